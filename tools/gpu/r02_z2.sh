#!/bin/bash
for rep in 1 2; do
for mode in default nochain nopair both; do
unset PB200_NO_CHAIN PB200_P8_NOPAIR
if [ $mode = nochain ] || [ $mode = both ]; then export PB200_NO_CHAIN=1; fi
if [ $mode = nopair ] || [ $mode = both ]; then export PB200_P8_NOPAIR=1; fi
timeout 200 python tools/latency_bench.py 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$mode', {k: round(v,4) for k,v in d.items() if isinstance(v,float) and ('bs1' in k or 'bs2_device' in k)})"
done
done
