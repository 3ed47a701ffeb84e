"""Summarise an `ncu --page source --csv` dump: top sampled SASS lines + all barrier waits / MMA / copy instructions."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ia, isamp, iex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
data = [(int(r[isamp] or 0), i, r[ia].strip(), r[iex]) for i, r in enumerate(rows[2:]) if len(r) > isamp]
tot = sum(d[0] for d in data)
print('total samples', tot)
for s, i, src, ex in sorted(data, reverse=True)[:top]:
    print(f"{s:6d} {100*s/tot:5.1f}% line{i:5d} ex={ex:>9s} {src[:100]}")
print('--- sync / async instructions in program order')
for s, i, src, ex in data:
    if any(k in src for k in ('SYNCS', 'UTCBAR', 'UBLKCP', 'UTMALDG', 'BAR.SYNC', 'LDTM', 'WARPSYNC', 'ARRIVE')) or (s >= 3 and 'BRA' in src):
        print(f"line{i:5d} samp={s:5d} ex={ex:>9s} {src[:100]}")
