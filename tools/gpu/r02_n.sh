#!/bin/bash
# multi-GPU bench as the driver launches it: bash tools/gpu/r02_n.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/n${N}_topo.txt 2>&1
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err; echo "bench N=$N exit $?"; tail -5 gpurun_out/n${N}_bench.err
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/n${N}_bench.json") if l.startswith("{")][-1])
print("N=%d value %.0f ms %.4f | e2e %.0f ms %.4f | h2d ceil %.0f (%.1f GB/s/gpu) frac %.3f | fp16x2 %.0f e2e %.0f" % (
    d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_ceiling_frames_per_s"],
    d["e2e"]["h2d_ceiling_GBps_per_gpu"], d["e2e"]["frac_of_h2d_ceiling"], d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"]))
print("limiter:", d["e2e"]["limiter"])
print("training:", {k: v for k, v in d["training_step"].items() if k != "what"})
print("factors sharded:", d["factors"].get("sharded"))
print("strong:", d["strong_scaling"])
PY
