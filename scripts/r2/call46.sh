#!/bin/bash
# round 2, call 46 (4 GPUs, tight timeout): the driver's N=4 bench line (never run this round)
mkdir -p gpurun_out; cd /root/repo
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r2_c46_bench_sphere_4gpu.json 2> gpurun_out/r2_c46_bench_4gpu.err; echo "bench 4gpu rc=$?"; tail -3 gpurun_out/r2_c46_bench_4gpu.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c46_bench_sphere_4gpu.json") if x.startswith("{")][-1])
    print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity","halo_us_per_exchange")}); print(d["config"]["workload"], "|", d["config"]["parallelism"])
except Exception as e: print("no result", e)
PY
