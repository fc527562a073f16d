#!/bin/bash
# round 2, call 51 (8 GPUs, tight timeout, final build of the round): the driver's N=8 bench line incl. mgpu_parity and the config_E sub-record (TGV 1024^3)
mkdir -p gpurun_out; cd /root/repo
nvidia-smi -L | wc -l
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2_c51_bench_sphere_8gpu.json 2> gpurun_out/r2_c51_bench_8gpu.err; echo "bench 8gpu rc=$?"; tail -4 gpurun_out/r2_c51_bench_8gpu.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c51_bench_sphere_8gpu.json") if x.startswith("{")][-1])
    print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity","halo_us_per_exchange")}); print(d["config"]["workload"], "|", d["config"]["parallelism"]); print(d.get("config_E"))
except Exception as e: print("no result", e)
PY
