#!/bin/bash
# round 2, call 11 (2 GPUs, tight timeouts): multi-GPU parity worker, then bench.py under torchrun (mgpu_parity key)
mkdir -p gpurun_out; cd /root/repo
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29541 tests/mgpu_worker.py tgv32 sphere sphere_exit cylinder torus > gpurun_out/r2_c11_mgpu_worker_2.log 2>&1; echo "worker rc=$?"; grep "MGPU" gpurun_out/r2_c11_mgpu_worker_2.log | cut -c1-330 | tail -12; tail -5 gpurun_out/r2_c11_mgpu_worker_2.log | cut -c1-300
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c11_bench_sphere_2gpu.json 2> gpurun_out/r2_c11_bench_2gpu.err; echo "bench 2gpu rc=$?"; tail -4 gpurun_out/r2_c11_bench_2gpu.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c11_bench_sphere_2gpu.json") if x.startswith("{")][-1])
    print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity","halo_us_per_exchange")}); print(d["config"]["workload"], "|", d["config"]["parallelism"])
except Exception as e: print("no result", e)
PY
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | cut -c1-500
