#!/bin/bash
# round 2, call 7 (2 GPUs): bench.py under torchrun -- mgpu_parity key, weak-scaling sphere line; multi-GPU tests
mkdir -p gpurun_out; cd /root/repo
( timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -x ) > gpurun_out/r2_c07_pytest_mgpu.log 2>&1; echo "pytest mgpu rc=$?"; tail -5 gpurun_out/r2_c07_pytest_mgpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c07_bench_sphere_2gpu.json 2> gpurun_out/r2_c07_bench_2gpu.err; echo "bench 2gpu rc=$?"; tail -5 gpurun_out/r2_c07_bench_2gpu.err
python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c07_bench_sphere_2gpu.json") if x.startswith("{")][-1])
    print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity")}); print(d["config"])
except Exception as e: print("no result", e)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 | cut -c1-400
