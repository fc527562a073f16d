#!/bin/bash
# 2-GPU visit: slab-decomposed parity (== single GPU, bit for bit) with the new K-MOM tiling, and a 2-GPU bench line
mkdir -p gpurun_out; cd /root/repo
nvidia-smi -L | head -4
( timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q ) > gpurun_out/pytest_mgpu.log 2>&1; echo "pytest mgpu rc=$?"; tail -4 gpurun_out/pytest_mgpu.log
( WL_M3_W32=0 timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q ) > gpurun_out/pytest_mgpu_w31.log 2>&1; echo "pytest mgpu w31 rc=$?"; tail -2 gpurun_out/pytest_mgpu_w31.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --no-cpu > gpurun_out/bench_tgv128_2gpu.json 2> gpurun_out/bench_tgv128_2gpu.err; echo "bench2 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus 2 --no-cpu --workload sphere --steps 10 --warmup 3 > gpurun_out/bench_sphere512_2gpu.json 2> gpurun_out/bench_sphere512_2gpu.err; echo "bench2 sphere rc=$?"
for f in bench_tgv128_2gpu bench_sphere512_2gpu; do python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/$f.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$f", d["value"], d["ms_per_step"], d["config"]["parallelism"], d["roofline"]["frac"])
except Exception as e: print("$f no result", e)
PY
done
