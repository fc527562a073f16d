#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q ) > gpurun_out/pytest_mgpu_final.log 2>&1; echo "pytest mgpu rc=$?"; tail -4 gpurun_out/pytest_mgpu_final.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus 2 --no-cpu --workload sphere --steps 10 --warmup 3 > gpurun_out/r1_v16_scale_sphere_2.json 2> gpurun_out/bench_sphere512_2gpu.err; echo "bench2 sphere rc=$?"
python - <<PY
import json
l=[x for x in open("gpurun_out/r1_v16_scale_sphere_2.json") if x.startswith("{")][-1]; d=json.loads(l)
print("sphere 2gpu", d["value"], d["ms_per_step"], d["config"]["parallelism"])
PY
