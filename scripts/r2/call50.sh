#!/bin/bash
# round 2, call 50 (2 GPUs): two rows per thread also in the fused-exchange half sweep (k_gs_half_cx<3,true>): multi-GPU tests (default and forced), bench 1 and 2 GPUs
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests/test_multigpu.py tests/test_variants_gpu.py -m gpu -x -q > gpurun_out/r2_c50_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c50_pytest.log | cut -c1-300
timeout 300 python bench.py --no-cpu --steps 20 > gpurun_out/r2_c50_bench_1gpu.json 2> gpurun_out/r2_c50_1gpu.err; echo "bench rc=$?"
for v in off on; do
  if [ $v = on ]; then unset WL_NO_GS_C2; else export WL_NO_GS_C2=1; fi
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c50_bench_2gpu_$v.json 2> gpurun_out/r2_c50_2gpu_$v.err; echo "bench 2gpu $v rc=$?"
done
unset WL_NO_GS_C2
python - <<'PY'
import json
for f in ("bench_1gpu","bench_2gpu_off","bench_2gpu_on"):
    try:
        d=json.loads([x for x in open(f"gpurun_out/r2_c50_{f}.json") if x.startswith("{")][-1]); print(f, d["value"], d["ms_per_step"], d.get("mgpu_parity"))
    except Exception as e: print(f, "ERR", e)
PY
