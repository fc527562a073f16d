#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 300 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_default.log 2>&1; echo "pytest default rc=$?"; tail -3 gpurun_out/pytest_gpu_default.log
( WL_COARSE_SMEM_KB=200 WL_SMALL_CELLS=8000 timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu_coarse_sm.log 2>&1; echo "pytest coarse_sm rc=$?"; tail -3 gpurun_out/pytest_gpu_coarse_sm.log
run() { name=$1; shift
  env "$@" timeout 120 python bench.py --no-cpu $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_$name.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$name", "ms/step", round(d["ms_per_step"],4), "launches", d["gpu_launches"])
except Exception as e: print("$name no result", e)
PY
}
BARGS="--steps 40 --warmup 5"
for sc in 2000 8000; do for kb in 0 48 200; do run tgv_sc${sc}_kb${kb} WL_SMALL_CELLS=$sc WL_COARSE_SMEM_KB=$kb; done; done
run tgv_sc300_kb48 WL_SMALL_CELLS=300 WL_COARSE_SMEM_KB=48
run tgv_sc2000_kb0_b WL_SMALL_CELLS=2000 WL_COARSE_SMEM_KB=0
BARGS="--workload sphere --steps 10 --warmup 3"
run sph_sc2000_kb0 WL_SMALL_CELLS=2000 WL_COARSE_SMEM_KB=0
run sph_sc2000_kb48 WL_SMALL_CELLS=2000 WL_COARSE_SMEM_KB=48
run sph_sc8000_kb200 WL_SMALL_CELLS=8000 WL_COARSE_SMEM_KB=200
BARGS="--workload cylinder --steps 5 --warmup 2 --breakdown gpurun_out/bd_cyl_measband.tsv"
run cyl_measband WL_X=0
grep -E "band|measure|patch" gpurun_out/bd_cyl_measband.tsv
