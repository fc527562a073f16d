#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
export OLD=/root/repo/variants/lib_finldg.so
( timeout 300 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_default.log 2>&1; echo "pytest default rc=$?"; tail -3 gpurun_out/pytest_gpu_default.log
run() { name=$1; shift
  env "$@" timeout 120 python bench.py --no-cpu $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_$name.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$name", "ms/step", round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],4), "kmom_ms", round(d["roofline"]["avg_ms"],4))
except Exception as e: print("$name no result", e)
PY
}
BARGS="--steps 40 --warmup 5"
run tgv_finsm WL_X=0; run tgv_finldg WL_B200_LIB=$OLD; run tgv_finsm2 WL_X=0; run tgv_finldg2 WL_B200_LIB=$OLD
BARGS="--workload sphere --steps 10 --warmup 3"
run sphere_finsm WL_X=0; run sphere_finldg WL_B200_LIB=$OLD
