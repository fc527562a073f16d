#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
export MB=/root/repo/variants/lib_mbar.so
( WL_B200_LIB=$MB timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu_mbar.log 2>&1; echo "pytest mbar rc=$?"; tail -3 gpurun_out/pytest_gpu_mbar.log
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
BARGS=""
run tgv_sync WL_X=0; run tgv_mbar WL_B200_LIB=$MB; run tgv_sync2 WL_X=0; run tgv_mbar2 WL_B200_LIB=$MB
BARGS="--workload sphere --steps 10 --warmup 3"
run sphere_sync WL_X=0; run sphere_mbar WL_B200_LIB=$MB
