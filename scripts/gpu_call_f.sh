#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_default.log 2>&1; echo "pytest default rc=$?"; tail -3 gpurun_out/pytest_gpu_default.log
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --breakdown gpurun_out/bd_$name.tsv $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_$name.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$name", "ms/step", round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],4), "kmom_ms", round(d["roofline"]["avg_ms"],4), "e2e", round(d["e2e"]["value"],4))
except Exception as e: print("$name no result", e)
PY
}
BARGS="--workload sphere --steps 10 --warmup 3"
run sphere_2pass WL_X=0
run sphere_1pass WL_PATCH_1PASS=1
BARGS="--workload cylinder --steps 5 --warmup 2"
run cylinder_2pass WL_X=0
run cylinder_1pass WL_PATCH_1PASS=1
for n in sphere_2pass sphere_1pass cylinder_2pass cylinder_1pass; do echo "== $n"; grep -E "patch|band|k_mom3" gpurun_out/bd_$n.tsv; done
