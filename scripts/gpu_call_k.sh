#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
export V=/root/repo/variants/lib_l2pf.so
( WL_B200_LIB=$V timeout 200 python -m pytest tests/test_parity_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu_l2pf.log 2>&1; echo "pytest l2pf rc=$?"; tail -2 gpurun_out/pytest_gpu_l2pf.log
run() { name=$1; shift
  env "$@" timeout 120 python bench.py --no-cpu --breakdown gpurun_out/bd_$name.tsv $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_$name.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$name", "ms/step", round(d["ms_per_step"],4))
except Exception as e: print("$name no result", e)
PY
}
BARGS="--workload sphere --steps 10 --warmup 3"
run sphere_base WL_X=0; run sphere_l2pf WL_B200_LIB=$V
BARGS="--steps 40 --warmup 5"
run tgv_base WL_X=0; run tgv_l2pf WL_B200_LIB=$V
for n in sphere_base sphere_l2pf; do echo "== $n"; grep -E "k_up4|k_down4|k_divres4|k_proj4|k_incr4|k_cfl4" gpurun_out/bd_$n.tsv; done
