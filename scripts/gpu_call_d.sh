#!/bin/bash
# march-length sweep of K-MOM for both tilings (WL_M3_KZ overrides the host heuristic)
mkdir -p gpurun_out; cd /root/repo
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu $BARGS > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/bench_$name.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$name", "ms/step", round(d["ms_per_step"],4), "kmom_ms", round(d["roofline"]["avg_ms"],4))
except Exception as e: print("$name no result", e)
PY
}
BARGS="--steps 20 --warmup 5"
for kz in 0 128 64 43 32 26 22; do run tgv_w31_kz$kz WL_M3_KZ=$kz; run tgv_w32_kz$kz WL_M3_W32=1 WL_M3_KZ=$kz; done
BARGS="--workload sphere --steps 6 --warmup 2"
for kz in 0 256 128 86 64 43 32 22; do run sph_w31_kz$kz WL_M3_KZ=$kz; run sph_w32_kz$kz WL_M3_W32=1 WL_M3_KZ=$kz; done
