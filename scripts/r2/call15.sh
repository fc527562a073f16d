#!/bin/bash
# round 2, call 15: multi-block fold; per-level breakdown (sphere, tgv); ncu --set full + source of k_mom3 and k_down4 (level 1)
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c15_pytest.log
timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c15_breakdown_sphere512.tsv > gpurun_out/r2_c15_bench_sphere.json 2> gpurun_out/r2_c15_bench_sphere.err; echo "bench rc=$?"
timeout 300 python bench.py --no-cpu --workload tgv --steps 50 --breakdown gpurun_out/r2_c15_breakdown_tgv128.tsv > gpurun_out/r2_c15_bench_tgv.json 2> gpurun_out/r2_c15_bench_tgv.err; echo "bench tgv rc=$?"
python - <<'PY'
import json
for f in ("sphere","tgv"):
    try:
        d=json.loads(open(f"gpurun_out/r2_c15_bench_{f}.json").read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_mom3 -s 2 -c 1 -f -o gpurun_out/r2_c15_mom3 python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c15_ncu_mom3.log 2>&1; echo "ncu mom3 rc=$?"
timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_down4 -s 10 -c 1 -f -o gpurun_out/r2_c15_down4 python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c15_ncu_down4.log 2>&1; echo "ncu down4 rc=$?"
ls -la gpurun_out/*.ncu-rep
