#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 600 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py --breakdown gpurun_out/r1_v16_breakdown_tgv128.tsv > gpurun_out/r1_v16_bench_tgv128.json 2> gpurun_out/bench_tgv128.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_v16_bench_reference_tgv128.json 2> gpurun_out/bench_reference.err; echo "bench ref rc=$?"
timeout 400 python bench.py --workload sphere --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r1_v16_breakdown_sphere512.tsv > gpurun_out/r1_v16_bench_sphere512.json 2> gpurun_out/bench_sphere.err; echo "bench sphere rc=$?"
timeout 400 python bench.py --workload cylinder --no-cpu --steps 5 --warmup 2 --breakdown gpurun_out/r1_v16_breakdown_cylinder.tsv > gpurun_out/r1_v16_bench_cylinder.json 2> gpurun_out/bench_cyl.err; echo "bench cyl rc=$?"
for f in r1_v16_bench_tgv128 r1_v16_bench_reference_tgv128 r1_v16_bench_sphere512 r1_v16_bench_cylinder; do python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/$f.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$f", d.get("value"), d.get("ms_per_step"), (d.get("roofline") or {}).get("frac"), (d.get("roofline") or {}).get("avg_ms"), (d.get("cpu_baseline") or {}).get("value"), (d.get("e2e") or {}).get("value"), d.get("gpu_launches"), (d.get("clocks") or {}).get("sm_mhz"))
except Exception as e: print("$f no result", e)
PY
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r1_v16_ncu_launches_tgv128.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
