#!/bin/bash
# round 2, call 23: final state -- smoke(), the default bench line (with cpu_baseline), the reference arm, then the ncu launch list of the default command
mkdir -p gpurun_out; cd /root/repo
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_c23_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_c23_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/r2_c23_bench_default.json 2> gpurun_out/r2_c23_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_c23_bench_reference.json 2> gpurun_out/r2_c23_bench_reference.err; echo "reference rc=$?"
python - <<'PY'
import json
for f in ("default","reference"):
    try:
        d=json.loads(open(f"gpurun_out/r2_c23_bench_{f}.json").read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(f, d["value"], d["ms_per_step"], d.get("gpu_launches"), r.get("kernel"), r.get("frac"), r.get("share_of_step"), (d.get("e2e") or {}).get("value"), (d.get("cpu_baseline") or {}).get("value"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2_c23_ncu_launches_default.csv python bench.py --no-cpu --steps 5 --warmup 3 > gpurun_out/r2_c23_ncu_bench.log 2>&1; echo "ncu rc=$?"; wc -l gpurun_out/r2_c23_ncu_launches_default.csv
