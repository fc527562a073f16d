#!/bin/bash
# round 2, call 16: class-independent loads hoisted above the tile-class branch; accel-field slot in K-MOM
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c16_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c16_pytest.log
for w in sphere tgv cylinder; do
timeout 300 python bench.py --no-cpu --workload $w --steps 20 --breakdown gpurun_out/r2_c16_breakdown_$w.tsv > gpurun_out/r2_c16_bench_$w.json 2> gpurun_out/r2_c16_bench_$w.err; echo "bench $w rc=$?"
done
python - <<'PY'
import json
for f in ("sphere","tgv","cylinder"):
    try:
        d=json.loads(open(f"gpurun_out/r2_c16_bench_{f}.json").read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["gpu_launches"], d["roofline"]["frac"])
    except Exception as e: print(f, "ERR", e)
PY
