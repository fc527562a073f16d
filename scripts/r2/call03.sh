#!/bin/bash
# round 2, call 3: paired red-black sweeps (K-GS2): whole GPU suite + sphere/TGV bench with breakdown
mkdir -p gpurun_out; cd /root/repo
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c03_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_c03_pytest_gpu.log
timeout 600 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c03_breakdown_sphere512.tsv > gpurun_out/r2_c03_bench_sphere512.json 2> gpurun_out/r2_c03_bench_sphere.err; echo "bench sphere rc=$?"; tail -3 gpurun_out/r2_c03_bench_sphere.err; python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c03_bench_sphere512.json") if x.startswith("{")][-1])
    print("sphere", d["value"], d["ms_per_step"], d["roofline"]["kernel"], d["roofline"]["frac"], d["gpu_launches"])
    for r in d["roofline"]["kernels"]: print("  ", r["kernel"], r["launches_per_step"], r["avg_us"], r["share"], r["frac"])
except Exception as e: print("no result", e)
PY
cat gpurun_out/r2_c03_breakdown_sphere512.tsv
WL_NO_GS_PAIR=1 timeout 600 python bench.py --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere no-pair', d['value'], d['ms_per_step'])"
timeout 600 python bench.py --workload tgv --no-cpu --breakdown gpurun_out/r2_c03_breakdown_tgv128.tsv 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tgv', d['value'], d['ms_per_step'], d['gpu_launches'])"
cat gpurun_out/r2_c03_breakdown_tgv128.tsv | head -20
