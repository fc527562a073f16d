#!/bin/bash
# one GPU box visit: parity suite, A/B of this round's kernel changes, launch list + full ncu capture
mkdir -p gpurun_out; cd /root/repo
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/gpu.txt 2>&1
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --breakdown gpurun_out/bd_tgv128.tsv > gpurun_out/bench_tgv128.json 2> gpurun_out/bench_tgv128.err; echo "bench rc=$?"
WL_NO_NOBODY=1 WL_SCALAR_CFL=1 timeout 200 python bench.py --no-cpu > gpurun_out/bench_tgv128_old.json 2>&1
timeout 300 python bench.py --workload sphere --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/bd_sphere512.tsv > gpurun_out/bench_sphere512.json 2> gpurun_out/bench_sphere512.err
for f in bench_tgv128 bench_tgv128_old bench_sphere512; do python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/$f.json") if x.startswith("{")][-1]; d=json.loads(l)
    print("$f", d.get("value"), d.get("ms_per_step"), (d.get("roofline") or {}).get("frac"), (d.get("roofline") or {}).get("avg_ms"), d.get("clocks"), (d.get("e2e") or {}).get("value"))
except Exception as e: print("$f no result", e)
PY
done
# launch list of the default command (cold-cache, serialised)
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/ncu_launches_tgv128.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
# full capture: one step of the sphere workload (skip construction + warm-up launches)
timeout 500 ncu --set full --clock-control none --launch-skip 600 --launch-count 100 -o gpurun_out/r1_v14_sphere512 python bench.py --workload sphere --no-cpu --steps 3 --warmup 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/r1_v14_sphere512.ncu-rep --page raw --csv 2>/dev/null | python scripts/ncu_summary.py > gpurun_out/ncu_full_sphere512.tsv
[ $(stat -c %s gpurun_out/r1_v14_sphere512.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ] && rm -f gpurun_out/r1_v14_sphere512.ncu-rep
ls -la gpurun_out | head -40
