#!/bin/bash
# round 2, call 41 (2 GPUs): final state after the launch-bounds and prefetch work -- full GPU suite (incl. the 2-rank tests), smoke(), default bench line with cpu_baseline, reference arm,
# bench on 2 GPUs, ncu launch list of the default workload
mkdir -p gpurun_out; cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c41_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c41_pytest.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_c41_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_c41_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/r2_c41_bench_default.json 2> gpurun_out/r2_c41_bench_default.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_c41_bench_reference.json 2> gpurun_out/r2_c41_bench_reference.err; echo "reference rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c41_bench_sphere_2gpu.json 2> gpurun_out/r2_c41_bench_2gpu.err; echo "bench 2gpu rc=$?"
python - <<'PY'
import json
for f in ("bench_default","bench_reference","bench_sphere_2gpu"):
    try:
        d=json.loads([x for x in open(f"gpurun_out/r2_c41_{f}.json") if x.startswith("{")][-1]); r=d.get("roofline") or {}
        print(f, d["value"], d["ms_per_step"], d.get("gpu_launches"), r.get("kernel"), r.get("frac"), r.get("share_of_step"), (d.get("e2e") or {}).get("value"), (d.get("cpu_baseline") or {}).get("value"), d.get("mgpu_parity"), (d.get("clocks") or {}).get("reasons"))
    except Exception as e: print(f, "ERR", e)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2_c41_ncu_launches_default.csv python bench.py --no-cpu --steps 5 --warmup 3 > gpurun_out/r2_c41_ncu_bench.log 2>&1; echo "ncu rc=$?"; wc -l gpurun_out/r2_c41_ncu_launches_default.csv
