#!/bin/bash
# round 2, call 52: HEAD of the round -- full single-GPU suite, smoke(), default bench line (with cpu_baseline)
mkdir -p gpurun_out; cd /root/repo
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c52_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2_c52_pytest.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_c52_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_c52_smoke.log
timeout 300 python bench.py > gpurun_out/r2_c52_bench_default.json 2> gpurun_out/r2_c52_bench_default.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_c52_bench_default.json').read().strip().splitlines()[-1]); r=d['roofline']; print(d['value'], d['ms_per_step'], d['gpu_launches'], r['kernel'], r['frac'], r['share_of_step'], d['e2e']['value'], d['cpu_baseline']['value'], d['clocks']['reasons'])"
