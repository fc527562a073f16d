#!/bin/bash
# round 2, call 6: the reworked bench.py end to end (N=1 default with cpu_baseline, reference arm), full GPU suite after the
# clean-up, and the whole-library -fmad=true experiment (tolerance parity + speed)
mkdir -p gpurun_out; cd /root/repo
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c06_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c06_pytest_gpu.log
timeout 900 python bench.py --breakdown gpurun_out/r2_c06_breakdown_sphere512.tsv > gpurun_out/r2_c06_bench_sphere512.json 2> gpurun_out/r2_c06_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_c06_bench.err
python - <<'PY'
import json
d=json.loads([x for x in open("gpurun_out/r2_c06_bench_sphere512.json") if x.startswith("{")][-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches","warmup","steps")}); print(d["config"]); print("cpu", d["cpu_baseline"]); print("e2e", d["e2e"]); print("clocks", d["clocks"])
r=d["roofline"]; print({k:r[k] for k in r if k!="kernels"})
PY
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_c06_bench_reference.json 2>> gpurun_out/r2_c06_bench.err; echo "ref rc=$?"; cat gpurun_out/r2_c06_bench_reference.json | cut -c1-600
WL_B200_LIB=gpurun_tmp/lib_fmad.so timeout 600 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c06_breakdown_sphere512_fmad.tsv 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere fmad=true', d['value'], d['ms_per_step'], d['roofline']['avg_ms'])"
head -8 gpurun_out/r2_c06_breakdown_sphere512_fmad.tsv
( WL_B200_LIB=/root/repo/gpurun_tmp/lib_fmad.so timeout 900 python -m pytest tests/test_fullsize_gpu.py tests/test_parity_gpu.py -m gpu -q ) > gpurun_out/r2_c06_pytest_fmad.log 2>&1; echo "pytest fmad rc=$?"; tail -15 gpurun_out/r2_c06_pytest_fmad.log | cut -c1-300
