#!/bin/bash
# round 2, call 1: the 5 builder-gated tests first, then the whole GPU suite on HEAD, then baseline bench lines
mkdir -p gpurun_out; cd /root/repo
( WL_RUN_UNVERIFIED=1 timeout 600 python -m pytest tests/test_zz_unverified_gpu.py -m gpu -q ) > gpurun_out/r2_c01_unverified.log 2>&1; echo "unverified rc=$?"; tail -30 gpurun_out/r2_c01_unverified.log
( timeout 900 python -m pytest tests -m gpu -q ) > gpurun_out/r2_c01_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_c01_pytest_gpu.log
timeout 400 python bench.py --workload sphere --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c01_breakdown_sphere512.tsv > gpurun_out/r2_c01_bench_sphere512.json 2> gpurun_out/r2_c01_bench_sphere.err; echo "bench sphere rc=$?"; tail -1 gpurun_out/r2_c01_bench_sphere512.json
timeout 400 python bench.py --no-cpu --breakdown gpurun_out/r2_c01_breakdown_tgv128.tsv > gpurun_out/r2_c01_bench_tgv128.json 2> gpurun_out/r2_c01_bench_tgv.err; echo "bench tgv rc=$?"; tail -1 gpurun_out/r2_c01_bench_tgv128.json
nproc; lscpu | grep "Model name"
