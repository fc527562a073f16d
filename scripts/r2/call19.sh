#!/bin/bash
# round 2, call 19: potential of class-1-only instantiations (fewer registers): TIMING ONLY, the build skips class 0/2 tiles (results wrong)
mkdir -p gpurun_out; cd /root/repo
export WL_B200_LIB=/root/repo/gpurun_tmp/lib_c12.so
timeout 300 python bench.py --no-cpu --steps 5 --warmup 3 --breakdown gpurun_out/r2_c19_breakdown_c12.tsv > gpurun_out/r2_c19_bench_c12.json 2> gpurun_out/r2_c19.err; echo "rc=$?"; tail -3 gpurun_out/r2_c19.err
grep "@32768\|@65536\|@2368\|@4096" gpurun_out/r2_c19_breakdown_c12.tsv | cut -f1,2,4 | tr '\n' ' '; echo
