#!/bin/bash
# round 2, call 24: TIMING PROBE ONLY (results wrong): pressure kernels compiled with the constant-coefficient paths only, class-0 tiles skipped
mkdir -p gpurun_out; cd /root/repo
for v in cur plain; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 5 --warmup 3 --breakdown gpurun_out/r2_c24_breakdown_$v.tsv > gpurun_out/r2_c24_bench_$v.json 2> gpurun_out/r2_c24_$v.err; echo "$v rc=$?"
  grep "@32768\|@65536\|@4096" gpurun_out/r2_c24_breakdown_$v.tsv | cut -f1,2,4 | tr '\n' ' '; echo
done
