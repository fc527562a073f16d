#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_down4|k_divres4|k_up4|k_incr4|k_proj4" --launch-count 19 -o gpurun_out/r1_v16_mg4_sphere512 python bench.py --workload sphere --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_mg4.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r1_v16_mg4_sphere512.ncu-rep --page raw --csv 2>/dev/null | python scripts/ncu_summary.py > gpurun_out/r1_v16_ncu_full_mg4_sphere512.tsv
cut -f1,2,5 gpurun_out/r1_v16_ncu_full_mg4_sphere512.tsv
ls -la gpurun_out/r1_v16_mg4_sphere512.ncu-rep
