#!/bin/bash
# round 2, call 57: bench.py at the head of the round (row labels of the one-launch small levels) (no CPU leg, short): the line must still print
mkdir -p gpurun_out; cd /root/repo
timeout 120 python bench.py --no-cpu --steps 10 > gpurun_out/r2_c57_bench.json 2> gpurun_out/r2_c57_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_c57_bench.json').read().strip().splitlines()[-1]); r=d['roofline']; print(d['value'], d['ms_per_step'], d['gpu_launches'], r['kernel'], r['frac']); print([ (k['kernel'], round(k['frac'],3), 'frac_note' in k) for k in r['kernels'] if k.get('frac')])"
tail -3 gpurun_out/r2_c57_bench.err
