#!/bin/bash
# round 2, call 29 (2 GPUs): kept variant of the split kernels (classes 1/3 lean, list = classes 0/2, 32-column x halo): full suite incl. the
# 2-rank worker, same-box A/B against call 26's build, bench at 2 GPUs
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c29_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c29_pytest.log | cut -c1-300
for v in prev cur prev cur; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c29_breakdown_sphere_$v.tsv > gpurun_out/r2_c29_bench_sphere_$v.json 2> gpurun_out/r2_c29_bench_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c29_bench_sphere_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"], d["roofline"]["frac"], d["config"]["pattern_tiles_level1"]["general"], d["config"]["pattern_tiles_level1"]["constants"], d["config"]["pattern_tiles_level1"]["constants_with_boundary"])
PY
  grep "@32768\|@65536\|@2368" gpurun_out/r2_c29_breakdown_sphere_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_B200_LIB
grep "k_down4<2>@\|k_divres4<2>@" gpurun_out/r2_c29_breakdown_sphere_cur.tsv | cut -f1,4 | tr '\n' ' '; echo
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c29_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c29_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c29_bench_sphere_2gpu.json 2> gpurun_out/r2_c29_bench_2gpu.err; echo "bench 2gpu rc=$?"
python - <<'PY'
import json
try:
    d=json.loads([x for x in open("gpurun_out/r2_c29_bench_sphere_2gpu.json") if x.startswith("{")][-1]); print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity")})
except Exception as e: print("no result", e)
PY
