#!/bin/bash
# round 2, call 18 (2 GPUs, tight timeouts): full GPU suite incl. the 2-rank worker (new: Function uBC/g and pcg! on slabs),
# bench on 1 GPU and on 2 GPUs
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c18_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_c18_pytest.log | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29541 tests/mgpu_worker.py fn_ubc fn_g pcg sphere > gpurun_out/r2_c18_mgpu_worker_2.log 2>&1; echo "worker rc=$?"; grep "MGPU" gpurun_out/r2_c18_mgpu_worker_2.log | cut -c1-420 | tail -10; tail -4 gpurun_out/r2_c18_mgpu_worker_2.log | cut -c1-300
timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c18_breakdown_sphere512.tsv > gpurun_out/r2_c18_bench_sphere.json 2> gpurun_out/r2_c18_bench_sphere.err; echo "bench rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_c18_bench_sphere_2gpu.json 2> gpurun_out/r2_c18_bench_2gpu.err; echo "bench 2gpu rc=$?"; tail -3 gpurun_out/r2_c18_bench_2gpu.err | cut -c1-300
python - <<'PY'
import json
for f in ("gpurun_out/r2_c18_bench_sphere.json","gpurun_out/r2_c18_bench_sphere_2gpu.json"):
    try:
        d=json.loads([x for x in open(f) if x.startswith("{")][-1])
        print({k:d.get(k) for k in ("value","ms_per_step","gpu_launches","n_gpus","mgpu_parity")}, d["roofline"]["frac"], d["roofline"]["avg_ms"])
    except Exception as e: print("no result", e)
PY
grep "@32768\|@65536\|@2368" gpurun_out/r2_c18_breakdown_sphere512.tsv | cut -f1,4 | tr '\n' ' '; echo
