#!/bin/bash
# round 2, call 55 (2 GPUs): slab-parallel VTK output from real slabs -- worker case "vtk"
mkdir -p gpurun_out; cd /root/repo
timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 tests/mgpu_worker.py vtk > gpurun_out/r2_c55_mgpu_vtk.log 2>&1; echo "worker rc=$?"
grep "MGPU" gpurun_out/r2_c55_mgpu_vtk.log | cut -c1-700; tail -3 gpurun_out/r2_c55_mgpu_vtk.log | cut -c1-300
