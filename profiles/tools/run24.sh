#!/bin/bash
# 2 GPUs: fused halo push in the CG step -- distributed tests, then the A/B at the per-GPU size of the 8-GPU strong run
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu > gpurun_out/r02_pytest_dist_n2_push.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_dist_n2_push.log
tail -8 gpurun_out/r02_pytest_dist_n2_push.log | cut -c1-300
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
tr 300 29801 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep -v "^W\|^\[W\|warn" | tail -3
tr 300 29802 profiles/tools/cg_push_probe.py 64 7 1 2>&1 | grep -v "^W\|^\[W\|warn" | tail -3
tr 300 29803 profiles/tools/cg_push_probe.py 256 27 0 2>&1 | grep -v "^W\|^\[W\|warn" | tail -3
