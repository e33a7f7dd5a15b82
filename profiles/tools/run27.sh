#!/bin/bash
# 2 GPUs: two-block dot epilogue (rowlist dot), evict_first by size: tests, then CG probes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_golden_gpu.py -x -q -m gpu -k "two_block or symv_dot or cg_" 2>&1 | tail -4 | cut -c1-300
timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu > gpurun_out/r02_pytest_dist_n2_dot.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_dist_n2_dot.log
tail -8 gpurun_out/r02_pytest_dist_n2_dot.log | cut -c1-300
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
for fd in 0 1; do
  echo "== B200_CG_FUSED_DOT_REMOTE=$fd"
  B200_CG_FUSED_DOT_REMOTE=$fd tr 300 2985$fd profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-160
  B200_CG_FUSED_DOT_REMOTE=$fd tr 300 2986$fd profiles/tools/cg_push_probe.py 512 27 0 2>&1 | grep "ranks," | cut -c1-160
  B200_CG_FUSED_DOT_REMOTE=$fd tr 300 2987$fd profiles/tools/cg_push_probe.py 512 7 1 2>&1 | grep "ranks," | cut -c1-160
  B200_CG_FUSED_DOT_REMOTE=$fd tr 300 2988$fd profiles/tools/cg_push_probe.py 64 7 1 2>&1 | grep "ranks," | cut -c1-160
done
