#!/bin/bash
# allreduce in the tail of the dot / update_xr kernels (unfused vector path)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
echo "== 2 GPUs: fused vec (default)"
tr 300 2 29921 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-130
echo "== 2 GPUs: separate kernels, allreduce launches"
B200_CG_FUSED_VEC=0 B200_CG_TAIL_ALLREDUCE=0 tr 300 2 29922 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-130
echo "== 2 GPUs: separate kernels, allreduce in the tails"
B200_CG_FUSED_VEC=0 tr 300 2 29923 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-130
B200_CG_FUSED_VEC=0 tr 300 2 29924 profiles/tools/cg_push_probe.py 64 7 1 2>&1 | grep "ranks," | cut -c1-130
B200_CG_FUSED_VEC=0 tr 300 2 29925 profiles/tools/cg_push_probe.py 512 27 0 2>&1 | grep "ranks," | cut -c1-130
echo "== 2 GPUs: fused, large"
tr 300 2 29926 profiles/tools/cg_push_probe.py 512 27 0 2>&1 | grep "ranks," | cut -c1-130
