#!/bin/bash
# two-block dot with the halo wait and the allreduce inside the remote kernel
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "two_block" 2>&1 | tail -2 | cut -c1-300
B200_CG_FUSED_VEC=0 B200_CG_FUSED_DOT_REMOTE=1 timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
export B200_CG_FUSED_VEC=0
for fd in 0 1; do
  echo "== separate kernels, B200_CG_FUSED_DOT_REMOTE=$fd"
  for a in "64 27 0" "64 7 1" "512 27 0" "512 7 1"; do
    B200_CG_FUSED_DOT_REMOTE=$fd tr 300 2 2993$fd profiles/tools/cg_push_probe.py $a 2>&1 | grep "ranks," | cut -c1-130
  done
done
echo "== separate kernels, dot remote on, wait as its own launch"
B200_CG_WAIT_IN_REMOTE=0 B200_CG_FUSED_DOT_REMOTE=1 tr 300 2 29941 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-130
