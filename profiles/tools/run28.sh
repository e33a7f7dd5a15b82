#!/bin/bash
# fused-step timeline at the per-GPU size of the 8-GPU strong run (2.1 M rows): 1 GPU (no peers) and 2 GPUs
mkdir -p gpurun_out
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
echo "== 1 GPU, fused vec forced on, timeline"
B200_CG_TIMELINE=1 B200_CG_FUSED_VEC=1 tr 300 1 29901 profiles/tools/cg_push_probe.py 32 27 0 2>&1 | grep "ranks,\|timeline" | tail -3 | cut -c1-260
echo "== 1 GPU, default (separate vector kernels)"
tr 300 1 29902 profiles/tools/cg_push_probe.py 32 27 0 2>&1 | grep "ranks," | cut -c1-160
echo "== 2 GPUs timeline"
B200_CG_TIMELINE=1 tr 300 2 29903 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks,\|timeline" | tail -5 | cut -c1-260
echo "== 2 GPUs fused vec off"
B200_CG_FUSED_VEC=0 tr 300 2 29904 profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-160
echo "== 2 GPUs fused CTAs per SM 1, 2, 4"
for c in 1 2 4; do B200_CG_FUSED_CTAS=$c tr 300 2 2991$c profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-160; done
