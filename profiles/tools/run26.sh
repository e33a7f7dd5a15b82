#!/bin/bash
# 1 GPU: evict_first on the general matrix stream, CG us/iteration by block size
mkdir -p gpurun_out
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
for ef in 0 1; do
  echo "== B200_SPMV_STREAM_EVICT_FIRST=$ef"
  for nz in 32 64 128 256; do
    B200_SPMV_STREAM_EVICT_FIRST=$ef tr 300 2983$ef profiles/tools/cg_push_probe.py $nz 27 0 2>&1 | grep "ranks," | cut -c1-130
  done
  B200_SPMV_STREAM_EVICT_FIRST=$ef tr 300 2984$ef profiles/tools/cg_push_probe.py 256 7 0 2>&1 | grep "ranks," | cut -c1-130
  B200_SPMV_STREAM_EVICT_FIRST=$ef tr 300 2984$ef profiles/tools/cg_push_probe.py 64 7 0 2>&1 | grep "ranks," | cut -c1-130
done
