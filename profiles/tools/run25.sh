#!/bin/bash
# 2 GPUs: evict_first on the general matrix stream at the per-GPU size of the 8-GPU strong run (vectors L2-resident?)
mkdir -p gpurun_out
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
for ef in 0 1; do
  echo "== B200_SPMV_STREAM_EVICT_FIRST=$ef"
  B200_SPMV_STREAM_EVICT_FIRST=$ef tr 300 2981$ef profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-400
  B200_SPMV_STREAM_EVICT_FIRST=$ef tr 300 2982$ef profiles/tools/cg_push_probe.py 128 27 0 2>&1 | grep "ranks," | cut -c1-400
done
