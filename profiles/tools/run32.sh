#!/bin/bash
# new defaults (separate vector kernels with tail allreduce, two-block dot, wait in the remote kernel); push grid cap
mkdir -p gpurun_out
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $3 "${@:4}"; }
for cap in 0 32 128; do
  echo "== B200_HALO_PUSH_CTAS=$cap"
  B200_HALO_PUSH_CTAS=$cap tr 300 2 2995${cap:0:1} profiles/tools/cg_push_probe.py 64 27 0 2>&1 | grep "ranks," | cut -c1-130
  B200_HALO_PUSH_CTAS=$cap tr 300 2 2996${cap:0:1} profiles/tools/cg_push_probe.py 64 7 1 2>&1 | grep "ranks," | cut -c1-130
done
