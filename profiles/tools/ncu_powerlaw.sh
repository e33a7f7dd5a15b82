#!/bin/bash
# one ncu --set full capture of the irregular-row walks on the 3 M-row power-law matrix
# usage: ncu_powerlaw.sh <variants> <shape> <out-name>     e.g. 3,6 24576x2x16x2 r02_powerlaw
V=${1:-3,6}; S=${2:-24576x2x16x2}; OUT=${3:-r02_powerlaw}
mkdir -p gpurun_out
export KB_VARIANTS=$V KB_SHAPES=$S
CMD="python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 2 --warmup 1"
$CMD > gpurun_out/${OUT}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -s 5 -c 5 -f -o gpurun_out/$OUT $CMD > gpurun_out/${OUT}_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/${OUT}_plain.log; tail -3 gpurun_out/${OUT}_ncu.log
