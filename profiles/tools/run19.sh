#!/bin/bash
mkdir -p gpurun_out
python profiles/tools/fused_probe.py 32 2>&1 | grep "nz=" | tee gpurun_out/r02_fused_probe.txt
python profiles/tools/fused_probe.py 256 2>&1 | grep "nz=" | tee -a gpurun_out/r02_fused_probe.txt
CMD="python bench.py --steps 3 --warmup 3 --no-extras"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -s 1 -c 1 -f -o gpurun_out/r02_sym7 $CMD > gpurun_out/r02_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/r02_ncu_full.log
