# full ncu capture (with source) of the segmented-scan walk on the power-law matrix
T="--case powerlaw --n 3000000 --tune 6,24576,2,16,2"
timeout 300 python profiles/tools/kbench.py $T 2>&1 | grep -v Warn | head -1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -c 1 -f -o gpurun_out/powerlaw_seg \
  python profiles/tools/kbench.py $T --iters 1 --warmup 0 > gpurun_out/ncu_powerlaw_seg.log 2>&1
echo ncu rc=$?
