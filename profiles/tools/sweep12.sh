#!/bin/bash
# r02: warp-private pipelined walk (variants 6 / 7) against the stream walk (3) on the power-law matrix
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "powerlaw or huge or tall or segmented or bit_exact or fixture_general" > gpurun_out/r02_seg_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_seg_tests.log
tail -5 gpurun_out/r02_seg_tests.log
KB_VARIANTS=3,6,7 timeout 600 python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 20 > gpurun_out/r02_sweep_powerlaw_seg.log 2>&1
B200_SPMV_SEG_EVICT_FIRST=1 KB_VARIANTS=6,7 KB_SHAPES=16384x3x16x2,24576x2x16x2,12288x4x16x2 timeout 300 python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 20 > gpurun_out/r02_sweep_powerlaw_seg_ef.log 2>&1
cat gpurun_out/r02_sweep_powerlaw_seg.log gpurun_out/r02_sweep_powerlaw_seg_ef.log
