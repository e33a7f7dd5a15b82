#!/bin/bash
# r02: split stream walk (variant 8) on the power-law matrix + its tests
mkdir -p gpurun_out
python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "powerlaw or huge or tall or segmented or bit_exact or fixture_general or dot_det or cg_kernels" > gpurun_out/r02_split_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_split_tests.log
tail -5 gpurun_out/r02_split_tests.log
KB_VARIANTS=3,8 timeout 600 python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 20 > gpurun_out/r02_sweep_powerlaw_split.log 2>&1
grep -v Warning gpurun_out/r02_sweep_powerlaw_split.log | grep -v "want ="
