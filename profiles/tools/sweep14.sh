#!/bin/bash
# r02: chunk walk (variant 9)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k "powerlaw or huge or tall or segmented or bit_exact or fixture_general or chunk_walk or general_alpha or zero_rows or dot_det or cg_kernels" > gpurun_out/r02_chunk_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_chunk_tests.log
tail -12 gpurun_out/r02_chunk_tests.log | cut -c1-300
KB_VARIANTS=8,9 timeout 600 python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 20 > gpurun_out/r02_sweep_powerlaw_chunk.log 2>&1
grep -v Warning gpurun_out/r02_sweep_powerlaw_chunk.log | grep -v "want ="
