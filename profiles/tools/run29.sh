#!/bin/bash
# after the whole-CTA fold of the grid hand-overs: tests, then the probes of run28
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_golden_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -k "dot or cg" 2>&1 | tail -4 | cut -c1-300
timeout 600 python -m pytest tests/test_dist.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
bash profiles/tools/run28.sh
