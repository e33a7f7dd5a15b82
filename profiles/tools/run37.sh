#!/bin/bash
# last GPU seconds of the round: distributed tests with every CG switch forced, then the CG / dot kernel tests
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_dist.py -x -q -m gpu > gpurun_out/r02_pytest_dist_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_dist_n2.log
tail -5 gpurun_out/r02_pytest_dist_n2.log | cut -c1-600
timeout 60 python -m pytest tests/test_golden_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
