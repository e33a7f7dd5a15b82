# segmented-scan walk: parity tests, then the power-law sweep at two sizes
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q 2>&1 | tail -5
K="timeout 300 python profiles/tools/kbench.py"
KB_VARIANTS=3,6 $K --case powerlaw --n 3000000 2>&1 | grep -v Warn | grep -v "want ="
KB_VARIANTS=6 $K --case powerlaw --n 6250000 2>&1 | grep -v Warn | grep -v "want ="
