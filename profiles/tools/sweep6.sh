K="timeout 120 python profiles/tools/kbench.py"
B200_SPMV_PREP=0 $K --case sym7
$K --case sym7
for c in 1 2 4 16 32; do B200_SPMV_PREP_CHUNK=$c $K --case sym7; done
$K --case sym27; B200_SPMV_PREP=0 $K --case sym27
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -4
