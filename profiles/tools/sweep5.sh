timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
K="timeout 120 python profiles/tools/kbench.py"
$K --case sym7; B200_SPMV_PREP=0 $K --case sym7
$K --case sym27; B200_SPMV_PREP=0 $K --case sym27
for ct in 2 3 4; do for cw in 8 16; do B200_SPMV_CTAS=$ct B200_SPMV_CWARPS=$cw $K --case sym7; done; done
B200_SPMV_TILE_BYTES=16384 B200_SPMV_CTAS=3 $K --case sym7
B200_SPMV_TILE_BYTES=32768 B200_SPMV_CTAS=2 $K --case sym7
timeout 100 python profiles/tools/dbg_cg.py 2>&1 | grep -E "max" | sed 's/dev first5.*max/max/'
