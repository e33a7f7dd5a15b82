K="timeout 120 python profiles/tools/kbench.py"
$K --case sym7; $K --case general7; $K --case sym27; $K --case general27
for tb in 24576 32768 49152; do for cw in 8 16; do for ct in 2 3 4; do
B200_SPMV_TILE_BYTES=$tb B200_SPMV_CWARPS=$cw B200_SPMV_CTAS=$ct $K --case sym7
done; done; done
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -3
