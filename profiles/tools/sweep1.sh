set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
K="python profiles/tools/kbench.py"
$K --case sym7
B200_SPMV_VARIANT=2 $K --case sym7
for tb in 24576 32768 49152 65536; do for cw in 8 16; do for ct in 1 2 3; do
B200_SPMV_TILE_BYTES=$tb B200_SPMV_CWARPS=$cw B200_SPMV_CTAS=$ct $K --case sym7
done; done; done
B200_SPMV_STAGES=3 B200_SPMV_TILE_BYTES=32768 $K --case sym7
$K --case general7
B200_SPMV_VARIANT=2 $K --case general7
B200_SPMV_CWARPS=16 $K --case general7
$K --case general27
B200_SPMV_VARIANT=3 $K --case general27
$K --case sym27
B200_SPMV_VARIANT=2 $K --case sym27
