python -m pytest tests -m gpu -x -q 2>&1 | tail -3
K="python profiles/tools/kbench.py"
for v in 2 3 4 5; do B200_SPMV_VARIANT=$v $K --case sym7; done
for tb in 24576 32768 49152; do for cw in 8 16; do for ct in 2 3 4; do
B200_SPMV_VARIANT=5 B200_SPMV_TILE_BYTES=$tb B200_SPMV_CWARPS=$cw B200_SPMV_CTAS=$ct $K --case sym7
done; done; done
for v in 2 3 4 5; do B200_SPMV_VARIANT=$v $K --case general7; done
for v in 2 3 4; do B200_SPMV_VARIANT=$v $K --case sym27; done
for v in 2 4; do B200_SPMV_VARIANT=$v $K --case general27; done
for ct in 3 4; do B200_SPMV_VARIANT=4 B200_SPMV_TILE_BYTES=24576 B200_SPMV_CTAS=$ct $K --case general7; done
for ct in 3 4; do B200_SPMV_VARIANT=4 B200_SPMV_TILE_BYTES=24576 B200_SPMV_CTAS=$ct $K --case sym27; done
