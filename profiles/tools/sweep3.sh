python -m pytest tests -m gpu -x -q 2>&1 | tail -3
K="python profiles/tools/kbench.py"
export B200_SPMV_TILE_BYTES=24576 B200_SPMV_CWARPS=16 B200_SPMV_CTAS=2
for v in 3 5; do B200_SPMV_VARIANT=$v $K --case sym7; B200SPMV_LIB=$PWD/spmv_b200/lib/libb200spmv_nored.so B200_SPMV_VARIANT=$v $K --case sym7; done
export B200_SPMV_TILE_BYTES=24576 B200_SPMV_CWARPS=8 B200_SPMV_CTAS=3
for v in 3 5; do B200_SPMV_VARIANT=$v $K --case sym7; B200SPMV_LIB=$PWD/spmv_b200/lib/libb200spmv_nored.so B200_SPMV_VARIANT=$v $K --case sym7; done
B200_SPMV_VARIANT=4 $K --case sym27; B200SPMV_LIB=$PWD/spmv_b200/lib/libb200spmv_nored.so B200_SPMV_VARIANT=4 $K --case sym27
