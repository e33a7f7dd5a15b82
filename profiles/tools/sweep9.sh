# power-law, segmented-scan walk: where does the time go?  exp1 = no x gathers, exp2 = consumers
# release every stage untouched (producer + bulk copies only).  Experiment libraries are built with
#   make -C spmv_b200/csrc OBJDIR=$PWD/build/expN EXTRA=-DB200_EXP=N LIBNAME=libexpN.so $PWD/spmv_b200/lib/libexpN.so
K="timeout 300 python profiles/tools/kbench.py"
for lib in libb200spmv.so libexp1.so libexp2.so; do
  echo "== $lib"
  B200SPMV_LIB=$PWD/spmv_b200/lib/$lib KB_VARIANTS=6 $K --case powerlaw --n 3000000 2>&1 | grep -v Warn | grep -v "want =" | grep -v "rel err"
done
