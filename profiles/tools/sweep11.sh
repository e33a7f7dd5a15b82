# after the incremental stage counters: the four stencil cases, then the power-law shapes incl. 3 CTAs/SM x 16 warps
K="timeout 300 python profiles/tools/kbench.py"
$K --case sym7; $K --case general7; $K --case sym27; $K --case general27
KB_VARIANTS=3,6 $K --case powerlaw --n 3000000 2>&1 | grep -v Warn | grep -v "want ="
