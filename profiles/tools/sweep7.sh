# power-law kernel: timing at two sizes, then one full ncu capture of the default plan
K="timeout 300 python profiles/tools/kbench.py"
$K --case powerlaw --n 3000000 2>&1 | grep -v Warn | grep -v "want ="
$K --case powerlaw --n 6250000 2>&1 | grep -v Warn | grep -v "want =" | head -3
timeout 300 ncu --set full --clock-control none --import-source on -k regex:spmv_pipe -c 1 -o gpurun_out/powerlaw_stream \
  python profiles/tools/kbench.py --case powerlaw --n 3000000 --iters 1 --warmup 0 > gpurun_out/ncu_powerlaw.log 2>&1
echo ncu rc=$?
