python -m pytest tests -m gpu -x -q 2>&1 | tail -3
K="python profiles/tools/kbench.py"
$K --case sym7; $K --case sym27; $K --case general7; $K --case general27
$K --case powerlaw --n 3000000
