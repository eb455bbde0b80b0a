python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/bench_step.py
python bench.py --cpu-seconds 3
