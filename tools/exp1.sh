for pf in 0 296 592 1184 2368; do
  echo "== prefetch $pf"; BB_PREFETCH_BLOCKS=$pf python tools/bench_step.py --worlds 1048576 65536 --reps 30
done
echo "== prefetch 592 nosync"; BB_PREFETCH_BLOCKS=592 BB_PHASE_SYNC=0 python tools/bench_step.py --worlds 1048576 --reps 30
echo "== prefetch 592 256thr"; BB_PREFETCH_BLOCKS=296 BB_BLOCK_THREADS=256 python tools/bench_step.py --worlds 1048576 --reps 30
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
