python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== default"; python tools/bench_step.py
echo "== 128 sync"; BB_PHASE_SYNC=1 python tools/bench_step.py --worlds 65536 1048576
for mb in 5 6; do echo "== minblocks $mb"; BB_LIB_PATH=$PWD/madrona_basketball_b200/libbball_mb$mb.so python tools/bench_step.py --worlds 65536 1048576;  BB_PHASE_SYNC=1 BB_LIB_PATH=$PWD/madrona_basketball_b200/libbball_mb$mb.so python tools/bench_step.py --worlds 1048576; done
