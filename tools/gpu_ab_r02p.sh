TAG=${1:-r02p}
O=gpurun_out
mkdir -p $O
L=/root/repo/build/libbball_D.so
timeout 400 python tools/ab_step.py $L:BB_LONE=0 $L:BB_LONE=1 2048 4096 8192 > $O/${TAG}_ab_lone_warps.txt 2>&1
cat $O/${TAG}_ab_lone_warps.txt
for rnd in 1 2 3; do for T in 32 128; do
echo "== BB_SERVE_THREADS=$T"; BB_LIB_PATH=$L BB_SERVE_THREADS=$T timeout 120 python tools/e2e_probe.py 2>&1 | grep -E "serve"
done; done > $O/${TAG}_ab_serve_threads.txt 2>&1
cat $O/${TAG}_ab_serve_threads.txt
BB_LIB_PATH=$L BB_LONE=1 BB_SERVE_THREADS=128 timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_lone.log 2>&1; tail -3 $O/${TAG}_pytest_lone.log
