# One gpurun call: GPU tests of the looping kernel and the served step, same-box A/B of bb_step_n (looping kernel vs graph of
# step kernels) and of the server's fence variants (BB_SERVE_FLAGS: 0 = acquire fence, 8 = round-2's acq_rel fence,
# 16 = system-scope fence per block, 24 = both).
#   gpurun --timeout 900 -- 'bash tools/gpu_ab_r02j.sh r02j'
TAG=${1:-r02j}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log
L=madrona_basketball_b200/libbball.so
timeout 400 python tools/ab_step.py $L:BB_STEP_N_GRAPH=1 $L 2048 8192 65536 1048576 > $O/${TAG}_ab_step_n_loop.txt 2>&1
cat $O/${TAG}_ab_step_n_loop.txt
for rnd in 1 2; do
for F in 0 8 16 24; do
echo "== BB_SERVE_FLAGS=$F"; BB_SERVE_FLAGS=$F timeout 120 python tools/e2e_probe.py 2>&1 | grep -E "serve"
done
done > $O/${TAG}_ab_serve_fences.txt 2>&1
cat $O/${TAG}_ab_serve_fences.txt
for F in 4 12; do echo "== BB_SERVE_FLAGS=$F"; BB_SERVE_FLAGS=$F timeout 120 python tools/e2e_probe.py 2>&1 | grep -E "marks|serve u8|serve no io"; done > $O/${TAG}_serve_marks.txt 2>&1
cat $O/${TAG}_serve_marks.txt
