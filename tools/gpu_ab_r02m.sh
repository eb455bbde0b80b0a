TAG=${1:-r02m}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
tail -4 $O/${TAG}_pytest_gpu.log
ENVS_LIST="2048 8192" bash tools/ab_rollout_loop.sh > $O/${TAG}_ab_rollout_loop.txt 2>&1; cat $O/${TAG}_ab_rollout_loop.txt
timeout 600 python bench.py --steps 20 --warmup 5 --cpu-seconds 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?" >> $O/${TAG}_bench.err
cat $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
