TAG=${1:-r02z}
O=gpurun_out
mkdir -p $O
H=/root/repo/build/libbball_H.so
run() { BB_LIB_PATH=$H BB_ROLLOUT_LOOP=$1 timeout 300 python tools/bench_rollout.py --envs 8192 --iters 30 --no-torch $2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('loop=$1 $2', 'fused ms/rollout (device) %.4f' % (d['fused']['s_per_rollout_device']*1e3), 'env-steps/s %.1f M' % (d['fused']['env_steps_per_s']/1e6))"; }
for rnd in 1 2; do run 0 --frozen; run 1 --frozen; run 3 --frozen; run 1 ""; run 3 ""; done > $O/${TAG}_ab_rollout_frozen_two_tiles.txt 2>&1
cat $O/${TAG}_ab_rollout_frozen_two_tiles.txt
BB_LIB_PATH=$H timeout 600 python -m pytest tests -m gpu -x -q -k "rollout or ppo or fixtures or reference" > $O/${TAG}_pytest_H.log 2>&1; tail -3 $O/${TAG}_pytest_H.log
