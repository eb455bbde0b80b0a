TAG=${1:-r02x}
O=gpurun_out
mkdir -p $O
A=/root/repo/build/libbball_A.so
G=/root/repo/build/libbball_G.so
for rnd in 1 2 3; do for L in $A $G; do
BB_LIB_PATH=$L timeout 300 python tools/bench_rollout.py --envs 8192 --iters 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$(basename $L)', 'fused ms/rollout (device) %.4f' % (d['fused']['s_per_rollout_device']*1e3), 'env-steps/s %.1f M' % (d['fused']['env_steps_per_s']/1e6))"
done; done > $O/${TAG}_ab_rollout_obs_image.txt 2>&1
cat $O/${TAG}_ab_rollout_obs_image.txt
BB_LIB_PATH=$G timeout 600 python -m pytest tests -m gpu -x -q -k "rollout or ppo or fixtures or reference" > $O/${TAG}_pytest_G.log 2>&1; tail -3 $O/${TAG}_pytest_G.log
