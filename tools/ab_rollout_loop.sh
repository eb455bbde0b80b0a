# same-box A/B of bb_rollout's two forms: BB_ROLLOUT_LOOP=1 the whole rollout as one launch of bb_rollout_loop_kernel,
# BB_ROLLOUT_LOOP=0 the chain of policy / step kernels (2 per step, programmatic dependent launches)
for ENVS in ${ENVS_LIST:-8192 65536}; do for rnd in 1 2; do for L in 0 1; do
BB_ROLLOUT_LOOP=$L timeout 300 python tools/bench_rollout.py --envs $ENVS --iters 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('envs=$ENVS loop=$L', 'fused ms/rollout (device) %.4f' % (d['fused']['s_per_rollout_device']*1e3), 'env-steps/s %.1f M' % (d['fused']['env_steps_per_s']/1e6))"
done; done; done
