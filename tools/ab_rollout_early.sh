# same-box A/B of griddepcontrol.launch_dependents in the rollout chain: BB_ROLLOUT_EARLY bit 0 = the step kernel releases the
# policy kernel early, bit 1 = the policy kernel releases the step kernel early
for rnd in 1 2; do for E in 0 1 2 3; do
BB_ROLLOUT_EARLY=$E timeout 200 python tools/bench_rollout.py --iters 30 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('early=$E', 'fused ms/rollout (device) %.4f' % (d['fused']['s_per_rollout_device']*1e3), 'env-steps/s %.1f M' % (d['fused']['env_steps_per_s']/1e6))"
done; done
