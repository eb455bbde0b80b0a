# One gpurun call of a round: GPU tests, both bench arms, floors, ncu launch list + --set full captures.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh r02a'
TAG=${1:-r02}
O=gpurun_out
set -x
# the .ncu-rep files (16 MB each with the source page) are summarised on the box and removed: gpurun brings back at most 64 MiB
summarise() { if [ -f $1.ncu-rep ]; then python tools/ncu_summary.py $1.ncu-rep $1_summary.csv > /dev/null 2>&1; python tools/ncu_by_func.py $1.ncu-rep > $1_by_function.txt 2>&1; python tools/ncu_opcodes.py $1.ncu-rep > $1_opcodes.txt 2>&1; rm -f $1.ncu-rep; fi; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
nproc > $O/${TAG}_nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log
cp $O/ref_ppo_py_sps.json $O/${TAG}_ref_ppo_py_sps.json 2>/dev/null
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err
timeout 600 python bench.py --steps 20 --warmup 5 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?" >> $O/${TAG}_bench.err
timeout 300 python tools/probe_floor.py > $O/${TAG}_floor.json 2> $O/${TAG}_floor.err
if [ -z "$NO_NCU" ]; then
timeout 300 python bench.py --steps 5 --warmup 3 --cpu-seconds 0.2 --no-serve --no-extras > $O/${TAG}_plain_bench.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${TAG}_launches_bench.csv python bench.py --steps 5 --warmup 3 --cpu-seconds 0.2 --no-serve --no-extras > $O/${TAG}_ncu_bench.log 2>&1
for W in 8192 65536 1048576; do
timeout 120 python tools/profile_step.py --worlds $W --steps 72 > $O/${TAG}_plain_prof_$W.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:bb_step_kernel -s 68 -c 1 -f -o $O/${TAG}_prof_step_$W python tools/profile_step.py --worlds $W --steps 72 > $O/${TAG}_ncu_prof_$W.log 2>&1
summarise $O/${TAG}_prof_step_$W
done
# the looping kernel (bb_step_n): 64 steps in one launch at 8192 worlds, and the one-launch rollout
timeout 120 python tools/profile_step.py --worlds 8192 --steps 8 --loop 64 > $O/${TAG}_plain_prof_loop.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:bb_step_loop_kernel -c 1 -f -o $O/${TAG}_prof_step_loop_8192 python tools/profile_step.py --worlds 8192 --steps 8 --loop 64 > $O/${TAG}_ncu_prof_loop.log 2>&1
summarise $O/${TAG}_prof_step_loop_8192
timeout 120 python tools/profile_rollout.py > $O/${TAG}_plain_prof_rollout.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:bb_rollout_loop_kernel -s 2 -c 1 -f -o $O/${TAG}_prof_rollout_loop_8192 python tools/profile_rollout.py > $O/${TAG}_ncu_prof_rollout.log 2>&1
summarise $O/${TAG}_prof_rollout_loop_8192
fi
timeout 300 python tools/ab_step.py madrona_basketball_b200/libbball.so:BB_STEP_N_GRAPH=1 madrona_basketball_b200/libbball.so 8192 65536 > $O/${TAG}_ab_step_n_loop.txt 2>&1
ENVS_LIST="8192" bash tools/ab_rollout_loop.sh > $O/${TAG}_ab_rollout_loop.txt 2>&1
timeout 200 python tools/e2e_probe.py > $O/${TAG}_e2e_probe.txt 2>&1
tail -3 $O/${TAG}_pytest_gpu.log; cat $O/${TAG}_bench.json; tail -2 $O/${TAG}_bench.err; cat $O/${TAG}_floor.json
