# Quick GPU iteration: parity tests, then step timings at three sizes (+ the graph replay figure).
#   gpurun --timeout 600 -- 'bash tools/gpu_quick.sh tag'
TAG=${1:-q}
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS} > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -4 $O/${TAG}_pytest.log
timeout 300 python tools/bench_step.py --reps 200 > $O/${TAG}_step.log 2>&1; cat $O/${TAG}_step.log
timeout 300 python tools/probe_floor.py > $O/${TAG}_floor.json 2> $O/${TAG}_floor.err; cat $O/${TAG}_floor.json
if [ -n "$NCU" ]; then
for W in ${NCU_WORLDS:-8192}; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bb_step_kernel -s 68 -c 1 -f -o $O/${TAG}_prof_step_$W python tools/profile_step.py --worlds $W --steps 72 > $O/${TAG}_ncu_prof_$W.log 2>&1
done
fi
if [ -n "$BENCH" ]; then
timeout 600 python bench.py --steps 20 --warmup 5 --cpu-seconds 3 > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?" >> $O/${TAG}_bench.err
cat $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
fi
if [ -n "$PREFETCH_TEST" ]; then
for P in 0 592 1184 2368; do echo "prefetch blocks $P"; BB_PREFETCH_BLOCKS=$P timeout 120 python tools/bench_step.py --worlds 32768 65536 131072 --reps 200; done > $O/${TAG}_prefetch.log 2>&1
cat $O/${TAG}_prefetch.log
fi
if [ -n "$E2E" ]; then timeout 300 python tools/e2e_probe.py > $O/${TAG}_e2e_probe.log 2>&1; cat $O/${TAG}_e2e_probe.log; fi
