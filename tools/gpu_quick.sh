# Quick GPU iteration: parity tests, then step timings at three sizes (+ the graph replay figure).
#   gpurun --timeout 600 -- 'bash tools/gpu_quick.sh tag'
TAG=${1:-q}
O=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS} > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -4 $O/${TAG}_pytest.log
timeout 300 python tools/bench_step.py --reps 200 > $O/${TAG}_step.log 2>&1; cat $O/${TAG}_step.log
timeout 300 python tools/probe_floor.py > $O/${TAG}_floor.json 2> $O/${TAG}_floor.err; cat $O/${TAG}_floor.json
