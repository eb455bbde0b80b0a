set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --steps 5 --warmup 3 --cpu-seconds 0.2 --no-serve > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 5 --warmup 3 --cpu-seconds 0.2 --no-serve > gpurun_out/ncu_bench.log 2>&1
python tools/profile_step.py --worlds 1048576 --steps 72 > gpurun_out/plain_prof.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bb_step_kernel -s 68 -c 2 -o gpurun_out/prof_step_1m python tools/profile_step.py --worlds 1048576 --steps 72 > gpurun_out/ncu_prof.log 2>&1
python tools/profile_step.py --worlds 8192 --steps 72 > gpurun_out/plain_prof8k.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bb_step_kernel -s 68 -c 2 -o gpurun_out/prof_step_8k python tools/profile_step.py --worlds 8192 --steps 72 > gpurun_out/ncu_prof8k.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; tail -2 gpurun_out/bench.err
