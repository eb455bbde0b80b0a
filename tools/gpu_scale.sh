# Scaling run on one multi-GPU box: bench.py at N = 1, 2, 4, 8 (the driver's launch lines) + the CPU arm.
#   gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_scale.sh r02c'
TAG=${1:-scale}
O=gpurun_out
nproc > $O/${TAG}_nproc.txt; nvidia-smi -L > $O/${TAG}_gpus.txt
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/${TAG}_bench_ref_n1.json 2> $O/${TAG}_bench_ref_n1.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err
for N in ${NS:-2 4 8}; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "N=$N rc=$?"
done
TAG_=$TAG python - <<'PY'
import json, glob, os
tag = os.environ.get("TAG_", "")
for f in sorted(glob.glob("gpurun_out/%s_bench_n*.json" % tag)):
    try:
        d = json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "value %.0f M  us/step %.2f  e2e %.0f M (%s)  ppo %s  strong %s  stress %s" % (
        d["value"] / 1e6, d["ms_per_step"] * 1e3, d["e2e"]["value"] / 1e6, d["e2e"]["path"],
        {k: (round(v / 1e6, 1) if isinstance(v, float) and v > 1e5 else v) for k, v in d.get("ppo", {}).items() if k.startswith(("env_steps", "p2p_max", "ranks"))},
        d.get("strong_65536", {}).get("value"), d.get("stress_2p20", {}).get("value")))
PY
