TAG=${1:-r02r}
O=gpurun_out
mkdir -p $O
L=/root/repo/build/libbball_E.so
timeout 500 python tools/ab_step.py $L:BB_PDL_EARLY=0,BB_PREFETCH_SELF=0 $L:BB_PDL_EARLY=1,BB_PREFETCH_SELF=0 16384 32768 65536 > $O/${TAG}_ab_early_mid_sizes.txt 2>&1
cat $O/${TAG}_ab_early_mid_sizes.txt
timeout 500 python tools/ab_step.py $L:BB_PDL_EARLY=1,BB_PREFETCH_SELF=1 $L:BB_PDL_EARLY=0,BB_PREFETCH_SELF=1 16384 32768 65536 > $O/${TAG}_ab_early_prefetch_mid_sizes.txt 2>&1
cat $O/${TAG}_ab_early_prefetch_mid_sizes.txt
