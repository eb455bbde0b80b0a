TAG=${1:-r02s}
O=gpurun_out
mkdir -p $O
L=/root/repo/build/libbball_F.so
timeout 600 python tools/ab_step.py $L:BB_PDL_SYNC=0 $L:BB_PDL_SYNC=1 131072 262144 1048576 > $O/${TAG}_ab_pdl_sync_shape.txt 2>&1
cat $O/${TAG}_ab_pdl_sync_shape.txt
timeout 300 python tools/ab_step.py $L:BB_PDL_EARLY=0,BB_PREFETCH_SELF=0 $L 4096 8192 16384 65536 > $O/${TAG}_ab_new_defaults.txt 2>&1
cat $O/${TAG}_ab_new_defaults.txt
BB_LIB_PATH=$L BB_PDL_SYNC=1 timeout 600 python -m pytest tests -m gpu -x -q -k "parity or golden or north_star or stress" > $O/${TAG}_pytest_F.log 2>&1; tail -3 $O/${TAG}_pytest_F.log
