TAG=${1:-r02t}
O=gpurun_out
mkdir -p $O
L=/root/repo/madrona_basketball_b200/libbball.so
timeout 600 python tools/ab_step.py $L $L:BB_PHASE_SYNC=0 98304 131072 262144 > $O/${TAG}_ab_nosync_128.txt 2>&1
cat $O/${TAG}_ab_nosync_128.txt
timeout 600 python tools/ab_step.py $L:BB_PHASE_SYNC=0,BB_BLOCK_THREADS=64 $L:BB_PHASE_SYNC=0,BB_OBS_ROUND=16 98304 131072 262144 > $O/${TAG}_ab_nosync_64_wpr16.txt 2>&1
cat $O/${TAG}_ab_nosync_64_wpr16.txt
