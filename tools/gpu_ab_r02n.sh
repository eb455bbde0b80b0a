TAG=${1:-r02n}
O=gpurun_out
mkdir -p $O
for rnd in 1 2 3; do for L in A C; do
echo "== lib $L"; BB_LIB_PATH=/root/repo/build/libbball_$L.so timeout 120 python tools/e2e_probe.py 2>&1 | grep -E "serve"
done; done > $O/${TAG}_ab_serve_one_word_doorbell.txt 2>&1
cat $O/${TAG}_ab_serve_one_word_doorbell.txt
BB_LIB_PATH=/root/repo/build/libbball_C.so BB_SERVE_FLAGS=4 timeout 120 python tools/e2e_probe.py 2>&1 | grep -E "marks|serve u8|serve no io" > $O/${TAG}_serve_marks_C.txt; cat $O/${TAG}_serve_marks_C.txt
BB_LIB_PATH=/root/repo/build/libbball_C.so timeout 300 python -m pytest tests -m gpu -x -q -k "serve or host or packed" > $O/${TAG}_pytest_serve_C.log 2>&1; tail -3 $O/${TAG}_pytest_serve_C.log
