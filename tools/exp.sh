# development experiment: block size x register budget of the fused kernel on the config-4 shape (needs -DMTB_EXPERIMENT)
run() { echo "== $*"; env "$@" python tools/quick_bench.py --only c4 2>&1 | cut -c1-100 | head -1; }
run MTB_BLOCK=256 MTB_MINB=3
run MTB_BLOCK=256 MTB_MINB=4
run MTB_BLOCK=128 MTB_MINB=6
run MTB_BLOCK=128 MTB_MINB=7
run MTB_BLOCK=128 MTB_MINB=8
run MTB_BLOCK=64 MTB_MINB=12
run MTB_BLOCK=64 MTB_MINB=14
run MTB_BLOCK=64 MTB_MINB=16
