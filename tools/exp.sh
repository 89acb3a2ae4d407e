# development experiment: threads per block (work items packed per block) of the Jacobian kernel (needs -DMTB_EXPERIMENT)
run() { echo "== $*"; env "$@" python tools/quick_bench.py --only jac 2>&1 | cut -c1-120 | sed -n '2p;4p;5p'; }
run MTB_JAC_THREADS=64
run MTB_JAC_THREADS=96
run MTB_JAC_THREADS=128
run MTB_JAC_THREADS=192
run MTB_JAC_THREADS=256
