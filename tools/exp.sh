# development experiment: register budget / direction chunk of the Jacobian kernel on the config-3 shape (needs -DMTB_EXPERIMENT)
run() { echo "== $*"; env "$@" python tools/quick_bench.py --only jac 2>&1 | cut -c1-120 | head -2 | tail -1; }
run MTB_JAC_VARIANT=1
run MTB_JAC_VARIANT=5
run MTB_JAC_VARIANT=6
run MTB_JAC_VARIANT=7
