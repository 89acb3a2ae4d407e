# development experiment: immediate vs register barrier ids in the fused kernel
run() { echo "== $*"; env "$@" python tools/quick_bench.py --only c4 2>&1 | cut -c1-100 | head -1; }
run A=immediate
MTB_NVCC_EXTRA="-DMTB_BAR_DYNAMIC" python -m mantrap_b200.build >/dev/null 2>&1
run A=dynamic
run A=dynamic
