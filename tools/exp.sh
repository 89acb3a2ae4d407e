run() { echo "== $*"; env "$@" python tools/quick_bench.py 2>&1 | cut -c1-90 | grep -v "grad=False" | head -1; }
MTB_NVCC_EXTRA="-DMTB_USE_CAS64=0" python -m mantrap_b200.build >/dev/null 2>&1
run MTB_NO_STAGE=1
run MTB_NO_STAGE=1 MTB_TAPE_GLOBAL=1 MTB_MINB=2
run MTB_NO_STAGE=1 MTB_TAPE_GLOBAL=1 MTB_MINB=3
run MTB_NO_STAGE=1 MTB_TAPE_GLOBAL=1 MTB_MINB=4
run MTB_NO_STAGE=0 MTB_MINB=2
