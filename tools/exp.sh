for cfg in "-DMTB_MIN_BLOCKS=2" "-DMTB_MIN_BLOCKS=3" "-DMTB_MIN_BLOCKS=4" "-DMTB_MIN_BLOCKS=3 -DMTB_USE_CAS64=0" "-DMTB_MIN_BLOCKS=2 -DMTB_USE_CAS64=0"; do
  echo "== $cfg"; MTB_NVCC_EXTRA="$cfg" python -m mantrap_b200.build > /dev/null 2>&1; python tools/quick_bench.py 2>&1 | cut -c1-120 | grep -v "grad=False"
done
