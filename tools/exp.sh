# development experiment: placement of the collapsed means (shared memory vs L2 scratch) on shared-memory-bound shapes
python - <<'PY'
import os, sys
sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import sweep
from mantrap_b200 import engine
eng = engine.Engine("cuda:0")
for t, n, p in ((50, 2, 1), (50, 2, 2), (50, 4, 2), (50, 8, 1), (50, 16, 2), (50, 32, 1), (50, 64, 1), (50, 64, 2), (50, 64, 8), (30, 16, 4), (20, 16, 4), (12, 20, 3), (10, 2, 1), (40, 16, 4), (50, 32, 4)):
    out = []
    for knob in ("2", "1", "0"):
        os.environ["MTB_MEANS_GLOBAL"] = knob
        r = sweep.measure(eng, t, n, p, 1 << 18)
        out.append(f"{r['ms']:8.2f} (bps {r['blocks_per_sm']} blk {r['block']})")
    print(f"T={t} N={n} P={p}: smem {out[0]} | global {out[1]} | auto {out[2]}", flush=True)
PY
