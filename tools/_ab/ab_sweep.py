"""A/B on one box: the round-1 library (tools/_ab/libmantrap_r01.so, built from commit 4f2853a) against the current one on a few
sweep shapes. Development probe."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from mantrap_b200 import _cabi, engine
import sweep
old = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(os.path.join(ROOT, "tools", "_ab", "libmantrap_r01.so"))), "cuda:0")
new = engine.Engine("cuda:0")
for shape in ((20, 4, 1), (50, 4, 2), (50, 2, 4), (40, 8, 1), (20, 4, 2), (10, 2, 2), (30, 4, 4), (12, 20, 3)):
    for rep in range(2):
        a = sweep.measure(old, *shape, 1 << 20)
        b = sweep.measure(new, *shape, 1 << 20)
        print(shape, "r01 ms", a["ms"], "new ms", b["ms"], "ratio", round(a["ms"] / b["ms"], 3), "kind", b["kernel_kind"], flush=True)
