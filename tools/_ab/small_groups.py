"""Development probe: the small-group sweep shapes that run evaluate_kernel, under the launch-plan knobs of the development library
(MTB_MINB = register budget variant, MTB_BLOCK = threads per block)."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from mantrap_b200 import _cabi, build, engine
import sweep
eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), "cuda:0")
shapes = ((50, 4, 2), (20, 4, 1), (50, 2, 4), (40, 8, 1), (50, 16, 1), (20, 2, 2), (20, 4, 2), (10, 4, 1), (50, 8, 2), (30, 4, 2))
for shape in shapes:
    row = []
    for knobs in ({}, {"MTB_MINB": "2"}, {"MTB_MINB": "4"}, {"MTB_BLOCK": "128"}, {"MTB_BLOCK": "128", "MTB_MINB": "4"}):
        for k in ("MTB_MINB", "MTB_BLOCK"):
            os.environ.pop(k, None)
        os.environ.update(knobs)
        r = sweep.measure(eng, *shape, 1 << 20)
        row.append(f"{knobs or 'default'}: {r['ms']} ms (bps {r['blocks_per_sm']}, blk {r['block']})")
    print(shape, " | ".join(row), flush=True)
