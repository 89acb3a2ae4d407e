"""Development probe: launch-geometry variants of the evaluation kernels on the config-4 shape through the -DMTB_DEV_KNOBS library."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from mantrap_b200 import _cabi, build, engine  # noqa: E402
import quick_bench as qb  # noqa: E402

eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), "cuda:0")
w = {"goal": 0.5, "acc": 1.0}
for env in ({"MTB_PAIR": "1"}, {}, {"MTB_BLOCK": "192", "MTB_MINB": "3"}, {"MTB_BLOCK": "128"}):
    for k in ("MTB_PAIR", "MTB_BLOCK", "MTB_MINB"):
        os.environ.pop(k, None)
    os.environ.update(env)
    print(env, flush=True)
    qb.run(eng, "sf", 1024, 256, 20, 3, 12, w)
