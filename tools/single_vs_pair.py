import os, sys, ctypes
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
os.environ["MTB_PAIR"] = "1"
from mantrap_b200 import _cabi, build, engine
import quick_bench as qb
eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), "cuda:0")
w = {"goal": 0.5, "acc": 1.0}
qb.run(eng, "sf", 1024, 256, 20, 3, 12, w)
qb.run(eng, "sf", 1024, 256, 20, 3, 12, w, grad=False)
os.environ["MTB_PAIR"] = "0"
qb.run(eng, "sf", 1024, 256, 20, 3, 12, w)
