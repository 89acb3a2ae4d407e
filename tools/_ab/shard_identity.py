"""Development probe: is a scene shard evaluated alone bit-identical to the same scenes inside the whole batch?"""
import ctypes, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mantrap_b200 import _cabi, build, engine, scenarios
dev = "cuda:0"
eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), dev)
s, c, n, p, t_hor = 37, 19, 6, 3, 8
scene = scenarios.synthetic_scenes(s, n, p, seed=11, param_seed=12, regular_sigma=True).to(dev)
controls = scenarios.candidate_controls(s, c, t_hor, seed=13, u_max=1.0).to(dev)
def spec_of(sc):
    wo = eng.rollout_wo_ego(sc, t_hor)["mus"].contiguous()
    return engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=wo, control_limits=True, speed_limits=True), wo
sp, wo_full = spec_of(scene)
full = eng.evaluate(scene, controls, sp, want=("total", "violation", "obj"))
for rep in range(3):
    again = eng.evaluate(scene, controls, sp, want=("total", "violation", "obj"))
    print("repeat identical:", torch.equal(again["total"], full["total"]))
for lo, hi in ((0, 19), (19, 37)):
    from mantrap_b200 import parallel
    shard = parallel.slice_scene(scene, lo, hi)
    sps, wo = spec_of(shard)
    out = eng.evaluate(shard, controls[lo:hi].contiguous(), sps, want=("total", "violation", "obj"))
    d = (out["total"] - full["total"][lo:hi]).abs()
    print(os.environ.get("MTB_MINB", "rule"), (lo, hi), "mus_wo identical:", torch.equal(wo, wo_full[lo:hi]), "total identical:", torch.equal(out["total"], full["total"][lo:hi]),
          "max diff", float(d.max()), "n diff", int((d > 0).sum()), "obj diff", (out["obj"] - full["obj"][lo:hi]).abs().amax(dim=(0, 1)).tolist())
