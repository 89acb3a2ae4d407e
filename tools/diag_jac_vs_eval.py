"""Diagnostic: min-distance values of the Jacobian kernel vs the rows of evaluate_kernel / evaluate_pair_kernel (config-3 shape)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, scenarios
eng = engine.Engine("cuda:0")
s, c, n, p, t_hor = 1024, 2, 10, 4, 20
scene = scenarios.synthetic_scenes(s, n, p, regular_sigma=True).to("cuda:0")
controls = scenarios.candidate_controls(s, c, t_hor, u_max=1.0)
controls[:, 1] = controls[:, 0]
controls = controls.to("cuda:0")
spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True, ellipsoid=True)
jac, values = eng.constraint_jacobian(scene, controls, spec, want_values=True)
n_ctl, n_spd, n_md = 2 * t_hor, 2 * (t_hor + 1), n * t_hor
out = {}
for cc in (2, 1):
    ctl = controls[:, :cc].contiguous()
    for want in (("cons",), ("cons", "grad"), ("cons", "mus")):
        res = eng.evaluate(scene, ctl, spec, want=want)
        cons = res["cons"]
        d = (values[:, 0].reshape(s, -1) - cons[:, 0, n_ctl + n_spd:n_ctl + n_spd + n_md]).abs().max(dim=1).values
        out[(cc, want)] = res
        print(cc, want, eng.launch_info(scene, ctl, "grad" in want)["kernel_kind"], "max", float(d.max()), "scenes > 1e-5:", int((d > 1e-5).sum()),
              torch.nonzero(d > 1e-5).flatten().tolist()[:8])
a, b = out[(2, ("cons", "mus"))]["mus"][:, 0], out[(1, ("cons", "mus"))]["mus"][:, 0]
dm = (a - b).abs().amax(dim=(1, 2, 3))
print("pair vs single mus: max", float(dm.max()), "scenes differing:", int((dm > 0).sum()), torch.nonzero(dm > 0).flatten().tolist()[:8])
if int((dm > 0).sum()) > 0:
    sidx = int(torch.nonzero(dm > 0).flatten()[0])
    dt_ = (a[sidx] - b[sidx]).abs().amax(dim=(0, 2))
    print("first differing scene", sidx, "per-level max diff", dt_.tolist())
