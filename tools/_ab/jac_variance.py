"""Per-launch times of the constraint-Jacobian kernel on one shape (development probe: run-to-run variance)."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mantrap_b200 import _cabi, build, engine, scenarios  # noqa: E402

if __name__ == "__main__":
    s, c, n, p, t_hor = (int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (1024, 16, 10, 4, 20)))
    prod = os.environ.get("MTB_AB_PROD", "") == "1"       # A/B: the production library (possibly built from another source state)
    eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.LIB_PATH if prod else build.build(dev_knobs=True))), "cuda:0")
    scene = scenarios.synthetic_scenes(s, n, p, env_kind="sf").to("cuda:0")
    controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True)
    eng.constraint_jacobian(scene, controls, spec)
    torch.cuda.synchronize()
    times = []
    for _ in range(12):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        eng.constraint_jacobian(scene, controls, spec)
        b.record()
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    print("prod" if prod else "dev", os.environ.get("MTB_JAC_BLOCK", "default"), " ".join(f"{x:.2f}" for x in times), flush=True)
