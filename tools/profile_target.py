"""Small fixed workload for Nsight Compute captures: config-4 shape, 1024 scenes x 256 candidates, 3 launches."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, scenarios  # noqa: E402

if __name__ == "__main__":
    env_kind = sys.argv[1] if len(sys.argv) > 1 else "sf"
    s = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
    eng = engine.Engine("cuda:0")
    n, p, t_hor, c = (20, 3, 12, 256) if env_kind == "sf" else (5, 5, 15, 4096)
    if env_kind == "pf":
        s = max(1, s // 16)
    scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind).to("cuda:0")
    controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
    mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
    spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True)
    out = None
    for _ in range(3):
        out = eng.evaluate(scene, controls, spec, want=("total", "grad", "violation"), out=out)
        eng.argmin(out["total"], out["violation"], controls)
    torch.cuda.synchronize()
    print("ok", float(out["total"].mean()))
