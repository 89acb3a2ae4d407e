"""Small fixed workload for Nsight Compute captures of the constraint-Jacobian kernel: config-3 shape (SocialForces N=10
P=4 T=20), 1024 scenes x 16 candidates, control + speed + min-distance rows, 3 launches."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, scenarios  # noqa: E402

if __name__ == "__main__":
    eng = engine.Engine("cuda:0")
    scene = scenarios.synthetic_scenes(1024, 10, 4, env_kind="sf").to("cuda:0")
    controls = scenarios.candidate_controls(1024, 16, 20).to("cuda:0")
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True)
    for _ in range(3):
        jac = eng.constraint_jacobian(scene, controls, spec)
    torch.cuda.synchronize()
    print("ok", float(jac.abs().mean()))
