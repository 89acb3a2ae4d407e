"""Tiny workloads of every kernel for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool racecheck python tools/sanitize_target.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, mpc, scenarios  # noqa: E402

if __name__ == "__main__":
    eng = engine.Engine("cuda:0")
    for env_kind, n, p, t_hor in (("sf", 20, 3, 12), ("sf", 10, 4, 20), ("pf", 5, 5, 15), ("sf", 3, 5, 10), ("kalman", 5, 1, 8)):
        s, c = 3, 5
        scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind, regular_sigma=True).to("cuda:0")
        controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
        mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
        spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0, "prob": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True,
                                 min_distance=True, ellipsoid=True)
        out = eng.evaluate(scene, controls, spec, want=("obj", "total", "grad", "cons", "violation", "mus", "sigmas"))
        eng.argmin(out["total"], out["violation"], controls)
        eng.constraint_jacobian(scene, controls, spec)
        mask = eng.attention("closest", scene.ego0, scene.ado0)
        loop = mpc.BatchedClosedLoop(eng, scene, t_planning=t_hor, num_candidates=c, attention="euclidean")
        loop.solve(2)
        torch.cuda.synchronize()
        print("ok", env_kind, n, p, t_hor, float(out["total"].mean()), int(mask.sum()), flush=True)
