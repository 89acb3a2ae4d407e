"""Wall clock of the UNMODIFIED reference (simon-schaefer/mantrap, imported from /root/reference) on the bench workload's shape:
one rollout + objective + gradient evaluation of goal + interaction-acceleration on the config-4 golden scene (SocialForces,
20 pedestrians x 3 particles, t_planning = 12), one process, one thread. Build container only (the reference cannot travel).

    python tools/time_reference.py [evaluations]
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import reference_harness as rh  # noqa: E402
import generate_golden as gg  # noqa: E402

if __name__ == "__main__":
    evaluations = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    torch.set_num_threads(1)
    mantrap = rh.import_reference()
    case = gg.CASES["c4_sf_n20_p3_t12"]
    from mantrap.environment.base.particle import ParticleEnvironment
    ParticleEnvironment._compute_distributions.__defaults__ = (case["P"], True)
    env = rh.build_env(case)
    goal = torch.tensor(case["ego_goal"], dtype=torch.float32)
    torch.manual_seed(case["seed_wo"])
    modules = [mantrap.modules.GoalNormModule(env=env, t_horizon=case["T"], goal=goal),
               mantrap.modules.baselines.InteractionAccelerationModule(env=env, t_horizon=case["T"], goal=goal)]
    rng = np.random.RandomState(0)
    times = []
    for k in range(evaluations + 1):
        controls = torch.tensor(rng.uniform(-1, 1, (case["T"], 2)), dtype=torch.float32, requires_grad=True)
        t0 = time.perf_counter()
        traj = env.ego.unroll_trajectory(controls, dt=env.dt)
        value = sum(m.objective(traj, ado_ids=env.ado_ids, tag="t") for m in modules)
        traj = env.ego.unroll_trajectory(controls, dt=env.dt)
        grad = sum(m.gradient(traj, grad_wrt=controls, ado_ids=env.ado_ids, tag="t") for m in modules)
        times.append(time.perf_counter() - t0)
        assert np.isfinite(value) and np.all(np.isfinite(grad))
    times = times[1:]                                    # first call: warm-up
    print(f"unmodified reference, config-4 shape (SF N=20 P=3 T=12), objective + gradient of goal + interaction acceleration: "
          f"{np.mean(times):.2f} s per evaluation (min {np.min(times):.2f}, max {np.max(times):.2f}, {evaluations} evaluations, "
          f"1 thread, asserts {"on" if __debug__ else "off (python -O)"}) = {1.0 / np.mean(times):.3f} evals/s/core")
