"""Latency of ONE solver callback through the drop-in API (BASELINE config 1: SocialForces, double-integrator ego, 3 ados,
t_planning = 10, goal + interaction-acceleration objective + control limits): what IPOPT pays per iteration.
Wall-clock per call including the host work of the call (particle draw on the CPU generator as the reference does, H2D of
the scene, one launch, D2H of the result). Development / measurement tool, not the bench contract."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mantrap_b200 as mantrap  # noqa: E402


def scene():
    env = mantrap.environment.SocialForcesEnvironment(ego_position=torch.tensor([-5.0, 2.0]), ego_velocity=torch.zeros(2),
                                                      ego_type=mantrap.agents.DoubleIntegratorDTAgent, dt=0.4)
    for pos, goal in (((3.0, 2.0), (-4.0, 2.0)), ((0.0, -2.0), (0.0, 4.0)), ((-1.0, 4.5), (-1.0, -4.0))):
        env.add_ado(position=torch.tensor(pos), velocity=torch.zeros(2), goal=torch.tensor(goal))
    return env, torch.tensor([1.0, 1.0])


if __name__ == "__main__":
    env, goal = scene()
    modules = [mantrap.modules.GoalNormModule, mantrap.modules.InteractionAccelerationModule, mantrap.modules.ControlLimitModule,
               mantrap.modules.MinDistanceModule]
    solver = mantrap.solver.IPOPTSolver(env, goal=goal, t_planning=10, modules=modules)
    rng = np.random.RandomState(0)
    zs = [rng.uniform(-1, 1, 20) for _ in range(220)]
    for name, fn in (("objective", solver.objective), ("gradient", solver.gradient), ("constraints", solver.constraints),
                     ("jacobian", solver.jacobian)):
        for z in zs[:20]:
            fn(z)
        solver._cache.clear()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for z in zs[20:]:
            solver._cache.clear()          # every call a fresh z: no per-z cache hit
            fn(z)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 200
        print(f"config 1 callback {name}: {dt * 1e6:.0f} us per call ({1.0 / dt:.0f} calls/s), one process, one GPU", flush=True)
    # where the time of one callback goes: particle draw (host, the reference's RNG stream) | staging into the pinned block |
    # upload + launch + read-back + synchronise
    prep = solver._prepared(solver.env.ado_ids)
    t_draw = t_stage = t_run = 0.0
    for z in zs[20:]:
        t0 = time.perf_counter()
        draw = solver.env.draw_particle_parameters(prep.num_particles, solver.planning_horizon - 1)
        t1 = time.perf_counter()
        prep.stage(np.asarray(z, dtype=np.float32), draw)
        t2 = time.perf_counter()
        prep.run()
        t3 = time.perf_counter()
        t_draw, t_stage, t_run = t_draw + t1 - t0, t_stage + t2 - t1, t_run + t3 - t2
    print(f"config 1 callback breakdown: particle draw {t_draw / 200 * 1e6:.0f} us | staging {t_stage / 200 * 1e6:.0f} us | "
          f"H2D + launch + D2H + sync {t_run / 200 * 1e6:.0f} us", flush=True)
    # all four callbacks at the same z (what one IPOPT iteration does): one launch + one Jacobian launch, cached per z
    t0 = time.perf_counter()
    for z in zs[20:]:
        solver.objective(z), solver.gradient(z), solver.constraints(z), solver.jacobian(z)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 200
    print(f"config 1 one IPOPT iteration (objective + gradient + constraints + jacobian at one z): {dt * 1e6:.0f} us", flush=True)
    zb = np.stack(zs[20:84])
    solver.evaluate_fused_batch(zb, want_jacobian=True)
    t0 = time.perf_counter()
    for _ in range(20):
        solver.evaluate_fused_batch(zb, want_jacobian=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 20
    print(f"config 1 batched callbacks: 64 callback sets (values + gradient + rows + Jacobian) in {dt * 1e6:.0f} us "
          f"({dt / 64 * 1e6:.1f} us per set)", flush=True)
