"""Golden-vector generator: runs the UNMODIFIED reference (imported from /root/reference through
`oracle/reference_harness.py`) with `torch.manual_seed` pinned and stores inputs (including every random number
the reference drew) and outputs as small .npz fixtures next to this file.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/generate_golden.py

Each case stores
  inputs : ego0[4], ego_goal[2], ado0[N,4], goal[N,2], controls[T,2], dt, v0/sigma/weights[N,P], noise[T-1,N,P,2]
           (the conditioned draw, seed `seed`), and the same for the ego-free baseline draw (`*_wo`, seed `seed_wo`)
  outputs: ego_traj[T+1,5] (`unroll_trajectory`), mus[N,T,4] / sigmas[N,T,2] (`compute_distributions`),
           mus_wo / sigmas_wo (`compute_distributions_wo_ego`), and per module the values returned by the
           reference's `objective`, `gradient`, `constraint`, `jacobian` (weighted + normalised, as IPOPT sees them).
The particle count P is forced through `ParticleEnvironment._compute_distributions.__defaults__` because the
reference modules do not forward `num_particles` (the default is frozen at def-time, SURVEY §5 "Config").
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import reference_harness as rh  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def _ados(rows):
    return [dict(pos=r[0:2], vel=r[2:4], goal=r[4:6]) for r in rows]


def random_ados(n, seed, spacing=2.2, speed=1.25, box=8.0):
    """Rejection-sampled ado states with a minimum spacing so the scene is well conditioned in float32."""
    rng = np.random.RandomState(seed)
    rows = []
    while len(rows) < n:
        pos = rng.uniform(-box, box, 2)
        if all(np.linalg.norm(pos - np.array(r[0:2])) >= spacing for r in rows):
            vel = rng.uniform(0, speed, 2) * rng.choice([-1, 1], 2)
            goal = rng.uniform(-box, box, 2)
            rows.append([*pos.round(3), *vel.round(3), *goal.round(3)])
    return _ados(rows)


CASES = {
    # BASELINE config 1: test_solver-style scene extended to three ados, goal + acceleration modules.
    # Seeds of the two headline fixtures were screened (search_cases.py) so that the reference's interaction gradient is
    # NON-trivial: min(sigma) > 0.05, raw_acc < 1, |grad_acc|max > 1e-4 (asserted in generate()). The round-1 seeds, which
    # draw a sigma <= 0 (autograd NaN -> nan_to_num -> zero gradient), stay as the `*_sigma_neg` edge cases.
    "c1_sf_n3_t10": dict(env="sf", ego_kind="double", ego_pos=(-5, 2), ego_vel=(0, 0), ego_goal=(1, 1), dt=0.4,
                         T=10, P=5, seed=107, seed_wo=107, useed=1, uscale=2.0, modules=("goal", "acc"), require=("acc",),
                         ados=_ados([[3, 2, 0, 0, -4, 2], [0, -2, 0, 0, 0, 4], [-1, 4.5, 0, 0, -1, -4]])),
    "c1_sf_n3_t10_sigma_neg": dict(env="sf", ego_kind="double", ego_pos=(-5, 2), ego_vel=(0, 0), ego_goal=(1, 1), dt=0.4,
                                   T=10, P=5, seed=101, seed_wo=101, useed=1, uscale=2.0, modules=("goal", "acc"),
                                   ados=_ados([[3, 2, 0, 0, -4, 2], [0, -2, 0, 0, 0, 4], [-1, 4.5, 0, 0, -1, -4]])),
    # same scene with moving pedestrians and the ego passing close by: all three interaction variants.
    "sf_n3_t10_moving": dict(env="sf", ego_kind="double", ego_pos=(-4, 1), ego_vel=(1.0, 0), ego_goal=(4, 1),
                             dt=0.4, T=10, P=5, seed=202, seed_wo=202, useed=2, uscale=1.0,
                             modules=("goal", "acc", "vel", "pos"),
                             ados=_ados([[3, 2, -1.0, 0.1, -4, 2], [0, -2, 0.2, 0.9, 0, 4],
                                         [-1, 4.5, 0.1, -1.1, -1, -4]])),
    # BASELINE config 2: potential field, five ados, T = 15.
    "c2_pf_n5_t15": dict(env="pf", ego_kind="double", ego_pos=(-3, 0), ego_vel=(0.5, 0), ego_goal=(5, 0), dt=0.4,
                         T=15, P=5, seed=303, seed_wo=304, useed=3, uscale=1.0, modules=("goal", "acc", "vel"),
                         ados=_ados([[2, 0.5, -1.0, 0.0, 0, 0], [0, -2, 0.1, 1.0, 0, 0], [-1, 3, 0.2, -1.0, 0, 0],
                                     [4, -1, -0.8, 0.3, 0, 0], [1, 2, -0.3, -0.6, 0, 0]])),
    # BASELINE config 3: ten ados, four particles ("modes"), T = 20, position module + min-distance rows.
    "c3_sf_n10_p4_t20": dict(env="sf", ego_kind="double", ego_pos=(-6, 0), ego_vel=(1.0, 0.2), ego_goal=(6, 1),
                             dt=0.4, T=20, P=4, seed=404, seed_wo=404, useed=4, uscale=0.6,
                             modules=("goal", "pos", "acc", "min_distance"), require=("acc",), ados=random_ados(10, 44)),
    # single-integrator ego, six ados.
    "sf_single_n6_t12": dict(env="sf", ego_kind="single", ego_pos=(-5, -1), ego_vel=(0, 0), ego_goal=(5, 2),
                             dt=0.4, T=12, P=5, seed=505, seed_wo=505, useed=5, uscale=1.5,
                             modules=("goal", "vel", "acc"), ados=random_ados(6, 55)),
    "pf_single_n4_t8": dict(env="pf", ego_kind="single", ego_pos=(0, 0), ego_vel=(0, 0), ego_goal=(5, 2),
                            dt=0.4, T=8, P=3, seed=606, seed_wo=607, useed=6, uscale=1.5,
                            modules=("goal", "acc", "pos"), ados=random_ados(4, 66, box=4.0)),
    # dense crowd, short horizon: ado-ado repulsion, force clamps and the L2 control clamp are active.
    "sf_dense_n5_t6": dict(env="sf", ego_kind="double", ego_pos=(-1.5, 0.3), ego_vel=(1.0, 0), ego_goal=(4, 0),
                           dt=0.4, T=6, P=5, seed=707, seed_wo=707, useed=7, uscale=1.0,
                           modules=("goal", "acc", "vel"),
                           ados=_ados([[0.6, 0.4, -0.8, 0, -4, 0], [0.2, -0.7, 0.1, 0.7, 0, 4],
                                       [1.4, 1.2, -0.3, -0.6, 0, -4], [1.1, -0.2, -0.9, 0.2, -5, 0],
                                       [-0.4, 1.1, 0.5, -0.5, 3, -3]])),
    # ETH-shaped (config 4): 20 pedestrians, three particles, T = 12.
    "c4_sf_n20_p3_t12": dict(env="sf", ego_kind="double", ego_pos=(-7, -6), ego_vel=(0.8, 0.6), ego_goal=(7, 6),
                             dt=0.4, T=12, P=3, seed=800, seed_wo=800, useed=8, uscale=1.0, require=("acc",),
                             modules=("goal", "acc"), ados=random_ados(20, 88, spacing=1.5, box=9.0)),
    "c4_sf_n20_p3_t12_sigma_neg": dict(env="sf", ego_kind="double", ego_pos=(-7, -6), ego_vel=(0.8, 0.6), ego_goal=(7, 6),
                                       dt=0.4, T=12, P=3, seed=808, seed_wo=808, useed=8, uscale=1.0,
                                       modules=("goal", "acc"), ados=random_ados(20, 88, spacing=1.5, box=9.0)),
    # Probability-objective fixtures: moving pedestrians (the particle weights spread the weighted
    # velocities, so the variances are ~1e-2 instead of 1e-15), short horizons and an ego a few metres away keep the raw
    # cost inside the +-120 clamp, i.e. the adjoint through means AND variances is compared with the reference's autograd.
    "sf_prob_n4_t6": dict(env="sf", ego_kind="double", ego_pos=(-4.5, -3.0), ego_vel=(0.8, 0.4), ego_goal=(5, 3), dt=0.4,
                          T=6, P=5, seed=1105, seed_wo=1105, useed=11, uscale=1.0, modules=("goal", "acc", "vel"), require=("acc", "prob"),
                          ados=_ados([[1.5, 2.0, -1.0, -0.3, -6, 0], [-1.0, 3.5, 0.6, -0.9, 2, -5],
                                      [3.0, -1.5, -0.7, 0.8, -2, 5], [0.0, -4.5, 0.9, 0.5, 6, 1]])),
    "sf_prob_single_n6_t8": dict(env="sf", ego_kind="single", ego_pos=(-6.0, 0.5), ego_vel=(0, 0), ego_goal=(6, 0), dt=0.4,
                                 T=8, P=4, seed=1102, seed_wo=1102, useed=12, uscale=1.2, modules=("goal", "acc", "pos"), require=("acc", "prob"),
                                 ados=_ados([[-1.0, 3.0, 0.9, -0.6, -5, -4], [2.0, 2.5, -0.8, -0.7, 6, -2], [0.5, -3.0, 0.7, 0.9, -5, 2],
                                             [3.5, -1.0, -1.0, 0.4, 3, 6], [-3.0, -2.5, 1.0, 0.5, -4, 5], [5.0, 3.5, -0.9, -0.8, 6, -3]])),
    "pf_prob_n4_t6": dict(env="pf", ego_kind="double", ego_pos=(7.0, 6.5), ego_vel=(-0.2, -0.1), ego_goal=(-5, -4), dt=0.4,
                          T=6, P=5, seed=1305, seed_wo=1305, useed=13, uscale=1.0, modules=("goal", "acc", "vel"), require=("acc", "prob"),
                          ados=_ados([[-0.5, 0.0, 0.8, 0.7, 0, 0], [0.5, -1.0, 0.6, 0.8, 0, 0], [-1.0, 1.0, 0.9, 0.6, 0, 0],
                                      [1.0, 0.5, 0.7, 0.7, 0, 0]])),
    "kalman_n5_t15": dict(env="kalman", ego_kind="double", ego_pos=(-3, 0), ego_vel=(0.5, 0), ego_goal=(5, 0),
                          dt=0.4, T=15, P=5, seed=909, seed_wo=909, useed=9, uscale=1.0, modules=("goal", "acc"),
                          ados=random_ados(5, 99)),
}


def generate(name, case):
    mantrap = rh.import_reference()
    from mantrap.environment.base.particle import ParticleEnvironment
    ParticleEnvironment._compute_distributions.__defaults__ = (case["P"], True)

    env = rh.build_env(case)
    n, t_hor, num_p = len(case["ados"]), case["T"], case["P"]
    rng = np.random.RandomState(case["useed"])
    controls = torch.tensor(rng.uniform(-case["uscale"], case["uscale"], (t_hor, 2)), dtype=torch.float32)
    goal = torch.tensor(case["ego_goal"], dtype=torch.float32)

    out = dict(env=case["env"], ego_kind=case["ego_kind"], dt=np.float32(case["dt"]), T=t_hor, P=num_p,
               ego0=np.array([*case["ego_pos"], *case["ego_vel"]], dtype=np.float32),
               ego_goal=np.array(case["ego_goal"], dtype=np.float32),
               ado0=np.array([[*a["pos"], *a["vel"]] for a in case["ados"]], dtype=np.float32),
               goal=np.array([a["goal"] for a in case["ados"]], dtype=np.float32),
               controls=rh.to_np(controls), seed=case["seed"], seed_wo=case["seed_wo"])

    ego_traj = env.ego.unroll_trajectory(controls, dt=env.dt)
    out["ego_traj"] = rh.to_np(ego_traj)
    out["dx_du"] = rh.to_np(env.ego.dx_du(controls, dt=env.dt))

    if case["env"] != "kalman":
        for tag, seed in (("", case["seed"]), ("_wo", case["seed_wo"])):
            v0, sigma, w, noise = rh.redraw_params(case["env"], n, num_p, t_hor - 1, seed)
            out["v0" + tag], out["sigma" + tag] = rh.to_np(v0), rh.to_np(sigma)
            out["weights" + tag], out["noise" + tag] = rh.to_np(w), rh.to_np(noise)
    mus, sig, _ = rh.ref_rollout(env, ego_traj, seed=case["seed"], num_particles=num_p)
    out["mus"], out["sigmas"] = rh.to_np(mus), rh.to_np(sig)
    mus_wo, sig_wo = rh.ref_rollout_wo_ego(env, t_hor, seed=case["seed_wo"], num_particles=num_p)
    out["mus_wo"], out["sigmas_wo"] = rh.to_np(mus_wo), rh.to_np(sig_wo)

    module_classes = {"goal": mantrap.modules.GoalNormModule,
                      "prob": mantrap.modules.InteractionProbabilityModule,
                      "acc": mantrap.modules.baselines.InteractionAccelerationModule,
                      "vel": mantrap.modules.baselines.InteractionVelocityModule,
                      "pos": mantrap.modules.baselines.InteractionPositionModule}
    ado_ids = env.ado_ids
    for key in (*case["modules"], "prob"):      # the probability module (TrajOptSolver.module_defaults) on every scene
        if key == "min_distance":
            continue
        torch.manual_seed(case["seed_wo"])
        module = module_classes[key](env=env, t_horizon=t_hor, goal=goal)
        # objective: fresh trajectory, pinned draw.
        traj = env.ego.unroll_trajectory(controls, dt=env.dt)
        torch.manual_seed(case["seed"])
        out[f"obj_{key}"] = np.float64(module.objective(traj, ado_ids=ado_ids, tag="g"))
        u = controls.clone().requires_grad_(True)
        traj = env.ego.unroll_trajectory(u, dt=env.dt)
        torch.manual_seed(case["seed"])
        out[f"grad_{key}"] = np.asarray(module.gradient(traj, grad_wrt=u, ado_ids=ado_ids, tag="g"), dtype=np.float64)
        if key == "prob":
            # un-clamped cost: minus the mean over ados of the summed log-probabilities (prob_interact.py:74-78).
            torch.manual_seed(case["seed"])
            dist = env.compute_distributions(traj)
            lp = sum(torch.sum(dist[i].log_prob(module._dist_un_conditioned[i].mean)) for i in ado_ids)
            out["raw_prob"] = np.float64((-lp / n).item())
        elif key != "goal":
            # un-clamped cost, to check gradients on unsaturated evaluations (SURVEY §8d "Parity gate").
            torch.manual_seed(case["seed"])
            summary = module.summarize_distribution(traj)
            wo = module._ado_accelerations_wo
            cost = sum(torch.mean(torch.norm(summary[m] - wo[m], dim=1)) for m in range(n)) / n
            out[f"raw_{key}"] = np.float64(cost.item())
        # attention subset: only the first two ados enter the cost (acc_interact.py:53-65).
        if key != "goal" and n >= 3:
            sub = ado_ids[:2]
            traj = env.ego.unroll_trajectory(controls, dt=env.dt)
            torch.manual_seed(case["seed"])
            out[f"obj_{key}_sub2"] = np.float64(module.objective(traj, ado_ids=sub, tag="g"))
            u = controls.clone().requires_grad_(True)
            traj = env.ego.unroll_trajectory(u, dt=env.dt)
            torch.manual_seed(case["seed"])
            out[f"grad_{key}_sub2"] = np.asarray(module.gradient(traj, grad_wrt=u, ado_ids=sub, tag="g"),
                                                 dtype=np.float64)

    # constraint modules (deterministic).
    for key, cls in (("control_limits", mantrap.modules.ControlLimitModule),
                     ("speed_limits", mantrap.modules.SpeedLimitModule)):
        module = cls(env=env, t_horizon=t_hor)
        u = controls.clone().requires_grad_(True)
        traj = env.ego.unroll_trajectory(u, dt=env.dt)
        out[f"con_{key}"] = np.asarray(module.constraint(traj, ado_ids=ado_ids, tag="g"), dtype=np.float64)
        out[f"jac_{key}"] = np.asarray(module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids, tag="g"), dtype=np.float64)
        out[f"vio_{key}"] = np.float64(module.compute_violation(traj, ado_ids=ado_ids, tag="g"))
        lower, upper = module.constraint_boundaries(ado_ids=ado_ids)
        out[f"lb_{key}"] = np.array([np.nan if x is None else x for x in lower], dtype=np.float64)
        out[f"ub_{key}"] = np.array([np.nan if x is None else x for x in upper], dtype=np.float64)

    # InteractionEllipsoidModule (modules/baselines/ellipsoid_interact.py): one particle draw per call, pinned.
    module = mantrap.modules.baselines.InteractionEllipsoidModule(env=env, t_horizon=t_hor)
    u = controls.clone().requires_grad_(True)
    traj = env.ego.unroll_trajectory(u, dt=env.dt)
    torch.manual_seed(case["seed"])
    out["con_ellipsoid"] = np.asarray(module.constraint(traj, ado_ids=ado_ids, tag="g"), dtype=np.float64)
    torch.manual_seed(case["seed"])
    out["jac_ellipsoid"] = np.asarray(module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids, tag="g"), dtype=np.float64)
    torch.manual_seed(case["seed"])
    out["vio_ellipsoid"] = np.float64(module.compute_violation(traj, ado_ids=ado_ids, tag="g"))
    lower, upper = module.constraint_boundaries(ado_ids=ado_ids)
    out["lb_ellipsoid"] = np.array([np.nan if x is None else x for x in lower], dtype=np.float64)
    out["ub_ellipsoid"] = np.array([np.nan if x is None else x for x in upper], dtype=np.float64)

    # attention filters (mantrap/attention): selected ado indices at the current states.
    for key, cls in (("euclidean", mantrap.attention.EuclideanModule), ("closest", mantrap.attention.ClosestModule)):
        out[f"att_{key}"] = np.asarray(cls(env=env, t_horizon=t_hor)._compute(), dtype=np.int64)

    # MinDistance: the module does not exist at this commit. Compose reference pieces only:
    # reference positional rollout + torch.norm + OptimizationModule.compute_gradient_auto_grad.
    if "min_distance" in case["modules"]:
        module = mantrap.modules.ControlLimitModule(env=env, t_horizon=t_hor)   # only for its autograd helper
        u = controls.clone().requires_grad_(True)
        traj = env.ego.unroll_trajectory(u, dt=env.dt)
        torch.manual_seed(case["seed"])
        dist = env.compute_distributions(traj, vel_dist=False, num_particles=num_p)
        mu_p = torch.stack([dist[i].mean[:, 0, :] for i in ado_ids])              # [N,T,2]
        g = torch.norm(traj[None, :t_hor, 0:2] - mu_p, dim=-1).flatten()
        out["con_min_distance"] = rh.to_np(g).astype(np.float64)
        out["jac_min_distance"] = np.asarray(module.compute_gradient_auto_grad(g, grad_wrt=u), dtype=np.float64)

    # A "main" fixture must pin a NON-trivial reference gradient of the modules it names in `require` (unsaturated cost,
    # no sampled sigma <= 0): refuse to write one that would compare zeros with zeros.
    for key in case.get("require", ()):
        limit = 120.0 if key == "prob" else {"acc": 1.0, "vel": 1.0, "pos": 3.0}[key]
        gmax = float(np.abs(out[f"grad_{key}"]).max())
        assert abs(float(out[f"raw_{key}"])) < limit, (name, key, "saturated", float(out[f"raw_{key}"]))
        assert gmax > 1e-4, (name, key, "reference gradient is (almost) zero", gmax)
        if case["env"] == "sf":
            assert float(min(out["sigma"].min(), out["sigma_wo"].min())) > 0.05, (name, "sampled sigma <= 0.05")
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
    print(f"{name}: " + ", ".join(f"{k}={out[k]:.6g}" for k in out if k.startswith(("obj_", "raw_"))) + " | " +
          ", ".join(f"|{k}|max={np.abs(out[k]).max():.3g}" for k in out if k.startswith("grad_") and not k.endswith("_sub2")))


def generate_agent_kats():
    """Replay of the reference's dynamics known-answer tests (test/test_agents.py:176-194, :216-235, :241-260)
    through the reference agents, stored so the GPU box can check the CUDA ego rollout against them."""
    mantrap = rh.import_reference()
    rows = []
    for kind, cls in (("single", mantrap.agents.IntegratorDTAgent), ("double", mantrap.agents.DoubleIntegratorDTAgent)):
        for pos, vel, ctrl in (((1, 0), (0, 0), (0, 0)), ((1, 0), (2, 3), (0, 0)), ((1, 0), (0, 0), (2, 3))):
            agent = cls(position=torch.tensor(pos), velocity=torch.tensor(vel))
            nxt = agent.dynamics(state=agent.state_with_time, action=torch.tensor(ctrl).float(), dt=1.0)
            rows.append(dict(kind=kind, x0=[*pos, *vel], u=ctrl, x1=rh.to_np(nxt[0:4]).tolist()))
    expected = {("single", 1): [1, 0, 0, 0], ("single", 2): [3, 3, 2, 3], ("double", 1): [3, 3, 2, 3],
                ("double", 2): [1, 0, 2, 3]}
    for (kind, idx), x1 in expected.items():
        row = [r for r in rows if r["kind"] == kind][idx]
        assert np.allclose(row["x1"], x1), (kind, idx, row)
    rng = np.random.RandomState(0)
    unroll = {}
    for kind, cls in (("single", mantrap.agents.IntegratorDTAgent), ("double", mantrap.agents.DoubleIntegratorDTAgent)):
        for t_hor in (1, 5, 20, 33):
            x0 = rng.uniform(-3, 3, 4).astype(np.float32)
            u = rng.uniform(-2, 2, (t_hor, 2)).astype(np.float32)
            agent = cls(position=torch.tensor(x0[0:2]), velocity=torch.tensor(x0[2:4]), dt=0.4)
            traj = agent.unroll_trajectory(torch.tensor(u), dt=0.4)
            unroll[f"{kind}_{t_hor}_x0"], unroll[f"{kind}_{t_hor}_u"] = x0, u
            unroll[f"{kind}_{t_hor}_traj"] = rh.to_np(traj)
            unroll[f"{kind}_{t_hor}_dxdu"] = rh.to_np(agent.dx_du(torch.tensor(u), dt=0.4))
    np.savez_compressed(os.path.join(HERE, "agent_kats.npz"),
                        kat_kind=np.array([r["kind"] for r in rows]), kat_x0=np.array([r["x0"] for r in rows], dtype=np.float32),
                        kat_u=np.array([r["u"] for r in rows], dtype=np.float32),
                        kat_x1=np.array([r["x1"] for r in rows], dtype=np.float32), **unroll)
    print("agent_kats: ok")


def generate_scene_ops():
    """Attention filters and the closed-loop environment step of the reference on random scenes -> scene_ops.npz.

    attention: 48 scenes (N = 1..8 ados within a few metres of the ego so that the 4 m radius and the `~argmin` quirk of
    closest.py:27 are exercised), outputs = 0/1 masks from `EuclideanModule._compute` / `ClosestModule._compute`.
    step: 6 scenes x 5 consecutive `env.step(action)` calls (graph_based.py:82-124) under a pinned seed; the standard-normal
    draws behind `Normal.sample` are re-drawn with the same generator sequence (particle parameters first)
    and stored, so the device step can be fed the identical numbers."""
    mantrap = rh.import_reference()
    rng = np.random.RandomState(2024)
    att = dict(att_n=[], att_ego=[], att_ado=[], att_euclidean=[], att_closest=[])
    for case_id in range(48):
        n = 1 + case_id % 8
        ego = np.array([*rng.uniform(-3, 3, 2), 0.0, 0.0], dtype=np.float32)
        ados = np.zeros((8, 4), dtype=np.float32)
        ados[:n, 0:2] = ego[0:2] + rng.uniform(-6, 6, (n, 2))
        if case_id % 6 == 5 and n >= 2:
            ados[1, 0:2] = 2 * ego[0:2] - ados[0, 0:2]      # exact tie of the two smallest distances
        scene = dict(env="pf", ego_kind="double", ego_pos=ego[0:2].tolist(), ego_vel=(0, 0), dt=0.4,
                     ados=[dict(pos=a[0:2].tolist(), vel=(0, 0), goal=(0, 0)) for a in ados[:n]])
        env = rh.build_env(scene)
        masks = {}
        for key, cls in (("euclidean", mantrap.attention.EuclideanModule), ("closest", mantrap.attention.ClosestModule)):
            idx = np.asarray(cls(env=env, t_horizon=5)._compute(), dtype=np.int64)
            masks[key] = np.zeros(8, dtype=np.uint8)
            masks[key][idx] = 1
        att["att_n"].append(n)
        att["att_ego"].append(ego)
        att["att_ado"].append(ados)
        att["att_euclidean"].append(masks["euclidean"])
        att["att_closest"].append(masks["closest"])
    out = {k: np.stack(v) if k != "att_n" else np.array(v, dtype=np.int64) for k, v in att.items()}

    steps = 5
    step = dict(step_env=[], step_ego_kind=[], step_n=[], step_seed=[], step_ego0=[], step_ado0=[], step_actions=[],
                step_eps=[], step_ego_states=[], step_ado_states=[])
    for case_id, (env_kind, ego_kind, n) in enumerate((("sf", "double", 3), ("sf", "single", 4), ("pf", "double", 5),
                                                       ("pf", "single", 2), ("kalman", "double", 3), ("sf", "double", 6))):
        ego = np.array([*rng.uniform(-4, 4, 2), *rng.uniform(-1, 1, 2)], dtype=np.float32)
        ados = np.zeros((n, 6), dtype=np.float32)
        ados[:, 0:2] = rng.uniform(-8, 8, (n, 2))
        ados[:, 2:4] = rng.uniform(-1.5, 1.5, (n, 2))
        ados[:, 4:6] = rng.uniform(-8, 8, (n, 2))
        if case_id == 5:
            ados[0, 2:4] = (2.4, 1.9)                          # faster than PED_SPEED_MAX: the update clamps the action
        scene = dict(env=env_kind, ego_kind=ego_kind, ego_pos=ego[0:2].tolist(), ego_vel=ego[2:4].tolist(), dt=0.4,
                     ados=[dict(pos=a[0:2].tolist(), vel=a[2:4].tolist(), goal=a[4:6].tolist()) for a in ados])
        env = rh.build_env(scene)
        actions = rng.uniform(-3.0 if case_id == 5 else -2.0, 3.0 if case_id == 5 else 2.0, (steps, 2)).astype(np.float32)
        seed = 4000 + case_id
        eps_all, ego_states, ado_states = [], [], []
        for k in range(steps):
            # the draws of this step: particle parameters of the one-control rollout first, then one Normal.sample per ado
            if env_kind != "kalman":
                rh.redraw_params(env_kind, n, 5, 0, seed + k)
            else:
                torch.manual_seed(seed + k)
            eps = torch.stack([torch.randn((1, 1, 1, 2)) for _ in range(n)]).view(n, 2)
            torch.manual_seed(seed + k)
            ado_next, ego_next = env.step(torch.tensor(actions[k]))
            eps_all.append(rh.to_np(eps))
            ego_states.append(rh.to_np(ego_next)[0:4])
            ado_states.append(rh.to_np(ado_next)[:, 0:4])
        for key, val in (("step_env", env_kind), ("step_ego_kind", ego_kind), ("step_n", n), ("step_seed", seed),
                         ("step_ego0", ego), ("step_ado0", np.pad(ados[:, 0:4], ((0, 6 - n), (0, 0)))),
                         ("step_actions", actions), ("step_eps", np.pad(np.stack(eps_all), ((0, 0), (0, 6 - n), (0, 0)))),
                         ("step_ego_states", np.stack(ego_states)),
                         ("step_ado_states", np.pad(np.stack(ado_states), ((0, 0), (0, 6 - n), (0, 0))))):
            step[key].append(val)
    out.update({k: np.array(v) if isinstance(v[0], (str, int)) else np.stack(v) for k, v in step.items()})

    # scene encoding + warm start from the reference's pre-computed database (trajopt.py:466-499, :308-331). The database
    # tensors (third_party/warm_start/encoding.pt [100,4], solution.pt [100,10,2]) are recorded as inputs of the fixture.
    ws = dict(ws_n=[], ws_ego=[], ws_goal=[], ws_ado=[], ws_T=[], ws_encoding=[], ws_match=[], ws_warm=[])
    enc_db = torch.load(os.path.join(rh.REFERENCE_ROOT, "third_party", "warm_start", "encoding.pt"))
    sol_db = torch.load(os.path.join(rh.REFERENCE_ROOT, "third_party", "warm_start", "solution.pt"))
    cwd = os.getcwd()
    for case_id in range(40):
        n, t_hor = 1 + case_id % 8, (3, 5, 8, 10)[case_id % 4]
        ego = np.array([*rng.uniform(-8, 8, 2), *rng.uniform(-2, 2, 2)], dtype=np.float32)
        goal = rng.uniform(-8, 8, 2).astype(np.float32)
        ados = np.zeros((8, 4), dtype=np.float32)
        ados[:n, 0:2] = rng.uniform(-9, 9, (n, 2))
        scene = dict(env="pf", ego_kind="double", ego_pos=ego[0:2].tolist(), ego_vel=ego[2:4].tolist(), dt=0.4,
                     ados=[dict(pos=a[0:2].tolist(), vel=(0, 0), goal=(0, 0)) for a in ados[:n]])
        env = rh.build_env(scene)
        solver = mantrap.solver.IPOPTSolver(env, goal=torch.tensor(goal), t_planning=t_hor, modules=[mantrap.modules.GoalNormModule])
        encoding = solver.encode()
        warm = solver.warm_start(method=mantrap.constants.WARM_START_ENCODING)
        match = int(torch.argmin(torch.norm(enc_db - encoding, dim=1)))
        assert torch.equal(warm, sol_db[match, :t_hor])
        warm_pad = np.zeros((10, 2), dtype=np.float32)
        warm_pad[:t_hor] = rh.to_np(warm)
        for key, val in (("ws_n", n), ("ws_ego", ego), ("ws_goal", goal), ("ws_ado", ados), ("ws_T", t_hor),
                         ("ws_encoding", rh.to_np(encoding)), ("ws_match", match), ("ws_warm", warm_pad)):
            ws[key].append(val)
    os.chdir(cwd)
    out.update({k: np.array(v) if isinstance(v[0], int) else np.stack(v) for k, v in ws.items()})
    out["ws_encoding_db"], out["ws_solution_db"] = rh.to_np(enc_db), rh.to_np(sol_db)
    np.savez_compressed(os.path.join(HERE, "scene_ops.npz"), **out)
    print("scene_ops: attention selected", int(out["att_euclidean"].sum()), "/", int(out["att_closest"].sum()),
          "steps", out["step_ego_states"].shape)


if __name__ == "__main__":
    only = sys.argv[1:]
    if not only or "scene_ops" in only:
        generate_scene_ops()
    if not only or "agent_kats" in only:
        generate_agent_kats()
    for case_name, case_spec in CASES.items():
        if not only or case_name in only:
            generate(case_name, case_spec)
