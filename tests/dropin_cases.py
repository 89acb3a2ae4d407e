"""Shared bodies of the drop-in API tests (agents / environment / modules / solver mirrors of the reference interface).
They run on the CPU against an Engine bound to the host emulation of the kernel source (tests/test_dropin_emulation.py)
and on the B200 against the real CUDA engine (tests/test_gpu_dropin.py, -m gpu). Expected values are the golden vectors
recorded from the unmodified reference with the torch RNG pinned (tests/golden/generate_golden.py)."""
import numpy as np
import torch

import golden_io as gio
from mantrap_b200 import agents, environment, modules, solver

ENVS = {"sf": environment.SocialForcesEnvironment, "pf": environment.PotentialFieldEnvironment,
        "kalman": environment.KalmanEnvironment}
EGOS = {"double": agents.DoubleIntegratorDTAgent, "single": agents.IntegratorDTAgent}
MODULES = {"goal": modules.GoalNormModule, "acc": modules.InteractionAccelerationModule,
           "vel": modules.InteractionVelocityModule, "pos": modules.InteractionPositionModule,
           "prob": modules.InteractionProbabilityModule}


def build_env(g):
    """The golden scene through the mirrored environment API (same calls as oracle/reference_harness.build_env)."""
    env = ENVS[str(g["env"])](ego_position=gio.t(g["ego0"][0:2]), ego_velocity=gio.t(g["ego0"][2:4]),
                             ego_type=EGOS[str(g["ego_kind"])], dt=float(g["dt"]))
    for i in range(g["ado0"].shape[0]):
        kwargs = {"goal": gio.t(g["goal"][i])} if str(g["env"]) == "sf" else {}
        env.add_ado(position=gio.t(g["ado0"][i, 0:2]), velocity=gio.t(g["ado0"][i, 2:4]), identifier=f"a{i:02d}", **kwargs)
    return env


def check_agents():
    k = gio.load("agent_kats")
    for kind, x0, u, x1 in zip(k["kat_kind"], k["kat_x0"], k["kat_u"], k["kat_x1"]):      # test_agents.py:176-260
        agent = EGOS[str(kind)](position=gio.t(x0[0:2]), velocity=gio.t(x0[2:4]))
        nxt = agent.dynamics(agent.state_with_time, gio.t(u), dt=1.0)
        gio.assert_parity(nxt[0:4].numpy(), x1, 1e-6, f"KAT {kind}")
        assert float(nxt[4]) == 1.0
    for kind in ("single", "double"):
        for t_hor in (1, 5, 20, 33):
            x0, u = k[f"{kind}_{t_hor}_x0"], gio.t(k[f"{kind}_{t_hor}_u"])
            agent = EGOS[kind](position=gio.t(x0[0:2]), velocity=gio.t(x0[2:4]), dt=0.4)
            traj = agent.unroll_trajectory(u, dt=0.4)
            gio.assert_parity(traj.numpy(), k[f"{kind}_{t_hor}_traj"], gio.RTOL_ROLLOUT, f"unroll {kind} {t_hor}")
            gio.assert_parity(agent.dx_du(u, dt=0.4).numpy(), k[f"{kind}_{t_hor}_dxdu"], 1e-6, f"dx_du {kind} {t_hor}")
            back = agent.roll_trajectory(traj, dt=0.4)                                        # test_agents.py:41-45
            gio.assert_parity(back.numpy(), u.numpy(), 1e-4, f"roll(unroll) {kind} {t_hor}", atol=1e-5)
            # autograd through the CUDA unroll == dx_du (test_agents.py:143-170)
            uu = u.clone().requires_grad_(True)
            tr = agent.unroll_trajectory(uu, dt=0.4)
            coeff = torch.linspace(-1, 1, tr.numel()).view_as(tr)
            (grad,) = torch.autograd.grad((tr * coeff).sum(), uu)
            expect = coeff.flatten() @ gio.t(k[f"{kind}_{t_hor}_dxdu"])
            gio.assert_parity(grad.flatten().numpy(), expect.numpy(), 1e-5, f"unroll adjoint {kind} {t_hor}", atol=1e-6)
    # update / update_inverse / feasibility
    agent = agents.IntegratorDTAgent(position=torch.zeros(2), velocity=torch.zeros(2))
    action, state = agent.update(torch.tensor([10.0, 0.0]), dt=0.5)                          # clamped to PED_SPEED_MAX
    assert np.allclose(action.numpy(), [2.5, 0.0]) and np.allclose(state.numpy(), [1.25, 0, 2.5, 0, 0.5])
    assert agent.history.shape == (2, 5)
    robot = agents.DoubleIntegratorDTAgent(position=torch.zeros(2), velocity=torch.zeros(2), is_robot=True)
    assert robot.control_limits() == (-2.0, 2.0) and robot.speed_limits == (-2.0, 2.0)
    assert robot.check_feasibility_controls(torch.tensor([[2.0, -2.0]])) and not robot.check_feasibility_controls(torch.tensor([[2.1, 0.0]]))


def check_environment(name):
    """compute_distributions (with / without ego, velocity / position) reproduces the reference after the same
    torch.manual_seed: the RNG mirror, the rollout kernel and the distribution packaging at once."""
    g = gio.load(name)
    env = build_env(g)
    num_p, t_hor = int(g["P"]), int(g["T"])
    kw = {"num_particles": num_p} if str(g["env"]) != "kalman" else {}
    traj = env.ego.unroll_trajectory(gio.t(g["controls"]), dt=env.dt)
    gio.assert_parity(traj.numpy(), g["ego_traj"], gio.RTOL_ROLLOUT, "ego trajectory")
    for vel_dist, cols in ((True, slice(2, 4)), (False, slice(0, 2))):
        torch.manual_seed(int(g["seed"]))
        dist = env.compute_distributions(traj, vel_dist=vel_dist, **kw)
        assert list(dist.keys()) == env.ado_ids
        mean = torch.stack([dist[i].mean for i in env.ado_ids])
        assert mean.shape == (env.num_ados, t_hor, 1, 2)
        gio.assert_parity(mean[:, :, 0].numpy(), g["mus"][..., cols], gio.RTOL_ROLLOUT, f"mean vel_dist={vel_dist}")
        if vel_dist:
            gio.assert_parity(torch.stack([dist[i].stddev for i in env.ado_ids])[:, :, 0].numpy(), g["sigmas"],
                              gio.RTOL_ROLLOUT, "scale")
    torch.manual_seed(int(g["seed_wo"]))
    dist = env.compute_distributions_wo_ego(t_hor, **kw)
    gio.assert_parity(torch.stack([dist[i].mean for i in env.ado_ids])[:, :, 0].numpy(), g["mus_wo"][..., 2:4],
                      gio.RTOL_ROLLOUT, "mean wo ego")
    # prediction API shapes (test_environment.py:16-179): positions incl. the current one
    torch.manual_seed(int(g["seed"]))
    pred = env.predict_w_controls(gio.t(g["controls"]))
    assert pred.shape == (env.num_ados, t_hor + 1, 1, 2)
    if num_p == 5 or str(g["env"]) == "kalman":     # predict_* always use the default particle count (graph_based.py:253)
        gio.assert_parity(pred[:, :t_hor, 0].numpy(), g["mus"][..., 0:2], gio.RTOL_ROLLOUT, "predicted positions", atol=1e-6)
    samples = env.sample_w_trajectory(traj, num_samples=3)
    assert samples.shape == (env.num_ados, 3, t_hor + 1, 1, 2)
    assert env.predict_wo_ego(t_hor).shape == (env.num_ados, t_hor + 1, 1, 2)
    assert env.copy().same_initial_conditions(env)


def check_modules(name):
    """module.objective / gradient / constraint / jacobian with the reference's call sequence and seeds."""
    g = gio.load(name)
    env = build_env(g)
    num_p, t_hor = int(g["P"]), int(g["T"])
    goal, controls, ado_ids = gio.t(g["ego_goal"]), gio.t(g["controls"]), env.ado_ids
    for key, cls in MODULES.items():
        if f"obj_{key}" not in g:
            continue
        torch.manual_seed(int(g["seed_wo"]))
        module = cls(env=env, t_horizon=t_hor, goal=goal, num_particles=num_p)
        for suffix, ids in (("", ado_ids), ("_sub2", ado_ids[:2])):
            if f"obj_{key}{suffix}" not in g:
                continue
            traj = env.ego.unroll_trajectory(controls, dt=env.dt)
            torch.manual_seed(int(g["seed"]))
            obj = module.objective(traj, ado_ids=ids, tag="g")
            f_obj, f_grad = (0.0, 0.0)
            if key == "prob":      # float32 conditioning of the variance-scaled log-probability (kernel_cases.prob_floor);
                import kernel_cases as kc   # here BOTH rollouts (with and without ego) are this implementation's: twice the floor
                f_obj, f_grad = kc.prob_floor(g, None if suffix == "" else (torch.arange(len(ado_ids))[None] < 2))
                f_obj, f_grad = 2 * f_obj, 2 * f_grad
            gio.assert_parity(obj, g[f"obj_{key}{suffix}"], gio.RTOL_ROLLOUT, f"objective {key}{suffix}", atol=gio.ATOL_OBJ + 2 * f_obj)
            u = controls.clone().requires_grad_(True)
            traj = env.ego.unroll_trajectory(u, dt=env.dt)
            torch.manual_seed(int(g["seed"]))
            grad = module.gradient(traj, grad_wrt=u, ado_ids=ids, tag="g")
            assert grad.shape == (2 * t_hor,)
            gio.assert_parity(grad, g[f"grad_{key}{suffix}"], gio.RTOL_GRAD, f"gradient {key}{suffix}", atol=gio.ATOL + 2 * f_grad)
        if key != "goal":
            assert module.objective(traj, ado_ids=[], tag="g") == 0.0                      # acc_interact.py:50-51
            assert not module.gradient(traj, grad_wrt=u, ado_ids=[], tag="g").any()
    for key, cls in (("control_limits", modules.ControlLimitModule), ("speed_limits", modules.SpeedLimitModule)):
        module = cls(env=env, t_horizon=t_hor)
        u = controls.clone().requires_grad_(True)
        traj = env.ego.unroll_trajectory(u, dt=env.dt)
        gio.assert_parity(module.constraint(traj, ado_ids=ado_ids, tag="g"), g[f"con_{key}"], gio.RTOL_ROLLOUT, f"constraint {key}")
        single_ctl = key == "control_limits" and str(g["ego_kind"]) == "single"    # reference: T rows, 2T x 2T Jacobian
        jac = module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids, tag="g")
        if single_ctl:
            un = controls / torch.norm(controls, dim=1, keepdim=True)
            expect = torch.zeros((t_hor, 2 * t_hor))
            for t in range(t_hor):
                expect[t, 2 * t:2 * t + 2] = un[t]
            gio.assert_parity(jac, expect.flatten().numpy(), gio.RTOL_GRAD, "jacobian control (single integrator)")
            continue
        gio.assert_parity(jac, g[f"jac_{key}"], gio.RTOL_GRAD, f"jacobian {key}")
        gio.assert_parity(module.compute_violation(traj, ado_ids=ado_ids, tag="g"), g[f"vio_{key}"], 1e-4, f"violation {key}", atol=1e-6)
        lower, upper = module.constraint_boundaries(ado_ids=ado_ids)
        for got, ref in ((lower, g[f"lb_{key}"]), (upper, g[f"ub_{key}"])):
            assert len(got) == ref.shape[0] == module.num_constraints(ado_ids)
            assert all((x is None and np.isnan(r)) or (x is not None and abs(x - r) < 1e-6) for x, r in zip(got, ref))
        structure = module.jacobian_structure(ado_ids=ado_ids, tag="g")                      # test_modules.py:290-306
        assert set(np.nonzero(g[f"jac_{key}"])[0]) <= set(structure.tolist())
    # InteractionEllipsoidModule (modules/baselines/ellipsoid_interact.py), the reference's call sequence and seeds
    module = modules.baselines.InteractionEllipsoidModule(env=env, t_horizon=t_hor, num_particles=num_p)
    u = controls.clone().requires_grad_(True)
    traj = env.ego.unroll_trajectory(u, dt=env.dt)
    torch.manual_seed(int(g["seed"]))
    gio.assert_parity(module.constraint(traj, ado_ids=ado_ids, tag="g"), g["con_ellipsoid"], gio.RTOL_ROLLOUT, "ellipsoid rows")
    torch.manual_seed(int(g["seed"]))
    gio.assert_parity(module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids, tag="g"), g["jac_ellipsoid"], gio.RTOL_GRAD, "jacobian ellipsoid")
    torch.manual_seed(int(g["seed"]))
    gio.assert_parity(module.compute_violation(traj, ado_ids=ado_ids, tag="g"), g["vio_ellipsoid"], 1e-4, "violation ellipsoid", atol=1e-6)
    lower, upper = module.constraint_boundaries(ado_ids=ado_ids)
    assert lower == g["lb_ellipsoid"].tolist() and all(x is None for x in upper) and len(upper) == module.num_constraints(ado_ids)
    assert module.constraint(traj, ado_ids=[], tag="g").size == 0 and module.jacobian(traj, grad_wrt=u, ado_ids=[], tag="g").size == 0
    assert module.name == "ellipsoid_interaction" and module.normalize(3.0) == 3.0
    if "con_min_distance" in g:
        module = modules.MinDistanceModule(env=env, t_horizon=t_hor, num_particles=num_p)
        u = controls.clone().requires_grad_(True)
        traj = env.ego.unroll_trajectory(u, dt=env.dt)
        torch.manual_seed(int(g["seed"]))
        gio.assert_parity(module.constraint(traj, ado_ids=ado_ids, tag="g"), g["con_min_distance"], gio.RTOL_ROLLOUT, "min distance")
        torch.manual_seed(int(g["seed"]))
        gio.assert_parity(module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids, tag="g"), g["jac_min_distance"], gio.RTOL_GRAD,
                          "jacobian min distance")
        torch.manual_seed(int(g["seed"]))
        sub = module.jacobian(traj, grad_wrt=u, ado_ids=ado_ids[1:3], tag="g").reshape(2 * t_hor, 2 * t_hor)
        gio.assert_parity(sub, g["jac_min_distance"].reshape(-1, t_hor, 2 * t_hor)[1:3].reshape(2 * t_hor, 2 * t_hor), gio.RTOL_GRAD,
                          "jacobian min distance (ado subset)")
        assert module.num_constraints(ado_ids[1:3]) == 2 * t_hor
        assert module.constraint_boundaries(ado_ids)[0][0] == 0.5 and module.constraint_boundaries(ado_ids)[1][0] is None


def check_attention():
    """mantrap_b200.attention.{EuclideanModule, ClosestModule}.compute() against the reference's filters on the recorded
    scenes (ids of the selected ados), and as the `attention_module` of a solver."""
    from mantrap_b200 import attention
    g = gio.load("scene_ops")
    for i in range(len(g["att_n"])):
        n = int(g["att_n"][i])
        env = environment.PotentialFieldEnvironment(ego_position=gio.t(g["att_ego"][i][0:2]), ego_velocity=torch.zeros(2),
                                                    ego_type=agents.DoubleIntegratorDTAgent, dt=0.4)
        for k in range(n):
            env.add_ado(position=gio.t(g["att_ado"][i][k, 0:2]), velocity=torch.zeros(2), identifier=f"a{k:02d}")
        for key, cls in (("euclidean", attention.EuclideanModule), ("closest", attention.ClosestModule)):
            module = cls(env=env, t_horizon=5)
            expect = [env.ado_ids[m] for m in np.nonzero(g[f"att_{key}"][i][:n])[0]]
            assert module.compute() == expect == module.ids_current and module.name == key
    env, goal = small_scene("sf")
    sol = solver.RandomSearch(env, goal=goal, t_planning=4, attention_module=attention.EuclideanModule,
                              modules=[modules.GoalNormModule, modules.InteractionAccelerationModule, modules.ControlLimitModule])
    z = sol.optimize(torch.zeros(8), tag="t", num_candidates=8, max_rounds=1)
    assert z.numel() == 8 and sol._attention_module.ids_current == attention.EuclideanModule(env=sol.env, t_horizon=4).compute()


def check_warm_start_encoding():
    """TrajOptSolver.encode() / warm_start("encoding") of the drop-in solver against the reference on the recorded scenes."""
    g = gio.load("scene_ops")
    db = (gio.t(g["ws_encoding_db"]), gio.t(g["ws_solution_db"]))
    for i in range(0, len(g["ws_n"]), 3):
        n, t_hor = int(g["ws_n"][i]), int(g["ws_T"][i])
        env = environment.PotentialFieldEnvironment(ego_position=gio.t(g["ws_ego"][i][0:2]), ego_velocity=gio.t(g["ws_ego"][i][2:4]),
                                                    ego_type=agents.DoubleIntegratorDTAgent, dt=0.4)
        for k in range(n):
            env.add_ado(position=gio.t(g["ws_ado"][i][k, 0:2]), velocity=torch.zeros(2), identifier=f"a{k:02d}")
        sol = solver.RandomSearch(env, goal=gio.t(g["ws_goal"][i]), t_planning=t_hor, modules=[modules.GoalNormModule], warm_start_db=db)
        gio.assert_parity(sol.encode().numpy(), g["ws_encoding"][i], gio.RTOL_ROLLOUT, f"encode {i}", atol=1e-6)
        z = sol.warm_start(method="encoding")
        assert z.shape == (2 * t_hor,) and np.array_equal(z.view(-1, 2).numpy(), g["ws_warm"][i][:t_hor])
    sol = solver.RandomSearch(env, goal=gio.t(g["ws_goal"][i]), t_planning=3, modules=[modules.GoalNormModule])
    import os
    if "MANTRAP_WARM_START_DIR" not in os.environ:
        try:
            sol.warm_start(method="encoding")
            raise AssertionError("expected FileNotFoundError")
        except FileNotFoundError:
            pass


def check_autograd_route(name):
    """A user-defined cost written with torch ops on `env.compute_distributions` (what a custom OptimizationModule does)
    gets its gradient from torch.autograd through the EgoUnroll / ParticleRollout nodes: it must equal the reference's
    gradient of the same cost (here: the interaction-acceleration cost, re-stated in torch)."""
    g = gio.load(name)
    if "grad_acc" not in g or str(g["env"]) == "kalman":
        return
    env = build_env(g)
    num_p, t_hor, n = int(g["P"]), int(g["T"]), g["ado0"].shape[0]
    wo = gio.t(g["mus_wo"][..., 2:4])
    acc_wo = (wo[:, 1:] - wo[:, :-1]) / env.dt
    u = gio.t(g["controls"]).clone().requires_grad_(True)
    traj = env.ego.unroll_trajectory(u, dt=env.dt)
    torch.manual_seed(int(g["seed"]))
    dist = env.compute_distributions(traj, num_particles=num_p)
    vel = torch.stack([dist[i].mean[:, 0, :] for i in env.ado_ids])
    acc = (vel[:, 1:] - vel[:, :-1]) / env.dt
    cost = torch.norm(acc - acc_wo, dim=-1).mean(dim=-1).mean().clamp_max(1.0)
    gio.assert_parity(float(cost.detach()), g["obj_acc"], gio.RTOL_ROLLOUT, "torch-side cost", atol=gio.ATOL_OBJ)
    (grad,) = torch.autograd.grad(cost, u)
    grad = np.nan_to_num(grad.flatten().numpy())                  # compute_gradient_auto_grad, optimization_module.py:446
    gio.assert_parity(grad, g["grad_acc"], gio.RTOL_GRAD, "autograd-route gradient")


def check_solver_callbacks(name):
    """The four IPOPT callbacks (solver/ipopt.py:201-229) at the golden z, fused (one launch per z) and module by module."""
    g = gio.load(name)
    env = build_env(g)
    num_p, t_hor = int(g["P"]), int(g["T"])
    keys = [k for k in ("goal", "acc", "vel", "pos") if f"obj_{k}" in g]
    mods_ = [MODULES[k] for k in keys] + [modules.ControlLimitModule, modules.SpeedLimitModule]
    if "con_min_distance" in g:
        mods_.append(modules.MinDistanceModule)
    z = g["controls"].astype(np.float64).flatten()
    expect_obj = sum(float(g[f"obj_{k}"]) for k in keys)
    expect_grad = sum(g[f"grad_{k}"] for k in keys)
    expect_cons = [g["con_control_limits"], g["con_speed_limits"]] + ([g["con_min_distance"]] if "con_min_distance" in g else [])
    expect_jac = [g["jac_control_limits"], g["jac_speed_limits"]] + ([g["jac_min_distance"]] if "con_min_distance" in g else [])
    for fused in (True, False):
        torch.manual_seed(int(g["seed_wo"]))
        interact = [k for k in keys if k != "goal"]
        # every interaction module draws its own ego-free baseline at construction: re-seed per module like the generator
        sol = solver.IPOPTSolver(env, goal=gio.t(g["ego_goal"]), t_planning=t_hor, modules=[], fused=fused, num_particles=num_p)
        for cls in mods_:
            torch.manual_seed(int(g["seed_wo"]))
            m = cls(t_horizon=t_hor, goal=sol.goal, env=sol.env, num_particles=num_p)
            sol._module_dict[m.name] = m
        seeds = 1 if fused else len(interact)

        def seeded(fn, *a, **k):
            torch.manual_seed(int(g["seed"]))
            return fn(*a, **k)
        if fused or len(interact) <= 1:       # module by module the k-th interaction module sees the k-th draw of the stream
            gio.assert_parity(seeded(sol.objective, z), expect_obj, gio.RTOL_ROLLOUT, f"objective fused={fused}", atol=gio.ATOL_OBJ)
            sol._cache.clear()
            gio.assert_parity(seeded(sol.gradient, z), expect_grad, gio.RTOL_GRAD, f"gradient fused={fused}")
            sol._cache.clear()
        cons, vio = seeded(sol.constraints, z, return_violation=True)
        gio.assert_parity(cons, np.concatenate(expect_cons), gio.RTOL_ROLLOUT, f"constraints fused={fused}")
        sol._cache.clear()
        jac = seeded(sol.jacobian, z)
        if str(g["ego_kind"]) == "double":
            gio.assert_parity(jac, np.concatenate(expect_jac), gio.RTOL_GRAD, f"jacobian fused={fused}")
        else:                                 # control rows: T x 2T here (the reference's block is inconsistent with its rows)
            gio.assert_parity(jac[2 * t_hor * t_hor:], np.concatenate(expect_jac[1:]), gio.RTOL_GRAD, f"jacobian fused={fused}")
        cl, cu = sol.constraint_bounds(sol.env.ado_ids)
        assert len(cl) == len(cu) == cons.shape[0]
        lb, ub = sol.optimization_variable_bounds()
        assert lb == [-2.0] * (2 * t_hor) and ub == [2.0] * (2 * t_hor)
        del seeds
    # per-z cache: the callbacks of one iteration share one launch
    from mantrap_b200 import runtime
    engine = runtime.default_engine()
    sol._fused = True
    sol._cache.clear()
    before = engine.launches
    sol.objective(z), sol.gradient(z), sol.constraints(z)
    assert engine.launches == before + 1
    sol.jacobian(z)
    assert engine.launches == before + 2


def small_scene(env_kind="sf", ego_kind="double"):
    """test_solver.py:24-35 style scene: ego at (-5, 2) heading to (1, 1), two pedestrians."""
    env = ENVS[env_kind](ego_position=torch.tensor([-5.0, 2.0]), ego_velocity=torch.zeros(2), ego_type=EGOS[ego_kind], dt=0.4)
    kw = (lambda goal: {"goal": torch.tensor(goal)}) if env_kind == "sf" else (lambda goal: {})
    env.add_ado(position=torch.tensor([3.0, 2.0]), velocity=torch.tensor([-1.0, 0.0]), identifier="p0", **kw([-4.0, 2.0]))
    env.add_ado(position=torch.tensor([0.0, -2.0]), velocity=torch.tensor([0.1, 0.9]), identifier="p1", **kw([0.0, 4.0]))
    return env, torch.tensor([1.0, 1.0])


def check_random_search(num_candidates):
    """Batched RandomSearch: one launch evaluates every candidate, the device argmin keeps the best FEASIBLE one
    (random_search.py:28-35). Kalman environment = no random numbers, so a brute-force host loop over the same candidates
    (np.random stream replayed) must pick the same z."""
    env, goal = small_scene("kalman")
    mods_ = [modules.GoalNormModule, modules.InteractionAccelerationModule, modules.ControlLimitModule, modules.SpeedLimitModule]
    sol = solver.RandomSearch(env, goal=goal, t_planning=4, modules=mods_)
    np.random.seed(7)
    z_opt, log = sol.optimize_core(torch.zeros(8), ado_ids=env.ado_ids, num_candidates=num_candidates, max_rounds=1)
    assert z_opt.shape == (4, 2) and log["candidates"] == num_candidates
    np.random.seed(7)
    lb, ub = sol.z_bounds
    np.random.uniform(lb, ub)                                          # the fall-back sample drawn first (search.py:49)
    cands = np.random.uniform(lb, ub, size=(num_candidates, lb.size))
    best, best_obj = None, np.inf
    for z in cands[:min(num_candidates, 48)]:
        obj, vio = sol.evaluate(z.astype(np.float32).astype(np.float64), ado_ids=env.ado_ids, tag="t")
        if vio < 1e-3 and obj < best_obj:
            best, best_obj = z, obj
    obj_opt, vio_opt = sol.evaluate(z_opt.numpy().flatten().astype(np.float64), ado_ids=env.ado_ids, tag="t")
    assert vio_opt < 1e-3                                             # speed limits prune most uniform samples
    assert obj_opt <= best_obj + 1e-6
    if num_candidates <= 48 and best is not None:
        assert np.allclose(z_opt.numpy().flatten(), best.astype(np.float32))
    # social forces: stochastic, so only feasibility + finiteness
    env, goal = small_scene("sf")
    sol = solver.RandomSearch(env, goal=goal, t_planning=4, modules=mods_)
    z_opt, log = sol.optimize_core(torch.zeros(8), ado_ids=env.ado_ids, num_candidates=num_candidates, max_rounds=2)
    assert torch.isfinite(z_opt).all() and np.isfinite(log["objective"])


def check_mcts(num_breadth, num_depth):
    """Batched MonteCarloTreeSearch against a host replay of the reference's loop (baselines/mcts.py:41-87) over the SAME
    samples (np.random stream replayed), each completion evaluated one by one through the solver callbacks. Kalman
    environment = no random numbers in the evaluation."""
    env, goal = small_scene("kalman")
    mods_ = [modules.GoalNormModule, modules.InteractionAccelerationModule, modules.ControlLimitModule, modules.SpeedLimitModule]
    t_hor = 3
    sol = solver.MonteCarloTreeSearch(env, goal=goal, t_planning=t_hor, modules=mods_)
    np.random.seed(11)
    z_opt, log = sol.optimize_core(torch.zeros(2 * t_hor), ado_ids=env.ado_ids, max_cpu_time=60.0,
                                   num_breadth_samples=num_breadth, num_depth_samples=num_depth)
    assert z_opt.shape == (t_hor, 2) and log["rounds"] <= t_hor and log["candidates"] == log["rounds"] * num_breadth * num_depth
    # host replay
    np.random.seed(11)
    lb, ub = sol.z_bounds
    np.random.uniform(lb, ub)                                          # fall-back sample (search.py:49)
    z_best, obj_best, iteration = None, np.inf, 0
    for _ in range(log["rounds"]):
        z = np.empty((num_breadth, num_depth, 2 * t_hor))
        if iteration > 0:
            z[:, :, :2 * iteration] = z_best[:2 * iteration]
        z[:, :, 2 * iteration:2 * iteration + 2] = np.random.uniform(lb[2 * iteration:2 * iteration + 2], ub[2 * iteration:2 * iteration + 2],
                                                                     size=(num_breadth, 1, 2))
        rest = 2 * (iteration + 1)
        if rest < 2 * t_hor:
            z[:, :, rest:] = np.random.uniform(lb[rest:], ub[rest:], size=(num_breadth, num_depth, 2 * t_hor - rest))
        means, bests = np.full(num_breadth, np.inf), [(None, np.inf)] * num_breadth
        for i in range(num_breadth):
            values = []
            for j in range(num_depth):
                zij = z[i, j].astype(np.float32).astype(np.float64)
                obj, vio = sol.evaluate(zij, ado_ids=env.ado_ids, tag="t")
                if vio < 1e-5:
                    values.append(obj)
                    if obj < bests[i][1]:
                        bests[i] = (z[i, j].copy(), obj)
            means[i] = np.mean(values) if values else np.inf
        i_best = int(np.argmin(means))
        cand, cand_obj = bests[i_best]
        if cand is not None and cand_obj <= obj_best:
            z_best, obj_best, iteration = cand, cand_obj, iteration + 1
    if z_best is not None:
        assert np.allclose(z_opt.numpy().flatten(), z_best.astype(np.float32), atol=1e-6)
        assert abs(log["objective"] - obj_best) <= 1e-5 * max(1.0, abs(obj_best))
    # social forces: stochastic evaluation, only feasibility + finiteness
    env, goal = small_scene("sf")
    sol = solver.MonteCarloTreeSearch(env, goal=goal, t_planning=t_hor, modules=mods_)
    z_opt, log = sol.optimize_core(torch.zeros(2 * t_hor), ado_ids=env.ado_ids, max_cpu_time=60.0,
                                   num_breadth_samples=num_breadth, num_depth_samples=num_depth)
    assert torch.isfinite(z_opt).all() and z_opt.abs().max() <= 2.0


def check_closed_loop(solver_cls, env_kind, steps=3, **kwargs):
    """`solve()`: warm start, optimise, step the evaluation environment, repeat (trajopt.py:109-193; test_solver.py:111-153)."""
    env, goal = small_scene(env_kind)
    torch.manual_seed(0)
    np.random.seed(0)
    sol = solver_cls(env, goal=goal, t_planning=3,
                     modules=[modules.GoalNormModule, modules.InteractionAccelerationModule, modules.SpeedLimitModule])
    start = float(torch.norm(env.ego.position - goal))
    ego_traj, ado_trajs = sol.solve(time_steps=steps, warm_start_method="zeros", **kwargs)
    assert ego_traj.shape[1] == 5 and 2 <= ego_traj.shape[0] <= steps + 1
    assert ado_trajs.shape == (env.num_ados, ego_traj.shape[0], 1, 5)
    assert torch.isfinite(ego_traj).all() and torch.isfinite(ado_trajs).all()
    assert float(torch.norm(ego_traj[-1, 0:2] - goal)) < start          # moved towards the goal
    assert np.allclose(ego_traj[:, 4].numpy(), 0.4 * np.arange(ego_traj.shape[0]), atol=1e-5)
    assert torch.allclose(sol.env.ego.position, env.ego.position)       # solver environment restored (trajopt.py:187-190)
    speeds = ego_traj[:, 2:4].abs().max()
    assert float(speeds) <= 2.0 + 1e-3                                  # speed limit respected along the executed path


def check_warm_start_potential():
    """`warm_start("potential")` (trajopt.py:266-268): the default formulation on a PotentialFieldEnvironment copy of the scene;
    the result is a control sequence inside the variable bounds (test_solver.py:156-172)."""
    from mantrap_b200 import constants
    env, goal = small_scene("sf")
    torch.manual_seed(0)
    sol = solver.SLSQPSolver(env, goal=goal, t_planning=4, modules=[modules.GoalNormModule, modules.InteractionAccelerationModule])
    z = sol.warm_start(method=constants.WARM_START_POTENTIAL)
    lb, ub = sol.optimization_variable_bounds()
    assert z.numel() == 8 and bool(torch.isfinite(z).all())
    assert np.all(np.asarray(z.flatten()) >= np.asarray(lb) - 1e-6) and np.all(np.asarray(z.flatten()) <= np.asarray(ub) + 1e-6)
    with np.testing.assert_raises(ValueError):
        sol.warm_start(method=constants.WARM_START_SOFT)


def check_goal_optimize_speed():
    """GoalNormModule(optimize_speed=True): the reference's speed term fails in torch.dot on the first objective call
    (goal_norm.py:63-65: a [T+1] vector dotted with a [T+1, 2] matrix); the mirror raises the same RuntimeError when evaluated."""
    env, goal = small_scene("pf")
    module = modules.GoalNormModule(goal=goal, env=env, optimize_speed=True, t_horizon=4)
    traj = env.ego.unroll_trajectory(torch.ones((4, 2)) * 0.3, dt=env.dt)
    try:
        module.objective(traj, ado_ids=env.ado_ids, tag="t")
    except RuntimeError as err:
        assert "1D tensors expected" in str(err)
    else:
        raise AssertionError("optimize_speed=True must raise like the reference")
    plain = modules.GoalNormModule(goal=goal, env=env, t_horizon=4)
    assert plain.objective(traj, ado_ids=env.ado_ids, tag="t") > 0.0


def check_warm_start_pre_compute(tmp_dir):
    """Batched `pre_compute` (third_party/warm_start/pre_compute.py:27-58): database shapes and file format, encodings equal to
    `TrajOptSolver.encode()` of the same scenes, solutions inside the control limits and zero after the goal was reached; the
    written files feed `warm_start("encoding")`."""
    from mantrap_b200 import constants, runtime, warm_start_db
    eng = runtime.default_engine()
    m, horizon = 6, 4
    encodings, solutions = warm_start_db.pre_compute(eng, num_scenarios=m, solution_horizon=horizon, num_candidates=16, seed=3)
    assert encodings.shape == (m, 4) and solutions.shape == (m, horizon, 2)
    assert bool(torch.isfinite(encodings).all()) and bool(torch.isfinite(solutions).all())
    assert float(solutions.abs().max()) <= constants.ROBOT_ACC_MAX + 1e-4
    scene = warm_start_db.random_scenes(m, seed=3)
    for i in range(m):
        env = environment.PotentialFieldEnvironment(ego_position=scene.ego0[i, 0:2], ego_velocity=scene.ego0[i, 2:4], dt=scene.dt)
        env.add_ado(position=scene.ado0[i, 0, 0:2], velocity=scene.ado0[i, 0, 2:4])
        sol = solver.IPOPTSolver(env, goal=scene.ego_goal[i], t_planning=horizon, modules=[modules.GoalNormModule])
        assert torch.allclose(sol.encode(), encodings[i], atol=1e-5)
    warm_start_db.save(tmp_dir, encodings, solutions)
    sol = solver.IPOPTSolver(env, goal=scene.ego_goal[m - 1], t_planning=horizon, modules=[modules.GoalNormModule], warm_start_db=tmp_dir)
    z = sol.warm_start(method=constants.WARM_START_ENCODING)
    assert torch.allclose(z.view(horizon, 2), solutions[m - 1])        # its own scene is its nearest entry


def check_plan_with_infeasible_warm_start():
    """`BatchedClosedLoop.plan` keeps the warm start as incumbent (candidate 0) -- a deliberate deviation from the reference's
    `SearchIntermediate.optimize_core`, which ignores z0 and falls back to a fresh random z (search.py:19-81):
    an infeasible warm start never wins while a feasible candidate exists, and when NO candidate is feasible the plan is the
    warm start again, reported with index -1 and the warm start's objective."""
    from mantrap_b200 import mpc, runtime, scenarios
    eng = runtime.default_engine()
    scene = scenarios.synthetic_scenes(3, 2, 3, env_kind="pf", seed=9, param_seed=10)
    loop = mpc.BatchedClosedLoop(eng, scene, t_planning=4, num_candidates=6, seed=1)
    warm_bad = torch.full((3, 4, 2), 5.0, device=eng.device)                 # |u| = 5 > ROBOT_ACC_MAX: violates the control limits
    cands = scenarios.candidate_controls(3, 6, 4, seed=2, u_max=1.0).to(eng.device)
    best, val, idx = loop.plan(warm_bad, cands)
    assert bool((idx > 0).all()) and not bool(torch.isclose(best, warm_bad).all())
    best, val, idx = loop.plan(warm_bad, cands * 10.0)                        # every candidate infeasible
    assert bool((idx == -1).all()) and torch.equal(best, warm_bad) and bool(torch.isfinite(val).all())


def check_jacobian_reuses_the_cached_draw():
    """A Jacobian requested for a z whose values are already cached uses the SAME particle draw as those values and leaves the
    global RNG untouched (one draw per z); a fresh z draws again."""
    env, goal = small_scene("sf")
    sol = solver.IPOPTSolver(env, goal=goal, t_planning=4, modules=[modules.GoalNormModule, modules.InteractionAccelerationModule,
                                                                  modules.ControlLimitModule, modules.MinDistanceModule])
    z = np.linspace(-0.5, 0.5, 8)
    torch.manual_seed(5)
    sol.constraints(z)
    state = torch.get_rng_state()
    jac_cached = sol.jacobian(z)
    assert torch.equal(state, torch.get_rng_state())
    sol._cache.clear()
    torch.manual_seed(5)
    jac_fresh = sol.jacobian(z)                      # same seed, cold cache: the identical draw -> the identical Jacobian
    assert np.array_equal(jac_cached, jac_fresh)


def check_evaluate_fused_batch():
    """K callback sets in one launch (`evaluate_fused_batch`) = K single callbacks under the same seed (one draw per scene)."""
    env, goal = small_scene("sf")
    sol = solver.IPOPTSolver(env, goal=goal, t_planning=4, modules=[modules.GoalNormModule, modules.InteractionAccelerationModule,
                                                                  modules.ControlLimitModule, modules.MinDistanceModule])
    rng = np.random.RandomState(3)
    zs = rng.uniform(-1, 1, (5, 8))
    torch.manual_seed(11)
    batch = sol.evaluate_fused_batch(zs, want_jacobian=True)
    assert batch["total"].shape == (5,) and batch["grad"].shape == (5, 8) and batch["jac"].shape[0] == 5
    for k in range(5):
        sol._cache.clear()
        torch.manual_seed(11)
        one = sol.evaluate_fused(zs[k], sol.env.ado_ids, want_jacobian=True)
        assert abs(one["total"] - float(batch["total"][k])) <= 1e-6 * max(1.0, abs(one["total"]))
        gio.assert_parity(batch["grad"][k], one["grad"], 1e-5, "batched gradient", atol=1e-7)
        assert abs(one["violation"] - float(batch["violation"][k])) <= 1e-6
        rows = np.concatenate([one["cons"][name] for name in ("control_limits", "min_distance")])
        gio.assert_parity(batch["cons"][k], rows, 1e-6, "batched rows")
        jac = np.concatenate([one["jac"][name] for name in ("control_limits", "min_distance")])
        gio.assert_parity(batch["jac"][k], jac, 1e-5, "batched Jacobian", atol=1e-7)
