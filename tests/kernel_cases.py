"""Shared body of the kernel parity tests. The same checks run
  * on the CPU against the host emulation of the kernel source (tests/test_kernel_emulation.py, no GPU needed),
  * on the B200 through the real C-ABI library (tests/test_gpu_parity.py, -m gpu).
`lib` is a bound C-ABI library, `device` where the tensors live."""
import numpy as np
import torch

import golden_io as gio
from mantrap_b200 import engine

MODULE_WEIGHT = {"goal": 0.5, "acc": 1.0, "vel": 1.0, "pos": 1.0, "prob": 1.0}


def scene_from_golden(g, device, wo=False):
    tag = "_wo" if wo else ""
    env = str(g["env"])
    kw = {}
    if env != "kalman":
        kw = dict(v0=gio.t(g["v0" + tag])[None], sigma=gio.t(g["sigma" + tag])[None],
                  weight=gio.t(g["weights" + tag])[None], noise=gio.t(g["noise" + tag])[None])
    scene = engine.SceneBatch(env_kind=env, ego_kind=str(g["ego_kind"]), dt=float(g["dt"]), ego0=gio.t(g["ego0"])[None],
                              ego_goal=gio.t(g["ego_goal"])[None], ado0=gio.t(g["ado0"])[None],
                              ado_goal=gio.t(g["goal"])[None], **kw)
    return scene.to(device)


def np_(x):
    return x.detach().cpu().numpy()


def oracle_scene_from_golden(g):
    from oracle import torch_oracle as to
    env = str(g["env"])
    n = g["ado0"].shape[0]
    if env == "kalman":
        v0 = sigma = w = torch.ones((1, n, int(g["P"])))
        noise = None
    else:
        v0, sigma, w, noise = gio.t(g["v0"])[None], gio.t(g["sigma"])[None], gio.t(g["weights"])[None], gio.t(g["noise"])[None]
    return to.Scene(env, str(g["ego_kind"]), float(g["dt"]), gio.t(g["ego0"])[None], gio.t(g["ego_goal"])[None],
                    gio.t(g["ado0"])[None], gio.t(g["goal"])[None], v0, sigma, w, noise)


def prob_floor(g, mask):
    """Float32 conditioning of the probability objective on a golden scene: |float32 oracle - float64 oracle| of value and
    gradient. The cost divides velocity differences by VARIANCES of nearly identical particle velocities (1e-15 .. 1e-3),
    so the float32 rounding of the variance alone moves it by 1e-5 .. 1e-4 relative; an independent float32
    implementation cannot agree with the reference's own float32 evaluation more closely than that."""
    from oracle import torch_oracle as to
    osc = oracle_scene_from_golden(g)
    controls, mus_wo = gio.t(g["controls"])[None], gio.t(g["mus_wo"])[None]
    m = None if mask is None else mask.cpu().bool()
    r32 = to.evaluate(osc, controls, mus_wo, {"prob": 1.0}, ado_mask=m)
    r64 = to.evaluate(to.scene_to(osc, torch.float64), controls.double(), mus_wo.double(), {"prob": 1.0}, ado_mask=m)
    # |f32 - f64| is ONE realisation of the rounding noise (norms rounded once, as in torch); the kernels evaluate norms as
    # x * rsqrt(x) (three roundings, MUFU-friendly), so their realisation is drawn from a distribution about twice as wide
    return 2 * float((r32["total"].double() - r64["total"]).abs().max()), 2 * float((r32["grad"].double() - r64["grad"]).abs().max())


def check_rollout(lib, device, name):
    g = gio.load(name)
    scene = scene_from_golden(g, device)
    controls = gio.t(g["controls"])[None, None].to(device).contiguous()
    res = engine.launch_evaluate(lib, scene, controls, None, want=("ego_traj", "mus", "sigmas"))
    if device != "cpu":
        torch.cuda.synchronize()
    gio.assert_parity(np_(res["ego_traj"][0, 0]), g["ego_traj"][:, :4], gio.RTOL_ROLLOUT, "ego_traj")
    gio.assert_parity(np_(res["mus"][0, 0, ..., 0:2]), g["mus"][..., 0:2], gio.RTOL_ROLLOUT, "mu_p")
    gio.assert_parity(np_(res["mus"][0, 0, ..., 2:4]), g["mus"][..., 2:4], gio.RTOL_ROLLOUT, "mu_v")
    gio.assert_parity(np_(res["sigmas"][0, 0]), g["sigmas"], gio.RTOL_ROLLOUT, "sigmas")
    # rollout without ego (compute_distributions_wo_ego)
    scene_wo = scene_from_golden(g, device, wo=True)
    res = engine.launch_evaluate(lib, scene_wo, None, None, want=("mus", "sigmas"), t_horizon=int(g["T"]))
    if device != "cpu":
        torch.cuda.synchronize()
    gio.assert_parity(np_(res["mus"][0, 0]), g["mus_wo"], gio.RTOL_ROLLOUT, "mus_wo")
    gio.assert_parity(np_(res["sigmas"][0, 0]), g["sigmas_wo"], gio.RTOL_ROLLOUT, "sigmas_wo")


def check_modules(lib, device, name):
    g = gio.load(name)
    scene = scene_from_golden(g, device)
    controls = gio.t(g["controls"])[None, None].to(device).contiguous()
    mus_wo = gio.t(g["mus_wo"])[None].to(device).contiguous()
    n = g["ado0"].shape[0]
    worst = {}
    for key in engine.OBJECTIVE_KEYS:
        if f"obj_{key}" not in g:
            continue
        for suffix in ("", "_sub2"):
            if f"obj_{key}{suffix}" not in g:
                continue
            mask = None
            if suffix:
                mask = (torch.arange(n)[None] < 2).to(torch.uint8).to(device).contiguous()
            spec = engine.ModuleSpec(weights={key: MODULE_WEIGHT[key]}, mus_wo=mus_wo, ado_mask=mask)
            res = engine.launch_evaluate(lib, scene, controls, spec, want=("obj", "obj_raw", "total", "grad"))
            if device != "cpu":
                torch.cuda.synchronize()
            col = engine.OBJECTIVE_KEYS.index(key)
            f_obj, f_grad = prob_floor(g, mask) if key == "prob" else (0.0, 0.0)
            e1 = gio.assert_parity(np_(res["obj"][0, 0, col]), g[f"obj_{key}{suffix}"], gio.RTOL_ROLLOUT, f"obj_{key}{suffix}", atol=gio.ATOL_OBJ + 2 * f_obj)
            gio.assert_parity(np_(res["total"][0, 0]), g[f"obj_{key}{suffix}"], gio.RTOL_ROLLOUT, f"total_{key}{suffix}", atol=gio.ATOL_OBJ + 2 * f_obj)
            e2 = gio.assert_parity(np_(res["grad"][0, 0]).flatten(), g[f"grad_{key}{suffix}"], gio.RTOL_GRAD, f"grad_{key}{suffix}",
                                   atol=gio.ATOL + 2 * f_grad)
            # (a saturated probability cost is 1e6 .. 1e13 from variances of 1e-15: only its clamped value is comparable)
            if suffix == "" and key != "goal" and not (key == "prob" and abs(float(g["raw_prob"])) > 120.0):
                gio.assert_parity(np_(res["obj_raw"][0, 0, col]), g[f"raw_{key}"], gio.RTOL_ROLLOUT, f"raw_{key}",
                                  atol=gio.ATOL_OBJ + 2 * 120.0 * f_obj)
            worst[key + suffix] = (e1, e2)
    return worst


def check_constraints(lib, device, name):
    g = gio.load(name)
    scene = scene_from_golden(g, device)
    t_hor = int(g["T"])
    controls = gio.t(g["controls"])[None, None].to(device).contiguous()
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance="con_min_distance" in g)
    res = engine.launch_evaluate(lib, scene, controls, spec, want=("cons", "violation"))
    if device != "cpu":
        torch.cuda.synchronize()
    cons = np_(res["cons"][0, 0])
    n_ctl = g["con_control_limits"].shape[0]
    gio.assert_parity(cons[:n_ctl], g["con_control_limits"], gio.RTOL_ROLLOUT, "control rows")
    gio.assert_parity(cons[n_ctl:n_ctl + 2 * (t_hor + 1)], g["con_speed_limits"], gio.RTOL_ROLLOUT, "speed rows")
    expect_vio = float(g["vio_control_limits"]) + float(g["vio_speed_limits"])
    if "con_min_distance" in g:
        gio.assert_parity(cons[n_ctl + 2 * (t_hor + 1):], g["con_min_distance"], gio.RTOL_ROLLOUT, "min-distance rows")
        expect_vio += float(np.maximum(0.5 - g["con_min_distance"], 0).sum())
    gio.assert_parity(np_(res["violation"][0, 0]), expect_vio, 1e-4, "violation", atol=1e-6)


def check_seeded_adjoint(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, seed=0):
    """Module interface of the kernel: the ego TRAJECTORY is given (not the controls) and the caller supplies adjoint
    seeds d(loss)/d(mus), d(loss)/d(sigmas) (what torch.autograd hands to the backward of compute_distributions).
    Checked against autograd of the CPU oracle for loss = <a, mus> + <b, sigmas>: gradient w.r.t. the trajectory rows
    (grad_traj) and folded through the ego dynamics w.r.t. the controls (grad)."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=700 + seed,
                                           param_seed=800 + seed, regular_sigma=True)
    controls_cpu = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=900 + seed, u_max=1.0)
    gen = torch.Generator().manual_seed(1000 + seed)
    b = num_scenes * num_candidates
    a_mus = torch.randn((b, n, t_hor, 4), generator=gen) * 0.1
    a_sig = torch.randn((b, n, t_hor, 2), generator=gen)
    osc = oracle_scene(scene_cpu, num_candidates)

    def oracle(dtype):
        sc = to.scene_to(osc, dtype)
        u = controls_cpu.reshape(b, t_hor, 2).to(dtype).requires_grad_(True)
        traj = to.ego_unroll(sc.ego0, u, sc.dt, ego_kind)
        traj.retain_grad()
        mus, sigmas = to.rollout(env_kind, sc.ado0, sc.goal, sc.v0, sc.sigma, sc.weights, sc.noise, traj, sc.dt)
        loss = (mus * a_mus.to(dtype)).sum() + (sigmas * a_sig.to(dtype)).sum()
        if not loss.requires_grad:                 # Kalman: the prediction does not depend on the ego (kalman.py:63)
            return traj.detach(), torch.zeros_like(u), torch.zeros_like(traj), mus.detach(), sigmas.detach()
        loss.backward()
        return traj.detach(), torch.nan_to_num(u.grad), torch.nan_to_num(traj.grad), mus.detach(), sigmas.detach()

    traj, g_u, g_traj, mus_ref, sig_ref = oracle(torch.float32)
    _, g_u64, g_traj64, mus64, sig64 = oracle(torch.float64)
    floor = max(float((g_u.double() - g_u64).abs().max()), float((g_traj.double() - g_traj64).abs().max()))
    # float32 conditioning of the rollout itself on random (dense) scenes, as in check_batched_vs_oracle
    floor_roll = max(float((mus_ref.double() - mus64).abs().max()), float((sig_ref.double() - sig64).abs().max()))
    shape5 = (num_scenes, num_candidates)
    spec = engine.ModuleSpec(seed_mus=a_mus.reshape(*shape5, n, t_hor, 4).to(device).contiguous(),
                             seed_sigmas=a_sig.reshape(*shape5, n, t_hor, 2).to(device).contiguous())
    ego_traj = traj.reshape(*shape5, t_hor + 1, 4).to(device).contiguous()
    res = engine.launch_evaluate(lib, scene_cpu.to(device), None, spec, want=("grad", "grad_traj", "mus", "sigmas"),
                                 ego_traj=ego_traj)
    if device != "cpu":
        torch.cuda.synchronize()
    gio.assert_parity(np_(res["mus"]).reshape(b, n, t_hor, 4), mus_ref.numpy(), gio.RTOL_ROLLOUT, "mus (trajectory given)",
                      atol=gio.ATOL + 1e-6 + 2 * floor_roll)
    if env_kind != "kalman":
        gio.assert_parity(np_(res["sigmas"]).reshape(b, n, t_hor, 2), sig_ref.numpy(), gio.RTOL_ROLLOUT, "sigmas",
                          atol=gio.ATOL + 1e-6 + 2 * floor_roll)
    e1 = gio.assert_parity(np_(res["grad_traj"]).reshape(b, t_hor + 1, 4), g_traj.numpy(), gio.RTOL_GRAD, "seeded grad_traj",
                           atol=gio.ATOL + 2 * floor)
    e2 = gio.assert_parity(np_(res["grad"]).reshape(b, t_hor, 2), g_u.numpy(), gio.RTOL_GRAD, "seeded grad (controls)",
                           atol=gio.ATOL + 2 * floor)
    if env_kind == "sf":                       # (potential field: the ego may be outside every field of view)
        assert float(g_u.abs().max()) > 0
    return e1, e2


def check_jacobian_golden(lib, device, name):
    """Constraint Jacobian kernel against the reference's `module.jacobian` vectors (control / speed limits: analytic,
    min-distance: row-by-row autograd through the reference rollout)."""
    g = gio.load(name)
    scene = scene_from_golden(g, device)
    t_hor, n = int(g["T"]), g["ado0"].shape[0]
    controls = gio.t(g["controls"])[None, None].to(device).contiguous()
    has_mind = "jac_min_distance" in g
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=has_mind)
    jac, values = engine.launch_jacobian(lib, scene, controls, spec, want_values=True)
    if device != "cpu":
        torch.cuda.synchronize()
    jac = np_(jac[0, 0])
    n_ctl = 2 * t_hor if str(g["ego_kind"]) == "double" else t_hor
    n_spd = 2 * (t_hor + 1)
    assert jac.shape == (n_ctl + n_spd + (n * t_hor if has_mind else 0), 2 * t_hor)
    if str(g["ego_kind"]) == "double":       # single integrator: the reference returns a 2T x 2T identity for T rows
        gio.assert_parity(jac[:n_ctl], g["jac_control_limits"].reshape(n_ctl, 2 * t_hor), gio.RTOL_GRAD, "jac control")
    gio.assert_parity(jac[n_ctl:n_ctl + n_spd], g["jac_speed_limits"].reshape(n_spd, 2 * t_hor), gio.RTOL_GRAD, "jac speed")
    # the speed rows are the velocity rows of dx_du (linear.py:169-196)
    dx_du = g["dx_du"].reshape(t_hor + 1, 5, 2 * t_hor)
    gio.assert_parity(jac[n_ctl:n_ctl + n_spd].reshape(t_hor + 1, 2, -1), dx_du[:, 2:4], gio.RTOL_GRAD, "dx_du velocity rows")
    if has_mind:
        gio.assert_parity(np_(values[0, 0]).flatten(), g["con_min_distance"], gio.RTOL_ROLLOUT, "min-distance values")
        gio.assert_parity(jac[n_ctl + n_spd:], g["jac_min_distance"].reshape(n * t_hor, 2 * t_hor), gio.RTOL_GRAD, "jac min-distance")


def check_jacobian_vs_oracle(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, seed=0):
    """Random batch through the tangent kernel vs. row-by-row autograd of the CPU oracle (optimization_module.py:428-434)."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=400 + seed,
                                           param_seed=500 + seed, regular_sigma=True)
    controls_cpu = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=600 + seed, u_max=1.0)
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True)
    jac, values = engine.launch_jacobian(lib, scene_cpu.to(device), controls_cpu.to(device).contiguous(), spec, want_values=True)
    if device != "cpu":
        torch.cuda.synchronize()
    jac, values = np_(jac), np_(values)
    n_ctl = 2 * t_hor if ego_kind == "double" else t_hor
    n_spd = 2 * (t_hor + 1)
    osc = oracle_scene(scene_cpu, num_candidates)
    worst = 0.0
    for b in range(num_scenes * num_candidates):
        one = to.Scene(*[(x[b:b + 1] if isinstance(x, torch.Tensor) else x) for x in osc])
        u = controls_cpu.reshape(-1, t_hor, 2)[b:b + 1]
        g_ref, j_ref = to.min_distance_jacobian(one, u)
        g64, j64 = to.min_distance_jacobian(to.scene_to(one, torch.float64), u.double())
        floor = float((j_ref.double() - j64).abs().max())
        sc, c = divmod(b, num_candidates)
        gio.assert_parity(values[sc, c].flatten(), g_ref.numpy(), gio.RTOL_ROLLOUT, "min-distance values", atol=gio.ATOL + 1e-6)
        worst = max(worst, gio.assert_parity(jac[sc, c, n_ctl + n_spd:], j_ref.numpy(), gio.RTOL_GRAD, "jac min-distance",
                                             atol=gio.ATOL + 2 * floor))
        # control / speed rows against autograd of the oracle's constraint functions
        uu = u.clone().requires_grad_(True)
        et = to.ego_unroll(one.ego0, uu, one.dt, ego_kind)
        rows = torch.cat((to.control_limit_raw(et, one.dt, ego_kind)[0], to.speed_limit_raw(et)[0]))
        j_cs = torch.stack([torch.autograd.grad(r, uu, retain_graph=True)[0].flatten() for r in rows]).numpy()
        if ego_kind == "double":                 # reference quirk: identity even for negative controls
            j_cs[:n_ctl] = np.eye(n_ctl)
        gio.assert_parity(jac[sc, c, :n_ctl + n_spd], j_cs, gio.RTOL_GRAD, "jac control+speed", atol=1e-6)
    return worst


def check_jacobian_block_vs_chunked(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, seed=0, setenv=None, ring=0):
    """The two Jacobian kernels on the same batch: jacobian_block_kernel (one evaluation per block, tangents in shared memory; the
    default whenever the shape fits) against jacobian_kernel (chunks of 8 directions in registers; `MTB_JAC=1` in a library built
    with the development knobs). Same forward arithmetic, hence identical constraint values; the Jacobians differ by the order in
    which the pair terms are accumulated (aggregated 2x2 maps vs. per-direction sums), bounded by the oracle's own
    float32-vs-float64 difference on the batch."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=1400 + seed,
                                           param_seed=1500 + seed, regular_sigma=True)
    controls_cpu = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=1600 + seed, u_max=1.0)
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True, ellipsoid=True)
    scene, controls = scene_cpu.to(device), controls_cpu.to(device).contiguous()
    out = {}
    setenv("MTB_JAC_RING", str(ring))                # 1: the block kernel with ONE ring slot (forward and tangent warps alternate)
    for knob in ("0", "1"):
        setenv("MTB_JAC", knob)
        jac, values = engine.launch_jacobian(lib, scene, controls, spec, want_values=True)
        if device != "cpu":
            torch.cuda.synchronize()
        out[knob] = (np_(jac), np_(values))
    setenv("MTB_JAC", "0")
    assert np.isfinite(out["0"][0]).all() and np.isfinite(out["1"][0]).all()
    if device == "cpu":
        np.testing.assert_array_equal(out["0"][1], out["1"][1])      # min-distance values: the same forward rollout (IEEE emulation)
    else:                                                            # nvcc --use_fast_math contracts the two kernels independently
        gio.assert_parity(out["0"][1], out["1"][1], gio.RTOL_ROLLOUT, "min-distance values: block kernel vs chunked kernel", atol=1e-5)
    floor = 0.0
    if env_kind != "kalman":
        osc = oracle_scene(scene_cpu, num_candidates)
        for b in range(min(num_scenes * num_candidates, 4)):
            one = to.Scene(*[(x[b:b + 1] if isinstance(x, torch.Tensor) else x) for x in osc])
            u = controls_cpu.reshape(-1, t_hor, 2)[b:b + 1]
            _, j_ref = to.min_distance_jacobian(one, u)
            _, j64 = to.min_distance_jacobian(to.scene_to(one, torch.float64), u.double())
            floor = max(floor, float((j_ref.double() - j64).abs().max()))
    n_const = (2 * t_hor if ego_kind == "double" else t_hor) + 2 * (t_hor + 1)
    n_md = n * t_hor
    a, b = out["0"][0], out["1"][0]
    np.testing.assert_array_equal(a[:, :, :n_const], b[:, :, :n_const])                        # control | speed rows: closed form
    gio.assert_parity(a[:, :, n_const + n_md:], b[:, :, n_const + n_md:], 1e-6, "ellipsoid rows: block kernel vs chunked kernel")
    assert env_kind == "kalman" or float(np.abs(b[:, :, n_const:n_const + n_md]).max()) > 1e-3
    return gio.assert_parity(a[:, :, n_const:n_const + n_md], b[:, :, n_const:n_const + n_md], gio.RTOL_GRAD,
                             "min-distance rows: block kernel vs chunked kernel", atol=gio.ATOL + 2 * floor)


def check_jacobian_sampled_sigma_nonpositive(lib, device):
    """Outside the parity domain: a particle whose sampled sigma is <= 0 (the reference draws sigma from a Normal, so this
    happens with small probability) has a non-finite local derivative; the reference's autograd then returns NaN and
    `np.nan_to_num` (optimization_module.py:440-446) reports zeros. The tangent kernel must never emit a non-finite entry
    either (an optimiser cannot use it), and scenes of the same batch without such a particle are unaffected."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    s, c, n, p, t_hor = 6, 2, 5, 4, 8
    scene_cpu = scenarios.synthetic_scenes(s, n, p, env_kind="sf", seed=41, param_seed=42, regular_sigma=True)
    scene_cpu.sigma[1, 2, 0] = -0.04                 # one particle each in scenes 1 and 4
    scene_cpu.sigma[4, 0, 3] = 0.0
    controls_cpu = scenarios.candidate_controls(s, c, t_hor, seed=43, u_max=1.0)
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True, ellipsoid=True)
    jac, values = engine.launch_jacobian(lib, scene_cpu.to(device), controls_cpu.to(device).contiguous(), spec, want_values=True)
    if device != "cpu":
        torch.cuda.synchronize()
    assert bool(torch.isfinite(jac).all()) and bool(torch.isfinite(values).all())
    jac = np_(jac)
    n_ctl, n_spd, n_md = 2 * t_hor, 2 * (t_hor + 1), n * t_hor
    osc = oracle_scene(scene_cpu, c)
    for sc in (0, 2, 3, 5):                          # regular scenes: same result as without the degenerate neighbours
        b = sc * c
        one = to.Scene(*[(x[b:b + 1] if isinstance(x, torch.Tensor) else x) for x in osc])
        _, j_ref = to.min_distance_jacobian(one, controls_cpu.reshape(-1, t_hor, 2)[b:b + 1])
        gio.assert_parity(jac[sc, 0, n_ctl + n_spd:n_ctl + n_spd + n_md], j_ref.numpy(), gio.RTOL_GRAD, "jac min-distance", atol=gio.ATOL + 1e-5)


def oracle_scene(scene, num_candidates):
    """Expand a SceneBatch (CPU copy) to the oracle's per-evaluation batch layout [S*C, ...]."""
    from oracle import torch_oracle as to

    def rep(x):
        return None if x is None else x.detach().cpu().repeat_interleave(num_candidates, dim=0)
    sc = scene.to("cpu")
    n, p = sc.num_ados, sc.num_particles
    ones = torch.ones((sc.num_scenes, n, p))
    return to.Scene(sc.env_kind, sc.ego_kind, sc.dt, rep(sc.ego0), rep(sc.ego_goal), rep(sc.ado0),
                    rep(sc.ado_goal if sc.ado_goal is not None else torch.zeros((sc.num_scenes, n, 2))),
                    rep(sc.v0 if sc.v0 is not None else ones), rep(sc.sigma if sc.sigma is not None else ones),
                    rep(sc.weight if sc.weight is not None else ones), rep(sc.noise), sc.axes_diag)


def conditioning_floor(osc, controls, mus_wo, weights):
    """max |float32 oracle - float64 oracle| of the total objective and the gradient on the same inputs: how much
    float32 rounding alone moves the answer (the step model amplifies perturbations). Parity cannot be
    tighter than this, so it is added (x2) to the absolute tolerance of oracle comparisons on random scenes."""
    from oracle import torch_oracle as to
    r32 = to.evaluate(osc, controls, mus_wo, weights)
    r64 = to.evaluate(to.scene_to(osc, torch.float64), controls.double(), mus_wo.double(), weights)
    return (float((r32["total"].double() - r64["total"]).abs().max()), float((r32["grad"].double() - r64["grad"]).abs().max()),
            float((r32["mus"].double() - r64["mus"]).abs().max()))


def check_batched_vs_oracle(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, weights,
                            seed=0, rtol_roll=gio.RTOL_ROLLOUT, rtol_grad=gio.RTOL_GRAD):
    """Random batch (S scenes x C candidates) through the kernel vs. the CPU oracle on the same inputs."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=100 + seed,
                                           param_seed=200 + seed, regular_sigma=True)
    controls_cpu = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=300 + seed, u_max=1.0)
    scene = scene_cpu.to(device)
    controls = controls_cpu.to(device).contiguous()
    # ego-free baseline from the kernel itself, checked against the oracle
    res_wo = engine.launch_evaluate(lib, scene.without_ego(), None, None, want=("mus",), t_horizon=t_hor)
    mus_wo = res_wo["mus"][:, 0].contiguous()
    osc1 = oracle_scene(scene_cpu, 1)
    if env_kind != "kalman":
        ref_wo, _ = to.rollout(env_kind, osc1.ado0, osc1.goal, osc1.v0, osc1.sigma, osc1.weights, None, t_hor, scene_cpu.dt)
        gio.assert_parity(np_(mus_wo), ref_wo.numpy(), rtol_roll, "mus_wo")
    spec = engine.ModuleSpec(weights=weights, mus_wo=mus_wo, control_limits=True, speed_limits=True, min_distance=True)
    stats = torch.zeros(4, dtype=torch.int64, device=device)
    res = engine.launch_evaluate(lib, scene, controls, spec, want=("obj", "obj_raw", "total", "grad", "mus", "sigmas", "violation", "cons"),
                                 stats=stats)
    if device != "cpu":
        torch.cuda.synchronize()
    osc = oracle_scene(scene_cpu, num_candidates)
    wo_rep = mus_wo.detach().cpu().repeat_interleave(num_candidates, dim=0)
    ref = to.evaluate(osc, controls_cpu.reshape(-1, t_hor, 2), wo_rep, weights)
    f_obj, f_grad, f_mus = conditioning_floor(osc, controls_cpu.reshape(-1, t_hor, 2), wo_rep, weights)
    b = num_scenes * num_candidates
    gio.assert_parity(np_(res["mus"]).reshape(b, n, t_hor, 4), ref["mus"].numpy(), rtol_roll, "mus", atol=gio.ATOL + 2 * f_mus)
    if env_kind != "kalman":
        gio.assert_parity(np_(res["sigmas"]).reshape(b, n, t_hor, 2), ref["sigmas"].numpy(), rtol_roll, "sigmas", atol=gio.ATOL + 2 * f_mus)
    gio.assert_parity(np_(res["obj"]).reshape(b, engine.NUM_OBJECTIVES), ref["obj"].numpy(), rtol_roll, "obj", atol=gio.ATOL_OBJ + 2 * f_obj)
    gio.assert_parity(np_(res["total"]).reshape(b), ref["total"].numpy(), rtol_roll, "total", atol=gio.ATOL_OBJ + 2 * f_obj)
    gio.assert_parity(np_(res["grad"]).reshape(b, t_hor, 2), ref["grad"].numpy(), rtol_grad, "grad", atol=gio.ATOL + 2 * f_grad)
    # constraint rows
    et = ref["ego_traj"]
    rows = [to.control_limit_raw(et, scene_cpu.dt, ego_kind) / to.ROBOT_ACC_MAX, to.speed_limit_raw(et) / to.ROBOT_SPEED_MAX,
            to.min_distance_raw(et, ref["mus"]).flatten(start_dim=1)]
    ref_cons = torch.cat(rows, dim=1).numpy()
    gio.assert_parity(np_(res["cons"]).reshape(b, -1), ref_cons, rtol_roll, "cons")
    n_ctl = rows[0].shape[1]
    n_spd = rows[1].shape[1]
    vio = to.violation(ref_cons[:, :n_ctl], None, 1.0) + to.violation(ref_cons[:, n_ctl:n_ctl + n_spd], -1.0, 1.0) + \
        to.violation(ref_cons[:, n_ctl + n_spd:], to.MIN_L2_DISTANCE, None)
    gio.assert_parity(np_(res["violation"]).reshape(b), vio, 1e-4, "violation", atol=1e-5)
    return res, ref, stats


def check_pair_vs_single(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, weights, seed=0, noise=True,
                         mask=False, constraints=True, given_traj=False):
    """The two-candidates-per-thread kernel (even candidate count) against the one-evaluation-per-group kernel (the same
    candidates evaluated one at a time, C = 1). Both spell out the same operation sequence (mtb_math.cuh / mtb_math2.cuh): on
    the host emulation (IEEE operations, no contraction) the forward rollout is BIT-IDENTICAL; under nvcc --use_fast_math the
    compiler contracts / rewrites the packed and the scalar expression trees differently (FMUL2 + FADD2 -> FFMA2, rcp(sqrt) ->
    rsqrt), so on the GPU the rollouts agree to a few ulp and objectives / gradients far inside the parity tolerances."""
    from mantrap_b200 import scenarios
    assert num_candidates % 2 == 0
    scene = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=500 + seed, param_seed=600 + seed,
                                       regular_sigma=True)
    if noise:
        gen = torch.Generator().manual_seed(700 + seed)
        scene.noise = torch.rand((num_scenes, t_hor - 1, n, p, 2), generator=gen)
    scene = scene.to(device)
    controls = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=800 + seed, u_max=1.5).to(device).contiguous()
    mus_wo = engine.launch_evaluate(lib, scene.without_ego(), None, None, want=("mus",), t_horizon=t_hor)["mus"][:, 0].contiguous()
    ado_mask = None
    if mask:
        ado_mask = (torch.arange(n)[None].expand(num_scenes, n) % 2 == 0).to(torch.uint8).to(device).contiguous()
    spec = engine.ModuleSpec(weights=weights, mus_wo=mus_wo, ado_mask=ado_mask, control_limits=constraints, speed_limits=constraints,
                             min_distance=constraints, ellipsoid=constraints)
    want = ["obj", "obj_raw", "total", "grad", "grad_traj", "mus", "sigmas", "violation", "ego_traj"] + (["cons"] if constraints else [])
    ego_traj = None
    if given_traj:
        ego_traj = engine.launch_evaluate(lib, scene, controls, None, want=("ego_traj",))["ego_traj"].contiguous()
    info = engine.query_launch(lib, scene, controls, True)
    assert info["kernel_kind"] in (1, 2) and info["evals_per_thread"] == 2, info
    stats = torch.zeros(4, dtype=torch.int64, device=device)
    pair = engine.launch_evaluate(lib, scene, None if given_traj else controls, spec, want=want, stats=stats, ego_traj=ego_traj)
    stats1 = torch.zeros(4, dtype=torch.int64, device=device)
    singles = []
    for c in range(num_candidates):
        ctl_c = controls[:, c:c + 1].contiguous()
        assert engine.query_launch(lib, scene, ctl_c, True)["kernel_kind"] == 0
        traj_c = None if ego_traj is None else ego_traj[:, c:c + 1].contiguous()
        singles.append(engine.launch_evaluate(lib, scene, None if given_traj else ctl_c, spec, want=want, stats=stats1, ego_traj=traj_c))
    if device != "cpu":
        torch.cuda.synchronize()
    assert stats[2].item() == stats1[2].item(), (stats.tolist(), stats1.tolist())    # saturated evaluations (the pair counters include padding lanes)
    # The probability objective divides mean differences by particle VARIANCES (1e-4 .. 1e-2): it amplifies the last-bit
    # differences of the two rollouts by 1e2 .. 1e3 (kernel_cases.prob_floor), everything else is compared tightly.
    loose = 100.0 if "prob" in weights else 1.0
    # On the GPU the two kernels' rollouts differ by a few ulp per step (see above), which dense scenes amplify like any other
    # float32 re-ordering: add the measured float32-vs-float64 conditioning of the oracle on the same batch (as every oracle test does).
    f_obj = f_grad = f_mus = 0.0
    if device != "cpu":
        osc = oracle_scene(scene.to("cpu"), num_candidates)
        wo_rep = mus_wo.detach().cpu().repeat_interleave(num_candidates, dim=0)
        m_rep = None if ado_mask is None else ado_mask.cpu().bool().repeat_interleave(num_candidates, dim=0)
        from oracle import torch_oracle as to
        ctl_flat = controls.cpu().reshape(-1, t_hor, 2)
        r32 = to.evaluate(osc, ctl_flat, wo_rep, weights, ado_mask=m_rep)
        r64 = to.evaluate(to.scene_to(osc, torch.float64), ctl_flat.double(), wo_rep.double(), weights, ado_mask=m_rep)
        f_obj = 2 * float((r32["obj"].double() - r64["obj"]).abs().max())
        f_grad = 2 * float((r32["grad"].double() - r64["grad"]).abs().max())
        f_mus = 2 * float((r32["mus"].double() - r64["mus"]).abs().max())
        # per evaluation: dense scenes hold the odd evaluation whose float32 gradient is off by 1e-2 from its own float64 value
        # (a force clamp or the interaction radius within an ulp): such an evaluation bounds nothing, the others are held tightly
        g_eval = 4 * (r32["grad"].double() - r64["grad"]).abs().amax(dim=(1, 2)).numpy()
    for key in want:
        a = np.nan_to_num(np_(pair[key]))
        b = np.nan_to_num(np_(torch.cat([r[key] for r in singles], dim=1)))
        assert np.array_equal(np.isnan(np_(pair[key])), np.isnan(np_(torch.cat([r[key] for r in singles], dim=1)))), key
        if key == "ego_traj" or (key in ("mus", "sigmas") and device == "cpu"):
            assert np.array_equal(a, b), (key, float(np.abs(a - b).max()))
            continue
        if key in ("obj", "obj_raw"):
            scale = 120.0 if key == "obj_raw" else 1.0
            gio.assert_parity(a[..., :4], b[..., :4], 2e-6, f"pair vs single: {key}[goal, acc, vel, pos]", atol=2e-7 + f_obj)
            gio.assert_parity(a[..., 4], b[..., 4], 2e-6 * loose, f"pair vs single: {key}[prob]", atol=2e-7 + scale * f_obj)
            continue
        rtol = 2e-5 if key in ("grad", "grad_traj") else 2e-6
        if key in ("grad", "grad_traj", "total"):
            rtol *= loose
        if key == "cons" and device != "cpu":
            rtol = gio.RTOL_ROLLOUT          # ellipsoid rows are (distance / 0.1)^2 ~ 1e4: a few ulp of the means are 3e-6 of the largest row
        floor = {"grad": f_grad, "grad_traj": 10 * f_grad, "total": f_obj, "mus": f_mus, "sigmas": f_mus, "cons": 10 * f_mus,
                 "violation": 100 * f_mus}.get(key, 0.0)
        if key == "grad" and device != "cpu":
            ga, gb = a.reshape(-1, a.shape[-2] * a.shape[-1]), b.reshape(-1, a.shape[-2] * a.shape[-1])
            err = np.abs(ga - gb).max(axis=1)
            bound = max(rtol, 0.5 * gio.RTOL_GRAD) * np.abs(gb).max() + 2e-7 + g_eval       # half the parity tolerance + conditioning
            assert np.all(err <= bound), (f"pair vs single: grad", float((err - bound).max()), int(np.argmax(err - bound)))
            continue
        gio.assert_parity(a, b, rtol, f"pair vs single: {key}", atol=2e-7 + floor)
    return pair


def check_lean_vs_general(lib, device, num_scenes, num_candidates, n, p, t_hor, seed=0):
    """The LEAN instantiation of the pair kernel (what the batched search launches: controls in, goal + acceleration objective,
    control / speed rows, totals / violation / gradient out) against the general instantiation on the same call (forced by asking
    for the pair counters): identical arithmetic, so identical forward results; gradients up to the order of the L2 float reductions."""
    from mantrap_b200 import scenarios
    scene = scenarios.synthetic_scenes(num_scenes, n, p, seed=40 + seed, param_seed=50 + seed, regular_sigma=True).to(device)
    controls = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=60 + seed, u_max=1.5).to(device).contiguous()
    mus_wo = engine.launch_evaluate(lib, scene.without_ego(), None, None, want=("mus",), t_horizon=t_hor)["mus"][:, 0].contiguous()
    spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True)
    want = ("obj", "total", "grad", "violation", "cons")
    stats = torch.zeros(4, dtype=torch.int64, device=device)
    general = engine.launch_evaluate(lib, scene, controls, spec, want=want, stats=stats)
    lean = engine.launch_evaluate(lib, scene, controls, spec, want=want)
    if device != "cpu":
        torch.cuda.synchronize()
    assert int(stats[1]) > 0
    for key in ("obj", "total", "violation", "cons"):
        assert torch.equal(general[key], lean[key]), key
    gio.assert_parity(np_(lean["grad"]), np_(general["grad"]), 1e-5, "lean vs general gradient", atol=1e-7)
    assert float(lean["grad"].abs().max()) > 0


def check_ellipsoid_golden(lib, device, name):
    """InteractionEllipsoidModule rows (values + violation from the evaluation kernel, Jacobian from the tangent kernel)
    against the reference module's `constraint`, `compute_violation` and `jacobian` on the pinned draw."""
    g = gio.load(name)
    scene = scene_from_golden(g, device)
    t_hor, n = int(g["T"]), g["ado0"].shape[0]
    controls = gio.t(g["controls"])[None, None].to(device).contiguous()
    spec = engine.ModuleSpec(ellipsoid=True)
    res = engine.launch_evaluate(lib, scene, controls, spec, want=("cons", "violation"))
    jac = engine.launch_jacobian(lib, scene, controls, spec)
    if device != "cpu":
        torch.cuda.synchronize()
    assert res["cons"].shape == (1, 1, n * t_hor) and jac.shape == (1, 1, n * t_hor, 2 * t_hor)
    gio.assert_parity(np_(res["cons"][0, 0]), g["con_ellipsoid"], gio.RTOL_ROLLOUT, "ellipsoid rows")
    gio.assert_parity(np_(res["violation"][0, 0]), g["vio_ellipsoid"], 1e-4, "ellipsoid violation", atol=1e-6)
    if str(g["env"]) == "kalman":
        # the reference reports zeros for a non-differentiable environment (optimization_module.py:298-301; the host module
        # does the same); the kernel's rows are the true derivative through the ego rollout -> compare with the oracle
        from oracle import torch_oracle as to
        assert not np.any(g["jac_ellipsoid"])
        _, j_ref = to.ellipsoid_jacobian(oracle_scene_from_golden(g), gio.t(g["controls"])[None])
        gio.assert_parity(np_(jac[0, 0]), j_ref.numpy(), gio.RTOL_GRAD, "jac ellipsoid (kalman, oracle)")
    else:
        gio.assert_parity(np_(jac[0, 0]).flatten(), g["jac_ellipsoid"], gio.RTOL_GRAD, "jac ellipsoid")
    # ... and after the other row blocks, in the documented order control | speed | min-distance | ellipsoid
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True, ellipsoid=True)
    res2 = engine.launch_evaluate(lib, scene, controls, spec, want=("cons",))
    jac2 = engine.launch_jacobian(lib, scene, controls, spec)
    if device != "cpu":
        torch.cuda.synchronize()
    assert torch.equal(res2["cons"][0, 0, -n * t_hor:], res["cons"][0, 0])
    assert torch.equal(jac2[0, 0, -n * t_hor:], jac[0, 0])


def check_ellipsoid_vs_oracle(lib, device, env_kind, ego_kind, num_scenes, num_candidates, n, p, t_hor, seed=0):
    """Random batch: ellipsoid rows, violation and Jacobian vs. the CPU oracle (row-by-row autograd). Controls are scaled so
    that some trajectories pass through ellipsoids (non-zero violation)."""
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=900 + seed,
                                           param_seed=950 + seed, regular_sigma=True)
    # one ado on top of the ego: rows below the bound (violation > 0). Not for social forces: at 7 cm the repulsion potential
    # is near-singular and the scene is outside the float32-conditioned parity domain; there the ado sits 0.8 m away.
    scene_cpu.ado0[:, 0, 0:2] = scene_cpu.ego0[:, 0:2] + (0.05 if env_kind != "sf" else 0.8)
    controls_cpu = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=970 + seed, u_max=1.0)
    spec = engine.ModuleSpec(ellipsoid=True)
    res = engine.launch_evaluate(lib, scene_cpu.to(device), controls_cpu.to(device).contiguous(), spec, want=("cons", "violation"))
    jac = engine.launch_jacobian(lib, scene_cpu.to(device), controls_cpu.to(device).contiguous(), spec)
    if device != "cpu":
        torch.cuda.synchronize()
    cons, vio, jac = np_(res["cons"]), np_(res["violation"]), np_(jac)
    osc = oracle_scene(scene_cpu, num_candidates)
    any_violation = False
    for b in range(num_scenes * num_candidates):
        one = to.Scene(*[(x[b:b + 1] if isinstance(x, torch.Tensor) else x) for x in osc])
        u = controls_cpu.reshape(-1, t_hor, 2)[b:b + 1]
        g_ref, j_ref = to.ellipsoid_jacobian(one, u)
        g64, j64 = to.ellipsoid_jacobian(to.scene_to(one, torch.float64), u.double())
        f_val, f_jac = float((g_ref.double() - g64).abs().max()), float((j_ref.double() - j64).abs().max())
        sc, c = divmod(b, num_candidates)
        gio.assert_parity(cons[sc, c], g_ref.numpy(), gio.RTOL_ROLLOUT, "ellipsoid rows", atol=gio.ATOL + 2 * f_val)
        gio.assert_parity(jac[sc, c], j_ref.numpy(), gio.RTOL_GRAD, "jac ellipsoid", atol=gio.ATOL + 2 * f_jac)
        v_ref = float(to.violation(g_ref.numpy()[None], 1.0, None)[0])
        any_violation |= v_ref > 0
        gio.assert_parity(vio[sc, c], v_ref, 1e-4, "ellipsoid violation", atol=1e-5 + 2 * f_val * n * t_hor)
    assert any_violation or env_kind == "sf"


def check_attention_golden(lib, device):
    """Attention masks against the reference's EuclideanModule / ClosestModule on the 48 recorded scenes (bit-exact), one
    launch per ado count, plus the oracle on a large random batch."""
    from oracle import torch_oracle as to
    g = gio.load("scene_ops")
    for n in sorted(set(g["att_n"].tolist())):
        sel = np.nonzero(g["att_n"] == n)[0]
        ego0 = gio.t(g["att_ego"][sel]).to(device).contiguous()
        ado0 = gio.t(g["att_ado"][sel][:, :n]).to(device).contiguous()
        for kind in ("euclidean", "closest"):
            mask = engine.launch_attention(lib, kind, ego0, ado0, 4.0)
            if device != "cpu":
                torch.cuda.synchronize()
            assert np.array_equal(np_(mask), g[f"att_{kind}"][sel][:, :n]), (kind, n)
    gen = torch.Generator().manual_seed(5)
    for s, n in ((257, 20), (64, 64), (33, 1), (40, 33)):
        ego0 = torch.cat((torch.rand((s, 2), generator=gen) * 10 - 5, torch.zeros((s, 2))), dim=1)
        ado0 = torch.cat((torch.rand((s, n, 2), generator=gen) * 14 - 7, torch.zeros((s, n, 2))), dim=2)
        for kind in ("euclidean", "closest"):
            mask = engine.launch_attention(lib, kind, ego0.to(device).contiguous(), ado0.to(device).contiguous(), 4.0)
            if device != "cpu":
                torch.cuda.synchronize()
            assert np.array_equal(np_(mask), to.attention_mask(kind, ego0, ado0).numpy()), (kind, s, n)


def check_env_step_golden(lib, device):
    """Closed-loop step kernel against five consecutive `env.step` calls of the reference per recorded scene (fed the same
    standard-normal draws), and the warm-start shift."""
    g = gio.load("scene_ops")
    for c in range(len(g["step_n"])):
        n, ego_kind = int(g["step_n"][c]), str(g["step_ego_kind"][c])
        scene = engine.SceneBatch(env_kind="kalman", ego_kind=ego_kind, dt=0.4, ego0=gio.t(g["step_ego0"][c])[None].clone(),
                                  ego_goal=None, ado0=gio.t(g["step_ado0"][c][:n])[None].clone()).to(device)
        for k in range(g["step_actions"].shape[1]):
            t_hor = 3
            controls = torch.zeros((1, t_hor, 2))
            controls[0, 0] = gio.t(g["step_actions"][c][k])
            controls[0, 1:] = torch.tensor([[1.0, 2.0], [3.0, 4.0]])
            warm = torch.full((1, t_hor, 2), -7.0).to(device)
            eps = gio.t(g["step_eps"][c][k][:n])[None].to(device).contiguous()
            engine.launch_env_step(lib, scene, controls.to(device).contiguous(), eps, warm)
            if device != "cpu":
                torch.cuda.synchronize()
            gio.assert_parity(np_(scene.ego0[0]), g["step_ego_states"][c][k], gio.RTOL_ROLLOUT, f"ego state {c}/{k}")
            gio.assert_parity(np_(scene.ado0[0]), g["step_ado_states"][c][k][:n], gio.RTOL_ROLLOUT, f"ado states {c}/{k}")
            assert np_(warm[0]).tolist() == [[1.0, 2.0], [3.0, 4.0], [0.0, 0.0]]


def check_closed_loop(lib, device, env_kind="sf", ego_kind="double", num_scenes=3, n=4, p=3, t_hor=5, num_candidates=6,
                      steps=3, attention=None):
    """Batched closed-loop MPC (`mantrap_b200.mpc.BatchedClosedLoop`: attention -> baseline -> candidates -> argmin -> step)
    against a per-scene loop over the CPU oracle fed the same candidates and standard-normal draws."""
    from oracle import torch_oracle as to
    from mantrap_b200 import mpc, scenarios
    eng = engine.Engine.from_library(lib, device) if device == "cpu" else engine.Engine(device)
    scene_cpu = scenarios.synthetic_scenes(num_scenes, n, p, env_kind=env_kind, ego_kind=ego_kind, seed=77, param_seed=78,
                                           regular_sigma=True)
    scene_cpu.ego_goal = scene_cpu.ego0[:, 0:2] + torch.tensor([[0.6, 0.3]])           # close goals: some scenes finish early
    scene_cpu.ego_goal[-1] += 6.0
    scene_cpu.ado0[:, 0, 0:2] = scene_cpu.ego0[:, 0:2] + 1.5                            # one ado inside the attention radius
    gen = torch.Generator().manual_seed(9)
    candidates = [(torch.rand((num_scenes, num_candidates, t_hor, 2), generator=gen) * 2 - 1) * 2.0 for _ in range(steps)]
    eps = [torch.randn((num_scenes, n, 2), generator=gen) for _ in range(steps)]
    weights = {"goal": 0.5, "acc": 1.0}
    loop = mpc.BatchedClosedLoop(eng, scene_cpu, t_planning=t_hor, weights=weights, num_candidates=num_candidates, attention=attention)
    out = loop.solve(steps, candidates=candidates, eps=eps)
    if device != "cpu":
        torch.cuda.synchronize()

    for sc in range(num_scenes):
        ego, ado = scene_cpu.ego0[sc:sc + 1].clone(), scene_cpu.ado0[sc:sc + 1].clone()
        warm = torch.zeros((t_hor, 2))
        done_at = steps
        for k in range(steps):
            cand = candidates[k][sc].clone()
            cand[0] = warm
            rep = lambda x: x[sc:sc + 1].expand(num_candidates, *x.shape[1:])
            osc = to.Scene(env_kind, ego_kind, scene_cpu.dt, ego.expand(num_candidates, -1), rep(scene_cpu.ego_goal),
                           ado.expand(num_candidates, -1, -1), rep(scene_cpu.ado_goal), rep(scene_cpu.v0), rep(scene_cpu.sigma),
                           rep(scene_cpu.weight), None)
            mask = None
            if attention is not None:
                mask = to.attention_mask(attention, ego, ado).bool().expand(num_candidates, -1)
            mus_wo, _ = to.rollout(env_kind, osc.ado0[:1], osc.goal[:1], osc.v0[:1], osc.sigma[:1], osc.weights[:1], None, t_hor, osc.dt)
            res = to.evaluate(osc, cand, mus_wo.expand(num_candidates, -1, -1, -1), weights, ado_mask=mask, want_grad=False)
            vio = to.violation(to.control_limit_raw(res["ego_traj"], osc.dt, ego_kind).numpy() / to.ROBOT_ACC_MAX, None, 1.0) + \
                to.violation(to.speed_limit_raw(res["ego_traj"]).numpy() / to.ROBOT_SPEED_MAX, -1.0, 1.0)
            total = res["total"].numpy().copy()
            feasible = vio < 1e-3
            idx = int(np.argmin(np.where(feasible, total, np.inf))) if feasible.any() else -1
            assert int(out.best_index[sc, k]) == idx, (sc, k, idx, total, vio)
            best = cand[idx] if idx >= 0 else warm
            gio.assert_parity(np_(out.objective[sc, k]), total[max(idx, 0)], gio.RTOL_ROLLOUT, "objective", atol=gio.ATOL_OBJ)
            ego, ado = to.env_step(ego_kind, osc.dt, ego, ado, best[0:1], eps[k][sc:sc + 1])
            warm = torch.cat((best[1:], torch.zeros((1, 2))))
            gio.assert_parity(np_(out.ego_trajectory[sc, k + 1]), ego[0].numpy(), gio.RTOL_ROLLOUT, f"ego {sc}/{k}")
            gio.assert_parity(np_(out.ado_trajectories[sc, :, k + 1]), ado[0].numpy(), gio.RTOL_ROLLOUT, f"ados {sc}/{k}")
            gio.assert_parity(np_(out.controls[sc, k]), best[0].numpy(), gio.RTOL_ROLLOUT, f"control {sc}/{k}")
            if float(torch.norm(ego[0, 0:2] - scene_cpu.ego_goal[sc])) < 0.5:
                done_at = k + 1
                break
        assert int(out.steps[sc]) == done_at, (sc, int(out.steps[sc]), done_at)
        # frozen after the goal
        for k in range(done_at, steps):
            assert torch.equal(out.ego_trajectory[sc, k + 1], out.ego_trajectory[sc, done_at])
    return out


def check_warm_start_encoding(lib, device):
    """Scene encoding + nearest pre-computed solution against `TrajOptSolver.encode` / `warm_start("encoding")` of the reference
    on 40 recorded scenes (database recorded with the fixture), and against the oracle on a random batch."""
    from oracle import torch_oracle as to
    g = gio.load("scene_ops")
    enc_db, sol_db = gio.t(g["ws_encoding_db"]).to(device).contiguous(), gio.t(g["ws_solution_db"]).to(device).contiguous()
    for i in range(len(g["ws_n"])):
        n, t_hor = int(g["ws_n"][i]), int(g["ws_T"][i])
        ego0, goal = gio.t(g["ws_ego"][i])[None].to(device), gio.t(g["ws_goal"][i])[None].to(device)
        ado0 = gio.t(g["ws_ado"][i][:n])[None].to(device).contiguous()
        enc, match, warm = engine.launch_warm_start_encoding(lib, ego0, goal, ado0, enc_db, sol_db, t_hor)
        if device != "cpu":
            torch.cuda.synchronize()
        gio.assert_parity(np_(enc[0]), g["ws_encoding"][i], gio.RTOL_ROLLOUT, f"encoding {i}", atol=1e-6)
        assert int(match[0]) == int(g["ws_match"][i]), i
        assert np.array_equal(np_(warm[0]), g["ws_warm"][i][:t_hor]), i
    gen = torch.Generator().manual_seed(17)
    s, n, t_hor = 300, 20, 7
    ego0 = torch.cat((torch.rand((s, 2), generator=gen) * 16 - 8, torch.rand((s, 2), generator=gen) * 4 - 2), dim=1)
    goal = torch.rand((s, 2), generator=gen) * 16 - 8
    ado0 = torch.cat((torch.rand((s, n, 2), generator=gen) * 18 - 9, torch.zeros((s, n, 2))), dim=2)
    enc, match, warm = engine.launch_warm_start_encoding(lib, ego0.to(device), goal.to(device), ado0.to(device).contiguous(), enc_db,
                                                         sol_db, t_hor)
    if device != "cpu":
        torch.cuda.synchronize()
    enc_ref = to.scene_encoding(ego0, goal, ado0)
    match_ref, warm_ref = to.warm_start_encoding(enc_ref, gio.t(g["ws_encoding_db"]), gio.t(g["ws_solution_db"]), t_hor)
    gio.assert_parity(np_(enc), enc_ref.numpy(), gio.RTOL_ROLLOUT, "encoding batch", atol=1e-6)
    # a near tie between two database rows may resolve differently under float32 rounding of the rotation: the distance to the
    # chosen row must equal the minimum distance to rounding
    dist = torch.cdist(enc_ref, gio.t(g["ws_encoding_db"]))
    chosen = dist[torch.arange(s), match.cpu().long()]
    assert torch.all(chosen <= dist.min(dim=1).values + 1e-4)
    same = match.cpu().long() == match_ref
    assert float(same.float().mean()) > 0.98
    assert torch.equal(warm.cpu()[same], warm_ref[same])


def check_closed_loop_encoding_warm_start(lib, device):
    """Closed loop started from the nearest pre-computed scenario of every scene: with ONE candidate per scene (= the warm
    start itself) the first executed control is row 0 of the database solution the oracle picks."""
    from oracle import torch_oracle as to
    from mantrap_b200 import mpc, scenarios
    eng = engine.Engine.from_library(lib, device) if device == "cpu" else engine.Engine(device)
    g = gio.load("scene_ops")
    db = (gio.t(g["ws_encoding_db"]), gio.t(g["ws_solution_db"]))
    scene = scenarios.synthetic_scenes(6, 4, 3, regular_sigma=True, seed=5, param_seed=6)
    scene.ego0[:, 2:4] = 0.3
    loop = mpc.BatchedClosedLoop(eng, scene, t_planning=5, num_candidates=1, control_limits=False, speed_limits=False)
    out = loop.solve(1, sample=False, warm_start_db=db)
    _, warm = to.warm_start_encoding(to.scene_encoding(scene.ego0, scene.ego_goal, scene.ado0), db[0], db[1], 5)
    assert torch.equal(out.best_index[:, 0].cpu(), torch.zeros(6, dtype=torch.int32))
    assert torch.equal(out.controls[:, 0].cpu(), warm[:, 0])
