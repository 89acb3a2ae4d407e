"""CPU oracle: a vectorised float32 torch restatement of mantrap's planning hot path.

TEST INFRASTRUCTURE ONLY — this module is the *checker*. Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import it; the product package `mantrap_b200` never
does (it fails loudly when its CUDA library is missing instead of falling back to anything here).

Parity status: PINNED. Every function below is checked in `tests/test_oracle_golden.py` against golden vectors
under `tests/golden/*.npz`, which `tests/golden/generate_golden.py` produced by importing the UNMODIFIED reference
from /root/reference with `torch.manual_seed` pinned (the reference's own tests hold no numeric vectors for the
particle rollouts; its dynamics known-answer tests `test/test_agents.py:176-279` are replayed as well).
Exception: `min_distance` (the module does not exist at the mounted reference commit) is pinned
only by composition of reference pieces (reference rollout + torch.norm + autograd), see the golden generator.

Gradients here come from torch.autograd applied to this restatement (closed-form social force), i.e. they are
independent of the hand-derived adjoint inside the CUDA kernels.

All tensors are float32 on the CPU. A leading batch axis B (independent evaluations) is supported everywhere so
that the same code serves as the timed CPU baseline.

Reference citations are relative to /root/reference.
"""
import math
import typing

import numpy as np
import torch

# mantrap/constants.py
PED_SPEED_MAX = 2.5                      # :28
ROBOT_SPEED_MAX = 2.0                    # :30
ROBOT_ACC_MAX = 2.0                      # :31
ENV_VAR_INITIAL = 1e-6                   # :43
ENV_PARTICLE_NOISE = 1e-6                # :46
KALMAN_ADDITIVE_NOISE = 0.2              # :48
SF_TAU = 0.4                             # :50
SF_V0 = (4.0, 1.0)                       # :51
SF_SIGMA = (0.9, 0.3)                    # :52
SF_MAX_GOAL_DISTANCE = 0.3               # :53
SF_MAX_INTERACTION_DISTANCE = 2.0        # :54
PF_V0 = (2.0, 1.0)                       # :56
PF_MAX_THETA = 30.0                      # :57
OBJ_POS_MAX, OBJ_VEL_MAX, OBJ_ACC_MAX = 3.0, 1.0, 1.0   # :101-103
OBJ_PROB_MAX = 120.0                                    # :100
MIN_L2_DISTANCE = 0.5                    # :107
VIOLATION_PRECISION = 1e-5               # :108


###########################################################################
# Ego dynamics ############################################################
###########################################################################
def ego_unroll(x0: torch.Tensor, controls: torch.Tensor, dt: float, ego_kind: str) -> torch.Tensor:
    """Closed form of `LinearDTAgent.unroll_trajectory` (mantrap/agents/base/linear.py:68-105) for the
    double integrator (agents/integrator_double.py:16-27: explicit Euler, position integrates the OLD velocity)
    and the single integrator (agents/integrator_single.py:16-27).

    :param x0: [..., 4] = (px, py, vx, vy), :param controls: [..., T, 2] -> trajectory [..., T+1, 4].
    """
    p0, v0 = x0[..., None, 0:2], x0[..., None, 2:4]
    if ego_kind == "double":
        v = torch.cat((v0, v0 + dt * torch.cumsum(controls, dim=-2)), dim=-2)          # [.., T+1, 2]
        p = torch.cat((p0, p0 + dt * torch.cumsum(v[..., :-1, :], dim=-2)), dim=-2)
    elif ego_kind == "single":
        v = torch.cat((v0, controls), dim=-2)
        p = torch.cat((p0, p0 + dt * torch.cumsum(controls, dim=-2)), dim=-2)
    else:
        raise ValueError(ego_kind)
    return torch.cat((p, v), dim=-1)


def ego_roll(trajectory: torch.Tensor, dt: float, ego_kind: str) -> torch.Tensor:
    """`_inverse_dynamics_batch` (integrator_double.py:40-41, integrator_single.py:39-40)."""
    if ego_kind == "double":
        return (trajectory[..., 1:, 2:4] - trajectory[..., :-1, 2:4]) / dt
    return (trajectory[..., 1:, 0:2] - trajectory[..., :-1, 0:2]) / dt


###########################################################################
# Particle parameters (RNG order) #########################################
###########################################################################
def draw_params(env_kind: str, num_ados: int, num_particles: int, num_steps: int,
                generator: typing.Optional[torch.Generator] = None):
    """Draw (v0, sigma, weights, noise) consuming the generator in the order of one reference
    `_compute_distributions` call (particle.py:40-79 then :165). With
    `generator=None` the global torch CPU generator is used, which reproduces the reference bit-for-bit after
    the same `torch.manual_seed`.
    Weights are the product of parameter cdfs (particle.py:62,79) normalised by the L2 norm over ADOS (:141)."""
    specs = [SF_V0, SF_SIGMA] if env_kind == "sf" else [PF_V0]
    samples = torch.zeros((num_ados, len(specs), num_particles))
    cdfs = torch.zeros_like(samples)
    for m in range(num_ados):
        for ip, (loc, scale) in enumerate(specs):
            s = torch.normal(torch.full((num_particles,), float(loc)), torch.full((num_particles,), float(scale)),
                             generator=generator)
            samples[m, ip] = s
            cdfs[m, ip] = 0.5 * (1 + torch.erf((s - loc) / scale / math.sqrt(2)))
    weights = torch.prod(cdfs, dim=1)
    weights = weights / torch.norm(weights, dim=0)
    noise = torch.rand((num_steps, num_ados, num_particles, 2), generator=generator) if num_steps > 0 \
        else torch.zeros((0, num_ados, num_particles, 2))
    v0 = samples[:, 0]
    sigma = samples[:, 1] if len(specs) > 1 else torch.ones_like(v0)
    return v0, sigma, weights, noise


###########################################################################
# Step models #############################################################
###########################################################################
def _safe_norm(x: torch.Tensor) -> torch.Tensor:
    return torch.norm(x, dim=-1, keepdim=True)   # backward at 0 is 0 in torch (probe verified)


def _repulsion(r, w, v0, sigma, dt, active):
    """Closed form of `_repulsive_force` (social_forces.py:94-126): F = dV/dr with V = v0 exp(-b/sigma),
    clamp to +-PED_SPEED_MAX element-wise (:122), NaN -> zeros(2) (:123-124).
    `active` masks pairs that the reference never evaluates (they must not contribute values or gradients)."""
    safe_r = r.new_tensor([1.0, 0.0])
    r = torch.where(active, r, safe_r.expand_as(r))
    w = torch.where(active, w, torch.zeros_like(w))

    def force(r_, w_):
        s = r_ - w_ * dt
        n1 = _safe_norm(r_)
        n2 = _safe_norm(s)
        n3sq = (w_ * w_).sum(-1, keepdim=True)
        a = n1 + n2
        big_b = a * a - (dt * dt) * n3sq
        b = 0.5 * torch.sqrt(big_b)
        pot = v0 * torch.exp(-b / sigma)
        # dV/dn1 = dV/dn2 =: k. The reference obtains F through autograd as k*(r/n1) + k*(s/n2), i.e. two separate
        # products: when V overflows to inf (sampled sigma <= 0) opposite-signed terms give inf - inf = NaN and the
        # whole force is zeroed (:123-124), whereas k*(r/n1 + s/n2) would give +-inf -> clamp. Keep the two products.
        k = -(pot / sigma) * (a / (4.0 * b))
        return k * (r_ / n1) + k * (s / n2)

    with torch.no_grad():
        bad = torch.isnan(force(r, w)).any(-1, keepdim=True)
    ok = active & ~bad
    r = torch.where(ok, r, safe_r.expand_as(r))
    w = torch.where(ok, w, torch.zeros_like(w))
    f = force(r, w).clamp(-PED_SPEED_MAX, PED_SPEED_MAX)
    return torch.where(ok, f, torch.zeros_like(f))


def _feasible(u):
    """`make_controls_feasible` for a pedestrian single integrator (discrete.py:358-370)."""
    nu = torch.norm(u, dim=-1, keepdim=True)
    return torch.div(u, nu.clamp(min=1e-6)) * nu.clamp(-PED_SPEED_MAX, PED_SPEED_MAX)


def sf_controls(p, v, v0, sigma, goal, mu_t, ego_t, dt, tau=SF_TAU):
    """`SocialForcesEnvironment.simulate_particle` (social_forces.py:128-165) for all particles at once.
    p, v: [B,N,P,2]; v0, sigma: [B,N,P,1]; goal: [B,N,1,2]; mu_t: [B,N,4]; ego_t: [B,4] or None."""
    n = p.shape[1]
    d = goal - p
    nd = _safe_norm(d)
    gate = nd >= SF_MAX_GOAL_DISTANCE
    nd_s = torch.where(gate, nd, torch.ones_like(nd))
    f_goal = torch.where(gate, (d / nd_s * _safe_norm(v) - v) / tau, torch.zeros_like(v))

    mp = mu_t[:, None, None, :, 0:2]                       # [B,1,1,N,2]
    mv = mu_t[:, None, None, :, 2:4]
    r = p[:, :, :, None, :] - mp                           # [B,N,P,N,2]
    w = v[:, :, :, None, :] - mv
    eye = torch.eye(n, dtype=torch.bool)[None, :, None, :, None]
    active = (torch.norm(r.detach(), dim=-1, keepdim=True) <= SF_MAX_INTERACTION_DISTANCE) & ~eye
    f_pair = _repulsion(r, w, v0[..., None, :], sigma[..., None, :], dt, active)
    f_rep = -f_pair.sum(dim=3)

    u = f_goal + f_rep
    if ego_t is not None:
        r = p - ego_t[:, None, None, 0:2]
        w = v - ego_t[:, None, None, 2:4]
        u = u - _repulsion(r, w, v0, sigma, dt, torch.ones_like(v0, dtype=torch.bool))
    return u


def pf_controls(p, v, v0, ego_t):
    """`PotentialFieldEnvironment.simulate_particle` (potential_field.py:70-86)."""
    if ego_t is None:
        return v
    v0c = v0.clamp(min=1e-3)
    delta = ego_t[:, None, None, 0:2] - p
    theta = torch.atan2(v[..., 1:2], v[..., 0:1]) - torch.atan2(delta[..., 1:2], delta[..., 0:1])
    gate = theta.detach().abs() < torch.tensor(PF_MAX_THETA / 180.0 * math.pi, dtype=theta.dtype)
    impact = -v0c * torch.sign(delta) * torch.exp(-delta.abs())
    return v + torch.where(gate, impact, torch.zeros_like(impact))


###########################################################################
# Environment rollouts ####################################################
###########################################################################
def rollout(env_kind: str, ado0, goal, v0, sigma, weights, noise, ego_traj, dt, tau=SF_TAU):
    """`ParticleEnvironment._compute_distributions` (particle.py:140-181) / Kalman (kalman.py:63-92).

    ado0 [B,N,4]; goal [B,N,2]; v0/sigma/weights [B,N,P]; noise [B,T-1,N,P,2] or None;
    ego_traj [B,T+1,4] or an int T (rollout without ego). Returns mus [B,N,T,4], sigmas [B,N,T,2].
    """
    if isinstance(ego_traj, int):
        t_horizon, ego_traj = ego_traj, None
    else:
        t_horizon = ego_traj.shape[-2] - 1
    b, n = ado0.shape[0], ado0.shape[1]
    if env_kind == "kalman":
        t = torch.arange(t_horizon, dtype=ado0.dtype)[None, None, :, None]
        mu_v = ado0[:, :, None, 2:4].expand(b, n, t_horizon, 2)
        mu_p = ado0[:, :, None, 0:2] + (t * dt) * ado0[:, :, None, 2:4]
        # P_k = F P F^T + Q with F = diag(1,1,0,0): velocity variance is reset to Q every step (kalman.py:79-92).
        sig_v = torch.full((b, n, t_horizon, 2), KALMAN_ADDITIVE_NOISE, dtype=ado0.dtype)
        sig_v[:, :, 0, :] = ENV_VAR_INITIAL
        return torch.cat((mu_p, mu_v), dim=-1), sig_v

    num_p = v0.shape[-1]
    p = ado0[:, :, None, 0:2].expand(b, n, num_p, 2)
    v = ado0[:, :, None, 2:4].expand(b, n, num_p, 2)
    v0_, sigma_, w_ = v0[..., None], sigma[..., None], weights[..., None]
    mu_t = ado0[..., 0:4]
    mus, sigmas = [mu_t], [torch.full((b, n, 2), ENV_VAR_INITIAL, dtype=ado0.dtype)]
    for t in range(t_horizon - 1):
        ego_t = None if ego_traj is None else ego_traj[:, t, :]
        if env_kind == "sf":
            u = sf_controls(p, v, v0_, sigma_, goal[:, :, None, :], mu_t, ego_t, dt, tau)
        else:
            u = pf_controls(p, v, v0_, ego_t)
        u = _feasible(u)
        v = u
        p = p + u * dt
        vel = v if noise is None else v + noise[:, t] * ENV_PARTICLE_NOISE
        q = vel * w_
        mu_v = torch.mean(q, dim=2)
        mu_p = mu_t[..., 0:2] + mu_t[..., 2:4] * dt
        mu_t = torch.cat((mu_p, mu_v), dim=-1)
        mus.append(mu_t)
        sigmas.append(torch.var(q, dim=2))
    return torch.stack(mus, dim=2), torch.stack(sigmas, dim=2)


def kalman_pos_sigma(b, n, t_horizon):
    """Positional variance of the Kalman prediction: 1e-6 + 0.2 t (kalman.py:84)."""
    t = torch.arange(t_horizon, dtype=torch.float32)
    return (ENV_VAR_INITIAL + KALMAN_ADDITIVE_NOISE * t)[None, None, :, None].expand(b, n, t_horizon, 2)


###########################################################################
# Modules #################################################################
###########################################################################
def goal_norm_raw(ego_traj, goal):
    """`GoalNormModule._objective_core` (modules/goal_norm.py:59-60). ego_traj [B,T+1,4], goal [B,2] -> [B]."""
    return (ego_traj[..., 0:2] - goal[:, None, :]).pow(2).mean(dim=(-1, -2))


def interaction_raw(kind: str, mus, mus_wo, dt, ado_mask=None):
    """`InteractionAcceleration/Velocity/PositionModule._objective_core` (modules/baselines/acc_interact.py:53-68,
    vel_interact.py:30-42, pos_interact.py:30-42) on the collapsed means. Returns (clamped, unclamped) cost [B].

    mus, mus_wo: [B,N,T,4]; ado_mask: [B,N] bool of `ado_ids` entering the mean (None = all).
    Quirk: the position module compares with-ego POSITIONS against without-ego VELOCITIES and its
    norm runs over the size-1 mode axis, i.e. it is a mean absolute difference."""
    if kind == "acc":
        x = (mus[..., 1:, 2:4] - mus[..., :-1, 2:4]) / dt
        x_wo = (mus_wo[..., 1:, 2:4] - mus_wo[..., :-1, 2:4]) / dt
        cost_ado = torch.norm(x - x_wo, dim=-1).mean(dim=-1)            # [B,N]
        max_value = OBJ_ACC_MAX
    elif kind == "vel":
        cost_ado = torch.norm(mus[..., 2:4] - mus_wo[..., 2:4], dim=-1).mean(dim=-1)
        max_value = OBJ_VEL_MAX
    elif kind == "pos":
        diff = mus[..., 0:2] - mus_wo[..., 2:4]
        cost_ado = torch.norm(diff[..., None, :], dim=-2).mean(dim=(-1, -2))
        max_value = OBJ_POS_MAX
    else:
        raise ValueError(kind)
    if ado_mask is None:
        cost = cost_ado.mean(dim=-1)
    else:
        m = ado_mask.to(cost_ado.dtype)
        cost = (cost_ado * m).sum(dim=-1) / m.sum(dim=-1).clamp(min=1.0)
    return cost.clamp_max(max_value), cost


INTERACT_MAX = {"acc": OBJ_ACC_MAX, "vel": OBJ_VEL_MAX, "pos": OBJ_POS_MAX, "prob": OBJ_PROB_MAX}


def prob_interact_raw(mus, sigmas, mus_wo, ado_mask=None):
    """`InteractionProbabilityModule._objective_core` (modules/prob_interact.py:57-86): minus the log-probability of the
    ego-free mean velocities in the conditioned distribution Normal(loc = mu_v, scale = sigmas) -- `sigmas` are VARIANCES
    used as scale (particle.py:175,180) --, summed over (t, modes, xy), averaged over the ados of the
    cost, clamped to +-OBJECTIVE_PROB_INTERACT_MAX. mus, mus_wo [B,N,T,4], sigmas [B,N,T,2] -> (clamped, unclamped) [B]."""
    loc, scale, x = mus[..., 2:4], sigmas, mus_wo[..., 2:4]
    if ado_mask is not None:
        # ados outside `ado_ids` are not part of the reference's graph at all: give them harmless inputs instead of
        # multiplying their (possibly overflowing) terms by zero, which would put 0 * inf = NaN into the backward pass.
        sel = ado_mask.to(torch.bool)[..., None, None]
        loc, scale = torch.where(sel, loc, x), torch.where(sel, scale, torch.ones_like(scale))
    dist = torch.distributions.Normal(loc=loc, scale=scale, validate_args=False)
    lp_ado = dist.log_prob(x).sum(dim=(-1, -2))                            # [B,N]
    if ado_mask is None:
        cost = -lp_ado.mean(dim=-1)
    else:
        m = ado_mask.to(lp_ado.dtype)
        cost = -(lp_ado * m).sum(dim=-1) / m.sum(dim=-1).clamp(min=1.0)
    return cost.clamp(-OBJ_PROB_MAX, OBJ_PROB_MAX), cost


def min_distance_raw(ego_traj, mus):
    """North-star MinDistance constraint (absent at the mounted commit; constants.py:107): g[k,t] =
    ||ego_p(t) - mu_p[k,t]||_2 for t = 0..T-1. ego_traj [B,T+1,4], mus [B,N,T,4] -> [B,N,T]."""
    t_horizon = mus.shape[2]
    return torch.norm(ego_traj[:, None, :t_horizon, 0:2] - mus[..., 0:2], dim=-1)


def control_limit_raw(ego_traj, dt, ego_kind):
    """`ControlLimitModule._constraint_core` (modules/control_limits.py:23-36): |u| per component for the double
    integrator (integrator_double.py:108-114), ||u||_2 for the single integrator (integrator_single.py:88-94)."""
    u = ego_roll(ego_traj, dt, ego_kind)
    if ego_kind == "double":
        return u.abs().flatten(start_dim=-2)
    return torch.norm(u, dim=-1)


def speed_limit_raw(ego_traj):
    """`SpeedLimitModule._constraint_core` (modules/speed_limits.py:26-37)."""
    return ego_traj[..., 2:4].flatten(start_dim=-2)


def violation(values: np.ndarray, lower, upper) -> np.ndarray:
    """`OptimizationModule._violation` (optimization_module.py:512-533) for normalised constraint rows [..., R]."""
    lo = -np.inf if lower is None else lower
    hi = np.inf if upper is None else upper
    v = np.maximum(lo - values, 0.0) + np.maximum(values - hi, 0.0)
    v = v.sum(axis=-1)
    return np.where(np.abs(v) < VIOLATION_PRECISION, 0.0, v)


###########################################################################
# Whole evaluations #######################################################
###########################################################################
def scene_to(scene: "Scene", dtype) -> "Scene":
    """Cast every tensor of a Scene (float64 copies are used to measure float32 conditioning in the tests)."""
    return scene._replace(**{k: v.to(dtype) for k, v in scene._asdict().items() if isinstance(v, torch.Tensor)})


class Scene(typing.NamedTuple):
    env_kind: str           # sf | pf | kalman
    ego_kind: str           # double | single
    dt: float
    ego0: torch.Tensor      # [B,4]
    ego_goal: torch.Tensor  # [B,2]
    ado0: torch.Tensor      # [B,N,4]
    goal: torch.Tensor      # [B,N,2]
    v0: torch.Tensor        # [B,N,P]
    sigma: torch.Tensor     # [B,N,P]
    weights: torch.Tensor   # [B,N,P]
    noise: typing.Optional[torch.Tensor]   # [B,T-1,N,P,2]
    axes_diag: float = math.sqrt(20.0 ** 2 + 20.0 ** 2)   # goal_norm.py:116-118 for the default +-10 m axes


def evaluate(scene: Scene, controls: torch.Tensor, mus_wo: typing.Optional[torch.Tensor],
             weights: typing.Dict[str, float], ado_mask=None, want_grad: bool = True):
    """One `IPOPTSolver.objective` + `gradient` evaluation (solver/base/trajopt.py:368-391, solver/ipopt.py:118-135)
    per batch entry: weighted, normalised module objectives and the gradient of their sum w.r.t. the controls.

    weights: {"goal": w, "acc": w, "vel": w, "pos": w, "prob": w} (missing = module absent).
    Returns dict(obj [B, 5] in the order goal/acc/vel/pos/prob, total [B], grad [B,T,2], mus, sigmas, ego_traj, raw).
    """
    controls = controls.detach().clone().requires_grad_(want_grad)
    ego_traj = ego_unroll(scene.ego0, controls, scene.dt, scene.ego_kind)
    mus, sigmas = rollout(scene.env_kind, scene.ado0, scene.goal, scene.v0, scene.sigma, scene.weights,
                          scene.noise, ego_traj, scene.dt)
    b = controls.shape[0]
    raw = torch.zeros((b, 5), dtype=controls.dtype)
    cols = []
    for i, key in enumerate(("goal", "acc", "vel", "pos", "prob")):
        wgt = weights.get(key, 0.0)
        if key == "goal":
            if key not in weights:
                cols.append(torch.zeros(b, dtype=controls.dtype))
                continue
            j = goal_norm_raw(ego_traj, scene.ego_goal)
            raw[:, i] = j.detach()
            cols.append(wgt * j / scene.axes_diag)
        else:
            if key not in weights or mus_wo is None:
                cols.append(torch.zeros(b, dtype=controls.dtype))
                continue
            if key == "prob":
                jc, ju = prob_interact_raw(mus, sigmas, mus_wo, ado_mask)
            else:
                jc, ju = interaction_raw(key, mus, mus_wo, scene.dt, ado_mask)
            raw[:, i] = ju.detach()
            cols.append(wgt * jc / INTERACT_MAX[key])
    obj = torch.stack(cols, dim=1)
    total = obj.sum(dim=1)
    grad = None
    if want_grad:
        if total.requires_grad:
            grad = torch.autograd.grad(total.sum(), controls, allow_unused=True)[0]
        if grad is None:
            grad = torch.zeros_like(controls)
        grad = torch.nan_to_num(grad)             # optimization_module.py:446
    return dict(obj=obj.detach(), total=total.detach(), grad=grad, mus=mus.detach(), sigmas=sigmas.detach(),
                ego_traj=ego_traj.detach(), raw=raw)


def min_distance_jacobian(scene: Scene, controls: torch.Tensor):
    """Row-by-row autograd Jacobian as `compute_gradient_auto_grad` does it (optimization_module.py:428-434).
    controls [1,T,2] -> g [N*T], jac [N*T, 2T]."""
    controls = controls.detach().clone().requires_grad_(True)
    ego_traj = ego_unroll(scene.ego0, controls, scene.dt, scene.ego_kind)
    mus, _ = rollout(scene.env_kind, scene.ado0, scene.goal, scene.v0, scene.sigma, scene.weights,
                     scene.noise, ego_traj, scene.dt)
    g = min_distance_raw(ego_traj, mus)[0].flatten()
    rows = []
    for i in range(g.numel()):
        gr = torch.autograd.grad(g[i], controls, retain_graph=True, allow_unused=True)[0]
        rows.append(torch.zeros_like(controls).flatten() if gr is None else gr.flatten())
    return g.detach(), torch.nan_to_num(torch.stack(rows))


###########################################################################
# Ellipsoid constraint, attention filters, closed-loop step ###############
###########################################################################
ELLIPSOID_MIN_SCALE = 0.01               # modules/baselines/ellipsoid_interact.py:52
ATTENTION_RADIUS = 4.0                   # constants.py:111-112


def ellipsoid_raw(ego_traj, mus, sigmas, env_kind: str = "sf"):
    """`InteractionEllipsoidModule._constraint_core` (modules/baselines/ellipsoid_interact.py:28-57) with the one "mode"
    the particle environments have: sum_xy (ego_p(t) - mu_p[k,t])^2 / clamp(scale[k,t], min=0.01), means and scales
    DETACHED (:74-75). scale = `distribution.stddev` of the positional distribution = the velocity variance for the
    particle environments (particle.py:177-183), var_initial + Q t for Kalman (kalman.py:79-92).
    ego_traj [B,T+1,4], mus [B,N,T,4], sigmas [B,N,T,2] -> [B,N,T]."""
    t_horizon = mus.shape[2]
    if env_kind == "kalman":
        sigmas = kalman_pos_sigma(mus.shape[0], mus.shape[1], t_horizon).to(mus.dtype)
    scale = sigmas.detach().clamp(min=ELLIPSOID_MIN_SCALE)
    dxy = ego_traj[:, None, :t_horizon, 0:2] - mus[..., 0:2].detach()
    return (dxy.pow(2) / scale).sum(dim=-1)


def ellipsoid_jacobian(scene: Scene, controls: torch.Tensor):
    """Row-by-row autograd Jacobian of the ellipsoid rows (optimization_module.py:428-434). controls [1,T,2] ->
    g [N*T], jac [N*T, 2T]."""
    controls = controls.detach().clone().requires_grad_(True)
    ego_traj = ego_unroll(scene.ego0, controls, scene.dt, scene.ego_kind)
    mus, sigmas = rollout(scene.env_kind, scene.ado0, scene.goal, scene.v0, scene.sigma, scene.weights,
                          scene.noise, ego_traj, scene.dt)
    g = ellipsoid_raw(ego_traj, mus, sigmas, scene.env_kind)[0].flatten()
    rows = []
    for i in range(g.numel()):
        gr = torch.autograd.grad(g[i], controls, retain_graph=True, allow_unused=True)[0]
        rows.append(torch.zeros_like(controls).flatten() if gr is None else gr.flatten())
    return g.detach(), torch.nan_to_num(torch.stack(rows))


def attention_mask(kind: str, ego0: torch.Tensor, ado0: torch.Tensor, radius: float = ATTENTION_RADIUS) -> torch.Tensor:
    """`EuclideanModule._compute` (attention/euclidean.py:22-29) / `ClosestModule._compute` (attention/closest.py:22-31)
    as 0/1 masks. ego0 [B,4], ado0 [B,N,4] -> uint8 [B,N]. The closest filter writes inf at `~index_min` with
    index_min a 0-dim integer tensor, i.e. at index -(index_min + 1) (bitwise not), exactly as the reference does."""
    out = torch.zeros(ado0.shape[0:2], dtype=torch.uint8)
    for b in range(ado0.shape[0]):
        distances = torch.norm(ado0[b, :, 0:2] - ego0[b, 0:2], dim=1)
        if kind == "closest":
            index_min = torch.argmin(distances)
            distances[~index_min] = np.inf
        out[b] = (distances < radius).to(torch.uint8)
    return out


def _clamp_action(u: torch.Tensor, limit: float, per_component: bool) -> torch.Tensor:
    """`DTAgent.make_controls_feasible` (agents/base/discrete.py:358-370)."""
    norm = u.abs() if per_component else torch.norm(u, dim=-1, keepdim=True)
    return u / norm.clamp(min=1e-6) * norm.clamp(-limit, limit)


def env_step(ego_kind: str, dt: float, ego0: torch.Tensor, ado0: torch.Tensor, action: torch.Tensor,
             eps: typing.Optional[torch.Tensor]):
    """`GraphBasedEnvironment.step` (environment/base/graph_based.py:82-124) for a batch of scenes.
    ego0 [B,4], ado0 [B,N,4], action [B,2], eps [B,N,2] standard-normal draws behind `Normal.sample` -> next (ego, ados).

    The ego executes the (clamped) action with its own dynamics (discrete.py:144-167). A one-control rollout performs no
    particle step (particle.py:140-181, t_horizon = 1), so every ado samples its velocity from Normal(loc = v, scale =
    ENV_VAR_INITIAL), moves to p + dt sample (`integrate_numerical`), and is updated through its inverse dynamics:
    action = (p' - p) / dt, clamped to the pedestrian speed limit, v' = action, p' = p + action dt."""
    if ego_kind == "double":
        u = _clamp_action(action, ROBOT_ACC_MAX, per_component=True)
        ego = torch.cat((ego0[:, 0:2] + dt * ego0[:, 2:4], ego0[:, 2:4] + dt * u), dim=1)
    else:
        u = _clamp_action(action, ROBOT_SPEED_MAX, per_component=False)
        ego = torch.cat((ego0[:, 0:2] + dt * u, u), dim=1)
    sample = ado0[..., 2:4] if eps is None else ado0[..., 2:4] + ENV_VAR_INITIAL * eps
    pos = ado0[..., 0:2] + sample * dt
    act = _clamp_action((pos - ado0[..., 0:2]) / dt, PED_SPEED_MAX, per_component=False)
    ados = torch.cat((ado0[..., 0:2] + act * dt, act), dim=-1)
    return ego, ados


def scene_encoding(ego0: torch.Tensor, ego_goal: torch.Tensor, ado0: torch.Tensor) -> torch.Tensor:
    """`TrajOptSolver.encode` (solver/base/trajopt.py:466-499) for a batch: position of the ado closest to the robot rotated
    into the robot -> goal frame (`maths.rotation_matrix`, utility/maths.py:226-233, applied to the absolute position) and
    the robot velocity. ego0 [B,4], ego_goal [B,2], ado0 [B,N,4] -> [B,4]."""
    out = torch.zeros((ego0.shape[0], 4))
    for b in range(ego0.shape[0]):
        xy = ego_goal[b] - ego0[b, 0:2]
        theta = torch.atan2(xy[1], xy[0])
        rot = torch.tensor([[torch.cos(theta), torch.sin(theta)], [-torch.sin(theta), torch.cos(theta)]])
        closest = torch.argmin(torch.norm(ado0[b, :, 0:2] - ego0[b, 0:2], dim=1))
        out[b] = torch.cat((torch.matmul(rot, ado0[b, closest, 0:2]), ego0[b, 2:4]))
    return out


def warm_start_encoding(encoding: torch.Tensor, encoding_db: torch.Tensor, solution_db: torch.Tensor, t_horizon: int):
    """`TrajOptSolver._warm_start_encoding` (trajopt.py:308-331): nearest database encoding (L2, first minimum) and the first
    `t_horizon` controls of its pre-computed solution. encoding [B,4] -> (match [B], warm [B,T,2])."""
    match = torch.stack([torch.argmin(torch.norm(encoding_db - e, dim=1)) for e in encoding])
    return match, solution_db[match, :t_horizon]
