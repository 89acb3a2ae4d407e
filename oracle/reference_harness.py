"""Reference harness: imports the UNMODIFIED mantrap reference from /root/reference and evaluates the hot path
with the torch CPU RNG pinned, recording every random number it consumed.

TEST INFRASTRUCTURE ONLY. This file is used in the build container to (a) generate the golden vectors under
`tests/golden/` (`tests/golden/generate_golden.py`) and (b) validate the CPU restatement in
`oracle/torch_oracle.py`. `/root/reference` does not exist on the GPU box, so nothing on the GPU test / smoke /
bench path imports this module. The product package `mantrap_b200` never imports anything from `oracle/`.

Reference entry points used (all relative to /root/reference):
  * `mantrap/environment/base/particle.py:98-183`  `ParticleEnvironment._compute_distributions`
  * `mantrap/environment/base/particle.py:17-79`   `create_particles` (RNG order)
  * `mantrap/environment/simplified/kalman.py:38-94`
  * `mantrap/modules/base/optimization_module.py:75-307` objective / gradient / constraint / jacobian
  * `mantrap/agents/base/linear.py:68-105` `unroll_trajectory`
"""
import os
import sys
import typing

import numpy as np
import torch

REFERENCE_ROOT = "/root/reference"
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_stubs")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "mantrap"))


def import_reference():
    """Import the reference `mantrap` package with the three stub modules."""
    if not reference_available():
        raise RuntimeError("reference not mounted at /root/reference (expected on the GPU box)")
    for path in (REFERENCE_ROOT, _STUBS):
        if path not in sys.path:
            sys.path.insert(0, path)
    import mantrap  # noqa
    return mantrap


ENV_KINDS = {"sf": "SocialForcesEnvironment", "pf": "PotentialFieldEnvironment", "kalman": "KalmanEnvironment"}
EGO_KINDS = {"double": "DoubleIntegratorDTAgent", "single": "IntegratorDTAgent"}


def build_env(scene: typing.Dict):
    """Build a reference environment from a plain scene dictionary.

    scene = {env: sf|pf|kalman, ego_kind: double|single, ego_pos: (2), ego_vel: (2), dt: float,
             ados: [{pos: (2), vel: (2), goal: (2)}, ...]}
    """
    mantrap = import_reference()
    env_class = getattr(mantrap.environment, ENV_KINDS[scene["env"]])
    ego_class = getattr(mantrap.agents, EGO_KINDS[scene["ego_kind"]])
    env = env_class(ego_position=torch.tensor(scene["ego_pos"], dtype=torch.float32),
                    ego_velocity=torch.tensor(scene["ego_vel"], dtype=torch.float32),
                    ego_type=ego_class, dt=scene["dt"])
    for i, ado in enumerate(scene["ados"]):
        kwargs = {}
        if scene["env"] == "sf":
            kwargs["goal"] = torch.tensor(ado["goal"], dtype=torch.float32)
        env.add_ado(position=torch.tensor(ado["pos"], dtype=torch.float32),
                    velocity=torch.tensor(ado["vel"], dtype=torch.float32),
                    identifier=f"a{i:02d}", **kwargs)
    return env


def redraw_params(env_kind: str, num_ados: int, num_particles: int, num_steps: int, seed: int):
    """Re-draw exactly the random numbers one `_compute_distributions` call consumes after
    `torch.manual_seed(seed)`: per ado, per parameter (SF: v0 then sigma; PF: v0) one
    `Normal(loc, scale).sample((P,))`; then per step one `torch.rand((N, P, 2))`.

    Returns v0[N,P], sigma[N,P] (ones for PF), cdf weights normalised over the ADO axis [N,P]
    (`particle.py:62,79,141`), noise[T-1,N,P,2].
    """
    import_reference()
    import mantrap.constants as mc
    if env_kind == "sf":
        specs = [mc.SOCIAL_FORCES_DEFAULT_V0, mc.SOCIAL_FORCES_DEFAULT_SIGMA]
    elif env_kind == "pf":
        specs = [mc.POTENTIAL_FIELD_V0_DEFAULT]
    else:
        raise ValueError(env_kind)
    torch.manual_seed(seed)
    samples = torch.zeros((num_ados, len(specs), num_particles))
    cdfs = torch.zeros((num_ados, len(specs), num_particles))
    for m in range(num_ados):
        for ip, (loc, scale) in enumerate(specs):
            dist = torch.distributions.Normal(loc=loc, scale=scale)
            s = dist.sample((num_particles,))
            samples[m, ip] = s
            cdfs[m, ip] = dist.cdf(s)
    weights = torch.prod(cdfs, dim=1)
    weights = weights / torch.norm(weights, dim=0)
    noise = torch.stack([torch.rand((num_ados, num_particles, 2)) for _ in range(num_steps)]) \
        if num_steps > 0 else torch.zeros((0, num_ados, num_particles, 2))
    v0 = samples[:, 0]
    sigma = samples[:, 1] if len(specs) > 1 else torch.ones_like(v0)
    return v0, sigma, weights, noise


def ref_rollout(env, ego_trajectory, seed: int, num_particles: int):
    """Run the reference rollout twice under the same seed (velocity and position distribution) and return
    mus[N,T,4] = (mu_p, mu_v) and sigmas[N,T,2] (variance used as scale)."""
    is_particle = hasattr(env, "create_particles")
    kwargs = {"num_particles": num_particles} if is_particle else {}
    out = {}
    for vel_dist in (True, False):
        torch.manual_seed(seed)
        out[vel_dist] = env.compute_distributions(ego_trajectory, vel_dist=vel_dist, **kwargs)
    mu_v = torch.stack([out[True][i].mean[:, 0, :] for i in env.ado_ids])
    mu_p = torch.stack([out[False][i].mean[:, 0, :] for i in env.ado_ids])
    sig = torch.stack([out[True][i].stddev[:, 0, :] for i in env.ado_ids])
    sig_p = torch.stack([out[False][i].stddev[:, 0, :] for i in env.ado_ids])
    return torch.cat((mu_p, mu_v), dim=-1), sig, sig_p


def ref_rollout_wo_ego(env, t_horizon: int, seed: int, num_particles: int):
    """Reference rollout without ego (`particle.py:185-198`), same output convention as `ref_rollout`."""
    is_particle = hasattr(env, "create_particles")
    kwargs = {"num_particles": num_particles} if is_particle else {}
    out = {}
    for vel_dist in (True, False):
        torch.manual_seed(seed)
        out[vel_dist] = env.compute_distributions_wo_ego(t_horizon, vel_dist=vel_dist, **kwargs)
    mu_v = torch.stack([out[True][i].mean[:, 0, :] for i in env.ado_ids])
    mu_p = torch.stack([out[False][i].mean[:, 0, :] for i in env.ado_ids])
    sig = torch.stack([out[True][i].stddev[:, 0, :] for i in env.ado_ids])
    return torch.cat((mu_p, mu_v), dim=-1), sig


def to_np(x) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    return np.asarray(x)
