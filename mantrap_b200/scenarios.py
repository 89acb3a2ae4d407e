"""Synthetic scene batches in the shape of the reference's evaluation scenarios.

Follows `mantrap_evaluation/scenarios/custom/random.py:9-39` (positions U(-10,10)^2, pedestrian velocities
U(0, PED_SPEED_MAX/2)^2, robot at rest) and the parameter sampling of `ParticleEnvironment.create_particles`
(mantrap/environment/base/particle.py:40-79): per (ado, particle) v0 ~ N(4,1), sigma ~ N(0.9,0.3) (social forces) or
v0 ~ N(2,1) (potential field), weight = product of parameter cdfs, L2-normalised over the ado axis (:141).
"ETH-shaped" = 20 pedestrians, dt = 0.4 s, 12-step horizon (`mantrap_evaluation/scenarios/eth/__init__.py:24`); the
ETH recordings themselves are not vendored by the reference.
Generation is host-side setup (it mirrors what the reference does with the torch CPU RNG) and is not timed.
"""
import math
import typing

import torch

from . import constants
from .engine import SceneBatch


def _cdf(x, loc, scale):
    return 0.5 * (1.0 + torch.erf((x - loc) / (scale * math.sqrt(2.0))))


def sample_particle_params(env_kind: str, num_scenes: int, num_ados: int, num_particles: int,
                           generator: torch.Generator = None):
    """(v0, sigma, weight) [S,N,P] — one `create_particles` draw per scene."""
    shape = (num_scenes, num_ados, num_particles)
    if env_kind == "sf":
        (m0, s0), (m1, s1) = constants.SOCIAL_FORCES_DEFAULT_V0, constants.SOCIAL_FORCES_DEFAULT_SIGMA
        v0 = torch.normal(torch.full(shape, m0), torch.full(shape, s0), generator=generator)
        sigma = torch.normal(torch.full(shape, m1), torch.full(shape, s1), generator=generator)
        weight = _cdf(v0, m0, s0) * _cdf(sigma, m1, s1)
    elif env_kind == "pf":
        m0, s0 = constants.POTENTIAL_FIELD_V0_DEFAULT
        v0 = torch.normal(torch.full(shape, m0), torch.full(shape, s0), generator=generator)
        sigma = torch.ones(shape)
        weight = _cdf(v0, m0, s0)
    else:
        return None, None, None
    weight = weight / torch.norm(weight, dim=1, keepdim=True)
    return v0, sigma, weight


def synthetic_scenes(num_scenes: int, num_ados: int, num_particles: int = constants.ENV_NUM_PARTICLES,
                     env_kind: str = "sf", ego_kind: str = "double", dt: float = constants.ENV_DT_DEFAULT,
                     seed: int = 1234, param_seed: int = 5678, regular_sigma: bool = False) -> SceneBatch:
    """Random scenes on the CPU (move with `.to(device)`). `regular_sigma` redraws sampled sigma <= 0.05, for which
    the reference potential overflows to inf and its gradient is NaN -> 0 (DESIGN.md "parity domain")."""
    gen = torch.Generator().manual_seed(seed)
    lo, hi = constants.ENV_X_AXIS_DEFAULT

    def uniform(*shape, a=lo, b=hi):
        return torch.rand(*shape, generator=gen) * (b - a) + a

    ego0 = torch.cat((uniform(num_scenes, 2), torch.zeros(num_scenes, 2)), dim=1)
    ego_goal = uniform(num_scenes, 2)
    ado_p = uniform(num_scenes, num_ados, 2)
    ado_v = uniform(num_scenes, num_ados, 2, a=0.0, b=constants.PED_SPEED_MAX / 2)
    ado_goal = uniform(num_scenes, num_ados, 2)
    pgen = torch.Generator().manual_seed(param_seed)
    v0, sigma, weight = sample_particle_params(env_kind, num_scenes, num_ados, num_particles, pgen)
    if regular_sigma and env_kind == "sf":
        m1, s1 = constants.SOCIAL_FORCES_DEFAULT_SIGMA
        for _ in range(16):
            bad = sigma <= 0.05
            if not bool(bad.any()):
                break
            redraw = torch.normal(torch.full(sigma.shape, m1), torch.full(sigma.shape, s1), generator=pgen)
            sigma = torch.where(bad, redraw, sigma)
        (m0, s0) = constants.SOCIAL_FORCES_DEFAULT_V0
        weight = _cdf(v0, m0, s0) * _cdf(sigma, m1, s1)
        weight = weight / torch.norm(weight, dim=1, keepdim=True)
    return SceneBatch(env_kind=env_kind, ego_kind=ego_kind, dt=dt, ego0=ego0.contiguous(), ego_goal=ego_goal.contiguous(),
                      ado0=torch.cat((ado_p, ado_v), dim=2).contiguous(), ado_goal=ado_goal.contiguous(),
                      v0=v0, sigma=sigma, weight=weight)


def candidate_controls(num_scenes: int, num_candidates: int, t_horizon: int, seed: int = 4321,
                       u_max: float = constants.ROBOT_ACC_MAX, pinned: bool = False) -> torch.Tensor:
    """z ~ U(-u_max, u_max)^{T x 2} per candidate (solver/base/trajopt.py:359-363, baselines/random_search.py:28)."""
    gen = torch.Generator().manual_seed(seed)
    z = (torch.rand((num_scenes, num_candidates, t_horizon, 2), generator=gen) * 2.0 - 1.0) * u_max
    return z.pin_memory() if pinned else z
