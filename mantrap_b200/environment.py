"""Environments of the hot path with the reference's interface (`mantrap/environment`).

An environment holds the scene (ego + ados, host side) and predicts the ados' velocity distributions over the planning
horizon, conditioned on an ego trajectory. The prediction itself -- the particle rollout of
`ParticleEnvironment._compute_distributions` (environment/base/particle.py:98-183) with the social-forces
(environment/social_forces.py:77-167) or potential-field (environment/simplified/potential_field.py:55-88) step model, or
the Kalman prediction (environment/simplified/kalman.py:38-94) -- runs in the fused CUDA kernel through
`autograd.ParticleRollout`, so the returned `torch.distributions.Normal` objects are differentiable w.r.t. the ego
trajectory exactly like the reference's.

Randomness: the reference re-samples the particle parameters on every call (particle.py:140, :59-62) and adds
`torch.rand * 1e-6` per step (:165) from torch's global CPU generator. `draw_particle_parameters` consumes the same
generator with the same calls in the same order, so after `torch.manual_seed(k)` this class and
the reference see identical random numbers.

Mirrors: GraphBasedEnvironment environment/base/graph_based.py:14-560, ParticleEnvironment environment/base/particle.py:14-205,
SocialForcesEnvironment environment/social_forces.py:11-210, PotentialFieldEnvironment environment/simplified/
potential_field.py:10-99, KalmanEnvironment environment/simplified/kalman.py:11-124.
"""
import dataclasses
import math
import typing

import torch

from . import agents, constants, runtime
from .autograd import ParticleRollout
from .engine import SceneBatch


def draw_particle_parameters(ado_ids: typing.List[str], param_defaults: typing.Dict[str, typing.Tuple[float, float]],
                             num_particles: int, num_noise_steps: int, **param_dicts) -> typing.Dict[str, torch.Tensor]:
    """(v0, sigma, weight) [1,N,P] and noise [1,T-1,N,P,2] with the reference's RNG consumption (particle.py:40-79,
    :141, :165). `param_dicts` may override the per-ado (mean, std) like `v0_dict` / `sigma_dict` do."""
    n = len(ado_ids)
    names = list(param_defaults.keys())
    # Normal(loc, scale).sample((P,)) per (ado, parameter) in the reference's order. torch's sampler is
    # `empty(P).normal_(0, 1).mul_(scale).add_(loc)` (distributions/normal.py:77-82 -> normal_out_impl), so one randn(P)
    # per (ado, parameter) consumes the generator identically and ONE mul/add and ONE cdf expression (normal.py:104-110)
    # over all draws give the same bits, without 2 N distribution objects per callback (this draw was the largest host
    # cost of a solver callback). The draws cannot be merged into one call: >= 16 elements switch torch to a different
    # (vectorised Box-Muller) sampler.
    if n > 0:
        locs, scales = [], []
        for ado_id in ado_ids:
            for key in names:
                override = param_dicts.get(f"{key}_dict")
                loc, scale = override[ado_id] if override is not None else param_defaults[key]
                locs.append(float(loc))
                scales.append(float(scale))
        loc_t = torch.tensor(locs, dtype=torch.float32).view(n, len(names), 1)
        scale_t = torch.tensor(scales, dtype=torch.float32).view(n, len(names), 1)
        unit = torch.stack([torch.randn(num_particles) for _ in range(n * len(names))]).view(n, len(names), num_particles)
        samples = unit.mul_(scale_t).add_(loc_t)
        cdfs = 0.5 * (1 + torch.erf((samples - loc_t) * scale_t.reciprocal() / math.sqrt(2)))
    else:
        samples = cdfs = torch.zeros((0, len(names), num_particles))
    weight = torch.prod(cdfs, dim=1)
    weight = weight / torch.norm(weight, dim=0)               # over the ADO axis (reference quirk, particle.py:141)
    out = {"v0": samples[None, :, 0].contiguous(), "weight": weight[None].contiguous(),
           "sigma": (samples[None, :, 1] if len(names) > 1 else torch.ones((1, n, num_particles))).contiguous()}
    if num_noise_steps > 0:
        # one torch.rand per step in the reference (particle.py:165); the CPU uniform sampler is serial, so one call of
        # [T-1, N, P, 2] is the same stream
        out["noise"] = torch.rand((1, num_noise_steps, n, num_particles, 2))
    return out


class GraphBasedEnvironment:
    env_kind: str = None

    def __init__(self, ego_position: torch.Tensor = None, ego_velocity: torch.Tensor = torch.zeros(2),
                 ego_history: torch.Tensor = None, ego_type=agents.DoubleIntegratorDTAgent,
                 x_axis: typing.Tuple[float, float] = constants.ENV_X_AXIS_DEFAULT,
                 y_axis: typing.Tuple[float, float] = constants.ENV_Y_AXIS_DEFAULT,
                 dt: float = constants.ENV_DT_DEFAULT, time: float = 0.0, config_name: str = "unknown"):
        assert x_axis[0] < x_axis[1] and y_axis[0] < y_axis[1] and dt > 0.0
        self._ego = None
        if ego_position is not None:
            assert ego_type is not None
            self._ego = ego_type(ego_position, velocity=ego_velocity, history=ego_history, time=time, is_robot=True,
                                 dt=dt, identifier=constants.ID_EGO)
        self._ados, self._ado_ids = [], []
        self._env_params = {"x_axis": x_axis, "y_axis": y_axis, "config_name": config_name}
        self._dt, self._time = dt, time

    # ---- scene -------------------------------------------------------------------------------------------------
    def add_ado(self, position: torch.Tensor, velocity: torch.Tensor = torch.zeros(2), history: torch.Tensor = None,
                **ado_kwargs) -> agents.IntegratorDTAgent:
        """Add a pedestrian as single integrator (graph_based.py:291-324)."""
        ado = agents.IntegratorDTAgent(position, velocity=velocity, history=history, dt=self.dt, **ado_kwargs)
        assert self.x_axis[0] <= ado.position[0] <= self.x_axis[1] and self.y_axis[0] <= ado.position[1] <= self.y_axis[1]
        assert ado.id not in self._ado_ids
        self._ado_ids.append(ado.id)
        self._ados.append(ado)
        return ado

    def states(self) -> typing.Tuple[typing.Optional[torch.Tensor], torch.Tensor]:
        ado_states = torch.zeros((self.num_ados, 5))
        for m, ado in enumerate(self.ados):
            ado_states[m] = ado.state_with_time
        return (self.ego.state_with_time if self.ego is not None else None), ado_states

    def agent_by_id(self, agent_id: str):
        return self.ego if agent_id == constants.ID_EGO else self.ados[self.index_ado_id(agent_id)]

    def index_ado_id(self, ado_id: str) -> int:
        return self._ado_ids.index(ado_id)

    # ---- simulation step ---------------------------------------------------------------------------------------
    def step(self, ego_action: torch.Tensor) -> typing.Tuple[torch.Tensor, torch.Tensor]:
        """Advance the scene by dt: ego executes `ego_action`, every ado moves to ONE sample of its conditioned
        distribution (graph_based.py:82-124)."""
        assert ego_action.shape == (2,)
        self._time = self._time + self.dt
        self._ego.update(ego_action, dt=self.dt)                      # the ego moves first, as in the reference
        ado_samples = self.sample_w_controls(ego_controls=ego_action.view(1, 2), num_samples=1)
        ado_states = torch.zeros((self.num_ados, 5))
        for m in range(self.num_ados):
            position = ado_samples[m, 0, 1, 0, :].detach().cpu()
            velocity = (position - self.ados[m].position) / self.dt
            self._ados[m].update_inverse(state_next=torch.cat((position, velocity)), dt=self.dt)
            ado_states[m] = self._ados[m].state_with_time
        self.detach()
        return ado_states.detach(), self.ego.state_with_time.detach()

    def step_reset(self, ego_next: typing.Optional[torch.Tensor], ado_next: typing.Optional[torch.Tensor]):
        """Set the next states directly (graph_based.py:126-151)."""
        self._time = self._time + self.dt
        if ego_next is not None:
            self._ego.reset(state=ego_next, history=None)
        if ado_next is not None:
            assert ado_next.shape == (self.num_ados, 5)
            for m in range(self.num_ados):
                self._ados[m].reset(ado_next[m], history=None)
        self.detach()

    # ---- distributions -----------------------------------------------------------------------------------------
    def compute_distributions(self, ego_trajectory: torch.Tensor, vel_dist: bool = True, **kwargs
                              ) -> typing.Dict[str, torch.distributions.Distribution]:
        """ado_id -> Normal with mean [T,1,2] of the velocity (or position) conditioned on the ego trajectory
        [T+1, >=4] (graph_based.py:362-381)."""
        assert ego_trajectory.dim() == 2 and ego_trajectory.shape[1] >= 4 and self.ego is not None
        return self._compute_distributions(ego_trajectory=ego_trajectory, vel_dist=vel_dist, **kwargs)

    def compute_distributions_wo_ego(self, t_horizon: int, vel_dist: bool = True, **kwargs
                                     ) -> typing.Dict[str, torch.distributions.Distribution]:
        assert t_horizon > 0
        return self._compute_distributions(ego_trajectory=None, t_horizon=t_horizon, vel_dist=vel_dist, **kwargs)

    def _compute_distributions(self, ego_trajectory, vel_dist: bool = True, t_horizon: int = None, **kwargs):
        raise NotImplementedError

    def _distributions_from(self, mus: torch.Tensor, sigmas: torch.Tensor, vel_dist: bool, device):
        """mus [N,T,4], sigmas [N,T,2] (engine device) -> the reference's dictionary of Normal([T,1,2])."""
        means = (mus[..., 2:4] if vel_dist else mus[..., 0:2]).unsqueeze(2).to(device)
        scales = sigmas.unsqueeze(2).to(device)
        return {ado_id: torch.distributions.Normal(loc=means[m], scale=scales[m], validate_args=False)
                for m, ado_id in enumerate(self.ado_ids)}

    # ---- predictions (means) and samples -----------------------------------------------------------------------
    def _integrate(self, velocities: torch.Tensor) -> torch.Tensor:
        """`integrate_numerical` (utility/maths.py:136-167): zero-pad, cumsum * dt, + x0 along the time axis (-3)."""
        _, ado_states = self.states()
        x0 = ado_states[:, 0:2].to(velocities.device)
        pad = torch.zeros_like(velocities[..., :1, :, :])
        steps = torch.cat((pad, velocities), dim=-3)
        x0 = x0.view(self.num_ados, *([1] * (velocities.dim() - 2)), 2)
        return torch.cumsum(steps, dim=-3) * self.dt + x0

    def predict_w_trajectory(self, ego_trajectory: torch.Tensor) -> typing.Optional[torch.Tensor]:
        """Mean ado paths [N, T+1, 1, 2] conditioned on the ego trajectory (graph_based.py:239-262)."""
        if self.num_ados == 0:
            return None
        dist = self.compute_distributions(ego_trajectory=ego_trajectory)
        return self._integrate(torch.stack([dist[i].mean for i in self.ado_ids]))

    def predict_w_controls(self, ego_controls: torch.Tensor) -> typing.Optional[torch.Tensor]:
        return self.predict_w_trajectory(self.ego.unroll_trajectory(controls=ego_controls, dt=self.dt))

    def predict_wo_ego(self, t_horizon: int) -> typing.Optional[torch.Tensor]:
        if self.num_ados == 0:
            return None
        dist = self.compute_distributions_wo_ego(t_horizon=t_horizon)
        return self._integrate(torch.stack([dist[i].mean for i in self.ado_ids]))

    def sample_w_trajectory(self, ego_trajectory: torch.Tensor, num_samples: int = 1) -> typing.Optional[torch.Tensor]:
        """Sampled ado paths [N, num_samples, T+1, 1, 2] (graph_based.py:171-196)."""
        if self.num_ados == 0:
            return None
        dist = self.compute_distributions(ego_trajectory=ego_trajectory)
        return self._integrate(torch.stack([dist[i].sample((num_samples,)) for i in self.ado_ids]))

    def sample_w_controls(self, ego_controls: torch.Tensor, num_samples: int = 1) -> typing.Optional[torch.Tensor]:
        return self.sample_w_trajectory(self.ego.unroll_trajectory(controls=ego_controls, dt=self.dt), num_samples)

    def sample_wo_ego(self, t_horizon: int, num_samples: int = 1) -> typing.Optional[torch.Tensor]:
        if self.num_ados == 0:
            return None
        dist = self.compute_distributions_wo_ego(t_horizon=t_horizon)
        return self._integrate(torch.stack([dist[i].sample((num_samples,)) for i in self.ado_ids]))

    # ---- copies ------------------------------------------------------------------------------------------------
    def detach(self):
        if self._ego is not None:
            self._ego.detach()
        for ado in self._ados:
            ado.detach()

    def copy(self, env_type=None) -> "GraphBasedEnvironment":
        """Re-create the environment (optionally as another type) with the same agents (graph_based.py:447-481)."""
        env_type = env_type if env_type is not None else self.__class__
        ego_kwargs = {}
        if self.ego is not None:
            ego_kwargs = dict(ego_position=self.ego.position.clone(), ego_velocity=self.ego.velocity.clone(),
                              ego_history=self.ego.history.clone(), ego_type=self.ego.__class__)
        env_copy = env_type(**ego_kwargs, dt=self.dt, time=self.time, **self._env_params)
        for ado in self.ados:
            env_copy.add_ado(position=ado.position.clone(), velocity=ado.velocity.clone(), history=ado.history.clone(),
                             time=self.time, color=ado.color, identifier=ado.id, **ado.params)
        return env_copy

    def same_initial_conditions(self, other: "GraphBasedEnvironment") -> bool:
        ok = self.dt == other.dt and self.num_ados == other.num_ados
        ok = ok and self.ego.__eq__(other.ego, check_class=True)
        for ado in self.ados:
            ok = ok and ado.__eq__(other.ados[other.index_ado_id(ado.id)], check_class=False, check_history=False)
        return bool(ok)

    def sanity_check(self, check_ego: bool = False) -> bool:
        assert self.num_ados == len(self.ado_ids)
        assert all(self.ados[m].id == self.ado_ids[m] for m in range(self.num_ados))
        if check_ego:
            assert self.ego is not None and self.ego.is_robot
        return True

    # ---- batched scene description (the kernels' input) -----------------------------------------------------------
    def scene_batch(self, num_particles: int = constants.ENV_NUM_PARTICLES, t_horizon: int = None,
                    with_noise: bool = True, **param_dicts) -> SceneBatch:
        """This scene as a `SceneBatch` of one scene on the host, with particle parameters freshly drawn from torch's
        global CPU generator in the reference's order (one `create_particles`, then one `torch.rand` per step)."""
        params = self.draw_particle_parameters(num_particles, (t_horizon - 1) if (with_noise and t_horizon) else 0,
                                               **param_dicts)
        return dataclasses.replace(self.scene_static(), v0=params.get("v0"), sigma=params.get("sigma"), weight=params.get("weight"),
                                   noise=params.get("noise"))

    def scene_static(self) -> SceneBatch:
        """The part of the scene that does not change between the callbacks of one solve (states, goals): no random numbers."""
        _, ado_states = self.states()
        goals = torch.stack([torch.as_tensor(a.params.get("goal", a.position)).float() for a in self.ados]) \
            if self.num_ados > 0 else torch.zeros((0, 2))
        ego = self.ego
        return SceneBatch(
            env_kind=self.env_kind, ego_kind=ego.ego_kind if ego is not None else None, dt=float(self.dt),
            ego0=ego.state[None].clone() if ego is not None else None, ego_goal=None,
            ado0=ado_states[None, :, 0:4].contiguous(), ado_goal=goals[None].contiguous(),
            axes_diag=float(((self.x_axis[1] - self.x_axis[0]) ** 2 + (self.y_axis[1] - self.y_axis[0]) ** 2) ** 0.5))

    def draw_particle_parameters(self, num_particles: int, num_noise_steps: int, **param_dicts
                                 ) -> typing.Dict[str, torch.Tensor]:
        return {}

    # ---- properties --------------------------------------------------------------------------------------------
    @property
    def ego(self):
        return self._ego

    @property
    def ados(self):
        return self._ados

    @property
    def ado_ids(self) -> typing.List[str]:
        return self._ado_ids

    @property
    def num_ados(self) -> int:
        return len(self._ados)

    @property
    def num_modes(self) -> int:
        return 1

    @property
    def dt(self) -> float:
        return self._dt

    @property
    def time(self) -> float:
        return round(self._time, 2)

    @property
    def x_axis(self):
        return self._env_params["x_axis"]

    @property
    def y_axis(self):
        return self._env_params["y_axis"]

    @property
    def axes(self):
        return self.x_axis, self.y_axis

    @property
    def name(self) -> str:
        raise NotImplementedError

    @property
    def log_name(self) -> str:
        return self.name

    @property
    def is_multi_modal(self) -> bool:
        return False

    @property
    def is_deterministic(self) -> bool:
        return True

    @property
    def is_differentiable_wrt_ego(self) -> bool:
        return True


class ParticleEnvironment(GraphBasedEnvironment):
    """Particle rollout collapsed to one Gaussian per ado and step (environment/base/particle.py:14-205)."""
    # parameter name -> (mean, "variance" used as std) in the reference's dict order (A.4)
    param_defaults: typing.Dict[str, typing.Tuple[float, float]] = {}

    def draw_particle_parameters(self, num_particles: int, num_noise_steps: int, **param_dicts):
        return draw_particle_parameters(self.ado_ids, self.param_defaults, num_particles, num_noise_steps, **param_dicts)

    def _compute_distributions(self, ego_trajectory, num_particles: int = constants.ENV_NUM_PARTICLES,
                               vel_dist: bool = True, t_horizon: int = None, **kwargs):
        if num_particles < 2:
            raise ValueError("num_particles must be >= 2 (the particle variance of one particle is undefined)")
        engine = runtime.default_engine()
        out_device = ego_trajectory.device if ego_trajectory is not None else torch.device("cpu")
        t_horizon = ego_trajectory.shape[0] - 1 if ego_trajectory is not None else t_horizon
        dicts = {k: v for k, v in kwargs.items() if k in ("v0_dict", "sigma_dict") and v is not None}
        scene = self.scene_batch(num_particles, t_horizon, **dicts).to(engine.device)
        if ego_trajectory is None:
            res = engine.rollout_wo_ego(scene, t_horizon)
            mus, sigmas = res["mus"][0], res["sigmas"][0]
        else:
            traj = runtime.to_device(ego_trajectory[:, 0:4], engine)[None, None]
            mus, sigmas = ParticleRollout.apply(traj, engine, scene)
            mus, sigmas = mus[0, 0], sigmas[0, 0]
        return self._distributions_from(mus, sigmas, vel_dist, out_device)


class SocialForcesEnvironment(ParticleEnvironment):
    env_kind = "sf"
    param_defaults = {"v0": constants.SOCIAL_FORCES_DEFAULT_V0, "sigma": constants.SOCIAL_FORCES_DEFAULT_SIGMA}

    def add_ado(self, position: torch.Tensor, velocity: torch.Tensor = torch.zeros(2), history: torch.Tensor = None,
                goal: torch.Tensor = torch.zeros(2), **ado_kwargs):
        """Pedestrian with a goal position, default (0, 0) (social_forces.py:171-196)."""
        goal = torch.as_tensor(goal).detach().float().cpu()
        return super().add_ado(position, velocity=velocity, history=history, goal=goal, **ado_kwargs)

    @property
    def name(self) -> str:
        return "social_forces"


class PotentialFieldEnvironment(ParticleEnvironment):
    env_kind = "pf"
    param_defaults = {"v0": constants.POTENTIAL_FIELD_V0_DEFAULT}

    @property
    def name(self) -> str:
        return "potential_field"


class KalmanEnvironment(GraphBasedEnvironment):
    """Constant-velocity prediction with linearly growing variance; not conditioned on the ego (kalman.py:11-124)."""
    env_kind = "kalman"

    def _compute_distributions(self, ego_trajectory, vel_dist: bool = True, t_horizon: int = None,
                               noise_additive: float = constants.KALMAN_ADDITIVE_NOISE, **kwargs):
        engine = runtime.default_engine()
        out_device = ego_trajectory.device if ego_trajectory is not None else torch.device("cpu")
        t_horizon = ego_trajectory.shape[0] - 1 if ego_trajectory is not None else t_horizon
        scene = self.scene_batch(1, t_horizon, with_noise=False).to(engine.device)
        res = engine.rollout_wo_ego(scene, t_horizon)                # the whole body is under no_grad (kalman.py:63)
        mus, sigmas = res["mus"][0], res["sigmas"][0]
        if noise_additive != constants.KALMAN_ADDITIVE_NOISE:
            sigmas = torch.where(sigmas == constants.KALMAN_ADDITIVE_NOISE, torch.full_like(sigmas, noise_additive), sigmas)
        if not vel_dist:       # positional variance 1e-6 + Q t: F = diag(1,1,0,0) keeps the position block (kalman.py:79-92)
            steps = torch.arange(t_horizon, dtype=torch.float32, device=sigmas.device).view(1, t_horizon, 1)
            sigmas = (constants.ENV_VAR_INITIAL + noise_additive * steps).expand(self.num_ados, t_horizon, 2)
        return self._distributions_from(mus, sigmas, vel_dist, out_device)

    @property
    def name(self) -> str:
        return "kalman"

    @property
    def is_differentiable_wrt_ego(self) -> bool:
        return False
