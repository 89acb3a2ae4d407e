"""Route B of INTEGRATION.md as code: the REFERENCE's own classes bound to the C ABI.

`bind_reference(mantrap)` takes the imported, unmodified reference package (simon-schaefer/mantrap) and returns
subclasses of ITS classes that override exactly the hooks the reference consults before it falls back to its
per-particle Python loop and torch.autograd:

  * `ParticleEnvironment._compute_distributions`   (mantrap/environment/base/particle.py:98-183,
    contract `mantrap/environment/base/graph_based.py:383-397`)  -> one `mtb_evaluate` launch behind the
    `autograd.ParticleRollout` node (backward = the fused kernel seeded with the incoming adjoints);
  * `OptimizationModule.compute_gradient_analytically` (mantrap/modules/base/optimization_module.py:154-156, :181-200)
    of the interaction objectives -> rollout + objective + hand-derived adjoint in ONE launch;
  * `OptimizationModule.compute_jacobian_analytically` (optimization_module.py:276-278, :334-353) of the ellipsoid
    constraint -> `mtb_constraint_jacobian`.

Everything else -- `TrajOptSolver.objective / constraints` (solver/base/trajopt.py:368-434), `IPOPTSolver.gradient /
jacobian` (solver/ipopt.py:118-154), the search solvers, logging, normalisation and weighting
(`_return_gradient`, optimization_module.py:565-573) -- stays the reference's own code and runs unmodified on top.

The particle parameters are drawn from torch's global CPU generator with the reference's calls in the reference's order
(`create_particles` then one `torch.rand` per step), so `torch.manual_seed(k)` pins the same
numbers in both implementations; no `IntegratorDTAgent` particle objects are built.

There is no CPU path here either: the engine is `runtime.default_engine()` (CUDA) unless the caller passes one.
"""
import types
import typing

import numpy as np
import torch

from . import constants, runtime
from .autograd import ParticleRollout
from .engine import Engine, ModuleSpec, SceneBatch
from .environment import draw_particle_parameters


def _ego_kind(env) -> typing.Optional[str]:
    if env.ego is None:
        return None
    return "double" if "Double" in type(env.ego).__name__ else "single"


def bind_reference(mantrap, engine: Engine = None) -> types.SimpleNamespace:
    """Subclasses of the reference's environments and modules running on `engine` (default: the process-wide CUDA engine).
    Returns a namespace with SocialForcesEnvironment, PotentialFieldEnvironment, InteractionAccelerationModule,
    InteractionVelocityModule, InteractionPositionModule, InteractionProbabilityModule, InteractionEllipsoidModule."""

    def eng() -> Engine:
        return engine if engine is not None else runtime.default_engine()

    class _DeviceParticleMixin:
        """`_compute_distributions` on the device; shared by the two particle environments."""
        _mtb_env_kind = None
        _mtb_param_defaults = None

        def _mtb_scene(self, num_particles: int, t_horizon: int, **kwargs) -> SceneBatch:
            """One-scene batch with a FRESH particle draw (the reference's RNG consumption of one call)."""
            _, ado_states = self.states()
            n = self.num_ados
            goals = torch.stack([torch.as_tensor(a.params.get("goal", a.position)).float() for a in self.ados]) \
                if n > 0 else torch.zeros((0, 2))
            dicts = {k: v for k, v in kwargs.items() if k in ("v0_dict", "sigma_dict") and v is not None}
            params = draw_particle_parameters(self.ado_ids, self._mtb_param_defaults, num_particles, t_horizon - 1, **dicts)
            ego = self.ego
            return SceneBatch(env_kind=self._mtb_env_kind, ego_kind=_ego_kind(self), dt=float(self.dt),
                              ego0=ego.state[None, 0:4].clone().float() if ego is not None else None, ego_goal=None,
                              ado0=ado_states[None, :, 0:4].float().contiguous(), ado_goal=goals[None].contiguous(),
                              v0=params.get("v0"), sigma=params.get("sigma"), weight=params.get("weight"),
                              noise=params.get("noise")).to(eng().device)

        def _compute_distributions(self, ego_trajectory, num_particles: int = constants.ENV_NUM_PARTICLES,
                                   vel_dist: bool = True, **kwargs):
            if num_particles < 2:      # the reference fails in torch.var -> Normal(scale = NaN) -> ValueError
                raise ValueError("num_particles must be >= 2 (the particle variance of one particle is undefined)")
            wo_ego = all(x is None for x in ego_trajectory)
            t_horizon = len(ego_trajectory) - 1
            scene = self._mtb_scene(num_particles, t_horizon, **kwargs)
            engine_ = eng()
            if wo_ego:
                res = engine_.rollout_wo_ego(scene, t_horizon)
                mus, sigmas = res["mus"][0], res["sigmas"][0]
                self._mtb_mus_wo = res["mus"].contiguous()                    # [1,N,T,4]: the modules' ego-free baseline
            else:
                traj = runtime.to_device(ego_trajectory[:, 0:4], engine_)[None, None]
                mus, sigmas = ParticleRollout.apply(traj, engine_, scene)
                mus, sigmas = mus[0, 0], sigmas[0, 0]
            mus, sigmas = mus.cpu(), sigmas.cpu()                              # the reference computes on CPU tensors
            means = mus[:, :, None, 2:4] if vel_dist else mus[:, :, None, 0:2]
            return {ado_id: torch.distributions.Normal(loc=means[m], scale=sigmas[m, :, None, :], validate_args=False)
                    for m, ado_id in enumerate(self.ado_ids)}

    class SocialForcesEnvironment(_DeviceParticleMixin, mantrap.environment.SocialForcesEnvironment):
        _mtb_env_kind = "sf"
        _mtb_param_defaults = {"v0": mantrap.constants.SOCIAL_FORCES_DEFAULT_V0,
                               "sigma": mantrap.constants.SOCIAL_FORCES_DEFAULT_SIGMA}

    class PotentialFieldEnvironment(_DeviceParticleMixin, mantrap.environment.PotentialFieldEnvironment):
        _mtb_env_kind = "pf"
        _mtb_param_defaults = {"v0": mantrap.constants.POTENTIAL_FIELD_V0_DEFAULT}

    class _FusedGradientMixin:
        """`compute_gradient_analytically` of an interaction objective: raw gradient [2T] from one fused launch. The
        reference weights and normalises it afterwards (`_return_gradient`)."""
        _mtb_key = None
        _mtb_max = None

        def __init__(self, env, t_horizon: int, **kwargs):
            super().__init__(env=env, t_horizon=t_horizon, **kwargs)          # runs the ego-free rollout (acc_interact.py:33-34)
            self._mtb_mus_wo = getattr(env, "_mtb_mus_wo", None)
            self._mtb_particles = kwargs.get("num_particles", None)

        def compute_gradient_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag):
            env = self._env
            if not isinstance(env, _DeviceParticleMixin) or self._mtb_mus_wo is None:
                return None                                                   # not a bound environment: reference path
            if len(ado_ids) == 0 or env.num_ados == 0 or not env.is_differentiable_wrt_ego:
                return np.zeros(grad_wrt.numel())                            # optimization_module.py:172-173
            engine_ = eng()
            t_horizon = ego_trajectory.shape[0] - 1
            num_particles = type(env)._compute_distributions.__defaults__[0] if self._mtb_particles is None else self._mtb_particles
            scene = env._mtb_scene(num_particles, t_horizon)
            mask = None
            if len(ado_ids) != env.num_ados:
                mask = torch.zeros((1, env.num_ados), dtype=torch.uint8)
                for ado_id in ado_ids:
                    mask[0, env.index_ado_id(ado_id)] = 1
                mask = mask.to(engine_.device)
            spec = ModuleSpec(weights={self._mtb_key: self._mtb_max}, mus_wo=self._mtb_mus_wo, ado_mask=mask)   # weight / max = 1: raw
            traj = runtime.to_device(ego_trajectory.detach()[:, 0:4], engine_)[None, None]
            res = engine_.evaluate(scene, None, spec, want=("grad",), ego_traj=traj)
            return res["grad"][0, 0].flatten().cpu().numpy().astype(np.float64)

    class InteractionAccelerationModule(_FusedGradientMixin, mantrap.modules.baselines.InteractionAccelerationModule):
        _mtb_key, _mtb_max = "acc", mantrap.constants.OBJECTIVE_ACC_INTERACT_MAX

    class InteractionVelocityModule(_FusedGradientMixin, mantrap.modules.baselines.InteractionVelocityModule):
        _mtb_key, _mtb_max = "vel", mantrap.constants.OBJECTIVE_VEL_INTERACT_MAX

    class InteractionPositionModule(_FusedGradientMixin, mantrap.modules.baselines.InteractionPositionModule):
        _mtb_key, _mtb_max = "pos", mantrap.constants.OBJECTIVE_POS_INTERACT_MAX

    class InteractionProbabilityModule(_FusedGradientMixin, mantrap.modules.InteractionProbabilityModule):
        _mtb_key, _mtb_max = "prob", mantrap.constants.OBJECTIVE_PROB_INTERACT_MAX

    class InteractionEllipsoidModule(mantrap.modules.baselines.InteractionEllipsoidModule):
        """`compute_jacobian_analytically`: rows [N*T] x 2T from `mtb_constraint_jacobian` (means and scales detached,
        ellipsoid_interact.py:74-75: the Jacobian goes through the ego rollout only)."""

        def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag):
            env = self._env
            if not isinstance(env, _DeviceParticleMixin):
                return None
            if env.num_ados == 0:
                return np.array([])
            engine_ = eng()
            t_horizon = ego_trajectory.shape[0] - 1
            scene = env._mtb_scene(type(env)._compute_distributions.__defaults__[0], t_horizon)
            traj = runtime.to_device(ego_trajectory.detach()[:, 0:4], engine_)[None, None]
            jac = engine_.constraint_jacobian(scene, None, ModuleSpec(ellipsoid=True), ego_traj=traj)
            return jac[0, 0].flatten().cpu().numpy().astype(np.float64)

    return types.SimpleNamespace(
        SocialForcesEnvironment=SocialForcesEnvironment, PotentialFieldEnvironment=PotentialFieldEnvironment,
        InteractionAccelerationModule=InteractionAccelerationModule, InteractionVelocityModule=InteractionVelocityModule,
        InteractionPositionModule=InteractionPositionModule, InteractionProbabilityModule=InteractionProbabilityModule,
        InteractionEllipsoidModule=InteractionEllipsoidModule)
