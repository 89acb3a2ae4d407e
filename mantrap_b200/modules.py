"""Objective / constraint modules of the hot path with the reference's interface (`mantrap/modules`).

Every module answers `objective / gradient / constraint / jacobian (ego_trajectory, [grad_wrt,] ado_ids, tag)` like
`mantrap/modules/base/optimization_module.py:11-702`. The built-in modules answer from the fused CUDA kernels:
ONE `mtb_evaluate` launch (ego trajectory given, `mtb_problem.ego_traj`) yields the weighted + normalised objective
and its gradient w.r.t. the controls; ONE `mtb_constraint_jacobian` launch yields a dense Jacobian. A user-defined
module that only implements `_objective_core` / `_constraint_core` with torch ops on `env.compute_distributions`
falls back to torch.autograd through the `EgoUnroll` / `ParticleRollout` nodes, exactly as in the reference.

Mirrors: GoalNormModule modules/goal_norm.py:12-136; InteractionAcceleration/Velocity/PositionModule
modules/baselines/{acc,vel,pos}_interact.py; ControlLimitModule modules/control_limits.py:12-125; SpeedLimitModule
modules/speed_limits.py:12-125; MinDistanceModule: absent at the reference commit (constants.py:107 only),
semantics from BASELINE.json's north star.
"""
import typing

import numpy as np
import torch

from . import constants, runtime
from .engine import OBJECTIVE_KEYS, ModuleSpec, SceneBatch


def _np(x: torch.Tensor) -> np.ndarray:
    return x.detach().cpu().numpy()


class OptimizationModule:
    """Base contract (optimization_module.py:11-702). `fused_key` names the kernel objective of a built-in module."""
    fused_key: typing.Optional[str] = None
    needs_rollout = False            # consumes one particle draw per call (interaction modules, min distance)

    def __init__(self, t_horizon: int, weight: float = 0.0, env=None, has_slack: bool = False, slack_weight: float = 0.0,
                 num_particles: int = constants.ENV_NUM_PARTICLES, **unused):
        assert t_horizon is None or t_horizon >= 1
        assert weight is None or weight >= 0.0
        assert slack_weight >= 0.0
        self._weight, self._t_horizon, self._env = weight, t_horizon, env
        if env is not None:
            assert env.ego is not None
        self._num_particles = num_particles
        self._has_slack, self._slack_weight, self._slack = has_slack, slack_weight, {}
        self._constraint_current, self._obj_current, self._grad_current, self._jacobian_current = {}, {}, {}, {}

    def reset_env(self, env):
        if self._env is not None:
            self._env = env

    # ---- fused kernel plumbing ---------------------------------------------------------------------------------
    def _scene(self) -> SceneBatch:
        """The module's environment as a one-scene batch on the engine device. Modules that need the environment rollout
        draw fresh particle parameters (the reference re-samples on every call); ego-only modules consume no random
        numbers and run against a single inert Kalman ado."""
        engine = runtime.default_engine()
        env = self._env
        if self.needs_rollout:
            num_p = self._num_particles if env.env_kind != "kalman" else 1
            scene = env.scene_batch(num_p, self.t_horizon, with_noise=env.env_kind != "kalman")
        else:
            scene = SceneBatch(env_kind="kalman", ego_kind=env.ego.ego_kind, dt=float(env.dt), ego0=env.ego.state[None].clone(),
                               ego_goal=None, ado0=torch.zeros((1, 1, 4)))
            scene.axes_diag = float(((env.x_axis[1] - env.x_axis[0]) ** 2 + (env.y_axis[1] - env.y_axis[0]) ** 2) ** 0.5)
        return scene.to(engine.device)

    def _ado_mask(self, ado_ids: typing.List[str]) -> typing.Optional[torch.Tensor]:
        if ado_ids is None or list(ado_ids) == list(self._env.ado_ids):
            return None
        mask = torch.tensor([[1 if i in ado_ids else 0 for i in self._env.ado_ids]], dtype=torch.uint8)
        return mask.to(runtime.default_engine().device)

    def _spec(self, ado_ids) -> ModuleSpec:
        """Which kernel objectives / constraint rows this module switches on."""
        raise NotImplementedError

    def _launch(self, ego_trajectory: torch.Tensor, ado_ids, want: typing.Sequence[str]):
        engine = runtime.default_engine()
        traj = runtime.to_device(ego_trajectory.detach()[:, 0:4], engine)[None, None]
        scene = self._scene()
        if scene.ego_goal is None and getattr(self, "_goal", None) is not None:
            scene.ego_goal = runtime.to_device(self._goal.view(1, 2), engine)
        return engine.evaluate(scene, None, self._spec(ado_ids), want=want, ego_traj=traj)

    # ---- objective ---------------------------------------------------------------------------------------------
    def objective(self, ego_trajectory: torch.Tensor, ado_ids: typing.List[str], tag: str) -> float:
        if self.fused_key is not None:
            value = 0.0
            if self._objective_defined(ado_ids):
                res = self._launch(ego_trajectory, ado_ids, want=("obj",))
                value = float(res["obj"][0, 0, OBJECTIVE_KEYS.index(self.fused_key)])
            self._obj_current[tag] = value                  # already weight * normalize(raw) (kernel epilogue)
            return value
        objective = self.compute_objective(ego_trajectory, ado_ids=ado_ids, tag=tag)
        obj_value = 0.0 if objective is None else float(objective.item())
        return self._return_objective(obj_value, tag=tag)

    def _objective_defined(self, ado_ids) -> bool:
        return True

    def compute_objective(self, ego_trajectory, ado_ids, tag):
        obj_value = self._objective_core(ego_trajectory, ado_ids=ado_ids, tag=tag)
        if self._has_slack:
            obj_value = torch.zeros(1) if obj_value is None else obj_value
            _ = self.compute_constraint(ego_trajectory, ado_ids=ado_ids, tag=tag)
            obj_value = obj_value + self._slack_weight * torch.clamp(self._slack[tag], min=0.0).sum()
        return obj_value

    def _objective_core(self, ego_trajectory, ado_ids, tag) -> typing.Optional[torch.Tensor]:
        return None

    def gradient(self, ego_trajectory: torch.Tensor, grad_wrt: torch.Tensor, ado_ids: typing.List[str], tag: str) -> np.ndarray:
        """Gradient of the weighted, normalised objective w.r.t. `grad_wrt` (= the controls the trajectory was unrolled
        from, as everywhere in the reference; optimization_module.py:138-179)."""
        if self.fused_key is not None:
            if not self._objective_defined(ado_ids) or not self._env.is_differentiable_wrt_ego and self.needs_rollout:
                gradient = np.zeros(grad_wrt.numel(), dtype=np.float32)
            else:
                res = self._launch(ego_trajectory, ado_ids, want=("grad",))
                gradient = _np(res["grad"][0, 0]).flatten()
            self._grad_current[tag] = gradient
            return gradient
        gradient = self.compute_gradient_analytically(ego_trajectory, grad_wrt, ado_ids=ado_ids, tag=tag)
        if gradient is None:
            assert grad_wrt.requires_grad and ego_trajectory.requires_grad
            objective = self._objective_core(ego_trajectory, ado_ids=ado_ids, tag=tag)
            if objective is None or not self.gradient_condition():
                gradient = np.zeros(grad_wrt.numel())
            else:
                gradient = self.compute_gradient_auto_grad(objective, grad_wrt=grad_wrt)
        return self._return_gradient(gradient, tag=tag)

    def compute_gradient_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> typing.Optional[np.ndarray]:
        return None

    # ---- constraint --------------------------------------------------------------------------------------------
    def constraint(self, ego_trajectory: torch.Tensor, ado_ids: typing.List[str], tag: str) -> np.ndarray:
        constraints = self.compute_constraint(ego_trajectory, ado_ids=ado_ids, tag=tag)
        constraints = np.array([]) if constraints is None else _np(constraints)
        return self._return_constraint(constraints, tag=tag)

    def compute_constraint(self, ego_trajectory, ado_ids, tag) -> typing.Optional[torch.Tensor]:
        constraints = self._constraint_core(ego_trajectory, ado_ids=ado_ids, tag=tag)
        if self._has_slack and constraints is not None:
            self._slack[tag] = -constraints
        return constraints

    def _constraint_core(self, ego_trajectory, ado_ids, tag) -> typing.Optional[torch.Tensor]:
        return None

    def jacobian(self, ego_trajectory: torch.Tensor, grad_wrt: torch.Tensor, ado_ids: typing.List[str], tag: str) -> np.ndarray:
        """Flattened [R * grad_wrt.numel()] Jacobian, NOT normalised (reference quirk, optimization_module.py:258-307)."""
        jacobian = self.compute_jacobian_analytically(ego_trajectory, grad_wrt, ado_ids=ado_ids, tag=tag)
        if jacobian is None:
            assert ego_trajectory.requires_grad
            constraints = self._constraint_core(ego_trajectory, ado_ids=ado_ids, tag=tag)
            if constraints is None:
                jacobian = np.array([])
            elif not self.gradient_condition():
                jacobian = np.zeros(int(grad_wrt.numel()) * int(constraints.numel()))
            else:
                jacobian = self.compute_gradient_auto_grad(constraints, grad_wrt=grad_wrt)
        return jacobian

    def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> typing.Optional[np.ndarray]:
        return None

    def jacobian_structure(self, ado_ids: typing.List[str], tag: str) -> typing.Optional[np.ndarray]:
        if self.num_constraints(ado_ids=ado_ids) == 0:
            return None
        controls = torch.rand((self.t_horizon, 2))
        ego_trajectory = self._env.ego.unroll_trajectory(controls, dt=self._env.dt)
        jacobian = self.compute_jacobian_analytically(ego_trajectory, grad_wrt=controls, ado_ids=ado_ids, tag=tag)
        return np.nonzero(jacobian)[0] if jacobian is not None else None

    def compute_gradient_auto_grad(self, x: torch.Tensor, grad_wrt: torch.Tensor) -> np.ndarray:
        """Row-by-row reverse mode (optimization_module.py:358-447), for user-defined modules."""
        grad_size, x_size = int(grad_wrt.numel()), int(x.numel())
        assert grad_wrt.requires_grad
        if not x.requires_grad:
            return np.zeros(grad_size)
        if x_size == 1:
            gradient = torch.autograd.grad(x, grad_wrt, retain_graph=True)[0]
        else:
            rows = [torch.autograd.grad(e, grad_wrt, retain_graph=True)[0].flatten() for e in x.flatten()]
            gradient = torch.cat(rows)
        return np.nan_to_num(_np(gradient.flatten()), copy=False)

    # ---- boundaries / violation --------------------------------------------------------------------------------
    def constraint_boundaries(self, ado_ids: typing.List[str]):
        lower, upper = self._constraint_limits()
        lower = self.normalize(lower) if lower is not None else None
        upper = self.normalize(upper) if upper is not None else None
        n = self._num_constraints(ado_ids=ado_ids)
        return ((lower * np.ones(n)).tolist() if lower is not None else [None] * n,
                (upper * np.ones(n)).tolist() if upper is not None else [None] * n)

    def _constraint_limits(self):
        raise NotImplementedError

    def num_constraints(self, ado_ids: typing.List[str]) -> int:
        return self._num_constraints(ado_ids=ado_ids)

    def _num_constraints(self, ado_ids: typing.List[str]) -> int:
        return 0

    def compute_violation(self, ego_trajectory: torch.Tensor, ado_ids: typing.List[str], tag: str) -> float:
        constraint = self.compute_constraint(ego_trajectory, ado_ids=ado_ids, tag=tag)
        if constraint is None:
            return self._violation(constraints=None, ado_ids=ado_ids)
        return self._violation(constraints=self.normalize(_np(constraint)), ado_ids=ado_ids)

    def compute_violation_internal(self, tag: str) -> float:
        return self._violation(constraints=self._constraint_current[tag])

    def _violation(self, constraints: typing.Optional[np.ndarray], ado_ids=None) -> float:
        """Summed bound violation of normalised rows, zero below 1e-5 (optimization_module.py:512-533)."""
        if constraints is None or constraints.size == 0:
            return 0.0
        lower, upper = self.constraint_boundaries(ado_ids=ado_ids if ado_ids is not None else self._env.ado_ids)
        lo = np.array([-np.inf if b is None else b for b in lower[:constraints.size]])
        hi = np.array([np.inf if b is None else b for b in upper[:constraints.size]])
        violation = float((np.maximum(lo - constraints, 0.0) + np.maximum(constraints - hi, 0.0)).sum())
        return 0.0 if abs(violation) < constants.CONSTRAINT_VIOLATION_PRECISION else violation

    # ---- conventions -------------------------------------------------------------------------------------------
    def normalize(self, x):
        raise NotImplementedError

    def gradient_condition(self) -> bool:
        return True

    def _return_constraint(self, value: np.ndarray, tag: str) -> np.ndarray:
        self._constraint_current[tag] = self.normalize(value)
        return self._constraint_current[tag]

    def _return_objective(self, value: float, tag: str) -> float:
        self._obj_current[tag] = self.weight * float(self.normalize(value))
        return self._obj_current[tag]

    def _return_gradient(self, gradient: np.ndarray, tag: str) -> np.ndarray:
        self._grad_current[tag] = self.weight * self.normalize(gradient)
        return self._grad_current[tag]

    def constraint_current(self, tag: str) -> np.ndarray:
        return self._constraint_current[tag]

    def inf_current(self, tag: str) -> float:
        return self.compute_violation_internal(tag=tag)

    def obj_current(self, tag: str) -> float:
        return self._obj_current[tag]

    def grad_current(self, tag: str) -> float:
        return float(np.linalg.norm(self._grad_current[tag]))

    def slack_variables(self, tag: str) -> torch.Tensor:
        return self._slack[tag]

    @property
    def weight(self) -> float:
        return self._weight

    @property
    def env(self):
        return self._env

    @property
    def t_horizon(self) -> int:
        return self._t_horizon

    @property
    def name(self) -> str:
        raise NotImplementedError


class PureObjectiveModule(OptimizationModule):
    """No constraint rows (modules/base/pure_objective.py)."""

    def __init__(self, t_horizon: int, weight: float = 1.0, env=None, **kwargs):
        super().__init__(t_horizon, weight=weight, env=env, has_slack=False, slack_weight=0.0, **kwargs)

    def _constraint_limits(self):
        return None, None


class PureConstraintModule(OptimizationModule):
    """No objective (modules/base/pure_constraint.py)."""

    def __init__(self, t_horizon: int, env=None, **kwargs):
        super().__init__(t_horizon, weight=0.0, env=env, has_slack=False, slack_weight=0.0, **kwargs)

    def normalize(self, x):
        return x


###########################################################################
# Objectives ##############################################################
###########################################################################
class GoalNormModule(PureObjectiveModule):
    """mean_{t, xy} (p_t - goal)^2 over all T+1 rows, / environment diagonal, weight 0.5 (goal_norm.py:37-118)."""
    fused_key = "goal"

    def __init__(self, goal: torch.Tensor, env, optimize_speed: bool = False, weight: float = 0.5, t_horizon: int = None,
                 **unused):
        super().__init__(env=env, t_horizon=t_horizon, weight=weight)
        self._goal = torch.as_tensor(goal).detach().float().cpu()
        assert self._goal.shape == (2,)
        # The reference's speed term `speeds.dot(torch.exp(-goal_distances * 0.5))` (goal_norm.py:63-65) multiplies a [T+1] with a
        # [T+1, 2] tensor: torch.dot raises "1D tensors expected" on the first objective call (probe-verified on the unmodified
        # reference), and its analytic gradient ignores the term (:99-105). Same behaviour here: constructible, raises when evaluated.
        self._optimize_speed = bool(optimize_speed)

    def _spec(self, ado_ids) -> ModuleSpec:
        if self._optimize_speed:
            raise RuntimeError("GoalNormModule(optimize_speed=True): 1D tensors expected, but got 1D and 2D tensors "
                               "(the reference's speed term, goal_norm.py:63-65, fails in torch.dot in exactly this way)")
        return ModuleSpec(weights={"goal": self.weight})

    def normalize(self, x):
        (x0, x1), (y0, y1) = self._env.axes
        return x / np.sqrt((x1 - x0) ** 2 + (y1 - y0) ** 2)

    @property
    def name(self) -> str:
        return "goal_norm"


class _InteractionModule(PureObjectiveModule):
    """Shared by the acceleration / velocity / position interaction costs (modules/baselines/acc_interact.py:13-106):
    distance between the ados' predicted means with and without the ego, averaged over time and `ado_ids`, clamped."""
    needs_rollout = True
    max_value: float = 1.0

    def __init__(self, env, t_horizon: int, weight: float = 1.0, **unused):
        super().__init__(env=env, t_horizon=t_horizon, weight=weight, **{k: v for k, v in unused.items() if k == "num_particles"})
        self._mus_wo = None
        if env is not None and env.num_ados > 0:      # baseline rollout without ego: one draw at construction (:29-34)
            engine = runtime.default_engine()
            num_p = self._num_particles if env.env_kind != "kalman" else 1
            scene = env.scene_batch(num_p, t_horizon, with_noise=env.env_kind != "kalman").to(engine.device)
            self._mus_wo = engine.rollout_wo_ego(scene, t_horizon)["mus"].contiguous()      # [1,N,T,4]

    def _objective_defined(self, ado_ids) -> bool:
        return self._mus_wo is not None and (ado_ids is None or len(ado_ids) > 0)          # acc_interact.py:50-51

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(weights={self.fused_key: self.weight}, mus_wo=self._mus_wo, ado_mask=self._ado_mask(ado_ids))

    def normalize(self, x):
        return x / self.max_value


class InteractionAccelerationModule(_InteractionModule):
    fused_key, max_value = "acc", constants.OBJECTIVE_ACC_INTERACT_MAX

    @property
    def name(self) -> str:
        return "interact_acc"


class InteractionVelocityModule(_InteractionModule):
    fused_key, max_value = "vel", constants.OBJECTIVE_VEL_INTERACT_MAX

    @property
    def name(self) -> str:
        return "interaction_vel"


class InteractionPositionModule(_InteractionModule):
    """Quirk kept: with-ego POSITIONS against without-ego VELOCITIES (pos_interact.py:34-36)."""
    fused_key, max_value = "pos", constants.OBJECTIVE_POS_INTERACT_MAX

    @property
    def name(self) -> str:
        return "interaction_pos"


class InteractionProbabilityModule(_InteractionModule):
    """Minus the log-probability of the ego-free mean velocities in the distribution conditioned on the ego trajectory
    (modules/prob_interact.py:13-122): `dist.log_prob(dist_un_conditioned.mean)` summed over (t, modes, xy), averaged over
    `ado_ids`, clamped to +-OBJECTIVE_PROB_INTERACT_MAX and divided by it. The conditioned Normal uses the VARIANCE of the
    weighted particle velocities as its scale (particle.py:175,180); the adjoint runs through both the means and the
    variances inside the fused kernel. This is the interaction objective of `TrajOptSolver.module_defaults()`."""
    fused_key, max_value = "prob", constants.OBJECTIVE_PROB_INTERACT_MAX

    @property
    def name(self) -> str:
        return "interaction_prob"


###########################################################################
# Constraints #############################################################
###########################################################################
class _FusedConstraintModule(PureConstraintModule):
    spec_flag: str = None

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(**{self.spec_flag: True}, ado_mask=self._ado_mask(ado_ids))

    def _rows(self, ego_trajectory, ado_ids) -> torch.Tensor:
        """Normalised rows straight from the kernel epilogue."""
        return self._launch(ego_trajectory, ado_ids, want=("cons",))["cons"][0, 0]

    def constraint(self, ego_trajectory: torch.Tensor, ado_ids: typing.List[str], tag: str) -> np.ndarray:
        rows = _np(self._select(self._rows(ego_trajectory, ado_ids), ado_ids))
        self._constraint_current[tag] = rows
        return rows

    def compute_constraint(self, ego_trajectory, ado_ids, tag):        # un-normalised, as the reference's core
        return self._select(self._rows(ego_trajectory, ado_ids), ado_ids).cpu() * self._norm_factor()

    def _select(self, rows: torch.Tensor, ado_ids) -> torch.Tensor:
        return rows

    def _norm_factor(self) -> float:
        return 1.0

    def _dense_jacobian(self, ego_trajectory, ado_ids) -> torch.Tensor:
        engine = runtime.default_engine()
        traj = runtime.to_device(ego_trajectory.detach()[:, 0:4], engine)[None, None]
        return engine.constraint_jacobian(self._scene(), None, self._spec(ado_ids), ego_traj=traj)[0, 0]


class ControlLimitModule(_FusedConstraintModule):
    """|u_t| <= u_max per component (double integrator) / ||u_t|| (single), rows / u_max (control_limits.py:12-125)."""
    spec_flag = "control_limits"

    def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> np.ndarray:
        if self._env.ego.ego_kind == "double":
            return np.eye(2 * self.t_horizon).flatten()          # reference quirk: identity (control_limits.py:62-68)
        # Single-integrator ego: the reference returns T rows ||u_t|| but a 2T x 2T identity and 2T bounds, i.e. it is
        # inconsistent with itself; here rows, bounds and Jacobian (u / ||u||, [T, 2T]) agree.
        return _np(self._dense_jacobian(ego_trajectory, ado_ids)).astype(np.float64).flatten()

    def jacobian_structure(self, ado_ids, tag) -> np.ndarray:
        if self._env.ego.ego_kind == "double":
            return np.array([(2 * self.t_horizon + 1) * n for n in range(2 * self.t_horizon)])
        return np.array([2 * self.t_horizon * t + 2 * t + c for t in range(self.t_horizon) for c in range(2)])

    def normalize(self, x):
        return x / self._env.ego.control_limits()[1]

    def _norm_factor(self) -> float:
        return self._env.ego.control_limits()[1]

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(control_limits=True, u_max=self._env.ego.control_limits()[1])

    def _constraint_limits(self):
        return None, self._env.ego.control_limits()[1]

    def _num_constraints(self, ado_ids) -> int:
        return 2 * self.t_horizon if self._env.ego.ego_kind == "double" else self.t_horizon

    @property
    def name(self) -> str:
        return "control_limits"


class SpeedLimitModule(_FusedConstraintModule):
    """v_min <= v_t <= v_max on all T+1 rows, rows / v_max (speed_limits.py:12-125)."""
    spec_flag = "speed_limits"

    def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> np.ndarray:
        return _np(self._dense_jacobian(ego_trajectory, ado_ids)).astype(np.float64).flatten()   # velocity rows of dx_du

    def normalize(self, x):
        return x / self._env.ego.speed_limits[1]

    def _norm_factor(self) -> float:
        return self._env.ego.speed_limits[1]

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(speed_limits=True, v_max=self._env.ego.speed_limits[1])

    def _constraint_limits(self):
        return self._env.ego.speed_limits

    def _num_constraints(self, ado_ids) -> int:
        return 2 * (self.t_horizon + 1)

    @property
    def name(self) -> str:
        return "speed_limits"


class MinDistanceModule(_FusedConstraintModule):
    """||ego_p(t) - mu_p[k, t]|| >= 0.5 m for k in ado_ids, t = 0..T-1: K*T rows, not normalised. The Jacobian goes through
    the ego rollout AND the environment rollout (forward-mode tangent kernel)."""
    spec_flag = "min_distance"
    needs_rollout = True

    def _select(self, rows: torch.Tensor, ado_ids) -> torch.Tensor:
        if ado_ids is None or list(ado_ids) == list(self._env.ado_ids):
            return rows
        idx = [self._env.index_ado_id(i) for i in ado_ids]
        return rows.view(self._env.num_ados, self.t_horizon, *rows.shape[1:])[idx].reshape(len(idx) * self.t_horizon, *rows.shape[1:])

    def constraint(self, ego_trajectory, ado_ids, tag) -> np.ndarray:
        if ado_ids is not None and len(ado_ids) == 0:
            return np.array([])
        return super().constraint(ego_trajectory, ado_ids, tag)

    def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> np.ndarray:
        if ado_ids is not None and len(ado_ids) == 0:
            return np.array([])
        jac = self._dense_jacobian(ego_trajectory, ado_ids)
        if not self._env.is_differentiable_wrt_ego:
            pass                                                  # Kalman: only the ego-position part is non-zero (in-kernel)
        return _np(self._select(jac, ado_ids)).astype(np.float64).flatten()

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(min_distance=True, ado_mask=self._ado_mask(ado_ids))

    def _constraint_limits(self):
        return constants.CONSTRAINT_MIN_L2_DISTANCE, None

    def _num_constraints(self, ado_ids) -> int:
        k = self._env.num_ados if ado_ids is None else len(ado_ids)
        return k * self.t_horizon

    @property
    def name(self) -> str:
        return "min_distance"


class InteractionEllipsoidModule(_FusedConstraintModule):
    """sum_xy (ego_p(t) - mu_p[k, t])^2 / max(scale[k, t], 0.01) >= 1 for EVERY ado of the scene (the reference's core ignores
    `ado_ids` once it is non-empty) and t = 0..T-1: N*T rows, not normalised; means and scales are detached, so the
    Jacobian goes through the ego rollout only (modules/baselines/ellipsoid_interact.py:13-125). One particle draw per
    call, like the reference's `env.compute_distributions`."""
    spec_flag = "ellipsoid"
    needs_rollout = True

    def __init__(self, env, t_horizon: int, **unused):
        super().__init__(env=env, t_horizon=t_horizon, **{k: v for k, v in unused.items() if k == "num_particles"})

    def _spec(self, ado_ids) -> ModuleSpec:
        return ModuleSpec(ellipsoid=True)

    def _defined(self, ado_ids) -> bool:
        return not (ado_ids is not None and len(ado_ids) == 0) and self._env.num_ados > 0      # ellipsoid_interact.py:37-38

    def constraint(self, ego_trajectory, ado_ids, tag) -> np.ndarray:
        if not self._defined(ado_ids):
            return self._return_constraint(np.array([]), tag=tag)
        return super().constraint(ego_trajectory, ado_ids, tag)

    def compute_constraint(self, ego_trajectory, ado_ids, tag):
        if not self._defined(ado_ids):
            return None
        return super().compute_constraint(ego_trajectory, ado_ids, tag)

    def compute_jacobian_analytically(self, ego_trajectory, grad_wrt, ado_ids, tag) -> np.ndarray:
        if not self._defined(ado_ids):
            return np.array([])
        if not self.gradient_condition():                         # optimization_module.py:298-301 (Kalman: zeros)
            return np.zeros(self._env.num_ados * self.t_horizon * int(grad_wrt.numel()))
        return _np(self._dense_jacobian(ego_trajectory, ado_ids)).astype(np.float64).flatten()

    def gradient_condition(self) -> bool:
        return self._env.is_differentiable_wrt_ego                # ellipsoid_interact.py:80-89

    def normalize(self, x):
        return x                                                  # ellipsoid_interact.py:91-99

    def _constraint_limits(self):
        return 1, None                                            # ellipsoid_interact.py:104-112

    def _num_constraints(self, ado_ids) -> int:
        num_modes = min(constants.CONSTRAINT_ELLIPSOID_NUM_MODES, self._env.num_modes)
        return num_modes * (self._env.num_ados if ado_ids is None else len(ado_ids)) * self.t_horizon

    @property
    def name(self) -> str:
        return "ellipsoid_interaction"


# `mantrap.modules.baselines.*` of the reference (modules/baselines/__init__.py:1-5)
import types as _types  # noqa: E402

baselines = _types.SimpleNamespace(InteractionAccelerationModule=InteractionAccelerationModule,
                                   InteractionVelocityModule=InteractionVelocityModule,
                                   InteractionPositionModule=InteractionPositionModule,
                                   InteractionEllipsoidModule=InteractionEllipsoidModule)
base = _types.SimpleNamespace(OptimizationModule=OptimizationModule, PureObjectiveModule=PureObjectiveModule,
                              PureConstraintModule=PureConstraintModule)
