"""Solvers of the hot path with the reference's interface (`mantrap/solver`).

The optimisation variable is z = the ego's controls, flattened (t, xy) (solver/base/trajopt.py:356-357, :439-444).
`TrajOptSolver` keeps the reference's frame -- module construction, `objective / constraints / gradient / jacobian (z)`,
warm start, the closed-loop `solve()` -- and answers the four callbacks from the CUDA kernels:

* `fused=True` (default): ONE `mtb_evaluate` launch per distinct z produces every module's objective, the summed
  gradient, all constraint rows and the violation (one particle draw shared by all modules); the four callbacks
  of an IPOPT iteration share it through a per-z cache (callbacks may be invoked in any order with the same z,
  SURVEY §8b). The dense Jacobian is one `mtb_constraint_jacobian` launch.
* `fused=False`: module by module, in the reference's call order, each drawing its own particles -- with
  `torch.manual_seed` pinned this reproduces the reference's numbers call for call (parity tests).

Mirrors: TrajOptSolver solver/base/trajopt.py:18-666; IPOPTSolver solver/ipopt.py:12-229 (cyipopt is third-party and not
shipped: `optimize_core` binds `ipopt` / `cyipopt` when importable and raises otherwise); SearchIntermediate
solver/base/search.py:11-108; RandomSearch solver/baselines/random_search.py:9-39 (here: thousands of candidates per
launch + device argmin instead of one candidate per Python iteration). `SLSQPSolver` drives the same four callbacks
with scipy and is NOT IPOPT; it exists for end-to-end checks of the callbacks.
"""
import dataclasses
import time
import typing

import numpy as np
import torch

from . import constants, modules as mods, runtime
from .engine import PreparedEvaluation, OBJECTIVE_KEYS, ModuleSpec, SceneBatch, to_device_packed


class TrajOptSolver:
    def __init__(self, env, goal: torch.Tensor, t_planning: int = constants.SOLVER_HORIZON_DEFAULT,
                 modules: typing.Union[typing.List[typing.Tuple], typing.List] = None, attention_module=None,
                 eval_env=None, config_name: str = "unknown", is_logging: bool = False, is_debug: bool = False,
                 fused: bool = True, num_particles: int = constants.ENV_NUM_PARTICLES, **solver_params):
        goal = torch.as_tensor(goal).detach().float().cpu()
        assert goal.shape == (2,)
        self._solver_params = dict(solver_params, t_planning=t_planning, config_name=config_name, goal=goal)
        self._env = env.copy()
        self._eval_env = env.copy(env_type=eval_env) if eval_env is not None else env.copy()
        assert self._env.ego is not None and self._env.same_initial_conditions(self._eval_env)
        self._fused, self._num_particles = fused, num_particles
        module_list = self.module_defaults() if modules is None else modules
        self._module_dict = {}
        for entry in module_list:
            cls, kwargs = (entry[0], (entry[1] if len(entry) > 1 and entry[1] is not None else {})) \
                if isinstance(entry, tuple) else (entry, {})
            module = cls(t_horizon=self.planning_horizon, goal=self.goal, env=self.env, num_particles=num_particles, **kwargs)
            self._module_dict[module.name] = module
        self._attention_module = attention_module(env=self.env, t_horizon=self.planning_horizon) \
            if attention_module is not None else None
        self._log, self._cache, self._prepared_calls = {}, {}, {}
        self._is_logging = is_logging
        assert self.num_optimization_variables() > 0

    # ---- closed loop -------------------------------------------------------------------------------------------
    def solve(self, time_steps: int, warm_start_method: str = constants.WARM_START_HARD, **kwargs
              ) -> typing.Tuple[torch.Tensor, torch.Tensor]:
        """MPC loop: optimise, execute the first control in the evaluation environment, shift the warm start, repeat
        until `time_steps` or the goal is reached (trajopt.py:109-193). The solver's environments are restored at the end."""
        ego_trajectory_opt = torch.zeros((time_steps + 1, 5))
        ado_trajectories = torch.zeros((self.env.num_ados, time_steps + 1, 1, 5))
        env_copy, eval_env_copy = self.env.copy(), self.eval_env.copy()
        ego_trajectory_opt[0] = self._env.ego.state_with_time
        for m, ado in enumerate(self.env.ados):
            ado_trajectories[m, 0, 0] = ado.state_with_time
        z_warm_start = self.warm_start(method=warm_start_method)
        for k in range(time_steps):
            z_k = self.optimize(z_warm_start, tag=constants.TAG_OPTIMIZATION, **kwargs)
            ego_controls_k = z_k.detach().float().view(-1, 2)
            z_warm_start = torch.cat((ego_controls_k[1:], torch.zeros((1, 2)))).flatten()
            ado_states, ego_state = self._eval_env.step(ego_action=ego_controls_k[0])
            self._env.step_reset(ego_next=ego_state, ado_next=ado_states)
            self._refresh_modules()
            ego_trajectory_opt[k + 1] = ego_state
            ado_trajectories[:, k + 1, 0] = ado_states
            if torch.norm(ego_state[0:2] - self.goal) < constants.SOLVER_GOAL_END_DISTANCE:
                ego_trajectory_opt = ego_trajectory_opt[:k + 2].detach()
                ado_trajectories = ado_trajectories[:, :k + 2].detach()
                break
        self._env, self._eval_env = env_copy, eval_env_copy
        self._refresh_modules()
        return ego_trajectory_opt, ado_trajectories

    def _refresh_modules(self):
        self._cache.clear()
        self._prepared_calls.clear()              # the static scene on the device belongs to the previous environment state
        for module in self.modules:
            module.reset_env(env=self.env)

    def optimize(self, z0: torch.Tensor, tag: str, **kwargs) -> torch.Tensor:
        ado_ids = self._attention_module.compute() if self._attention_module is not None else self.env.ado_ids
        z_opt, self._log = self.optimize_core(z0, ado_ids=ado_ids, tag=tag, **kwargs)
        return z_opt

    def optimize_core(self, z0: torch.Tensor, tag: str, ado_ids: typing.List[str], **kwargs):
        raise NotImplementedError

    # ---- warm start --------------------------------------------------------------------------------------------
    def warm_start(self, method: str = constants.WARM_START_HARD) -> torch.Tensor:
        """"hard": the same problem with the hard modules only, started from full throttle towards the goal
        (trajopt.py:248-306); "potential": the full default formulation on a `PotentialFieldEnvironment` copy of the scene
        (trajopt.py:266-268); "zeros": zero controls; "encoding": nearest pre-computed scenario (trajopt.py:308-331; needs a
        database, see `warm_start_database` / `mantrap_b200.warm_start_db.pre_compute`). "soft" needs `HJReachabilityModule`,
        whose value grid the reference does not ship."""
        if method == constants.WARM_START_HARD:
            return self._warm_start_optimization(self.env, self.module_hard())
        if method == constants.WARM_START_POTENTIAL:
            from .environment import PotentialFieldEnvironment
            return self._warm_start_optimization(self.env.copy(env_type=PotentialFieldEnvironment), self.module_defaults())
        if method == constants.WARM_START_SOFT:
            raise ValueError("warm start 'soft' needs HJReachabilityModule (third_party/reachability/2D.mat is not shipped by the reference)")
        if method == constants.WARM_START_ZEROS:
            return torch.zeros(2 * self.planning_horizon)
        if method == constants.WARM_START_ENCODING:
            return self._warm_start_encoding()
        raise ValueError(f"Invalid warm starting-method {method} !")

    def _warm_start_optimization(self, env, modules) -> torch.Tensor:
        """The same solver class on `env` with `modules` only, started from full throttle towards the goal (trajopt.py:280-306)."""
        part = self.__class__(env=env, goal=self.goal, modules=modules, t_planning=self.planning_horizon,
                              config_name=self.config_name, fused=self._fused, num_particles=self._num_particles)
        _, u_max = self.env.ego.control_limits()
        dx_goal = self.goal - self.env.ego.position
        init = torch.stack([dx_goal / torch.norm(dx_goal).clamp(min=1e-6) * u_max] * self.planning_horizon)
        return part.optimize(z0=init.flatten(), tag=constants.TAG_WARM_START)

    # ---- encoding / pre-computed warm starts ---------------------------------------------------------------------
    def _scene_arrays(self):
        engine = runtime.default_engine()
        ego_state, ado_states = self.env.states()
        return (engine, runtime.to_device(ego_state[0:4].view(1, 4), engine), runtime.to_device(self.goal.view(1, 2), engine),
                runtime.to_device(ado_states[:, 0:4].reshape(1, self.env.num_ados, 4), engine))

    def warm_start_database(self) -> typing.Tuple[torch.Tensor, torch.Tensor]:
        """(encoding_db [M,4], solution_db [M,H,2]): the `warm_start_db` solver parameter -- a pair of tensors or a directory
        holding the reference's `encoding.pt` / `solution.pt` (constants.py:79, third_party/warm_start/) -- else the directory
        named by $MANTRAP_WARM_START_DIR. The database is data of the reference and is not shipped with this package."""
        import os
        db = self._solver_params.get("warm_start_db", os.environ.get("MANTRAP_WARM_START_DIR"))
        if db is None:
            raise FileNotFoundError("warm start 'encoding' needs the pre-computed database: pass warm_start_db=(encodings, solutions) "
                                    "or a directory with encoding.pt / solution.pt, or set MANTRAP_WARM_START_DIR")
        if isinstance(db, (str, os.PathLike)):
            enc_file, sol_file = constants.WARM_START_PRE_COMPUTATION_FILE
            db = (torch.load(os.path.join(db, enc_file)), torch.load(os.path.join(db, sol_file)))
        enc_db, sol_db = db
        assert enc_db.ndim == 2 and enc_db.shape[1] == 4 and sol_db.ndim == 3 and sol_db.shape[2] == 2
        return enc_db, sol_db

    def encode(self) -> torch.Tensor:
        """Four-dimensional scene encoding (trajopt.py:466-499): closest ado in the robot -> goal frame + robot velocity."""
        engine, ego0, goal, ado0 = self._scene_arrays()
        dummy = torch.zeros((1, 4), dtype=torch.float32, device=engine.device)
        encoding, _, _ = engine.warm_start_encoding(ego0, goal, ado0, dummy, None, self.planning_horizon)
        return encoding[0].cpu()

    def _warm_start_encoding(self) -> torch.Tensor:
        """Solution of the nearest pre-computed scenario (trajopt.py:308-331), flattened z."""
        assert self.env.dt == constants.ENV_DT_DEFAULT            # pre-computation only for the default time step (:313)
        enc_db, sol_db = self.warm_start_database()
        assert sol_db.shape[1] >= self.planning_horizon
        engine, ego0, goal, ado0 = self._scene_arrays()
        _, _, warm = engine.warm_start_encoding(ego0, goal, ado0, runtime.to_device(enc_db, engine), runtime.to_device(sol_db, engine),
                                                self.planning_horizon)
        return warm[0].cpu().flatten()

    @staticmethod
    def module_defaults():
        """The reference's defaults: goal, interaction-probability, speed limits and HJ reachability (trajopt.py:336-343).
        The HJ value grid (`third_party/reachability/2D.mat`) is not shipped with the reference, so that module is absent."""
        return [mods.GoalNormModule, mods.InteractionProbabilityModule, mods.SpeedLimitModule]

    @staticmethod
    def module_hard():
        return [mods.GoalNormModule, mods.ControlLimitModule, mods.SpeedLimitModule]

    # ---- problem formulation -------------------------------------------------------------------------------------
    def num_optimization_variables(self) -> int:
        return 2 * self.planning_horizon

    def optimization_variable_bounds(self) -> typing.Tuple[typing.List, typing.List]:
        lower, upper = self._env.ego.control_limits()
        n = self.num_optimization_variables()
        return (np.ones(n) * lower).tolist(), (np.ones(n) * upper).tolist()

    def constraint_bounds(self, ado_ids) -> typing.Tuple[typing.List, typing.List]:
        cl, cu = [], []
        for module in self.modules:
            lower, upper = module.constraint_boundaries(ado_ids=ado_ids)
            cl += list(lower)
            cu += list(upper)
        return cl, cu

    def _fusable(self) -> bool:
        return self._fused and all(getattr(m, "fused_key", None) is not None or getattr(m, "spec_flag", None) is not None
                                   for m in self.modules)

    def _fused_spec(self, ado_ids) -> ModuleSpec:
        spec = ModuleSpec(weights={}, ado_mask=None)
        for m in self.modules:
            one = m._spec(ado_ids)
            spec.weights.update(one.weights)
            spec.mus_wo = one.mus_wo if one.mus_wo is not None else spec.mus_wo
            spec.ado_mask = one.ado_mask if one.ado_mask is not None else spec.ado_mask
            spec.control_limits |= one.control_limits
            spec.speed_limits |= one.speed_limits
            spec.min_distance |= one.min_distance
            spec.ellipsoid |= one.ellipsoid
            if one.control_limits:
                spec.u_max = one.u_max
            if one.speed_limits:
                spec.v_max = one.v_max
        if len(ado_ids) == 0:                       # interaction objectives are undefined without ados (-> 0)
            spec.ellipsoid = False                  # ellipsoid_interact.py:37-38: no rows
            for key in ("acc", "vel", "pos", "prob"):
                spec.weights.pop(key, None)
        return spec

    def _prepared(self, ado_ids: typing.List[str], num_candidates: int = 1) -> PreparedEvaluation:
        """The pre-built launch of this solver's callback (engine.PreparedEvaluation): static scene on the device, C-ABI structs,
        result buffers and scratch are set up once per (environment state, ado_ids, candidate count)."""
        key = (tuple(ado_ids), num_candidates)
        prep = self._prepared_calls.get(key)
        if prep is None:
            engine, env = runtime.default_engine(), self.env
            rollout = any(m.needs_rollout for m in self.modules)
            if rollout:
                scene = env.scene_static()
                num_p = self._num_particles if env.env_kind != "kalman" else 1
            else:                                   # ego-only modules: no random numbers, one inert Kalman ado
                scene, num_p = self.modules[0]._scene().to("cpu"), 1
            scene.ego_goal = self.goal.view(1, 2).clone()
            prep = PreparedEvaluation(engine, scene, self._fused_spec(ado_ids), self.planning_horizon, num_p, num_candidates,
                                      with_noise=rollout and env.env_kind != "kalman")
            prep.draws = rollout and env.env_kind != "kalman"
            prep.num_particles = num_p
            self._prepared_calls[key] = prep
        return prep

    def evaluate_fused(self, z: np.ndarray, ado_ids: typing.List[str], want_jacobian: bool = False) -> typing.Dict:
        """One launch for everything the callbacks need at this z (cached). Keys: obj {module: value}, total, grad [2T],
        cons {module: rows}, violation, optionally jac {module: [rows, 2T]}. Per call only z and the fresh particle draw travel to
        the device (one pinned block, one copy) and one block of results comes back (engine.PreparedEvaluation)."""
        key = (np.asarray(z, dtype=np.float64).tobytes(), tuple(ado_ids))
        hit = self._cache.get(key)
        if hit is not None and (not want_jacobian or "jac" in hit):
            return hit
        prep = self._prepared(ado_ids)
        spec = prep.spec
        if hit is None:
            draw = self.env.draw_particle_parameters(prep.num_particles, self.planning_horizon - 1) if prep.draws else None
            prep.stage(np.asarray(z, dtype=np.float32), draw)
            res = prep.run(want_jacobian=want_jacobian)
            obj, cons = res["obj"][0], res["cons"][0]
            out = dict(obj={m.name: float(obj[OBJECTIVE_KEYS.index(m.fused_key)]) for m in self.modules
                            if getattr(m, "fused_key", None) is not None},
                       total=float(res["total"][0]), grad=res["grad"][0].astype(np.float64), violation=float(res["violation"][0]),
                       cons=self._split_rows(cons, ado_ids, spec), _staged=prep.staged_inputs())
        else:
            # Jacobian requested for a z whose values are cached: the SAME particle draw (staged inputs kept with the entry), so that
            # cons / grad / violation / jac of one z belong together ("one draw per callback set") and the global
            # RNG does not advance on a cache hit
            out = hit
            prep.restore(hit["_staged"])
            res = prep.run(want_jacobian=True, want_values=False)
        if want_jacobian:
            out["jac"] = self._split_rows(res["jac"][0].astype(np.float64), ado_ids, spec) if res["jac"] is not None else {}
        if len(self._cache) > 64:
            self._cache.clear()
        self._cache[key] = out
        return out

    def evaluate_fused_batch(self, z: np.ndarray, ado_ids: typing.List[str] = None, want_jacobian: bool = False) -> typing.Dict:
        """K independent callback sets in ONE launch (SURVEY §8e "replicas only": K IPOPT replicas, line-search points or restarts
        of the same scene): z [K, 2T] -> total [K], grad [K, 2T], cons [K, R] (kernel row order), violation [K] and, on request, the
        dense Jacobians jac [K, R, 2T]. All K share one particle draw (they are candidates of one scene)."""
        ado_ids = ado_ids if ado_ids is not None else self.env.ado_ids
        assert self._fusable(), "evaluate_fused_batch needs the built-in (fused) modules"
        z = np.asarray(z, dtype=np.float32)
        prep = self._prepared(ado_ids, z.shape[0])
        draw = self.env.draw_particle_parameters(prep.num_particles, self.planning_horizon - 1) if prep.draws else None
        prep.stage(z, draw)
        return prep.run(want_jacobian=want_jacobian)

    def _split_rows(self, rows: np.ndarray, ado_ids, spec: ModuleSpec) -> typing.Dict[str, np.ndarray]:
        """Kernel row order (control | speed | min-distance | ellipsoid) -> per module, min-distance restricted to
        `ado_ids` (the ellipsoid rows cover every ado, like the reference's core)."""
        t_hor, n = self.planning_horizon, self.env.num_ados
        out, base = {}, 0
        for flag, count, name in (("control_limits", 2 * t_hor if self.env.ego.ego_kind == "double" else t_hor, "control_limits"),
                                  ("speed_limits", 2 * (t_hor + 1), "speed_limits"), ("min_distance", n * t_hor, "min_distance"),
                                  ("ellipsoid", n * t_hor, "ellipsoid_interaction")):
            if not getattr(spec, flag):
                continue
            block = rows[base:base + count]
            if name == "min_distance" and list(ado_ids) != list(self.env.ado_ids):
                idx = [self.env.index_ado_id(i) for i in ado_ids]
                block = block.reshape(n, t_hor, *block.shape[1:])[idx].reshape(len(idx) * t_hor, *block.shape[1:])
            out[name] = block
            base += count
        return out

    # ---- the four callbacks --------------------------------------------------------------------------------------
    def objective(self, z: np.ndarray, ado_ids: typing.List[str] = None, tag: str = constants.TAG_OPTIMIZATION) -> float:
        ado_ids = ado_ids if ado_ids is not None else self.env.ado_ids
        if self._fusable():
            return self.evaluate_fused(z, ado_ids)["total"]
        ego_trajectory = self.z_to_ego_trajectory(z)
        return float(np.sum([m.objective(ego_trajectory, ado_ids=ado_ids, tag=tag) for m in self.modules]))

    def constraints(self, z: np.ndarray, ado_ids: typing.List[str] = None, tag: str = constants.TAG_OPTIMIZATION,
                    return_violation: bool = False):
        ado_ids = ado_ids if ado_ids is not None else self.env.ado_ids
        if self._fusable():
            res = self.evaluate_fused(z, ado_ids)
            rows = [res["cons"][m.name] for m in self.modules if m.name in res["cons"]]
            constraints = np.concatenate(rows) if rows else np.array([])
            violation = res["violation"]
            if "min_distance" in res["cons"] and list(ado_ids) != list(self.env.ado_ids):   # violation over ado_ids only
                violation = float(sum(self.module_dict[k]._violation(v, ado_ids) for k, v in res["cons"].items()))
        else:
            ego_trajectory = self.z_to_ego_trajectory(z)
            constraints = np.concatenate([m.constraint(ego_trajectory, tag=tag, ado_ids=ado_ids) for m in self.modules])
            violation = float(np.sum([m.compute_violation_internal(tag=tag) for m in self.modules]))
        return constraints if not return_violation else (constraints, violation)

    def gradient(self, z: np.ndarray, ado_ids: typing.List[str] = None, tag: str = constants.TAG_OPTIMIZATION) -> np.ndarray:
        ado_ids = ado_ids if ado_ids is not None else self.env.ado_ids
        if self._fusable():
            return self.evaluate_fused(z, ado_ids)["grad"]
        ego_trajectory, grad_wrt = self.z_to_ego_trajectory(z, return_leaf=True)
        gradient = [m.gradient(ego_trajectory, grad_wrt=grad_wrt, tag=tag, ado_ids=ado_ids) for m in self.modules]
        return np.sum(gradient, axis=0)

    def jacobian(self, z: np.ndarray, ado_ids: typing.List[str] = None, tag: str = constants.TAG_OPTIMIZATION) -> np.ndarray:
        """Concatenated, flattened, un-normalised module Jacobians (solver/ipopt.py:140-154)."""
        ado_ids = ado_ids if ado_ids is not None else self.env.ado_ids
        if self._fusable():
            res = self.evaluate_fused(z, ado_ids, want_jacobian=True)
            blocks = []
            for m in self.modules:
                if m.name == "control_limits" and self.env.ego.ego_kind == "double":     # reference quirk: identity
                    blocks.append(np.eye(2 * self.planning_horizon).flatten())
                elif m.name in res["jac"]:
                    blocks.append(res["jac"][m.name].flatten())
            return np.concatenate(blocks) if blocks else np.array([])
        ego_trajectory, grad_wrt = self.z_to_ego_trajectory(z, return_leaf=True)
        jacobian = [m.jacobian(ego_trajectory, grad_wrt=grad_wrt, tag=tag, ado_ids=ado_ids) for m in self.modules]
        jacobian = [x.flatten() for x in jacobian if x.size > 0]
        return np.concatenate(jacobian) if jacobian else np.array([])

    def evaluate(self, z: np.ndarray, ado_ids: typing.List[str], tag: str) -> typing.Tuple[float, float]:
        """objective + summed constraint violation (solver/base/search.py:98-103)."""
        objective = self.objective(z, tag=tag, ado_ids=ado_ids)
        _, violation = self.constraints(z, ado_ids=ado_ids, return_violation=True, tag=tag)
        return objective, violation

    # ---- z <-> controls <-> trajectory -----------------------------------------------------------------------------
    def z_to_ego_trajectory(self, z: np.ndarray, return_leaf: bool = False):
        ego_controls = torch.from_numpy(np.asarray(z)).view(-1, 2).float()
        ego_controls.requires_grad = True
        ego_trajectory = self.env.ego.unroll_trajectory(controls=ego_controls, dt=self.env.dt)
        return ego_trajectory if not return_leaf else (ego_trajectory, ego_controls)

    @staticmethod
    def z_to_ego_controls(z: np.ndarray, return_leaf: bool = False):
        ego_controls = torch.from_numpy(np.asarray(z)).view(-1, 2).float()
        ego_controls.requires_grad = True
        return ego_controls if not return_leaf else (ego_controls, ego_controls)

    def ego_trajectory_to_z(self, ego_trajectory: torch.Tensor) -> np.ndarray:
        return self.env.ego.roll_trajectory(ego_trajectory, dt=self.env.dt).flatten().detach().numpy()

    @staticmethod
    def ego_controls_to_z(ego_controls: torch.Tensor) -> np.ndarray:
        return ego_controls.flatten().detach().numpy()

    # ---- properties ------------------------------------------------------------------------------------------------
    @property
    def env(self):
        return self._env

    @property
    def eval_env(self):
        return self._eval_env

    @property
    def goal(self) -> torch.Tensor:
        return self._solver_params["goal"]

    @property
    def planning_horizon(self) -> int:
        return self._solver_params["t_planning"]

    @property
    def module_dict(self):
        return self._module_dict

    @property
    def modules(self):
        return list(self._module_dict.values())

    @property
    def module_names(self) -> typing.List[str]:
        return list(self._module_dict.keys())

    @property
    def config_name(self) -> str:
        return self._solver_params["config_name"]

    @property
    def log(self):
        return self._log

    @property
    def name(self) -> str:
        raise NotImplementedError


class IPOPTProblem:
    """The object cyipopt calls back into (solver/ipopt.py:201-229)."""

    def __init__(self, problem: "IPOPTSolver", ado_ids: typing.List[str], tag: str = constants.TAG_OPTIMIZATION):
        self.problem, self.tag, self.ado_ids = problem, tag, ado_ids

    def objective(self, z: np.ndarray) -> float:
        return self.problem.objective(z, tag=self.tag, ado_ids=self.ado_ids)

    def gradient(self, z: np.ndarray) -> np.ndarray:
        return self.problem.gradient(z, tag=self.tag, ado_ids=self.ado_ids)

    def constraints(self, z: np.ndarray) -> np.ndarray:
        return self.problem.constraints(z, tag=self.tag, ado_ids=self.ado_ids)

    def jacobian(self, z: np.ndarray) -> np.ndarray:
        return self.problem.jacobian(z, tag=self.tag, ado_ids=self.ado_ids)

    def intermediate(self, *args):
        pass


class IPOPTSolver(TrajOptSolver):
    def optimize_core(self, z0: torch.Tensor, ado_ids: typing.List[str], tag: str = constants.TAG_OPTIMIZATION,
                      max_cpu_time: float = constants.IPOPT_MAX_CPU_TIME_DEFAULT, approx_jacobian: bool = False, **solver_kwargs):
        """NLP assembly + IPOPT call with the reference's options (solver/ipopt.py:15-99)."""
        lb, ub = self.optimization_variable_bounds()
        cl, cu = self.constraint_bounds(ado_ids)
        z0_flat = z0.detach().flatten().double().numpy().tolist()
        problem = IPOPTProblem(self, ado_ids=ado_ids, tag=tag)
        options = {"max_cpu_time": max_cpu_time, "tol": constants.IPOPT_OPTIMALITY_TOLERANCE, "mu_strategy": "adaptive",
                   "mehrotra_algorithm": "yes", "hessian_approximation": "limited-memory", "print_level": 0}
        if approx_jacobian:
            options["jacobian_approximation"] = "finite-difference-values"
        big = 2.0e19
        cl = [(-big if x is None else x) for x in cl]
        cu = [(big if x is None else x) for x in cu]
        try:
            import ipopt                                            # cyipopt 0.1.9, the version the reference pins
            nlp = ipopt.problem(n=len(z0_flat), m=len(cl), problem_obj=problem, lb=lb, ub=ub, cl=cl, cu=cu)
            add = nlp.addOption
        except ImportError:
            try:
                import cyipopt                                      # cyipopt >= 1.0: same callbacks, renamed entry points
            except ImportError as err:
                raise ImportError("IPOPTSolver.optimize_core needs cyipopt (`ipopt` or `cyipopt` module), which is a "
                                  "third-party dependency of the reference and is not installed. The four callbacks "
                                  "(objective / gradient / constraints / jacobian) work without it; SLSQPSolver drives "
                                  "them with scipy instead (not IPOPT).") from err
            nlp = cyipopt.Problem(n=len(z0_flat), m=len(cl), problem_obj=problem, lb=lb, ub=ub, cl=cl, cu=cu)
            add = nlp.add_option
        for key, value in options.items():
            add(key, value)
        z_opt, _ = nlp.solve(z0_flat)
        nlp.close()
        return torch.from_numpy(np.asarray(z_opt)).view(-1, 2), self._log

    @staticmethod
    def module_hard():
        """Variable bounds already are the control limits (solver/ipopt.py:105-113)."""
        return [mods.GoalNormModule, mods.SpeedLimitModule]

    @property
    def name(self) -> str:
        return "ipopt"


class SLSQPSolver(TrajOptSolver):
    """NOT IPOPT: scipy's SLSQP over the same four callbacks and bounds. For end-to-end checks where cyipopt is absent."""

    def optimize_core(self, z0: torch.Tensor, ado_ids: typing.List[str], tag: str = constants.TAG_OPTIMIZATION,
                      max_iter: int = 30, **solver_kwargs):
        import scipy.optimize
        lb, ub = self.optimization_variable_bounds()
        cl, cu = self.constraint_bounds(ado_ids)
        n = self.num_optimization_variables()
        cons = []
        if len(cl) > 0:
            lo = np.array([-np.inf if x is None else x for x in cl])
            hi = np.array([np.inf if x is None else x for x in cu])

            def g(z):
                c = self.constraints(z, ado_ids=ado_ids, tag=tag)
                return np.concatenate((c[np.isfinite(lo)] - lo[np.isfinite(lo)], hi[np.isfinite(hi)] - c[np.isfinite(hi)]))

            def g_jac(z):
                j = self.jacobian(z, ado_ids=ado_ids, tag=tag).reshape(len(cl), n)
                j = j / self._row_normalisers(ado_ids)[:, None]           # constraints are normalised, Jacobians are not
                return np.concatenate((j[np.isfinite(lo)], -j[np.isfinite(hi)]))

            cons = [{"type": "ineq", "fun": g, "jac": g_jac}]
        res = scipy.optimize.minimize(lambda z: self.objective(z, ado_ids=ado_ids, tag=tag), z0.detach().flatten().double().numpy(),
                                      jac=lambda z: self.gradient(z, ado_ids=ado_ids, tag=tag), bounds=list(zip(lb, ub)),
                                      constraints=cons, method="SLSQP", options={"maxiter": max_iter, "ftol": 1e-4})
        return torch.from_numpy(res.x).view(-1, 2), self._log

    def _row_normalisers(self, ado_ids) -> np.ndarray:
        factors = []
        for m in self.modules:
            k = m.num_constraints(ado_ids=ado_ids)
            if k > 0:
                factors.append(np.ones(k) * float(1.0 / m.normalize(1.0)))
        return np.concatenate(factors)

    @property
    def name(self) -> str:
        return "slsqp"


class SearchIntermediate(TrajOptSolver):
    """Sampling solvers. Same frame as the reference (solver/base/search.py:19-81): an inner search step is repeated until it
    reports completion or the time budget is spent, and its candidate replaces the incumbent only if it is not worse. The
    reference evaluates ONE candidate per Python call of `evaluate` (search.py:98-103); here an inner step evaluates a whole
    batch of candidates in one launch (`evaluate_batch`) and reduces it on the device."""

    def optimize_core(self, z0: torch.Tensor, ado_ids: typing.List[str], tag: str = constants.TAG_OPTIMIZATION,
                      max_cpu_time: float = constants.SEARCH_MAX_CPU_TIME, max_rounds: int = None, **solver_kwargs):
        start = time.time()
        z_fallback = np.random.uniform(*self.z_bounds)
        z_best, obj_best, iteration, rounds, evaluated = None, np.inf, 0, 0, 0
        while True:
            z_cand, obj_cand, iteration_cand, finished, count = self._optimize_inner(z_best, obj_best, iteration, tag, ado_ids,
                                                                                     **solver_kwargs)
            evaluated += count
            if obj_cand <= obj_best and z_cand is not None:
                z_best, obj_best, iteration = z_cand, obj_cand, iteration_cand
            rounds += 1
            if finished or (max_rounds is not None and rounds >= max_rounds) or time.time() - start > max_cpu_time:
                break
        z_best = z_fallback if z_best is None else z_best
        self._log = {"rounds": rounds, "candidates": evaluated, "objective": obj_best, "iteration": iteration}
        return torch.from_numpy(np.asarray(z_best, dtype=np.float32)).view(-1, 2), self._log

    def _optimize_inner(self, z_best, obj_best, iteration, tag, ado_ids, **kwargs):
        """-> (candidate z or None, its objective, the search's iteration counter, finished?, candidates evaluated)"""
        raise NotImplementedError

    def evaluate_batch(self, z: np.ndarray, ado_ids: typing.List[str]):
        """z [C, 2T] -> device tensors total [1,C], violation [1,C] and the controls [1,C,T,2] (one launch)."""
        assert self._fusable(), "batched search needs the built-in (fused) modules"
        engine = runtime.default_engine()
        controls = torch.from_numpy(np.asarray(z, dtype=np.float32)).view(1, z.shape[0], -1, 2).to(engine.device).contiguous()
        env = self.env
        if any(m.needs_rollout for m in self.modules):
            num_p = self._num_particles if env.env_kind != "kalman" else 1
            scene = env.scene_batch(num_p, self.planning_horizon, with_noise=False)
        else:
            scene = self.modules[0]._scene().to("cpu")
        scene.ego_goal = self.goal.view(1, 2).clone()
        res = engine.evaluate(scene.to(engine.device), controls, self._fused_spec(ado_ids), want=("total", "violation"))
        return res["total"], res["violation"], controls

    @property
    def z_bounds(self) -> typing.Tuple[np.ndarray, np.ndarray]:
        lb, ub = self.optimization_variable_bounds()
        return np.asarray(lb), np.asarray(ub)


class RandomSearch(SearchIntermediate):
    """Uniform candidates within the control bounds, keep the best feasible (baselines/random_search.py:12-35)."""

    def _optimize_inner(self, z_best, obj_best, iteration, tag, ado_ids, num_candidates: int = 4096, **unused):
        lb, ub = self.z_bounds
        z = np.random.uniform(lb, ub, size=(num_candidates, lb.size))
        total, violation, controls = self.evaluate_batch(z, ado_ids)
        idx, val, best = runtime.default_engine().argmin(total, violation, controls, constants.SOLVER_CONSTRAINT_LIMIT)
        if int(idx[0]) < 0:
            return None, np.inf, iteration, False, num_candidates
        return best[0].flatten().cpu().numpy().astype(np.float64), float(val[0]), iteration, False, num_candidates

    @property
    def name(self) -> str:
        return "random_search"


class MonteCarloTreeSearch(SearchIntermediate):
    """Monte-Carlo tree search over the control sequence (baselines/mcts.py:12-98). Search step i fixes the controls of the
    steps before i to the incumbent, samples `num_breadth_samples` values for the control of step i ("priors") and, for each
    prior, `num_depth_samples` random completions of the remaining steps. The value of a prior is the mean objective of its
    FEASIBLE completions (violation < CONSTRAINT_VIOLATION_PRECISION, mcts.py:56); the best completion of the best prior is
    returned and the search moves on to step i + 1, finishing after the last step.

    The reference evaluates the breadth x depth completions one by one (2 x 2 by default, constants.py:87-88, because each
    costs a Python rollout); here they are ONE launch, so the defaults are 64 x 64."""

    def _optimize_inner(self, z_best, obj_best, iteration, tag, ado_ids, num_breadth_samples: int = 64,
                        num_depth_samples: int = 64, **unused):
        lb, ub = self.z_bounds
        t_hor, nb, nd = self.planning_horizon, num_breadth_samples, num_depth_samples
        assert len(lb) == len(ub) == 2 * t_hor and 0 <= iteration <= t_hor
        if iteration >= t_hor:                      # the whole horizon has been searched
            return z_best, obj_best, iteration, True, 0
        z = np.empty((nb, nd, 2 * t_hor))
        if iteration > 0:                           # controls before step i: the incumbent (mcts.py:46-49)
            assert z_best is not None and len(z_best) == 2 * t_hor
            z[:, :, :2 * iteration] = np.asarray(z_best)[:2 * iteration]
        lo_i, hi_i = lb[2 * iteration:2 * iteration + 2], ub[2 * iteration:2 * iteration + 2]
        z[:, :, 2 * iteration:2 * iteration + 2] = np.random.uniform(lo_i, hi_i, size=(nb, 1, 2))   # one value per prior
        rest = 2 * (iteration + 1)
        if rest < 2 * t_hor:
            z[:, :, rest:] = np.random.uniform(lb[rest:], ub[rest:], size=(nb, nd, 2 * t_hor - rest))
        total, violation, _ = self.evaluate_batch(z.reshape(nb * nd, 2 * t_hor), ado_ids)
        total, violation = total.view(nb, nd), violation.view(nb, nd)
        feasible = (violation < constants.CONSTRAINT_VIOLATION_PRECISION) & ~torch.isnan(total)
        inf = torch.full_like(total, float("inf"))
        count = feasible.sum(dim=1)
        mean = torch.where(count > 0, torch.where(feasible, total, torch.zeros_like(total)).sum(dim=1) / count.clamp(min=1),
                           inf[:, 0])                # np.mean of the feasible values, inf without any (mcts.py:68)
        best_val, best_idx = torch.where(feasible, total, inf).min(dim=1)
        prior = int(torch.argmin(mean))              # mcts.py:75
        should_terminate = iteration + 1 == t_hor
        if int(count[prior]) == 0:                   # no feasible completion: keep the incumbent (mcts.py:80-81)
            return z_best, obj_best, iteration, should_terminate, nb * nd
        return z[prior, int(best_idx[prior])].copy(), float(best_val[prior]), iteration + 1, should_terminate, nb * nd

    @property
    def name(self) -> str:
        return "mcts"


# `mantrap.solver.baselines.*` of the reference (solver/baselines/__init__.py:1-4)
import types as _types  # noqa: E402

baselines = _types.SimpleNamespace(RandomSearch=RandomSearch, MonteCarloTreeSearch=MonteCarloTreeSearch)
base = _types.SimpleNamespace(TrajOptSolver=TrajOptSolver, SearchIntermediate=SearchIntermediate)
