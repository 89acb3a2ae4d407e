"""Batched closed-loop planning: `TrajOptSolver.solve` (mantrap/solver/base/trajopt.py:109-193) for S scenes in lock step.

The reference plans ONE scene: optimise -> execute the first control in the evaluation environment
(`GraphBasedEnvironment.step`, graph_based.py:82-124) -> shift the warm start -> repeat until `time_steps` or the goal is
reached. Here S independent scenes advance together and stay on the device between the solves:

    per step k:  attention mask            mtb_attention           (attention/euclidean.py, closest.py; optional)
                 ego-free baseline         mtb_evaluate, no ego     (module __init__, acc_interact.py:33-34)
                 C candidates per scene    mtb_evaluate + mtb_argmin (search.py:98-103, random_search.py:28-35); the warm
                                           start is kept as incumbent (candidate 0) -- a deliberate deviation, see `plan`
                 closed-loop step          mtb_env_step            (graph_based.py:82-124, warm-start shift trajopt.py:150-155)

This feeds `mantrap_evaluation.evaluate`-style sweeps (metrics.py:12-63) and warm-start data generation
(third_party/warm_start/pre_compute.py:27-58) without a host round trip per scene. Random numbers (candidate controls,
the step's standard-normal draws) come from a `torch.Generator` on the engine device or are supplied by the caller
(parity tests feed the oracle the same numbers). A scene that reached its goal (`SOLVER_GOAL_END_DISTANCE`) is frozen,
as the reference's loop stops there.
"""
import typing

import torch

from . import constants
from .engine import Engine, ModuleSpec, SceneBatch


class ClosedLoopResult(typing.NamedTuple):
    ego_trajectory: torch.Tensor     # [S, K+1, 4]  (states after k executed steps; frozen once the goal is reached)
    ado_trajectories: torch.Tensor   # [S, N, K+1, 4]
    controls: torch.Tensor           # [S, K, 2]    first control of every solve (before the agent's own clamp)
    objective: torch.Tensor          # [S, K]       objective of the chosen plan at step k
    best_index: torch.Tensor         # [S, K] int32 chosen candidate (-1: no feasible candidate, warm start kept)
    steps: torch.Tensor              # [S] int32    executed steps until the goal was reached (K if never)


class BatchedClosedLoop:
    """Random-search model-predictive control of a whole `SceneBatch` (the batched `RandomSearch.solve`)."""

    def __init__(self, engine: Engine, scene: SceneBatch, t_planning: int = constants.SOLVER_HORIZON_DEFAULT,
                 weights: typing.Dict[str, float] = None, num_candidates: int = 256, attention: typing.Optional[str] = None,
                 control_limits: bool = True, speed_limits: bool = True, min_distance: bool = False, ellipsoid: bool = False,
                 feas_tol: float = constants.SOLVER_CONSTRAINT_LIMIT, seed: int = 0):
        assert scene.ego_goal is not None and attention in (None, "euclidean", "closest")
        self.engine, self.t_planning, self.num_candidates = engine, t_planning, num_candidates
        self.weights = dict(weights) if weights is not None else {"goal": 0.5, "acc": 1.0}
        self.attention, self.feas_tol = attention, feas_tol
        self.flags = dict(control_limits=control_limits, speed_limits=speed_limits, min_distance=min_distance, ellipsoid=ellipsoid)
        self.scene = scene.to(engine.device)
        self.scene.ego0, self.scene.ado0 = self.scene.ego0.clone(), self.scene.ado0.clone()   # advanced in place
        self.generator = torch.Generator(device=engine.device).manual_seed(seed)
        self._needs_baseline = any(self.weights.get(k, 0.0) != 0.0 for k in ("acc", "vel", "pos", "prob"))

    def draw_candidates(self) -> torch.Tensor:
        """z ~ U(-u_max, u_max) per candidate (trajopt.py:359-363, random_search.py:28), on the device."""
        s, c, t = self.scene.num_scenes, self.num_candidates, self.t_planning
        z = torch.rand((s, c, t, 2), generator=self.generator, device=self.engine.device)
        return (z * 2.0 - 1.0) * constants.ROBOT_ACC_MAX

    def plan(self, warm: torch.Tensor, candidates: torch.Tensor = None):
        """One solve for every scene: candidates [S,C,T,2] (drawn when None) with candidate 0 replaced by the warm start
        -> (controls [S,T,2], objective [S], index [S]).

        DELIBERATE DEVIATION from the reference's search frame: `SearchIntermediate.optimize_core`
        (solver/base/search.py:19-81) ignores z0, starts from obj = inf and falls back to a fresh random z when nothing feasible
        was found. A closed loop that re-plans every step from scratch throws its previous plan away, so here the shifted warm
        start competes as candidate 0 (it can only win when it is feasible), and when NO candidate is feasible the plan is the
        warm start again, reported with index -1 and the warm start's objective (`tests/dropin_cases.py:
        check_plan_with_infeasible_warm_start`)."""
        eng, scene = self.engine, self.scene
        candidates = self.draw_candidates() if candidates is None else candidates.to(eng.device).clone()
        candidates[:, 0] = warm
        mask = eng.attention(self.attention, scene.ego0, scene.ado0) if self.attention is not None else None
        mus_wo = eng.rollout_wo_ego(scene, self.t_planning, want=("mus",))["mus"].contiguous() if self._needs_baseline else None
        spec = ModuleSpec(weights=self.weights, mus_wo=mus_wo, ado_mask=mask, **self.flags)
        res = eng.evaluate(scene, candidates.contiguous(), spec, want=("total", "violation"))
        idx, val, best = eng.argmin(res["total"], res["violation"], candidates, self.feas_tol)
        none = (idx < 0)
        # no feasible candidate: keep the incumbent (the reference would draw a fresh random z here, search.py:69-70)
        best = torch.where(none[:, None, None], warm, best)
        val = torch.where(none, res["total"][:, 0], val)
        return best, val, idx

    def solve(self, time_steps: int, candidates: typing.Sequence[torch.Tensor] = None, eps: typing.Sequence[torch.Tensor] = None,
              sample: bool = True, warm_start_db: typing.Tuple[torch.Tensor, torch.Tensor] = None) -> ClosedLoopResult:
        """`time_steps` closed-loop steps. `candidates[k]` / `eps[k]` override the device draws of step k (tests); `sample
        = False` moves the ados along the mean of their one-step distribution; `warm_start_db` = (encodings [M,4], solutions
        [M,H,2]) starts every scene from its nearest pre-computed scenario instead of zeros."""
        eng, scene = self.engine, self.scene
        s, n, dev = scene.num_scenes, scene.num_ados, eng.device
        ego_traj = torch.empty((s, time_steps + 1, 4), device=dev)
        ado_traj = torch.empty((s, n, time_steps + 1, 4), device=dev)
        taken = torch.zeros((s, time_steps, 2), device=dev)
        objective = torch.zeros((s, time_steps), device=dev)
        chosen = torch.full((s, time_steps), -1, dtype=torch.int32, device=dev)
        steps = torch.full((s,), time_steps, dtype=torch.int32, device=dev)
        done = torch.zeros(s, dtype=torch.bool, device=dev)
        ego_traj[:, 0], ado_traj[:, :, 0] = scene.ego0, scene.ado0
        if warm_start_db is not None:                                      # WARM_START_ENCODING for every scene (trajopt.py:308-331)
            enc_db, sol_db = (x.to(dev).float().contiguous() for x in warm_start_db)
            _, _, warm = eng.warm_start_encoding(scene.ego0, scene.ego_goal, scene.ado0, enc_db, sol_db, self.t_planning)
        else:
            warm = torch.zeros((s, self.t_planning, 2), device=dev)        # WARM_START_ZEROS (constants.py:76)
        next_warm = torch.empty_like(warm)
        for k in range(time_steps):
            best, val, idx = self.plan(warm, None if candidates is None else candidates[k])
            ego_prev, ado_prev = scene.ego0.clone(), scene.ado0.clone()
            if eps is not None:
                eps_k = eps[k].to(dev).contiguous()
            else:
                eps_k = torch.randn((s, n, 2), generator=self.generator, device=dev) if sample else None
            eng.env_step(scene, best.contiguous(), eps_k, next_warm)
            # scenes that already reached their goal stay where they are (the reference's loop has ended for them)
            scene.ego0.copy_(torch.where(done[:, None], ego_prev, scene.ego0))
            scene.ado0.copy_(torch.where(done[:, None, None], ado_prev, scene.ado0))
            ego_traj[:, k + 1], ado_traj[:, :, k + 1] = scene.ego0, scene.ado0
            taken[:, k] = torch.where(done[:, None], torch.zeros_like(best[:, 0]), best[:, 0])
            objective[:, k], chosen[:, k] = val, idx
            reached = torch.norm(scene.ego0[:, 0:2] - scene.ego_goal, dim=1) < constants.SOLVER_GOAL_END_DISTANCE
            steps = torch.where(reached & ~done, torch.full_like(steps, k + 1), steps)
            done = done | reached
            warm, next_warm = next_warm, warm
        return ClosedLoopResult(ego_traj, ado_traj, taken, objective, chosen, steps)
