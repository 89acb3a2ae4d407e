"""Batched generator of the pre-computed warm-start database (third_party/warm_start/pre_compute.py:27-58 of the reference).

The reference samples `WARM_START_PRE_COMPUTATION_NUM` (100) random one-pedestrian `PotentialFieldEnvironment` scenes
(pre_compute.py:9-24), solves each with `IPOPTSolver.solve(time_steps=10)` one after the other, and stores
`encoding.pt` [M,4] (`TrajOptSolver.encode`, trajopt.py:466-499) and `solution.pt` [M,H,2] (the executed controls, zero-padded when
the goal is reached early, pre_compute.py:49-55); `TrajOptSolver.warm_start("encoding")` later looks up the nearest encoding
(trajopt.py:308-331).

Here all M scenes are one `SceneBatch` solved in lock step by `mpc.BatchedClosedLoop` (attention -> baseline -> candidates ->
argmin -> `mtb_env_step`, everything on the device); the encodings come from `mtb_warm_start_encoding`. The planner inside
the loop is the batched random search (the sampling baseline of the reference), NOT IPOPT -- cyipopt is a third-party
dependency the reference does not vendor -- so the stored solutions are the search's closed-loop controls; the file format
and the consumer are the reference's.
"""
import os
import typing

import torch

from . import constants
from .engine import Engine, SceneBatch
from .mpc import BatchedClosedLoop
from .scenarios import sample_particle_params


def random_scenes(num_scenarios: int, seed: int = 0) -> SceneBatch:
    """`generate_random_scene` (pre_compute.py:9-24) for M scenes at once: robot position / goal and pedestrian position
    ~ U(axis)^2, robot velocity ~ U(0, ROBOT_SPEED_MAX)^2, pedestrian velocity ~ U(0, 1)^2, potential-field environment."""
    gen = torch.Generator().manual_seed(seed)
    lo, hi = constants.ENV_X_AXIS_DEFAULT

    def uniform(*shape, a=lo, b=hi):
        return torch.rand(*shape, generator=gen) * (b - a) + a

    m = num_scenarios
    ego0 = torch.cat((uniform(m, 2), uniform(m, 2, a=0.0, b=constants.ROBOT_SPEED_MAX)), dim=1)
    ado0 = torch.cat((uniform(m, 1, 2), uniform(m, 1, 2, a=0.0, b=1.0)), dim=2)
    v0, sigma, weight = sample_particle_params("pf", m, 1, constants.ENV_NUM_PARTICLES, torch.Generator().manual_seed(seed + 1))
    return SceneBatch(env_kind="pf", ego_kind="double", dt=constants.ENV_DT_DEFAULT, ego0=ego0.contiguous(), ego_goal=uniform(m, 2),
                      ado0=ado0.contiguous(), ado_goal=ado0[:, :, 0:2].contiguous(), v0=v0, sigma=sigma, weight=weight)


def pre_compute(engine: Engine, num_scenarios: int = constants.WARM_START_PRE_COMPUTATION_NUM,
                solution_horizon: int = constants.WARM_START_PRE_COMPUTATION_HORIZON, num_candidates: int = 1024,
                weights: typing.Dict[str, float] = None, seed: int = 0) -> typing.Tuple[torch.Tensor, torch.Tensor]:
    """`solve_random_scenario` (pre_compute.py:27-58), batched: (encodings [M,4], solutions [M,H,2]) on the host.
    The objective defaults to the reference's default modules on this path: goal + interaction probability, speed limits."""
    scene = random_scenes(num_scenarios, seed).to(engine.device)
    dummy = torch.zeros((1, 4), dtype=torch.float32, device=engine.device)
    encodings, _, _ = engine.warm_start_encoding(scene.ego0.contiguous(), scene.ego_goal.contiguous(), scene.ado0.contiguous(), dummy, None,
                                                 solution_horizon)
    loop = BatchedClosedLoop(engine, scene, t_planning=solution_horizon, num_candidates=num_candidates,
                             weights=weights if weights is not None else {"goal": 0.5, "prob": 1.0}, seed=seed + 2)
    result = loop.solve(solution_horizon)
    # executed controls (what `roll_trajectory` of the closed-loop ego trajectory returns for a double integrator,
    # integrator_double.py:40-41), zero after the goal was reached (pre_compute.py:51-54)
    vel = result.ego_trajectory[:, :, 2:4]
    solutions = (vel[:, 1:] - vel[:, :-1]) / scene.dt
    steps = result.steps.to(torch.int64)[:, None]
    solutions = torch.where(torch.arange(solution_horizon, device=solutions.device)[None, :, None] < steps[:, :, None], solutions,
                            torch.zeros_like(solutions))
    return encodings.cpu(), solutions.contiguous().cpu()


def save(directory: str, encodings: torch.Tensor, solutions: torch.Tensor) -> typing.Tuple[str, str]:
    """Write the database in the reference's file format (constants.py:79): `encoding.pt`, `solution.pt`."""
    os.makedirs(directory, exist_ok=True)
    enc_file, sol_file = constants.WARM_START_PRE_COMPUTATION_FILE
    paths = os.path.join(directory, enc_file), os.path.join(directory, sol_file)
    torch.save(encodings, paths[0])
    torch.save(solutions, paths[1])
    return paths
