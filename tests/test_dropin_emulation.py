"""CPU run of the drop-in API tests: the host classes (agents / environment / modules / solver) drive an Engine that is
bound to the host emulation of the kernel source (tests/host_emul) instead of the CUDA library. The GPU run of the same
bodies is tests/test_gpu_dropin.py."""
import ctypes
import os
import sys

import pytest

import dropin_cases as dc
import golden_io as gio
from mantrap_b200 import _cabi, engine, runtime

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul"))
import build_emul  # noqa: E402

SMALL = ["c1_sf_n3_t10", "c2_pf_n5_t15", "kalman_n5_t15", "sf_single_n6_t12", "pf_single_n4_t8", "sf_dense_n5_t6",
         "c4_sf_n20_p3_t12"]


@pytest.fixture(scope="module", autouse=True)
def emul_engine():
    lib = _cabi.bind(ctypes.CDLL(build_emul.build()))
    previous = runtime.set_default_engine(engine.Engine.from_library(lib, "cpu"))
    yield
    runtime.set_default_engine(previous)


def test_agents():
    dc.check_agents()


@pytest.mark.parametrize("name", gio.case_names())
def test_environment(name):
    dc.check_environment(name)


@pytest.mark.parametrize("name", gio.case_names())
def test_modules(name):
    dc.check_modules(name)


@pytest.mark.parametrize("name", SMALL)
def test_autograd_route(name):
    dc.check_autograd_route(name)


@pytest.mark.parametrize("name", ["c1_sf_n3_t10", "c2_pf_n5_t15", "c3_sf_n10_p4_t20", "kalman_n5_t15", "sf_single_n6_t12"])
def test_solver_callbacks(name):
    dc.check_solver_callbacks(name)


def test_attention():
    dc.check_attention()


def test_warm_start_encoding():
    dc.check_warm_start_encoding()


def test_random_search():
    dc.check_random_search(32)


@pytest.mark.parametrize("env_kind", ["sf", "pf", "kalman"])
def test_closed_loop_slsqp(env_kind):
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.SLSQPSolver, env_kind, steps=3, max_iter=15)


def test_closed_loop_random_search():
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.RandomSearch, "sf", steps=2, num_candidates=32, max_rounds=1)


def test_mcts():
    dc.check_mcts(4, 6)


def test_closed_loop_mcts():
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.MonteCarloTreeSearch, "sf", steps=2, num_breadth_samples=4, num_depth_samples=8, max_cpu_time=60.0)


def test_ipopt_solver_reports_missing_cyipopt():
    """cyipopt is a third-party dependency of the reference (ipopt==0.1.9) and not installed: the callbacks work, the
    interior-point call raises a clear ImportError instead of silently doing something else."""
    import importlib.util
    import torch
    from mantrap_b200 import solver
    if importlib.util.find_spec("ipopt") or importlib.util.find_spec("cyipopt"):
        pytest.skip("cyipopt installed")
    env, goal = dc.small_scene("kalman")
    sol = solver.IPOPTSolver(env, goal=goal, t_planning=3, modules=dc.solver.IPOPTSolver.module_hard())
    with pytest.raises(ImportError):
        sol.optimize(torch.zeros(6), tag="t")
