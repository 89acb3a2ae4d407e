"""GPU run (-m gpu) of the drop-in API tests: agents / environment / modules / solver mirrors of the reference interface
on the real CUDA engine, against the golden vectors recorded from the unmodified reference."""
import pytest

import dropin_cases as dc
import golden_io as gio
from mantrap_b200 import runtime

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def cuda_engine():
    from mantrap_b200 import build, engine
    build.build()
    previous = runtime.set_default_engine(engine.Engine("cuda:0"))
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


@pytest.mark.parametrize("name", gio.case_names())
def test_autograd_route(name):
    dc.check_autograd_route(name)


@pytest.mark.parametrize("name", gio.case_names())
def test_solver_callbacks(name):
    dc.check_solver_callbacks(name)


def test_attention():
    dc.check_attention()


def test_warm_start_encoding():
    dc.check_warm_start_encoding()


def test_random_search():
    dc.check_random_search(4096)


@pytest.mark.parametrize("env_kind", ["sf", "pf", "kalman"])
def test_closed_loop_slsqp(env_kind):
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.SLSQPSolver, env_kind, steps=5, max_iter=20)


def test_closed_loop_random_search():
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.RandomSearch, "sf", steps=4, num_candidates=4096, max_rounds=2)


def test_mcts():
    dc.check_mcts(8, 16)


def test_closed_loop_mcts():
    from mantrap_b200 import solver
    dc.check_closed_loop(solver.MonteCarloTreeSearch, "sf", steps=4, num_breadth_samples=64, num_depth_samples=64, max_cpu_time=60.0)
