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


@pytest.mark.parametrize("env_kind", ["sf", "pf"])
@pytest.mark.parametrize("num_ados,num_particles", [(0, 5), (1, 2), (3, 5), (8, 10), (2, 17)])
def test_particle_draw_follows_the_reference_rng_stream(env_kind, num_ados, num_particles):
    """`draw_particle_parameters` consumes torch's global CPU generator exactly as the reference's `create_particles`
    (particle.py:40-79: one Normal(loc, scale).sample((P,)) + .cdf per (ado, parameter)) followed by one torch.rand per
    step (particle.py:165) does, bit for bit, including above 16 particles where torch switches its normal sampler."""
    import torch
    from mantrap_b200 import environment
    cls = {"sf": environment.SocialForcesEnvironment, "pf": environment.PotentialFieldEnvironment}[env_kind]
    env = cls(ego_position=torch.tensor([-5.0, 0.0]), ego_velocity=torch.tensor([1.0, 0.0]))
    for i in range(num_ados):
        env.add_ado(position=torch.tensor([float(i), 2.0]), velocity=torch.tensor([0.0, -1.0]), goal=torch.tensor([0.0, -5.0]))
    torch.manual_seed(7)
    got = env.draw_particle_parameters(num_particles, 9)
    after = torch.rand(1)
    torch.manual_seed(7)
    k = len(env.param_defaults)
    samples, cdfs = torch.zeros(num_ados, k, num_particles), torch.zeros(num_ados, k, num_particles)
    for m in range(num_ados):
        for ip, (loc, scale) in enumerate(env.param_defaults.values()):
            dist = torch.distributions.Normal(loc, scale)
            samples[m, ip] = dist.sample((num_particles,))
            cdfs[m, ip] = dist.cdf(samples[m, ip])
    weight = torch.prod(cdfs, dim=1)
    weight = weight / torch.norm(weight, dim=0)
    noise = torch.stack([torch.rand((num_ados, num_particles, 2)) for _ in range(9)])[None]
    assert torch.equal(got["v0"][0], samples[:, 0]) and torch.equal(got["weight"][0], weight)
    assert k == 1 or torch.equal(got["sigma"][0], samples[:, 1])
    assert torch.equal(got["noise"], noise) and torch.equal(after, torch.rand(1))


def test_scene_copy_to_host_keeps_values_and_none():
    """`to_device_packed` towards a non-GPU device is the plain per-tensor fp32 copy (the staged single copy is GPU-only and
    covered by tests/test_gpu_dropin.py::test_packed_scene_copy)."""
    import torch
    from mantrap_b200.engine import to_device_packed
    src = [None, torch.arange(6.0).double().view(2, 3), torch.ones(5)[::2]]
    out = to_device_packed(src, "cpu")
    assert out[0] is None and out[1].dtype == torch.float32 and torch.equal(out[1], src[1].float())
    assert out[2].is_contiguous() and out[2].tolist() == [1.0, 1.0, 1.0]


def test_warm_start_potential():
    dc.check_warm_start_potential()


def test_goal_optimize_speed():
    dc.check_goal_optimize_speed()


def test_warm_start_pre_compute(tmp_path):
    dc.check_warm_start_pre_compute(str(tmp_path))


def test_plan_with_infeasible_warm_start():
    dc.check_plan_with_infeasible_warm_start()


def test_jacobian_reuses_the_cached_draw():
    dc.check_jacobian_reuses_the_cached_draw()


def test_evaluate_fused_batch():
    dc.check_evaluate_fused_batch()
