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


def test_packed_scene_copy():
    """`SceneBatch.to` of a small host scene (one staged host->device copy): every tensor arrives bit-identical, fp32,
    contiguous and 16-byte aligned (the kernels read states as float4), None stays None, and the large-batch path (one
    copy per tensor) gives the same values."""
    import torch
    from mantrap_b200 import engine, scenarios
    for s, n, p in ((1, 3, 5), (1, 1, 2), (3, 7, 3), (400, 20, 3)):
        scene = scenarios.synthetic_scenes(s, n, p, env_kind="sf")
        scene.noise = torch.rand((s, 9, n, p, 2))
        scene.ego_goal = torch.rand((s, 2)).double()            # converted on the way
        packed = sum(-(-t.numel() // 4) * 4 for t in (getattr(scene, f) for f in scene._TENSOR_FIELDS) if t is not None)
        assert (packed <= engine.PACKED_COPY_MAX_FLOATS) == (s < 400)      # both paths are exercised
        moved = scene.to("cuda:0")
        for f in scene._TENSOR_FIELDS:
            src, dst = getattr(scene, f), getattr(moved, f)
            if src is None:
                assert dst is None
                continue
            assert dst.device.type == "cuda" and dst.dtype == torch.float32 and dst.is_contiguous() and dst.shape == src.shape
            assert dst.data_ptr() % 16 == 0
            assert torch.equal(dst.cpu(), src.float())
        again = moved.to("cuda:0")                               # already on the device: no staging, same values
        assert torch.equal(again.ado0, moved.ado0)
    out = engine.to_device_packed([None, torch.ones(3), None, torch.arange(5.0)], "cuda:0")
    assert out[0] is None and out[2] is None and out[1].tolist() == [1.0] * 3 and out[3].tolist() == [0.0, 1.0, 2.0, 3.0, 4.0]


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
