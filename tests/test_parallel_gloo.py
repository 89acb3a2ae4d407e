"""Multi-process tests of the sharded search (mantrap_b200/parallel.py) on the CPU: world_size 2 over gloo, every rank
drives an Engine bound to the host emulation of the kernel source. Checks that the scene-partitioned and the
candidate-partitioned search give, on every rank, exactly the winners of a single-process run over the whole problem."""
import ctypes
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(num_scenes, num_candidates, n, p, t_hor):
    from mantrap_b200 import scenarios
    scene = scenarios.synthetic_scenes(num_scenes, n, p, seed=11, param_seed=12, regular_sigma=True)
    controls = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=13, u_max=1.0)
    return scene, controls


def _worker(rank, world, port, shape):
    for path in (ROOT, HERE, os.path.join(HERE, "host_emul")):
        if path not in sys.path:
            sys.path.insert(0, path)
    import build_emul
    from mantrap_b200 import _cabi, engine, parallel
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build_emul.build())), "cpu")
        scene, controls = _problem(*shape)
        t_hor = shape[4]

        def make_spec(shard):
            wo = eng.rollout_wo_ego(shard, t_hor)["mus"].contiguous()
            return engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=wo, control_limits=True, speed_limits=True)

        # single-process answer over the whole problem (computed redundantly on every rank)
        # the single-GPU reference asks for the same outputs as sharded_search (total, violation, grad): the gradient and the
        # forward-only instantiation of a kernel are different binaries, whose forward values may differ in the last bit on the
        # GPU (ptxas contracts them independently under --use_fast_math)
        full = eng.evaluate(scene, controls, make_spec(scene), want=("total", "violation", "grad"))
        idx0, val0, best0 = eng.argmin(full["total"], full["violation"], controls, 1e-3)
        assert int((idx0 >= 0).sum()) > 0, "test problem has no feasible candidate"
        for mode in ("scenes", "candidates"):
            val, idx, best, local = parallel.sharded_search(eng, scene, controls, make_spec, 1e-3, mode=mode)
            assert torch.equal(idx, idx0), (mode, idx, idx0)
            assert torch.equal(val, val0) and torch.equal(best, best0), mode
            lo, hi = parallel.shard_range(controls.shape[0 if mode == "scenes" else 1], rank, world)
            expect = (hi - lo, controls.shape[1]) if mode == "scenes" else (controls.shape[0], hi - lo)
            assert tuple(local["total"].shape) == expect                      # each rank evaluated only its shard
        # every rank holds the same answer
        flat = torch.cat((val, idx.float(), best.flatten()))
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(flat, ref)
    finally:
        dist.destroy_process_group()


def _loop_worker(rank, world, port):
    for path in (ROOT, HERE, os.path.join(HERE, "host_emul")):
        if path not in sys.path:
            sys.path.insert(0, path)
    import build_emul
    from mantrap_b200 import _cabi, engine, mpc, parallel
    torch.set_num_threads(1)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build_emul.build())), "cpu")
        s, c, n, p, t_hor, steps = 5, 4, 3, 2, 4, 2                            # ragged: 5 scenes over 2 ranks
        scene, _ = _problem(s, c, n, p, t_hor)
        gen = torch.Generator().manual_seed(21)
        candidates = [(torch.rand((s, c, t_hor, 2), generator=gen) * 2 - 1) * 2.0 for _ in range(steps)]
        eps = [torch.randn((s, n, 2), generator=gen) for _ in range(steps)]
        kwargs = dict(t_planning=t_hor, num_candidates=c, attention="euclidean")
        ref = mpc.BatchedClosedLoop(eng, scene, **kwargs).solve(steps, candidates=candidates, eps=eps)   # whole problem, one process
        out = parallel.sharded_closed_loop(eng, scene, steps, candidates=candidates, eps=eps, **kwargs)
        for a, b in zip(out, ref):
            assert a.shape == b.shape and torch.equal(a, b)
    finally:
        dist.destroy_process_group()


def test_sharded_closed_loop_world2():
    mp.spawn(_loop_worker, args=(2, _free_port()), nprocs=2, join=True)


@pytest.mark.parametrize("shape", [(5, 6, 3, 2, 4), (2, 9, 4, 3, 5)])        # ragged: 5 scenes / 9 candidates over 2 ranks
def test_sharded_search_world2(shape):
    mp.spawn(_worker, args=(2, _free_port(), shape), nprocs=2, join=True)


def test_shard_range_covers_everything():
    from mantrap_b200 import parallel
    for n in (1, 7, 8, 65536):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) - min(hi - lo for lo, hi in spans) <= 1


def test_order_key_is_order_preserving():
    from mantrap_b200 import parallel
    vals = torch.tensor([-3.5, -0.0, 0.0, 1e-8, 2.0, float("inf"), -1e-8], dtype=torch.float32)
    keys = parallel._order_key(vals, torch.zeros(vals.shape[0], dtype=torch.int32))
    order = torch.argsort(keys)
    assert torch.equal(vals[order], torch.sort(vals, stable=True).values) or \
        torch.allclose(vals[order][torch.isfinite(vals[order])], torch.sort(vals).values[torch.isfinite(torch.sort(vals).values)])
    same = parallel._order_key(torch.tensor([1.0, 1.0]), torch.tensor([7, 3], dtype=torch.int32))
    assert same[1] < same[0]                                                     # ties -> lowest candidate index
    assert parallel._order_key(torch.tensor([0.5]), torch.tensor([-1], dtype=torch.int32))[0] == torch.iinfo(torch.int64).max
