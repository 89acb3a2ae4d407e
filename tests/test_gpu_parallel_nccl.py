"""GPU run (-m gpu) of the sharded paths over NCCL: one process per GPU (world_size 2) on the CUDA engine. The same
assertions as tests/test_parallel_gloo.py (which covers the host logic on the CPU): every rank ends with exactly the
winners / trajectories of a single-GPU run over the whole problem. Skipped on a box with fewer than two GPUs."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port):
    for path in (ROOT, HERE):
        if path not in sys.path:
            sys.path.insert(0, path)
    from mantrap_b200 import engine, mpc, parallel, scenarios
    torch.cuda.set_device(rank)
    dev = f"cuda:{rank}"
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device(dev))
    try:
        eng = engine.Engine(dev)
        # ---- sharded search, scene- and candidate-partitioned (ragged: 37 scenes / 19 candidates over 2 ranks)
        s, c, n, p, t_hor = 37, 19, 6, 3, 8
        scene = scenarios.synthetic_scenes(s, n, p, seed=11, param_seed=12, regular_sigma=True).to(dev)
        controls = scenarios.candidate_controls(s, c, t_hor, seed=13, u_max=1.0).to(dev)

        def make_spec(shard):
            wo = eng.rollout_wo_ego(shard, t_hor)["mus"].contiguous()
            return engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=wo, control_limits=True, speed_limits=True)

        # the single-GPU reference asks for the same outputs as sharded_search (total, violation, grad): the gradient and the
        # forward-only instantiation of a kernel are different binaries, whose forward values may differ in the last bit on the
        # GPU (ptxas contracts them independently under --use_fast_math)
        full = eng.evaluate(scene, controls, make_spec(scene), want=("total", "violation", "grad"))
        idx0, val0, best0 = eng.argmin(full["total"], full["violation"], controls, 1e-3)
        assert int((idx0 >= 0).sum()) > 0
        for mode in ("scenes", "candidates"):
            val, idx, best, local = parallel.sharded_search(eng, scene, controls, make_spec, 1e-3, mode=mode)
            assert torch.equal(idx, idx0), (mode, idx, idx0)
            assert torch.equal(best, best0), mode
            if mode == "scenes":
                assert torch.equal(val, val0), mode
            else:
                # 19 candidates over 2 ranks = shards of 10 (even: evaluate_pair_kernel) and 9 (odd: evaluate_kernel), the whole
                # batch (odd) runs evaluate_kernel. Under --use_fast_math the scalar and the packed expression trees are contracted
                # differently, so on the GPU the two kernels agree to a few ulp per step (bit-identical on the host emulation,
                # tests/test_parallel_gloo.py): same winners, values within the objective tolerance.
                torch.testing.assert_close(val, val0, rtol=1e-5, atol=1e-6)
        # ---- sharded closed loop (5 scenes over 2 ranks), given candidates and step noise
        s, c, n, p, t_hor, steps = 5, 16, 4, 3, 6, 3
        scene = scenarios.synthetic_scenes(s, n, p, seed=31, param_seed=32, regular_sigma=True)
        gen = torch.Generator().manual_seed(21)
        candidates = [(torch.rand((s, c, t_hor, 2), generator=gen) * 2 - 1) * 2.0 for _ in range(steps)]
        eps = [torch.randn((s, n, 2), generator=gen) for _ in range(steps)]
        kwargs = dict(t_planning=t_hor, num_candidates=c, attention="euclidean")
        ref = mpc.BatchedClosedLoop(eng, scene, **kwargs).solve(steps, candidates=candidates, eps=eps)
        out = parallel.sharded_closed_loop(eng, scene, steps, candidates=candidates, eps=eps, **kwargs)
        for a, b in zip(out, ref):
            assert a.shape == b.shape and torch.equal(a, b)
        torch.cuda.synchronize()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_sharded_paths_nccl_world2():
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)
