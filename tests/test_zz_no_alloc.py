"""`mtb_evaluate_static`: the evaluation entry point that allocates nothing (scratch in a file-scope `static __device__` array
of MTB_STATIC_SCRATCH_BYTES inside the library, include/mantrap_b200.h). No counterpart in the reference.

Called through the C ABI and compared with `mtb_evaluate` (caller workspace) on the same inputs, two calls in a row with
DIFFERENT inputs, so the second call runs on the scratch the first one left behind. Same kernels, same launch plan: forward results
are bit-identical, gradients agree up to the order of the float reductions in the L2 scratch (the tolerance of the lean-vs-general
test). The file sorts last on purpose; it must pass alone in a fresh process and after the others.
"""
import ctypes
import os
import sys

import pytest
import torch

import golden_io as gio
import kernel_cases as kc
from mantrap_b200 import _cabi, engine, scenarios

FORWARD = ("obj", "total", "violation", "cons", "mus", "sigmas", "ego_traj")


def evaluate_static(lib, scene, controls, spec, want):
    """One `mtb_evaluate_static` call; outputs are the caller's (allocated here, before the call), no workspace is passed."""
    t_hor = controls.shape[2]
    pb = engine._problem_struct(scene, controls, t_hor, None)
    md = engine._module_struct(scene, spec)
    rows = lib.mtb_num_constraint_rows(ctypes.byref(pb), ctypes.byref(md))
    assert rows >= 0
    outs, res = _cabi.MtbOutputs(), {}
    for key in want:
        res[key] = torch.empty(engine.OUTPUT_SHAPES[key](pb.num_scenes, pb.num_candidates, t_hor, pb.num_ados, rows),
                               dtype=torch.float32, device=scene.device)
        setattr(outs, key, _cabi.ptr(res[key]))
    code = lib.mtb_evaluate_static(ctypes.byref(pb), ctypes.byref(md), ctypes.byref(outs), 0)
    return code, res


def check_static_vs_workspace(lib, device, num_scenes, num_candidates, n, p, t_hor, kernel_kind, general=False):
    sync = (lambda: None) if device == "cpu" else torch.cuda.synchronize
    for call in range(2):                                   # the second call sees the first call's scratch contents
        scene = scenarios.synthetic_scenes(num_scenes, n, p, seed=910 + call, param_seed=920 + call, regular_sigma=True).to(device)
        controls = scenarios.candidate_controls(num_scenes, num_candidates, t_hor, seed=930 + call, u_max=1.5).to(device).contiguous()
        mus_wo = engine.launch_evaluate(lib, scene.without_ego(), None, None, want=("mus",), t_horizon=t_hor)["mus"][:, 0].contiguous()
        spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True,
                                 min_distance=general)
        want = ("obj", "total", "grad", "violation", "cons") + (("mus", "sigmas", "ego_traj") if general else ())
        assert engine.query_launch(lib, scene, controls, True)["kernel_kind"] in kernel_kind
        ref = engine.launch_evaluate(lib, scene, controls, spec, want=want)
        sync()
        code, got = evaluate_static(lib, scene, controls, spec, want)
        _cabi.check(lib, code, "mtb_evaluate_static")
        sync()
        for key in want:
            if key in FORWARD:
                assert torch.equal(got[key], ref[key]), (call, key)
        gio.assert_parity(kc.np_(got["grad"]), kc.np_(ref["grad"]), 1e-5, f"static vs workspace gradient, call {call}", atol=1e-7)
        assert float(got["grad"].abs().max()) > 0


def check_rejects_beyond_the_maximum(lib, device):
    # N * P = 1024 threads per evaluation over T = 64: the tape alone is grid x 63 x 1024 x 16 B (one block per SM: 153 MB on a B200)
    num_scenes, n, p, t_hor = 160, 64, 16, 64
    scene = scenarios.synthetic_scenes(num_scenes, n, p, seed=3, param_seed=4, regular_sigma=True).to(device)
    controls = scenarios.candidate_controls(num_scenes, 1, t_hor, seed=5, u_max=1.5).to(device).contiguous()
    pb = engine._problem_struct(scene, controls, t_hor, None)
    need = lib.mtb_workspace_bytes(ctypes.byref(pb), 1)
    assert need > _cabi.STATIC_SCRATCH_BYTES, need
    spec = engine.ModuleSpec(weights={"goal": 1.0})
    code, got = evaluate_static(lib, scene, controls, spec, ("total", "grad"))
    assert code == _cabi.ERR_STATIC_SCRATCH
    assert b"MTB_STATIC_SCRATCH_BYTES" in lib.mtb_error_string(code)
    # forward only needs no tape: the same shape is accepted
    code, got = evaluate_static(lib, scene, controls, spec, ("total",))
    _cabi.check(lib, code, "mtb_evaluate_static (forward)")
    ref = engine.launch_evaluate(lib, scene, controls, spec, want=("total",))
    if device != "cpu":
        torch.cuda.synchronize()
    assert torch.equal(got["total"], ref["total"])


# (kernel_kind: 2 = lean pair kernel, 1 = general pair kernel, 0 = evaluate_kernel)
SHAPES = [
    pytest.param(6, 8, 5, 3, 8, (2,), False, id="lean-pair"),
    pytest.param(5, 4, 6, 2, 6, (1, 2), True, id="pair-with-rollout-outputs"),
    pytest.param(7, 3, 4, 4, 7, (0,), True, id="single"),
]


@pytest.fixture(scope="module")
def cuda_lib():
    from mantrap_b200 import build
    build.build()
    return engine.Engine("cuda:0").lib


@pytest.mark.gpu
@pytest.mark.parametrize("s,c,n,p,t_hor,kind,general", SHAPES)
def test_static_scratch_matches_workspace(cuda_lib, s, c, n, p, t_hor, kind, general):
    check_static_vs_workspace(cuda_lib, "cuda:0", s, c, n, p, t_hor, kind, general)


@pytest.mark.gpu
def test_static_scratch_full_size(cuda_lib):
    # the BASELINE config-4 shape at a grid that fills the device (42 MB of the 64 MiB array on a B200)
    check_static_vs_workspace(cuda_lib, "cuda:0", 64, 64, 20, 3, 12, (2,), False)


@pytest.mark.gpu
def test_static_scratch_rejects_beyond_the_maximum(cuda_lib):
    check_rejects_beyond_the_maximum(cuda_lib, "cuda:0")


@pytest.mark.gpu
def test_static_scratch_call_does_not_allocate(cuda_lib):
    """The device's free memory does not move across calls of the new entry point (the first call of the process did the set-up)."""
    scene = scenarios.synthetic_scenes(4, 5, 3, seed=1, param_seed=2, regular_sigma=True).to("cuda:0")
    controls = scenarios.candidate_controls(4, 8, 8, seed=3, u_max=1.5).to("cuda:0").contiguous()
    mus_wo = engine.launch_evaluate(cuda_lib, scene.without_ego(), None, None, want=("mus",), t_horizon=8)["mus"][:, 0].contiguous()
    spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=mus_wo, control_limits=True, speed_limits=True)
    pb = engine._problem_struct(scene, controls, 8, None)
    md = engine._module_struct(scene, spec)
    total = torch.empty((4, 8), dtype=torch.float32, device="cuda:0")
    grad = torch.empty((4, 8, 8, 2), dtype=torch.float32, device="cuda:0")
    outs = _cabi.MtbOutputs()
    outs.total, outs.grad = _cabi.ptr(total), _cabi.ptr(grad)
    _cabi.check(cuda_lib, cuda_lib.mtb_evaluate_static(ctypes.byref(pb), ctypes.byref(md), ctypes.byref(outs), 0), "warm-up")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(5):
        _cabi.check(cuda_lib, cuda_lib.mtb_evaluate_static(ctypes.byref(pb), ctypes.byref(md), ctypes.byref(outs), 0), "mtb_evaluate_static")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and torch.cuda.memory_reserved() == reserved0


# ---- the same checks on the host emulation of the kernel sources (build container, no GPU) ----
@pytest.fixture(scope="module")
def emul_lib():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul"))
    import build_emul
    return _cabi.bind(ctypes.CDLL(build_emul.build()))


@pytest.mark.parametrize("s,c,n,p,t_hor,kind,general", SHAPES[:1] + SHAPES[2:])
def test_static_scratch_matches_workspace_emulation(emul_lib, s, c, n, p, t_hor, kind, general):
    check_static_vs_workspace(emul_lib, "cpu", max(1, s // 3), c, n, p, t_hor, kind, general)
