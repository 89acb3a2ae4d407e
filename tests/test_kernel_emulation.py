"""CPU tests of the ACTUAL kernel source: `mantrap_b200/csrc/mtb_kernels.cu` compiled with g++ against the host
emulation shim (tests/host_emul/cuda_shim.h; every CUDA thread = one OS thread) and driven through the same
C ABI as the CUDA library. Catches indexing / synchronisation / adjoint bugs without a GPU; the GPU run of the same
checks is tests/test_gpu_parity.py."""
import ctypes
import os
import sys

import pytest
import torch

import golden_io as gio
import kernel_cases as kc
from mantrap_b200 import _cabi, engine

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_emul"))
import build_emul  # noqa: E402


@pytest.fixture(scope="module")
def emul_lib():
    return _cabi.bind(ctypes.CDLL(build_emul.build()))


def test_exports(emul_lib):
    assert emul_lib.mtb_abi_version() == _cabi.ABI_VERSION


@pytest.mark.parametrize("name", gio.case_names())
def test_rollout(emul_lib, name):
    kc.check_rollout(emul_lib, "cpu", name)


@pytest.mark.parametrize("name", gio.case_names())
def test_modules(emul_lib, name):
    kc.check_modules(emul_lib, "cpu", name)


@pytest.mark.parametrize("name", gio.case_names())
def test_constraints(emul_lib, name):
    kc.check_constraints(emul_lib, "cpu", name)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor,weights", [
    ("sf", "double", 3, 5, 4, 3, 6, {"goal": 0.5, "acc": 1.0}),            # sub-warp groups, partial last item
    ("sf", "single", 2, 3, 12, 3, 5, {"goal": 0.5, "vel": 1.0, "pos": 1.0}),  # two warps per evaluation
    ("pf", "double", 2, 4, 5, 5, 7, {"goal": 0.5, "acc": 1.0, "vel": 1.0}),
    ("kalman", "double", 2, 3, 5, 1, 6, {"goal": 0.5, "acc": 1.0}),
    ("sf", "double", 2, 3, 2, 2, 9, {"goal": 0.5, "acc": 1.0}),            # groups of 4 lanes, horizon > group: serial fold
    ("sf", "single", 1, 5, 1, 2, 5, {"goal": 0.5, "acc": 1.0, "vel": 1.0}),  # groups of 2 lanes, one ado (no pair terms)
    ("sf", "double", 2, 3, 5, 4, 6, {"goal": 0.5, "prob": 1.0, "acc": 1.0}),  # probability objective (variance adjoint)
    ("kalman", "double", 1, 2, 4, 1, 7, {"goal": 0.5, "prob": 1.0}),
])
def test_batched_vs_oracle(emul_lib, env_kind, ego_kind, s, c, n, p, t_hor, weights):
    kc.check_batched_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, weights)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor,weights", [
    ("sf", "double", 3, 4, 4, 3, 6, {"goal": 0.5, "acc": 1.0}),            # sub-warp groups (16 lanes), two groups per warp
    ("sf", "single", 2, 2, 12, 3, 5, {"goal": 0.5, "vel": 1.0, "pos": 1.0}),  # two warps per group
    ("pf", "double", 2, 4, 5, 5, 7, {"goal": 0.5, "acc": 1.0, "vel": 1.0}),
    ("sf", "double", 2, 2, 2, 2, 9, {"goal": 0.5, "acc": 1.0}),            # groups of 4 lanes, horizon > group: serial fold
    ("sf", "double", 1, 6, 5, 4, 6, {"goal": 0.5, "prob": 1.0, "acc": 1.0}),  # probability objective (variance adjoint)
    ("pf", "single", 1, 2, 3, 2, 4, {"goal": 0.5, "prob": 1.0}),
])
def test_batched_vs_oracle_pair_kernel(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor, weights):
    """Even candidate counts run the two-candidates-per-thread kernel (potential field: only with the development knob)."""
    monkeypatch.setenv("MTB_PAIR", "2")
    kc.check_batched_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, weights)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor,weights,kw", [
    ("sf", "double", 2, 4, 4, 3, 6, {"goal": 0.5, "acc": 1.0}, {}),
    ("sf", "double", 1, 2, 6, 3, 5, {"goal": 0.5, "acc": 1.0, "vel": 1.0, "pos": 1.0, "prob": 1.0}, {"mask": True}),
    ("sf", "single", 1, 2, 12, 3, 4, {"goal": 0.5, "acc": 1.0, "prob": 1.0}, {"given_traj": True}),
    ("pf", "double", 2, 2, 5, 4, 6, {"goal": 0.5, "acc": 1.0, "prob": 1.0}, {}),
    ("sf", "double", 1, 2, 34, 2, 3, {"goal": 0.5, "acc": 1.0}, {"constraints": False}),      # N > 32: two mask words
    ("sf", "double", 3, 2, 1, 2, 5, {"goal": 0.5, "acc": 1.0, "vel": 1.0}, {"noise": False}),  # one ado: groups of 2 lanes
    ("sf", "double", 1, 2, 3, 2, 1, {"goal": 0.5}, {}),                                        # T = 1
])
def test_pair_kernel_vs_single_kernel(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor, weights, kw):
    monkeypatch.setenv("MTB_PAIR", "2")
    kc.check_pair_vs_single(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, weights, **kw)


@pytest.mark.parametrize("s,c,n,p,t_hor", [(2, 4, 4, 3, 6), (1, 2, 12, 3, 5), (3, 2, 2, 2, 9)])
def test_lean_vs_general_pair_kernel(emul_lib, s, c, n, p, t_hor):
    kc.check_lean_vs_general(emul_lib, "cpu", s, c, n, p, t_hor)


@pytest.mark.parametrize("name", gio.case_names())
def test_jacobian_golden(emul_lib, name):
    kc.check_jacobian_golden(emul_lib, "cpu", name)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [
    ("sf", "double", 2, 2, 4, 3, 6),
    ("sf", "single", 1, 2, 5, 2, 5),
    ("pf", "double", 1, 2, 5, 3, 7),
    ("kalman", "double", 1, 1, 3, 1, 5),
    ("sf", "double", 1, 1, 1, 2, 1),            # T = 1: no particle step, one ado
    ("pf", "single", 2, 2, 1, 2, 1),
    ("sf", "double", 1, 1, 34, 2, 3),           # N > 32: two interaction mask words
])
def test_jacobian_vs_oracle(emul_lib, env_kind, ego_kind, s, c, n, p, t_hor):
    kc.check_jacobian_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [
    ("sf", "double", 2, 2, 4, 3, 6),
    ("sf", "single", 1, 2, 5, 2, 5),
    ("sf", "double", 1, 2, 10, 4, 9),           # several live control steps per particle, neighbours in most steps
    ("pf", "double", 1, 2, 5, 3, 7),
    ("kalman", "double", 1, 1, 3, 1, 5),
    ("sf", "double", 1, 1, 34, 2, 3),           # N > 32: two interaction mask words
])
def test_jacobian_block_vs_chunked(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor):
    kc.check_jacobian_block_vs_chunked(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, setenv=monkeypatch.setenv)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [("sf", "double", 2, 2, 4, 3, 6), ("sf", "double", 1, 2, 10, 4, 9), ("pf", "single", 1, 2, 5, 3, 7)])
def test_jacobian_block_one_ring_slot(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor):
    """The block kernel with one ring slot (shapes where a second slot would cost a resident block per SM)."""
    kc.check_jacobian_block_vs_chunked(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, setenv=monkeypatch.setenv, ring=1)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [
    ("sf", "double", 2, 2, 4, 3, 6),
    ("pf", "double", 1, 2, 5, 3, 7),
    ("sf", "double", 1, 1, 34, 2, 3),
])
def test_jacobian_vs_oracle_chunked_kernel(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor):
    """jacobian_kernel (the fall-back for shapes whose tangents do not fit into shared memory) on its own against the oracle."""
    monkeypatch.setenv("MTB_JAC", "1")
    kc.check_jacobian_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [
    ("sf", "double", 2, 2, 4, 3, 6),
    ("sf", "single", 1, 2, 12, 3, 5),
    ("pf", "double", 1, 3, 5, 5, 7),
    ("kalman", "single", 1, 1, 3, 1, 5),
])
def test_seeded_adjoint(emul_lib, env_kind, ego_kind, s, c, n, p, t_hor):
    kc.check_seeded_adjoint(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, seed=3)


@pytest.mark.parametrize("name", gio.case_names())
def test_ellipsoid_golden(emul_lib, name):
    kc.check_ellipsoid_golden(emul_lib, "cpu", name)


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor", [
    ("sf", "double", 2, 2, 4, 3, 6),
    ("pf", "single", 1, 3, 5, 2, 5),
    ("kalman", "double", 1, 2, 3, 1, 5),
])
def test_ellipsoid_vs_oracle(emul_lib, env_kind, ego_kind, s, c, n, p, t_hor):
    kc.check_ellipsoid_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor)


def test_attention(emul_lib):
    kc.check_attention_golden(emul_lib, "cpu")


def test_closed_loop_encoding_warm_start(emul_lib):
    kc.check_closed_loop_encoding_warm_start(emul_lib, "cpu")


def test_warm_start_encoding(emul_lib):
    kc.check_warm_start_encoding(emul_lib, "cpu")


def test_env_step(emul_lib):
    kc.check_env_step_golden(emul_lib, "cpu")


@pytest.mark.parametrize("env_kind,ego_kind,attention", [("sf", "double", None), ("pf", "single", "euclidean"),
                                                        ("sf", "double", "closest")])
def test_closed_loop(emul_lib, env_kind, ego_kind, attention):
    kc.check_closed_loop(emul_lib, "cpu", env_kind=env_kind, ego_kind=ego_kind, attention=attention)


def test_argmin(emul_lib):
    total = torch.tensor([[3.0, 1.0, 2.0, 0.5], [4.0, -1.0, 5.0, 6.0], [1.0, 2.0, 3.0, 4.0]])
    vio = torch.tensor([[0.0, 0.0, 0.0, 1.0], [0.0, 0.5, 0.0, 0.0], [1.0, 1.0, 1.0, 1.0]])
    controls = torch.arange(3 * 4 * 2 * 2, dtype=torch.float32).reshape(3, 4, 2, 2).contiguous()
    idx, val, best = engine.launch_argmin(emul_lib, total, vio, controls, 1e-3)
    assert idx.tolist() == [1, 0, -1]
    assert val[:2].tolist() == [1.0, 4.0] and torch.isinf(val[2])
    assert torch.equal(best[0], controls[0, 1]) and torch.equal(best[1], controls[1, 0])


@pytest.mark.parametrize("env_kind,ego_kind,s,c,n,p,t_hor,weights", [
    ("sf", "double", 2, 3, 4, 3, 6, {"goal": 0.5, "acc": 1.0}),
    ("sf", "single", 1, 3, 12, 3, 5, {"goal": 0.5, "vel": 1.0, "pos": 1.0}),
    ("sf", "double", 2, 3, 5, 4, 6, {"goal": 0.5, "prob": 1.0, "acc": 1.0}),
    ("pf", "double", 2, 2, 5, 5, 7, {"goal": 0.5, "acc": 1.0}),
])
def test_means_in_l2_scratch(emul_lib, monkeypatch, env_kind, ego_kind, s, c, n, p, t_hor, weights):
    """The alternative placement of the collapsed means (L2-resident scratch instead of shared memory, chosen by the launch plan
    for long horizons / many ados) gives the same results: forced through the development knob."""
    monkeypatch.setenv("MTB_MEANS_GLOBAL", "1")
    kc.check_batched_vs_oracle(emul_lib, "cpu", env_kind, ego_kind, s, c, n, p, t_hor, weights)
    kc.check_modules(emul_lib, "cpu", "c4_sf_n20_p3_t12")
    kc.check_constraints(emul_lib, "cpu", "c3_sf_n10_p4_t20")


def test_jacobian_sampled_sigma_nonpositive(emul_lib):
    kc.check_jacobian_sampled_sigma_nonpositive(emul_lib, "cpu")
