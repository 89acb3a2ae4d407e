"""Helpers shared by the CPU (oracle) and GPU (CUDA path) parity tests: golden loading and the parity metric."""
import glob
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerances: 1e-5 relative on rollouts / objectives, 1e-4 on gradients / Jacobians, measured per
# tensor as max|a-b| <= rtol * max|ref| + atol (SURVEY §8d "Parity gate", atol for ~1e-4-sized gradients).
RTOL_ROLLOUT = 1e-5
RTOL_GRAD = 1e-4
ATOL = 1e-7
# Interaction objectives are means of |differences of O(1) m/s means| / dt: one float32 ulp of a mean (1.2e-7)
# already moves them by 3e-7 absolute, so objective VALUES get this absolute floor next to the 1e-5 relative bound
# (the rollouts they are computed from are themselves only required to agree to 1e-5 * max|mu| ~ 1e-5 absolute).
ATOL_OBJ = 1e-6


def case_names():
    names = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return [n for n in names if n not in ("agent_kats", "scene_ops")]


def load(name):
    with np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"), allow_pickle=False) as data:
        return {k: data[k] for k in data.files}


def parity_error(a, ref):
    """Return (max|a-ref|, max|ref|)."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    if a.size == 0:
        return 0.0, 0.0
    return float(np.max(np.abs(a - ref))), float(np.max(np.abs(ref)))


def assert_parity(a, ref, rtol, what, atol=ATOL):
    err, scale = parity_error(a, ref)
    assert err <= rtol * scale + atol, f"{what}: max|diff| {err:.3e} > {rtol:g} * {scale:.3e} + {atol:g}"
    return err / max(scale, 1e-30)


def t(x):
    return torch.as_tensor(np.asarray(x), dtype=torch.float32)
