"""CPU tests of the boundary: the CUDA library builds for sm_100a, loads, and exports every symbol that
include/mantrap_b200.h declares (no compute calls here: there is no GPU in the build container)."""
import ctypes
import os
import re
import subprocess

import pytest

from mantrap_b200 import _cabi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    return build.build()


def test_header_symbols_exported(lib_path):
    header = open(os.path.join(ROOT, "include", "mantrap_b200.h")).read()
    declared = set(re.findall(r"\b(mtb_[a-z_]+)\s*\(", header))
    assert declared == set(_cabi.EXPORTED_SYMBOLS), declared ^ set(_cabi.EXPORTED_SYMBOLS)
    lib = _cabi.bind(ctypes.CDLL(lib_path))
    assert lib.mtb_abi_version() == _cabi.ABI_VERSION
    assert b"bad argument" in lib.mtb_error_string(-1)


def test_struct_sizes_match_header(lib_path, tmp_path):
    """sizeof of the ctypes mirrors == sizeof of the C structs (compiled with gcc from the public header)."""
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "mantrap_b200.h"\nint main(){printf("%zu %zu %zu %zu\\n",'
                   "sizeof(mtb_problem),sizeof(mtb_modules),sizeof(mtb_outputs),sizeof(mtb_launch_info));return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert sizes == [ctypes.sizeof(_cabi.MtbProblem), ctypes.sizeof(_cabi.MtbModules), ctypes.sizeof(_cabi.MtbOutputs),
                     ctypes.sizeof(_cabi.MtbLaunchInfo)]


def test_sass_is_sm100(lib_path):
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


def test_engine_requires_gpu():
    import torch
    from mantrap_b200 import engine
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        engine.Engine()


def test_argument_validation_returns_error_codes(lib_path):
    """Bad arguments are rejected with MTB_ERR_* before any device work (no GPU needed): the error behaviour of the boundary."""
    lib = _cabi.bind(ctypes.CDLL(lib_path))
    dummy = ctypes.c_void_p(16)          # never dereferenced: validation fails first
    assert lib.mtb_attention(_cabi.ATTENTION_EUCLIDEAN, None, dummy, 1, 1, 4.0, dummy, None) == -1
    assert lib.mtb_attention(7, dummy, dummy, 1, 1, 4.0, dummy, None) == -1
    assert lib.mtb_attention(_cabi.ATTENTION_CLOSEST, dummy, dummy, 1, _cabi.MAX_ADOS + 1, 4.0, dummy, None) == -2
    assert lib.mtb_env_step(_cabi.EGO_DOUBLE_INTEGRATOR, 0.4, 1e-6, 2.0, 2.5, None, dummy, dummy, None, 1, 1, 1, None, None) == -1
    assert lib.mtb_env_step(_cabi.EGO_NONE, 0.4, 1e-6, 2.0, 2.5, dummy, dummy, dummy, None, 1, 1, 1, None, None) == -1
    assert lib.mtb_env_step(_cabi.EGO_DOUBLE_INTEGRATOR, 0.4, 1e-6, 2.0, 2.5, dummy, dummy, dummy, None, 1, 1, 0, None, None) == -1
    pb, md = _cabi.MtbProblem(), _cabi.MtbModules()
    assert lib.mtb_num_constraint_rows(ctypes.byref(pb), ctypes.byref(md)) == -1          # empty problem
    pb.num_scenes = pb.num_candidates = pb.t_horizon = 1
    pb.num_ados, pb.num_particles, pb.ado0 = _cabi.MAX_ADOS + 1, 2, 16
    assert lib.mtb_num_constraint_rows(ctypes.byref(pb), ctypes.byref(md)) == -2          # too many ados
    assert lib.mtb_argmin(None, None, 1, 1, 1e-3, None, 0, None, None, None, None) == -1
    assert lib.mtb_constraint_jacobian(ctypes.byref(pb), ctypes.byref(md), None, None, None, 0, None) == -1
    for code, text in ((-1, b"bad argument"), (-2, b"exceeds"), (-3, b"workspace")):
        assert text in lib.mtb_error_string(code)
