"""ctypes mirror of `include/mantrap_b200.h` (the C-ABI boundary of the hot path).

Only struct definitions, prototype binding and pointer plumbing live here; there is no arithmetic. The functions
take torch tensors purely as owners of memory (`data_ptr()`); which library is bound decides where the work runs.
The product binds `mantrap_b200/libmantrap_b200.so` (CUDA, see `_lib.py`).
"""
import ctypes
import typing

import torch

ENV_SOCIAL_FORCES, ENV_POTENTIAL_FIELD, ENV_KALMAN = 0, 1, 2
EGO_DOUBLE_INTEGRATOR, EGO_SINGLE_INTEGRATOR, EGO_NONE = 0, 1, 2
MAX_ADOS, MAX_HORIZON = 64, 64
ABI_VERSION = 7
STATIC_SCRATCH_BYTES = 64 << 20     # MTB_STATIC_SCRATCH_BYTES (mtb_evaluate_static)
ERR_STATIC_SCRATCH = -4            # MTB_ERR_STATIC_SCRATCH
ATTENTION_EUCLIDEAN, ATTENTION_CLOSEST = 0, 1

_f, _i, _p = ctypes.c_float, ctypes.c_int32, ctypes.c_void_p


class MtbProblem(ctypes.Structure):
    _fields_ = [("num_scenes", _i), ("num_candidates", _i), ("t_horizon", _i), ("num_ados", _i),
                ("num_particles", _i), ("env_kind", _i), ("ego_kind", _i),
                ("dt", _f), ("sf_tau", _f), ("ped_speed_max", _f), ("sf_goal_radius", _f),
                ("sf_interaction_radius", _f), ("pf_theta_max", _f), ("particle_noise", _f), ("var_initial", _f),
                ("kalman_noise", _f),
                ("ego0", _p), ("ado0", _p), ("ado_goal", _p), ("v0", _p), ("sigma", _p), ("weight", _p),
                ("noise", _p), ("controls", _p), ("ego_traj", _p)]


class MtbModules(ctypes.Structure):
    _fields_ = [("w_goal", _f), ("w_acc", _f), ("w_vel", _f), ("w_pos", _f),
                ("max_acc", _f), ("max_vel", _f), ("max_pos", _f),
                ("ego_goal", _p), ("mus_wo", _p), ("ado_mask", _p),
                ("con_control", _i), ("con_speed", _i), ("con_min_distance", _i),
                ("u_max", _f), ("v_max", _f), ("min_distance", _f), ("violation_eps", _f),
                ("seed_mus", _p), ("seed_sigmas", _p), ("w_prob", _f), ("max_prob", _f),
                ("con_ellipsoid", _i), ("ellipsoid_min_scale", _f)]


class MtbOutputs(ctypes.Structure):
    _fields_ = [("obj", _p), ("obj_raw", _p), ("total", _p), ("grad", _p), ("grad_traj", _p), ("cons", _p), ("violation", _p),
                ("ego_traj", _p), ("mus", _p), ("sigmas", _p), ("stats", _p)]


class MtbLaunchInfo(ctypes.Structure):
    _fields_ = [("grid", _i), ("block", _i), ("evals_per_block", _i), ("group_size", _i), ("smem_bytes", _i),
                ("tape_in_smem", _i), ("num_items", _i), ("blocks_per_sm", _i), ("variant_maxt", _i), ("variant_minb", _i),
                ("kernel_kind", _i), ("evals_per_thread", _i)]


EXPORTED_SYMBOLS = ("mtb_abi_version", "mtb_error_string", "mtb_num_constraint_rows", "mtb_workspace_bytes",
                    "mtb_evaluate", "mtb_evaluate_static", "mtb_query_launch", "mtb_constraint_jacobian", "mtb_jacobian_workspace_bytes",
                    "mtb_argmin", "mtb_attention", "mtb_warm_start_encoding", "mtb_env_step", "mtb_probe_launch", "mtb_ego_unroll", "mtb_ego_roll", "mtb_ego_unroll_adjoint",
                    "mtb_ego_dx_du")


def bind(lib: ctypes.CDLL) -> ctypes.CDLL:
    """Attach prototypes to a loaded library and check every symbol the header declares is exported."""
    missing = [name for name in EXPORTED_SYMBOLS if not hasattr(lib, name)]
    if missing:
        raise RuntimeError(f"mantrap_b200 library does not export {missing}")
    pp, pm, po = ctypes.POINTER(MtbProblem), ctypes.POINTER(MtbModules), ctypes.POINTER(MtbOutputs)
    lib.mtb_abi_version.restype = ctypes.c_int
    lib.mtb_error_string.restype, lib.mtb_error_string.argtypes = ctypes.c_char_p, [ctypes.c_int]
    lib.mtb_num_constraint_rows.restype, lib.mtb_num_constraint_rows.argtypes = ctypes.c_int, [pp, pm]
    lib.mtb_workspace_bytes.restype, lib.mtb_workspace_bytes.argtypes = ctypes.c_size_t, [pp, ctypes.c_int]
    lib.mtb_evaluate.restype = ctypes.c_int
    lib.mtb_evaluate.argtypes = [pp, pm, po, _p, ctypes.c_size_t, _p]
    lib.mtb_evaluate_static.restype = ctypes.c_int
    lib.mtb_evaluate_static.argtypes = [pp, pm, po, _p]
    lib.mtb_query_launch.restype = ctypes.c_int
    lib.mtb_query_launch.argtypes = [pp, ctypes.c_int, ctypes.POINTER(MtbLaunchInfo)]
    lib.mtb_constraint_jacobian.restype = ctypes.c_int
    lib.mtb_constraint_jacobian.argtypes = [pp, pm, _p, _p, _p, ctypes.c_size_t, _p]
    lib.mtb_jacobian_workspace_bytes.restype, lib.mtb_jacobian_workspace_bytes.argtypes = ctypes.c_size_t, [pp]
    lib.mtb_probe_launch.restype = ctypes.c_int
    lib.mtb_probe_launch.argtypes = [ctypes.c_int, _p, ctypes.c_int, ctypes.c_int, _p]
    lib.mtb_ego_unroll.restype, lib.mtb_ego_unroll.argtypes = ctypes.c_int, [ctypes.c_int, _f, _p, _p, _i, _i, _p, _p]
    lib.mtb_ego_roll.restype, lib.mtb_ego_roll.argtypes = ctypes.c_int, [ctypes.c_int, _f, _p, _i, _i, _p, _p]
    lib.mtb_ego_unroll_adjoint.restype = ctypes.c_int
    lib.mtb_ego_unroll_adjoint.argtypes = [ctypes.c_int, _f, _p, _i, _i, _p, _p]
    lib.mtb_ego_dx_du.restype, lib.mtb_ego_dx_du.argtypes = ctypes.c_int, [ctypes.c_int, _f, _i, _p, _p]
    lib.mtb_attention.restype = ctypes.c_int
    lib.mtb_attention.argtypes = [ctypes.c_int, _p, _p, _i, _i, _f, _p, _p]
    lib.mtb_warm_start_encoding.restype = ctypes.c_int
    lib.mtb_warm_start_encoding.argtypes = [_p, _p, _p, _i, _i, _p, _p, _i, _i, _i, _p, _p, _p, _p]
    lib.mtb_env_step.restype = ctypes.c_int
    lib.mtb_env_step.argtypes = [ctypes.c_int, _f, _f, _f, _f, _p, _p, _p, _p, _i, _i, _i, _p, _p]
    lib.mtb_argmin.restype = ctypes.c_int
    lib.mtb_argmin.argtypes = [_p, _p, _i, _i, _f, _p, _i, _p, _p, _p, _p]
    return lib


def ptr(t: typing.Optional[torch.Tensor], dtype=torch.float32) -> typing.Optional[int]:
    """Raw address of a contiguous tensor (None stays NULL)."""
    if t is None:
        return None
    if t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"expected contiguous {dtype} tensor, got {t.dtype}, contiguous={t.is_contiguous()}")
    return t.data_ptr()


def check(lib: ctypes.CDLL, code: int, what: str):
    if code != 0:
        raise RuntimeError(f"{what} failed with code {code}: {lib.mtb_error_string(code).decode()}")
