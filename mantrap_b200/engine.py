"""Batched scene data model and the launcher of the fused CUDA hot path.

`SceneBatch` is the HBM layout of S scenes (structure of arrays, float32); `launch_evaluate` fills
the C-ABI structs of `include/mantrap_b200.h` from it and launches `mtb_evaluate` on the current stream. It is
pointer plumbing only. `Engine` is what the product uses: it binds the CUDA library and refuses anything that is
not a CUDA tensor (there is no CPU path).
"""
import contextlib
import ctypes
import dataclasses
import math
import typing

import torch

from . import _cabi, constants

ENV_KINDS = {"sf": _cabi.ENV_SOCIAL_FORCES, "pf": _cabi.ENV_POTENTIAL_FIELD, "kalman": _cabi.ENV_KALMAN}
EGO_KINDS = {"double": _cabi.EGO_DOUBLE_INTEGRATOR, "single": _cabi.EGO_SINGLE_INTEGRATOR, None: _cabi.EGO_NONE}
ATTENTION_KINDS = {"euclidean": _cabi.ATTENTION_EUCLIDEAN, "closest": _cabi.ATTENTION_CLOSEST}
OBJECTIVE_KEYS = ("goal", "acc", "vel", "pos", "prob")     # columns of `obj` / `obj_raw` (MTB_NUM_OBJECTIVES)
NUM_OBJECTIVES = len(OBJECTIVE_KEYS)


PACKED_COPY_MAX_FLOATS = 1 << 16


def to_device_packed(tensors: typing.Sequence[typing.Optional[torch.Tensor]], device) -> typing.List[typing.Optional[torch.Tensor]]:
    """fp32 contiguous copies of `tensors` (None stays None) on `device`. A small all-host set going to a GPU (the one-scene
    batch of a solver callback: eight tensors of a few hundred bytes) is staged in ONE host buffer and moved with ONE
    host->device copy, every tensor a 16-byte aligned view of it: per-copy driver latency, not bandwidth, is the cost there.
    Large sets (batched scenes) are copied one by one, without the extra host pass."""
    device = torch.device(device)
    live = [t for t in tensors if t is not None]
    total = sum(-(-t.numel() // 4) * 4 for t in live)
    if device.type != "cuda" or len(live) < 2 or total > PACKED_COPY_MAX_FLOATS or any(t.device.type != "cpu" for t in live):
        return [None if t is None else t.to(device=device, dtype=torch.float32).contiguous() for t in tensors]
    stage = torch.empty(total, dtype=torch.float32)
    spans, offset = [], 0
    for t in live:
        stage[offset:offset + t.numel()].copy_(t.reshape(-1))
        spans.append((offset, t.numel(), t.shape))
        offset += -(-t.numel() // 4) * 4
    moved = stage.to(device)
    views = iter(moved[o:o + n].view(shape) for o, n, shape in spans)
    return [None if t is None else next(views) for t in tensors]


@dataclasses.dataclass
class SceneBatch:
    """S scenes with N ados and P particles per ado. All tensors float32, contiguous, on one device.

    ego0 [S,4], ego_goal [S,2], ado0 [S,N,4], ado_goal [S,N,2] (social forces), v0 / sigma / weight [S,N,P]
    (one `create_particles` draw per scene, mantrap/environment/base/particle.py:17-79; weights normalised over
    the ado axis, :141), noise [S,T-1,N,P,2] or None.
    """
    env_kind: str
    ego_kind: typing.Optional[str]
    dt: float
    ego0: typing.Optional[torch.Tensor]
    ego_goal: typing.Optional[torch.Tensor]
    ado0: torch.Tensor
    ado_goal: typing.Optional[torch.Tensor] = None
    v0: typing.Optional[torch.Tensor] = None
    sigma: typing.Optional[torch.Tensor] = None
    weight: typing.Optional[torch.Tensor] = None
    noise: typing.Optional[torch.Tensor] = None
    axes_diag: float = math.sqrt(20.0 ** 2 + 20.0 ** 2)

    @property
    def num_scenes(self) -> int:
        return self.ado0.shape[0]

    @property
    def num_ados(self) -> int:
        return self.ado0.shape[1]

    @property
    def num_particles(self) -> int:
        return 1 if self.v0 is None else self.v0.shape[2]

    @property
    def device(self) -> torch.device:
        return self.ado0.device

    _TENSOR_FIELDS = ("ego0", "ego_goal", "ado0", "ado_goal", "v0", "sigma", "weight", "noise")

    def to(self, device) -> "SceneBatch":
        vals = [getattr(self, f) for f in self._TENSOR_FIELDS]
        return dataclasses.replace(self, **dict(zip(self._TENSOR_FIELDS, to_device_packed(vals, device))))

    def without_ego(self) -> "SceneBatch":
        return dataclasses.replace(self, ego_kind=None)


@dataclasses.dataclass
class ModuleSpec:
    """Which objective / constraint modules enter the evaluation, in the reference's conventions:
    `weights[key]` is the module weight (goal_norm.py:37 -> 0.5, interaction modules 1.0); normalisation
    (goal: / environment diagonal, interaction: / clamp maximum) is applied by the launcher."""
    weights: typing.Dict[str, float] = dataclasses.field(default_factory=dict)
    mus_wo: typing.Optional[torch.Tensor] = None          # [S,N,T,4]
    ado_mask: typing.Optional[torch.Tensor] = None        # [S,N] uint8
    control_limits: bool = False
    speed_limits: bool = False
    min_distance: bool = False
    ellipsoid: bool = False
    u_max: float = constants.ROBOT_ACC_MAX
    v_max: float = constants.ROBOT_SPEED_MAX
    d_min: float = constants.CONSTRAINT_MIN_L2_DISTANCE
    seed_mus: typing.Optional[torch.Tensor] = None        # [S,C,N,T,4] external adjoint seeds (autograd backward)
    seed_sigmas: typing.Optional[torch.Tensor] = None     # [S,C,N,T,2]


def _problem_struct(scene: SceneBatch, controls: typing.Optional[torch.Tensor], t_horizon: int,
                    ego_traj: typing.Optional[torch.Tensor] = None) -> _cabi.MtbProblem:
    pb = _cabi.MtbProblem()
    pb.num_scenes = scene.num_scenes
    with_ego = controls is not None or ego_traj is not None
    pb.num_candidates = controls.shape[1] if controls is not None else (ego_traj.shape[1] if ego_traj is not None else 1)
    pb.t_horizon = t_horizon
    pb.num_ados = scene.num_ados
    pb.num_particles = scene.num_particles
    pb.env_kind = ENV_KINDS[scene.env_kind]
    pb.ego_kind = EGO_KINDS[scene.ego_kind if with_ego else None]
    pb.ego_traj = _cabi.ptr(ego_traj)
    pb.dt = scene.dt
    pb.sf_tau = constants.SOCIAL_FORCES_DEFAULT_TAU
    pb.ped_speed_max = constants.PED_SPEED_MAX
    pb.sf_goal_radius = constants.SOCIAL_FORCES_MAX_GOAL_DISTANCE
    pb.sf_interaction_radius = constants.SOCIAL_FORCES_MAX_INTERACTION_DISTANCE
    pb.pf_theta_max = constants.POTENTIAL_FIELD_MAX_THETA / 180.0 * math.pi
    pb.particle_noise = constants.ENV_PARTICLE_NOISE
    pb.var_initial = constants.ENV_VAR_INITIAL
    pb.kalman_noise = constants.KALMAN_ADDITIVE_NOISE
    pb.ego0 = _cabi.ptr(scene.ego0)
    pb.ado0 = _cabi.ptr(scene.ado0)
    pb.ado_goal = _cabi.ptr(scene.ado_goal)
    pb.v0 = _cabi.ptr(scene.v0)
    pb.sigma = _cabi.ptr(scene.sigma)
    pb.weight = _cabi.ptr(scene.weight)
    pb.noise = _cabi.ptr(scene.noise)
    pb.controls = _cabi.ptr(controls)
    return pb


def _module_struct(scene: SceneBatch, spec: ModuleSpec) -> _cabi.MtbModules:
    md = _cabi.MtbModules()
    w = spec.weights
    md.w_goal = w.get("goal", 0.0) / scene.axes_diag
    md.w_acc = w.get("acc", 0.0) / constants.OBJECTIVE_ACC_INTERACT_MAX
    md.w_vel = w.get("vel", 0.0) / constants.OBJECTIVE_VEL_INTERACT_MAX
    md.w_pos = w.get("pos", 0.0) / constants.OBJECTIVE_POS_INTERACT_MAX
    md.max_acc = constants.OBJECTIVE_ACC_INTERACT_MAX
    md.max_vel = constants.OBJECTIVE_VEL_INTERACT_MAX
    md.max_pos = constants.OBJECTIVE_POS_INTERACT_MAX
    md.w_prob = w.get("prob", 0.0) / constants.OBJECTIVE_PROB_INTERACT_MAX
    md.max_prob = constants.OBJECTIVE_PROB_INTERACT_MAX
    md.ego_goal = _cabi.ptr(scene.ego_goal)
    md.mus_wo = _cabi.ptr(spec.mus_wo)
    md.ado_mask = _cabi.ptr(spec.ado_mask, dtype=torch.uint8)
    md.con_control = int(spec.control_limits)
    md.con_speed = int(spec.speed_limits)
    md.con_min_distance = int(spec.min_distance)
    md.con_ellipsoid = int(spec.ellipsoid)
    md.ellipsoid_min_scale = constants.CONSTRAINT_ELLIPSOID_MIN_SCALE
    md.u_max, md.v_max, md.min_distance = spec.u_max, spec.v_max, spec.d_min
    md.violation_eps = constants.CONSTRAINT_VIOLATION_PRECISION
    md.seed_mus = _cabi.ptr(spec.seed_mus)
    md.seed_sigmas = _cabi.ptr(spec.seed_sigmas)
    return md


OUTPUT_SHAPES = {
    "obj": lambda s, c, t, n, r: (s, c, NUM_OBJECTIVES), "obj_raw": lambda s, c, t, n, r: (s, c, NUM_OBJECTIVES),
    "total": lambda s, c, t, n, r: (s, c), "grad": lambda s, c, t, n, r: (s, c, t, 2),
    "grad_traj": lambda s, c, t, n, r: (s, c, t + 1, 4),
    "cons": lambda s, c, t, n, r: (s, c, r), "violation": lambda s, c, t, n, r: (s, c),
    "ego_traj": lambda s, c, t, n, r: (s, c, t + 1, 4), "mus": lambda s, c, t, n, r: (s, c, n, t, 4),
    "sigmas": lambda s, c, t, n, r: (s, c, n, t, 2),
}


def launch_evaluate(lib: ctypes.CDLL, scene: SceneBatch, controls: typing.Optional[torch.Tensor],
                    spec: typing.Optional[ModuleSpec], want: typing.Sequence[str], t_horizon: int = None,
                    stream: int = 0, out: typing.Dict[str, torch.Tensor] = None, stats: torch.Tensor = None,
                    workspace: torch.Tensor = None, ego_traj: torch.Tensor = None) -> typing.Dict[str, torch.Tensor]:
    """One `mtb_evaluate` launch. controls [S,C,T,2] (None = rollout without ego, needs `t_horizon`), or a given ego
    trajectory `ego_traj` [S,C,T+1,4] (module interface). `want` names the output arrays to produce (keys of
    OUTPUT_SHAPES); `out` may supply pre-allocated tensors."""
    spec = spec if spec is not None else ModuleSpec()
    if controls is not None:
        t_horizon = controls.shape[2]
    elif ego_traj is not None:
        t_horizon = ego_traj.shape[2] - 1
    pb = _problem_struct(scene, controls, t_horizon, ego_traj)
    md = _module_struct(scene, spec)
    rows = lib.mtb_num_constraint_rows(ctypes.byref(pb), ctypes.byref(md))
    if rows < 0:
        _cabi.check(lib, rows, "mtb_num_constraint_rows")
    s, c, n = pb.num_scenes, pb.num_candidates, pb.num_ados
    res = {} if out is None else dict(out)
    outs = _cabi.MtbOutputs()
    for key in want:
        if key not in res:
            res[key] = torch.empty(OUTPUT_SHAPES[key](s, c, t_horizon, n, rows), dtype=torch.float32, device=scene.device)
        setattr(outs, key, _cabi.ptr(res[key]))
    if stats is not None:
        outs.stats = _cabi.ptr(stats, dtype=torch.int64)
    need = lib.mtb_workspace_bytes(ctypes.byref(pb), int("grad" in want or "grad_traj" in want))
    ws_ptr, ws_bytes = None, 0
    if need > 0:
        if workspace is None or workspace.numel() * workspace.element_size() < need:
            workspace = torch.empty(need, dtype=torch.uint8, device=scene.device)
        res["_workspace"] = workspace
        ws_ptr, ws_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    code = lib.mtb_evaluate(ctypes.byref(pb), ctypes.byref(md), ctypes.byref(outs), ws_ptr, ws_bytes, stream)
    _cabi.check(lib, code, "mtb_evaluate")
    return res


def launch_jacobian(lib: ctypes.CDLL, scene: SceneBatch, controls: typing.Optional[torch.Tensor], spec: ModuleSpec,
                    stream: int = 0, want_values: bool = False, ego_traj: torch.Tensor = None):
    """One `mtb_constraint_jacobian` launch: dense UN-normalised Jacobian [S,C,R,2T] of every constraint row enabled
    in `spec` (control | speed | min-distance | ellipsoid, the row order of `cons`) w.r.t. the flattened controls; optionally the
    un-normalised min-distance values [S,C,N,T]."""
    t_horizon = controls.shape[2] if controls is not None else ego_traj.shape[2] - 1
    pb = _problem_struct(scene, controls, t_horizon, ego_traj)
    md = _module_struct(scene, spec)
    rows = lib.mtb_num_constraint_rows(ctypes.byref(pb), ctypes.byref(md))
    if rows <= 0:
        _cabi.check(lib, rows if rows < 0 else -1, "mtb_num_constraint_rows")
    s, c, n = pb.num_scenes, pb.num_candidates, pb.num_ados
    jac = torch.empty((s, c, rows, 2 * t_horizon), dtype=torch.float32, device=scene.device)
    values = torch.empty((s, c, n, t_horizon), dtype=torch.float32, device=scene.device) \
        if (want_values and spec.min_distance) else None
    code = lib.mtb_constraint_jacobian(ctypes.byref(pb), ctypes.byref(md), _cabi.ptr(values), _cabi.ptr(jac), None, 0, stream)
    _cabi.check(lib, code, "mtb_constraint_jacobian")
    return (jac, values) if want_values else jac


def launch_attention(lib: ctypes.CDLL, kind: str, ego0: torch.Tensor, ado0: torch.Tensor, radius: float, stream: int = 0
                     ) -> torch.Tensor:
    """`mtb_attention`: mask [S,N] uint8 of the ados an attention filter keeps (mantrap/attention)."""
    s, n = ado0.shape[0], ado0.shape[1]
    mask = torch.empty((s, n), dtype=torch.uint8, device=ado0.device)
    code = lib.mtb_attention(ATTENTION_KINDS[kind], _cabi.ptr(ego0), _cabi.ptr(ado0), s, n, radius,
                             _cabi.ptr(mask, dtype=torch.uint8), stream)
    _cabi.check(lib, code, "mtb_attention")
    return mask


def launch_warm_start_encoding(lib: ctypes.CDLL, ego0: torch.Tensor, ego_goal: torch.Tensor, ado0: torch.Tensor,
                               encoding_db: torch.Tensor, solution_db: typing.Optional[torch.Tensor], t_horizon: int, stream: int = 0):
    """`mtb_warm_start_encoding`: (encoding [S,4], match [S] int32, warm [S,T,2] or None without a solution database)."""
    s, n, dev = ado0.shape[0], ado0.shape[1], ado0.device
    encoding = torch.empty((s, 4), dtype=torch.float32, device=dev)
    match = torch.empty(s, dtype=torch.int32, device=dev)
    warm = torch.empty((s, t_horizon, 2), dtype=torch.float32, device=dev) if solution_db is not None else None
    code = lib.mtb_warm_start_encoding(_cabi.ptr(ego0), _cabi.ptr(ego_goal), _cabi.ptr(ado0), s, n, _cabi.ptr(encoding_db),
                                       _cabi.ptr(solution_db), encoding_db.shape[0], 0 if solution_db is None else solution_db.shape[1],
                                       t_horizon, _cabi.ptr(encoding), _cabi.ptr(match, dtype=torch.int32), _cabi.ptr(warm), stream)
    _cabi.check(lib, code, "mtb_warm_start_encoding")
    return encoding, match, warm


def launch_env_step(lib: ctypes.CDLL, scene: SceneBatch, controls: torch.Tensor, eps: typing.Optional[torch.Tensor],
                    warm: typing.Optional[torch.Tensor] = None, stream: int = 0):
    """`mtb_env_step`: advance `scene.ego0` / `scene.ado0` IN PLACE by one closed-loop step with controls[:, 0]
    (controls [S,T,2]); eps [S,N,2] standard-normal draws or None (mean step); `warm` [S,T,2] receives the shifted
    warm start."""
    s, t_horizon = controls.shape[0], controls.shape[1]
    ego_u_max = constants.ROBOT_ACC_MAX if scene.ego_kind == "double" else constants.ROBOT_SPEED_MAX
    code = lib.mtb_env_step(EGO_KINDS[scene.ego_kind], scene.dt, constants.ENV_VAR_INITIAL, ego_u_max,
                            constants.PED_SPEED_MAX, _cabi.ptr(scene.ego0), _cabi.ptr(scene.ado0), _cabi.ptr(controls),
                            _cabi.ptr(eps), s, scene.num_ados, t_horizon, _cabi.ptr(warm), stream)
    _cabi.check(lib, code, "mtb_env_step")


def query_launch(lib: ctypes.CDLL, scene: SceneBatch, controls: torch.Tensor, want_grad: bool) -> typing.Dict[str, int]:
    pb = _problem_struct(scene, controls, controls.shape[2])
    info = _cabi.MtbLaunchInfo()
    _cabi.check(lib, lib.mtb_query_launch(ctypes.byref(pb), int(want_grad), ctypes.byref(info)), "mtb_query_launch")
    return {name: getattr(info, name) for name, _ in info._fields_}


def launch_argmin(lib: ctypes.CDLL, total: torch.Tensor, violation: typing.Optional[torch.Tensor],
                  controls: typing.Optional[torch.Tensor], feas_tol: float, stream: int = 0):
    """Per scene: best feasible candidate (index -1 when none is feasible)."""
    s, c = total.shape
    best_idx = torch.empty(s, dtype=torch.int32, device=total.device)
    best_val = torch.empty(s, dtype=torch.float32, device=total.device)
    t_horizon = 0 if controls is None else controls.shape[2]
    best_controls = None if controls is None else torch.empty((s, t_horizon, 2), dtype=torch.float32, device=total.device)
    code = lib.mtb_argmin(_cabi.ptr(total), _cabi.ptr(violation), s, c, feas_tol, _cabi.ptr(controls), t_horizon,
                          _cabi.ptr(best_idx, dtype=torch.int32), _cabi.ptr(best_val), _cabi.ptr(best_controls), stream)
    _cabi.check(lib, code, "mtb_argmin")
    return best_idx, best_val, best_controls


class PreparedEvaluation:
    """A solver callback as a PRE-BUILT launch (the per-callback host work of `launch_evaluate` done once): the C-ABI structs,
    the output buffers, the kernel scratch and a device copy of the static scene (states, goals, ego-free baseline) are created
    when the solver is set up; a call only writes what changes between callbacks -- K candidate control sequences and the fresh
    particle draw (`create_particles`, particle.py:40-79, + the per-step noise, :165) -- into ONE pinned host block, copies it
    with ONE host->device transfer, launches `mtb_evaluate` (and `mtb_constraint_jacobian` on request) and reads every result
    back with ONE device->host transfer. This is what `TrajOptSolver.objective / gradient / constraints / jacobian`
    (solver/base/trajopt.py:368-434, solver/ipopt.py:118-154) cost per IPOPT iteration.

    `scene` holds the static fields on the host (ego0, ado0, ado_goal, ego_goal); `spec.mus_wo` / `spec.ado_mask` are device tensors.
    """
    RESULT_KEYS = ("obj", "cons", "grad", "total", "violation")

    def __init__(self, engine: "Engine", scene: SceneBatch, spec: ModuleSpec, t_horizon: int, num_particles: int,
                 num_candidates: int = 1, with_noise: bool = True):
        self.engine, self.spec, self.t_horizon, self.k = engine, spec, t_horizon, num_candidates
        dev, n, p = engine.device, scene.num_ados, num_particles
        particle_env = scene.env_kind != "kalman"
        # dynamic inputs: one block [controls | v0 | sigma | weight | noise], 16-byte aligned segments
        shapes = [("controls", (1, num_candidates, t_horizon, 2))]
        if particle_env:
            shapes += [("v0", (1, n, p)), ("sigma", (1, n, p)), ("weight", (1, n, p))]
            if with_noise and t_horizon > 1:
                shapes.append(("noise", (1, t_horizon - 1, n, p, 2)))
        self._spans, total = {}, 0
        for name, shape in shapes:
            count = int(math.prod(shape))
            self._spans[name] = (total, count, shape)
            total += -(-count // 4) * 4
        pin = dev.type == "cuda"
        self.hin = torch.zeros(total, dtype=torch.float32, pin_memory=pin)
        self.din = torch.zeros(total, dtype=torch.float32, device=dev)
        self._hin_np = self.hin.numpy()
        dyn = {name: self.din[o:o + c].view(shape) for name, (o, c, shape) in self._spans.items()}
        static = dataclasses.replace(scene, v0=None, sigma=None, weight=None, noise=None).to(dev)
        self.scene = dataclasses.replace(static, v0=dyn.get("v0"), sigma=dyn.get("sigma"), weight=dyn.get("weight"), noise=dyn.get("noise"))
        self.controls = dyn["controls"]
        self._pb = _problem_struct(self.scene, self.controls, t_horizon)
        self._md = _module_struct(self.scene, spec)
        lib = engine.lib
        rows = lib.mtb_num_constraint_rows(ctypes.byref(self._pb), ctypes.byref(self._md))
        if rows < 0:
            _cabi.check(lib, rows, "mtb_num_constraint_rows")
        self.rows = rows
        # results: one block [obj | cons | grad | total | violation] (+ a second one for the dense Jacobian)
        sizes = {"obj": num_candidates * NUM_OBJECTIVES, "cons": num_candidates * rows, "grad": num_candidates * 2 * t_horizon,
                 "total": num_candidates, "violation": num_candidates}
        self._out_spans, total = {}, 0
        for name in self.RESULT_KEYS:
            self._out_spans[name] = (total, sizes[name])
            total += -(-sizes[name] // 4) * 4
        self.dout = torch.zeros(max(total, 4), dtype=torch.float32, device=dev)
        self.hout = torch.zeros(max(total, 4), dtype=torch.float32, pin_memory=pin)
        self._hout_np = self.hout.numpy()
        self._outs = _cabi.MtbOutputs()
        for name in self.RESULT_KEYS:
            o, c = self._out_spans[name]
            if c > 0:
                setattr(self._outs, name, self.dout[o:o + c].data_ptr())
        self.djac = torch.zeros(max(num_candidates * rows * 2 * t_horizon, 1), dtype=torch.float32, device=dev)
        self.hjac = torch.zeros(max(num_candidates * rows * 2 * t_horizon, 1), dtype=torch.float32, pin_memory=pin)
        need = lib.mtb_workspace_bytes(ctypes.byref(self._pb), 1)
        self._ws = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
        self._ws_bytes = need

    def stage(self, controls, draw: typing.Optional[typing.Dict[str, torch.Tensor]]):
        """Write this call's inputs into the pinned host block: controls [K, 2T] (numpy or tensor) and the particle draw."""
        o, c, _ = self._spans["controls"]
        self._hin_np[o:o + c] = controls.reshape(-1) if not isinstance(controls, torch.Tensor) else controls.reshape(-1).numpy()
        if draw is not None:
            for name in ("v0", "sigma", "weight", "noise"):
                if name in self._spans and draw.get(name) is not None:
                    o, c, _ = self._spans[name]
                    self.hin[o:o + c].copy_(draw[name].reshape(-1))

    def staged_inputs(self):
        """Copy of the staged host block (kept with a cache entry so that a later Jacobian request re-uses the same draw)."""
        return self._hin_np.copy()

    def restore(self, staged):
        self._hin_np[:] = staged

    def run(self, want_jacobian: bool = False, want_values: bool = True) -> typing.Dict:
        """Upload the staged inputs, launch, read back. Returns numpy arrays: obj [K,5], cons [K,R], grad [K,2T], total [K],
        violation [K] (when `want_values`) and jac [K,R,2T] (when `want_jacobian`)."""
        eng, lib = self.engine, self.engine.lib
        stream = eng._stream()
        with eng._guard():
            self.din.copy_(self.hin, non_blocking=True)
            if want_values:
                code = lib.mtb_evaluate(ctypes.byref(self._pb), ctypes.byref(self._md), ctypes.byref(self._outs),
                                        self._ws.data_ptr() if self._ws_bytes > 0 else None, self._ws_bytes, stream)
                _cabi.check(lib, code, "mtb_evaluate")
                self.hout.copy_(self.dout, non_blocking=True)
                eng.launches += 1
            if want_jacobian and self.rows > 0:
                code = lib.mtb_constraint_jacobian(ctypes.byref(self._pb), ctypes.byref(self._md), None, self.djac.data_ptr(), None, 0, stream)
                _cabi.check(lib, code, "mtb_constraint_jacobian")
                self.hjac.copy_(self.djac, non_blocking=True)
                eng.launches += 1
            if eng.device.type == "cuda":
                torch.cuda.current_stream(eng.device).synchronize()
        out = {}
        k, t2 = self.k, 2 * self.t_horizon
        if want_values:
            shapes = {"obj": (k, NUM_OBJECTIVES), "cons": (k, self.rows), "grad": (k, t2), "total": (k,), "violation": (k,)}
            for name in self.RESULT_KEYS:
                o, c = self._out_spans[name]
                out[name] = self._hout_np[o:o + c].reshape(shapes[name]).copy()
        if want_jacobian:
            out["jac"] = self.hjac.numpy()[:k * self.rows * t2].reshape(k, self.rows, t2).copy() if self.rows > 0 else None
        return out


class Engine:
    """The product entry point of the batched hot path: binds the CUDA library (no fallback) and counts launches."""

    def __init__(self, device: typing.Union[str, torch.device] = "cuda:0"):
        from . import _lib
        if not torch.cuda.is_available():
            raise RuntimeError("mantrap_b200 needs a CUDA device (there is no CPU path)")
        self.lib = _lib.load()
        self.device = torch.device(device)
        self.launches = 0
        self._workspace = {}                              # kernel scratch per CUDA stream

    @classmethod
    def from_library(cls, lib: ctypes.CDLL, device: typing.Union[str, torch.device]) -> "Engine":
        """TEST HOOK: an engine over an explicitly given, already bound C-ABI library. The CPU test-suite uses it with
        the host emulation of the kernel source (tests/host_emul) to exercise the host classes without a GPU; nothing
        in the package calls it."""
        self = cls.__new__(cls)
        self.lib, self.device, self.launches, self._workspace = lib, torch.device(device), 0, {}
        return self

    def _check(self, *tensors):
        for t in tensors:
            if t is not None and t.device != self.device and not (t.is_cuda and self.device.type == "cuda" and
                                                                  self.device.index in (None, t.device.index)):
                raise RuntimeError(f"mantrap_b200.Engine only accepts tensors on {self.device} (got {t.device})")

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream if self.device.type == "cuda" else 0

    def _guard(self):
        return torch.cuda.device(self.device) if self.device.type == "cuda" else contextlib.nullcontext()

    # ---- ego dynamics (agents interface) ---------------------------------------------------------------------
    def ego_unroll(self, ego_kind: str, dt: float, x0: torch.Tensor, controls: torch.Tensor) -> torch.Tensor:
        """x0 [B,5] (px,py,vx,vy,time), controls [B,T,2] -> trajectory [B,T+1,5] (linear.py:68-105)."""
        self._check(x0, controls)
        b, t_horizon = controls.shape[0], controls.shape[1]
        out = torch.empty((b, t_horizon + 1, 5), dtype=torch.float32, device=self.device)
        with self._guard():
            code = self.lib.mtb_ego_unroll(EGO_KINDS[ego_kind], dt, _cabi.ptr(x0), _cabi.ptr(controls), b, t_horizon,
                                           _cabi.ptr(out), self._stream())
        _cabi.check(self.lib, code, "mtb_ego_unroll")
        self.launches += 1
        return out

    def ego_roll(self, ego_kind: str, dt: float, trajectory: torch.Tensor) -> torch.Tensor:
        """trajectory [B,T+1,5] -> controls [B,T,2] (linear.py:107-130)."""
        self._check(trajectory)
        b, t_horizon = trajectory.shape[0], trajectory.shape[1] - 1
        out = torch.empty((b, t_horizon, 2), dtype=torch.float32, device=self.device)
        with self._guard():
            code = self.lib.mtb_ego_roll(EGO_KINDS[ego_kind], dt, _cabi.ptr(trajectory), b, t_horizon, _cabi.ptr(out), self._stream())
        _cabi.check(self.lib, code, "mtb_ego_roll")
        self.launches += 1
        return out

    def ego_unroll_adjoint(self, ego_kind: str, dt: float, grad_trajectory: torch.Tensor) -> torch.Tensor:
        """grad_trajectory [B,T+1,5] -> grad_controls [B,T,2] (reverse mode of unroll_trajectory)."""
        self._check(grad_trajectory)
        b, t_horizon = grad_trajectory.shape[0], grad_trajectory.shape[1] - 1
        out = torch.empty((b, t_horizon, 2), dtype=torch.float32, device=self.device)
        with self._guard():
            code = self.lib.mtb_ego_unroll_adjoint(EGO_KINDS[ego_kind], dt, _cabi.ptr(grad_trajectory), b, t_horizon,
                                                   _cabi.ptr(out), self._stream())
        _cabi.check(self.lib, code, "mtb_ego_unroll_adjoint")
        self.launches += 1
        return out

    def ego_dx_du(self, ego_kind: str, dt: float, t_horizon: int) -> torch.Tensor:
        """[5(T+1), 2T] Jacobian of the trajectory w.r.t. the controls (linear.py:169-196)."""
        out = torch.empty((5 * (t_horizon + 1), 2 * t_horizon), dtype=torch.float32, device=self.device)
        with self._guard():
            code = self.lib.mtb_ego_dx_du(EGO_KINDS[ego_kind], dt, t_horizon, _cabi.ptr(out), self._stream())
        _cabi.check(self.lib, code, "mtb_ego_dx_du")
        self.launches += 1
        return out

    def evaluate(self, scene: SceneBatch, controls: typing.Optional[torch.Tensor], spec: ModuleSpec = None,
                 want: typing.Sequence[str] = ("obj", "total", "grad"), out=None, stats=None, ego_traj=None):
        """One fused launch on the CURRENT CUDA stream. The kernel's scratch (adjoint rows, tape, masks) is kept per stream: two
        streams that evaluate on one engine (e.g. `api.HostPipeline.s_run` next to the caller's default stream) never share it,
        and a buffer that has to grow is replaced only for its own stream, whose launches are ordered."""
        self._check(scene.ado0, controls, ego_traj)
        stream = self._stream()
        with self._guard():
            res = launch_evaluate(self.lib, scene, controls, spec, want, stream=stream, out=out, stats=stats,
                                  workspace=self._workspace.get(stream), ego_traj=ego_traj)
        ws = res.pop("_workspace", None)
        if ws is not None:
            self._workspace[stream] = ws
        self.launches += 1
        return res

    def rollout_wo_ego(self, scene: SceneBatch, t_horizon: int, want=("mus", "sigmas")):
        """Ego-free prediction (`compute_distributions_wo_ego`): mus [S,N,T,4], sigmas [S,N,T,2]."""
        self._check(scene.ado0)
        stream = self._stream()
        with self._guard():
            res = launch_evaluate(self.lib, scene.without_ego(), None, None, want, t_horizon=t_horizon, stream=stream,
                                  workspace=self._workspace.get(stream))
        ws = res.pop("_workspace", None)
        if ws is not None:
            self._workspace[stream] = ws
        self.launches += 1
        return {k: v[:, 0] for k, v in res.items() if not k.startswith("_")}

    def constraint_jacobian(self, scene: SceneBatch, controls: typing.Optional[torch.Tensor], spec: ModuleSpec,
                            want_values: bool = False, ego_traj=None):
        """Dense constraint Jacobian [S,C,R,2T] (forward-mode tangent kernel)."""
        self._check(scene.ado0, controls, ego_traj)
        with self._guard():
            out = launch_jacobian(self.lib, scene, controls, spec, stream=self._stream(), want_values=want_values,
                                  ego_traj=ego_traj)
        self.launches += 1
        return out

    def argmin(self, total, violation, controls, feas_tol: float = constants.SOLVER_CONSTRAINT_LIMIT):
        self._check(total, violation, controls)
        with self._guard():
            out = launch_argmin(self.lib, total, violation, controls, feas_tol, stream=self._stream())
        self.launches += 1
        return out

    def attention(self, kind: str, ego0: torch.Tensor, ado0: torch.Tensor, radius: float = None) -> torch.Tensor:
        """Attention filter of S scenes -> mask [S,N] uint8 (`kind` = "euclidean" | "closest")."""
        self._check(ego0, ado0)
        if radius is None:
            radius = constants.ATTENTION_EUCLIDEAN_RADIUS if kind == "euclidean" else constants.ATTENTION_CLOSEST_RADIUS
        with self._guard():
            mask = launch_attention(self.lib, kind, ego0, ado0, radius, stream=self._stream())
        self.launches += 1
        return mask

    def warm_start_encoding(self, ego0, ego_goal, ado0, encoding_db, solution_db, t_horizon: int):
        """Scene encodings of S scenes, their nearest pre-computed scenario and its solution as warm start
        (`TrajOptSolver.encode` + `_warm_start_encoding`)."""
        self._check(ego0, ego_goal, ado0, encoding_db, solution_db)
        with self._guard():
            out = launch_warm_start_encoding(self.lib, ego0, ego_goal, ado0, encoding_db, solution_db, t_horizon, stream=self._stream())
        self.launches += 1
        return out

    def env_step(self, scene: SceneBatch, controls: torch.Tensor, eps: torch.Tensor = None, warm: torch.Tensor = None):
        """One closed-loop step of every scene, in place (GraphBasedEnvironment.step for S scenes in lock step)."""
        self._check(scene.ego0, scene.ado0, controls, eps, warm)
        with self._guard():
            launch_env_step(self.lib, scene, controls, eps, warm, stream=self._stream())
        self.launches += 1

    def launch_info(self, scene: SceneBatch, controls: torch.Tensor, want_grad: bool = True):
        return query_launch(self.lib, scene, controls, want_grad)
