"""Public batched API with HOST buffers: what a sampling solver (RandomSearch / MCTS driver) or a batch of IPOPT
callbacks calls. Inputs come from pinned host memory, results go back to pinned host memory; the host<->device copies
are part of the call (this is the `e2e` number of bench.py)."""
import typing

import torch

from . import constants
from .engine import Engine, ModuleSpec, SceneBatch


class HostBuffers:
    """Pinned host outputs of one batched evaluation: objective, violation and control gradient of every
    (scene, candidate) plus the best feasible candidate per scene."""

    def __init__(self, num_scenes: int, num_candidates: int, t_horizon: int):
        pin = dict(dtype=torch.float32, pin_memory=True)
        self.total = torch.empty((num_scenes, num_candidates), **pin)
        self.violation = torch.empty((num_scenes, num_candidates), **pin)
        self.grad = torch.empty((num_scenes, num_candidates, t_horizon, 2), **pin)
        self.best_idx = torch.empty(num_scenes, dtype=torch.int32, pin_memory=True)
        self.best_val = torch.empty(num_scenes, **pin)
        self.best_controls = torch.empty((num_scenes, t_horizon, 2), **pin)
        self.last_packed_device = None

    def d2h_bytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.total, self.violation, self.grad, self.best_idx,
                                                          self.best_val, self.best_controls))

    @staticmethod
    def h2d_bytes(controls_host: torch.Tensor) -> int:
        return controls_host.numel() * controls_host.element_size()


def evaluate_candidates_host(engine: Engine, scene: SceneBatch, controls_host: torch.Tensor, spec: ModuleSpec,
                             out: HostBuffers, device_buffers: typing.Dict[str, torch.Tensor] = None,
                             device_controls: torch.Tensor = None,
                             feas_tol: float = constants.SOLVER_CONSTRAINT_LIMIT) -> HostBuffers:
    """controls_host [S,C,T,2] (pinned) -> H2D -> fused rollout/objective/adjoint kernel -> per-scene argmin -> D2H of
    everything into `out`; returns after the results are readable on the host (one stream synchronisation).
    The scene batch (initial states, particle parameters, ego-free baseline) stays resident on the device."""
    dev = engine.device
    if device_controls is None:
        device_controls = torch.empty(controls_host.shape, dtype=torch.float32, device=dev)
    device_controls.copy_(controls_host, non_blocking=True)
    res = engine.evaluate(scene, device_controls, spec, want=("total", "grad", "violation"), out=device_buffers)
    idx, val, best = engine.argmin(res["total"], res["violation"], device_controls, feas_tol)
    out.total.copy_(res["total"], non_blocking=True)
    out.violation.copy_(res["violation"], non_blocking=True)
    out.grad.copy_(res["grad"], non_blocking=True)
    out.best_idx.copy_(idx, non_blocking=True)
    out.best_val.copy_(val, non_blocking=True)
    out.best_controls.copy_(best, non_blocking=True)
    out.last_packed_device = torch.cat((val[:, None], idx.float()[:, None], best.reshape(best.shape[0], -1)), dim=1)
    torch.cuda.current_stream(dev).synchronize()
    return out


class HostPipeline:
    """The same host-buffer evaluation as `evaluate_candidates_host`, software-pipelined over consecutive batches: while the
    kernels of batch i run, the controls of batch i+1 are copied to the device and the results of batch i-1 are copied back
    (three streams, two device buffer sets, CUDA events between them; the copy engines and the SMs work concurrently).
    Every batch still moves all of its inputs host -> device and all of its results device -> host; only the waiting is
    overlapped. What a sampling solver does when it evaluates one candidate batch after the other (`search.py:47-67`).

        pipe = HostPipeline(engine, scene, spec, S, C, T)
        for k, controls_host in enumerate(batches):          # pinned [S,C,T,2]
            slot = pipe.submit(controls_host)                  # returns immediately; slot = k % depth
        pipe.wait(slot) -> HostBuffers of that batch           # results readable on the host
    """

    def __init__(self, engine: Engine, scene: SceneBatch, spec: ModuleSpec, num_scenes: int, num_candidates: int,
                 t_horizon: int, depth: int = 2, feas_tol: float = constants.SOLVER_CONSTRAINT_LIMIT):
        self.engine, self.scene, self.spec, self.depth, self.feas_tol = engine, scene, spec, depth, feas_tol
        dev = engine.device
        self.s_h2d, self.s_run, self.s_d2h = (torch.cuda.Stream(device=dev) for _ in range(3))
        self.controls = [torch.empty((num_scenes, num_candidates, t_horizon, 2), dtype=torch.float32, device=dev) for _ in range(depth)]
        self.bufs: typing.List[typing.Optional[typing.Dict[str, torch.Tensor]]] = [None] * depth
        self.best: typing.List[typing.Optional[tuple]] = [None] * depth
        self.host = [HostBuffers(num_scenes, num_candidates, t_horizon) for _ in range(depth)]
        self.ev_h2d = [torch.cuda.Event() for _ in range(depth)]
        self.ev_run = [torch.cuda.Event() for _ in range(depth)]
        self.ev_d2h = [torch.cuda.Event() for _ in range(depth)]
        self.count = 0
        for s in (self.s_h2d, self.s_run, self.s_d2h):       # order the side streams after whatever prepared the scene
            s.wait_stream(torch.cuda.current_stream(dev))

    def submit(self, controls_host: torch.Tensor) -> int:
        slot = self.count % self.depth
        reuse = self.count >= self.depth
        with torch.cuda.stream(self.s_h2d):
            if reuse:
                self.s_h2d.wait_event(self.ev_run[slot])         # the kernels that read this control buffer are done
            self.controls[slot].copy_(controls_host, non_blocking=True)
            self.ev_h2d[slot].record(self.s_h2d)
        with torch.cuda.stream(self.s_run):
            self.s_run.wait_event(self.ev_h2d[slot])
            if reuse:
                self.s_run.wait_event(self.ev_d2h[slot])         # the previous results of this slot have left the device
            res = self.engine.evaluate(self.scene, self.controls[slot], self.spec, want=("total", "grad", "violation"),
                                       out=self.bufs[slot])
            self.bufs[slot] = {k: v for k, v in res.items() if not k.startswith("_")}
            self.best[slot] = self.engine.argmin(res["total"], res["violation"], self.controls[slot], self.feas_tol)
            self.ev_run[slot].record(self.s_run)
        with torch.cuda.stream(self.s_d2h):
            self.s_d2h.wait_event(self.ev_run[slot])
            out, res = self.host[slot], self.bufs[slot]
            idx, val, best = self.best[slot]
            out.total.copy_(res["total"], non_blocking=True)
            out.violation.copy_(res["violation"], non_blocking=True)
            out.grad.copy_(res["grad"], non_blocking=True)
            out.best_idx.copy_(idx, non_blocking=True)
            out.best_val.copy_(val, non_blocking=True)
            out.best_controls.copy_(best, non_blocking=True)
            self.ev_d2h[slot].record(self.s_d2h)
        self.count += 1
        return slot

    def wait(self, slot: int) -> HostBuffers:
        """Block the host until the results of the batch in `slot` are readable."""
        self.ev_d2h[slot].synchronize()
        return self.host[slot]

    def best_device(self, slot: int):
        """(idx, val, controls) of the batch in `slot` on the device (valid once ev_run[slot] has completed)."""
        return self.best[slot]

    def drain(self):
        for s in (self.s_h2d, self.s_run, self.s_d2h):
            s.synchronize()
