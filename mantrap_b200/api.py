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
