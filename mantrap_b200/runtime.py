"""Process-wide engine used by the drop-in host classes (`agents`, `environment`, `modules`, `solver`).

The reference's objects compute on whatever torch gives them; here every computation of the hot path goes through
ONE `Engine` (the CUDA C-ABI library bound on one device). `default_engine()` creates it lazily on `cuda:<LOCAL_RANK>`
and raises when there is no GPU or no built library -- there is no CPU path in the product.

`set_default_engine` exists so that (a) multi-GPU drivers pin one engine per rank and (b) the CPU test-suite can
install an engine bound to the host emulation of the kernel source (tests/host_emul, test infrastructure).
"""
import os
import typing

import torch

from .engine import Engine

_ENGINE: typing.Optional[Engine] = None


def default_engine() -> Engine:
    global _ENGINE
    if _ENGINE is None:
        _ENGINE = Engine(f"cuda:{int(os.environ.get('LOCAL_RANK', '0'))}")
    return _ENGINE


def set_default_engine(engine: typing.Optional[Engine]) -> typing.Optional[Engine]:
    """Install `engine` as the process-wide engine; returns the previous one."""
    global _ENGINE
    previous, _ENGINE = _ENGINE, engine
    return previous


def to_device(x: typing.Optional[torch.Tensor], engine: Engine = None) -> typing.Optional[torch.Tensor]:
    """float32, contiguous, on the engine's device (autograd-transparent)."""
    if x is None:
        return None
    engine = engine if engine is not None else default_engine()
    return x.to(device=engine.device, dtype=torch.float32).contiguous()
