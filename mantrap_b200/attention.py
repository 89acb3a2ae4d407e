"""Attention filters of the reference (`mantrap/attention`) on the device: which ados enter the planning cost.

Same class names, constructor and `compute()` contract as the reference (attention_module.py:9-77): `compute()` returns
the list of selected ado ids of the module's environment. The selection itself is one `mtb_attention` launch (one warp
per scene); `compute_mask_batch` exposes the batched form ([S,N] uint8, the `ado_mask` input of the fused kernels) that
the batched solvers use for S scenes at once.
"""
import typing

import numpy as np
import torch

from . import constants, runtime


class AttentionModule:
    """attention_module.py:9-77."""
    kind: str = None
    radius: float = None

    def __init__(self, env, t_horizon: int, **unused):
        self._env = env
        self._t_horizon = t_horizon
        self._ids_current = []

    def compute(self) -> typing.List[str]:
        if self._env.num_ados == 0:
            filtered_ids = []
        else:
            filtered_ids = [self._env.ado_ids[m] for m in self._compute()]
        return self._return_filtered(filtered_ids)

    def _compute(self) -> np.ndarray:
        """Indices of the selected ados (euclidean.py:22-29, closest.py:22-31), computed on the device."""
        engine = runtime.default_engine()
        ego_state, ado_states = self._env.states()
        ego0 = runtime.to_device(ego_state[0:4].view(1, 4), engine)
        ado0 = runtime.to_device(ado_states[:, 0:4].reshape(1, self._env.num_ados, 4), engine)
        mask = engine.attention(self.kind, ego0, ado0, self.radius)
        return np.nonzero(mask[0].cpu().numpy())[0]

    @classmethod
    def compute_mask_batch(cls, ego0: torch.Tensor, ado0: torch.Tensor, engine=None) -> torch.Tensor:
        """Batched form: ego0 [S,4], ado0 [S,N,4] on the engine device -> mask [S,N] uint8."""
        engine = engine if engine is not None else runtime.default_engine()
        return engine.attention(cls.kind, ego0, ado0, cls.radius)

    def _return_filtered(self, ids_filtered: typing.List[str]) -> typing.List[str]:
        self._ids_current = ids_filtered
        return self._ids_current

    @property
    def ids_current(self) -> typing.List[str]:
        return self._ids_current

    @property
    def env(self):
        return self._env

    @property
    def t_horizon(self) -> int:
        return self._t_horizon

    @property
    def name(self) -> str:
        raise NotImplementedError


class EuclideanModule(AttentionModule):
    """All ados closer than the attention radius to the ego at t = t_current (euclidean.py:9-37)."""
    kind = "euclidean"
    radius = constants.ATTENTION_EUCLIDEAN_RADIUS

    @property
    def name(self) -> str:
        return "euclidean"


class ClosestModule(AttentionModule):
    """closest.py:9-40. As in the reference, `distances[~index_min] = inf` uses the BITWISE not of the 0-dim argmin tensor,
    i.e. it removes the ado at index N - 1 - argmin (not "all but the closest"); every other ado inside the radius stays."""
    kind = "closest"
    radius = constants.ATTENTION_CLOSEST_RADIUS

    @property
    def name(self) -> str:
        return "closest"
