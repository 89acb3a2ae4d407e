"""`torch.autograd.Function` wrappers around the C-ABI kernels.

The reference differentiates the hot path with torch.autograd through thousands of tiny CPU nodes
(`mantrap/modules/base/optimization_module.py:358-447`). Here each differentiable stage is ONE node whose forward and
backward are CUDA kernels with hand-derived adjoints, so user code that composes its own torch cost on top of the
predicted distributions (as a custom `OptimizationModule._objective_core` would) still gets exact gradients w.r.t.
the controls:

    controls --EgoUnroll--> ego trajectory --ParticleRollout--> (mus, sigmas) --any torch ops--> cost

The fused module path (`Engine.evaluate`) does not go through these nodes: it produces objective AND gradient in a
single launch. All tensors are float32 on the engine's device; the host classes move data in and out.
"""
import dataclasses

import torch

from .engine import Engine, ModuleSpec, SceneBatch


class EgoUnroll(torch.autograd.Function):
    """`LinearDTAgent.unroll_trajectory` (agents/base/linear.py:68-105): controls [B,T,2], x0 [B,5] -> [B,T+1,5]."""

    @staticmethod
    def forward(ctx, controls: torch.Tensor, x0: torch.Tensor, engine: Engine, ego_kind: str, dt: float):
        ctx.engine, ctx.ego_kind, ctx.dt = engine, ego_kind, dt
        return engine.ego_unroll(ego_kind, dt, x0.contiguous(), controls.contiguous())

    @staticmethod
    def backward(ctx, grad_trajectory: torch.Tensor):
        grad = ctx.engine.ego_unroll_adjoint(ctx.ego_kind, ctx.dt, grad_trajectory.contiguous())
        return grad, None, None, None, None


class ParticleRollout(torch.autograd.Function):
    """`GraphBasedEnvironment.compute_distributions` (environment/base/graph_based.py:362-381) for a batch: ego
    trajectories [S,C,T+1,4] -> collapsed means mus [S,C,N,T,4] (mu_p, mu_v) and velocity "scales" sigmas [S,C,N,T,2]
    (= variances). Backward re-runs the fused kernel with the incoming adjoints as seeds and returns
    the gradient w.r.t. the trajectory rows."""

    @staticmethod
    def forward(ctx, ego_traj: torch.Tensor, engine: Engine, scene: SceneBatch):
        ego_traj = ego_traj.contiguous()
        res = engine.evaluate(scene, None, None, want=("mus", "sigmas"), ego_traj=ego_traj)
        ctx.engine, ctx.scene = engine, scene
        ctx.save_for_backward(ego_traj)
        return res["mus"], res["sigmas"]

    @staticmethod
    def backward(ctx, grad_mus: torch.Tensor, grad_sigmas: torch.Tensor):
        (ego_traj,) = ctx.saved_tensors
        spec = ModuleSpec(seed_mus=grad_mus.contiguous(), seed_sigmas=grad_sigmas.contiguous())
        res = ctx.engine.evaluate(ctx.scene, None, spec, want=("grad_traj",), ego_traj=ego_traj)
        return res["grad_traj"], None, None


def rollout_wo_ego(engine: Engine, scene: SceneBatch, t_horizon: int):
    """`compute_distributions_wo_ego` (particle.py:185-198): no ego, hence no gradient: mus [S,N,T,4], sigmas [S,N,T,2]."""
    res = engine.rollout_wo_ego(dataclasses.replace(scene), t_horizon)
    return res["mus"], res["sigmas"]
