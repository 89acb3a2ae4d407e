"""Device-side timing of the batched closed-loop MPC (mantrap_b200.mpc.BatchedClosedLoop) on ETH-shaped scenes:
attention -> ego-free baseline -> C candidates (rollout + objectives + violation) -> argmin -> environment step, per
closed-loop step, all scenes in lock step. Development / measurement tool, not the bench contract."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, mpc, scenarios  # noqa: E402

if __name__ == "__main__":
    s = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    c = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    eng = engine.Engine("cuda:0")
    scene = scenarios.synthetic_scenes(s, 20, 3, regular_sigma=True)
    for attention in (None, "euclidean"):
        loop = mpc.BatchedClosedLoop(eng, scene, t_planning=12, num_candidates=c, attention=attention, seed=1)
        loop.solve(1)
        torch.cuda.synchronize()
        launches0 = eng.launches
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        out = loop.solve(steps)
        stop.record()
        torch.cuda.synchronize()
        ms = start.elapsed_time(stop) / steps
        print(f"closed loop S={s} C={c} N=20 P=3 T=12 attention={attention}: {ms:.2f} ms/step, {s / ms * 1e3:.3e} scene-steps/s, "
              f"{s * c / ms * 1e3:.3e} candidate evals/s, {(eng.launches - launches0) / steps:.0f} kernel launches/step, "
              f"feasible picks {float((out.best_index >= 0).float().mean()):.3f}, reached goal {int((out.steps < steps).sum())}", flush=True)
