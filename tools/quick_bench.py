"""Quick device-side timing probe of the fused kernel on a few shapes (development tool, not the bench contract)."""
import argparse
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, scenarios  # noqa: E402


def run(eng, env_kind, s, c, n, p, t_hor, weights, iters=5, grad=True):
    scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind).to("cuda:0")
    controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
    mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
    spec = engine.ModuleSpec(weights=weights, mus_wo=mus_wo)
    want = ("obj", "total", "grad") if grad else ("obj", "total")
    stats = torch.zeros(4, dtype=torch.int64, device="cuda:0")
    out = eng.evaluate(scene, controls, spec, want=want, stats=stats)
    torch.cuda.synchronize()
    st = stats.tolist()
    info = eng.launch_info(scene, controls, grad)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(2):
        eng.evaluate(scene, controls, spec, want=want, out=out)
    start.record()
    for _ in range(iters):
        eng.evaluate(scene, controls, spec, want=want, out=out)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / iters
    evals = s * c
    sat = st[2] / evals
    print(f"{env_kind} S={s} C={c} N={n} P={p} T={t_hor} grad={grad}: {ms:.3f} ms/launch, {evals / ms * 1e3:.3e} evals/s, "
          f"ado-pairs/eval={st[0] / evals:.1f} ego-pairs/eval={st[1] / evals:.1f} saturated={sat:.3f} {info}", flush=True)
    return ms


def run_jacobian(eng, env_kind, s, c, n, p, t_hor, iters=5, ellipsoid=False):
    """Constraint-Jacobian kernel (control | speed | min-distance [| ellipsoid] rows) on one shape."""
    scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind).to("cuda:0")
    controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True, ellipsoid=ellipsoid)
    jac = eng.constraint_jacobian(scene, controls, spec)
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(iters):
        eng.constraint_jacobian(scene, controls, spec)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / iters
    rows = jac.shape[2]
    print(f"jacobian {env_kind} S={s} C={c} N={n} P={p} T={t_hor} rows={rows} cols={2 * t_hor}: {ms:.3f} ms/launch, "
          f"{s * c / ms * 1e3:.3e} Jacobians/s, {jac.numel() * 4 / ms / 1e6:.1f} GB/s written", flush=True)
    return ms


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--dev", action="store_true", help="load the -DMTB_DEV_KNOBS library (launch overrides from the environment, e.g. MTB_JAC=1)")
    args = ap.parse_args()
    if args.dev:
        import ctypes
        from mantrap_b200 import _cabi, build
        eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), "cuda:0")
    else:
        eng = engine.Engine("cuda:0")
    w = {"goal": 0.5, "acc": 1.0}
    if args.only == "c4":
        run(eng, "sf", 2048, 256, 20, 3, 12, w)
        sys.exit(0)
    if args.only == "jac":
        run_jacobian(eng, "sf", 1024, 1, 10, 4, 20)
        run_jacobian(eng, "sf", 1024, 16, 10, 4, 20)
        run_jacobian(eng, "sf", 1024, 16, 10, 4, 20, ellipsoid=True)
        run_jacobian(eng, "sf", 256, 16, 20, 3, 12)
        run_jacobian(eng, "pf", 256, 64, 5, 5, 15)
        run_jacobian(eng, "sf", 4096, 16, 3, 5, 10)
        run_jacobian(eng, "sf", 64, 16, 20, 3, 40)
        sys.exit(0)
    run(eng, "sf", 1024, 256, 20, 3, 12, w)
    run(eng, "sf", 1024, 256, 20, 3, 12, w, grad=False)
    run(eng, "sf", 8192, 256, 20, 3, 12, w)
    run(eng, "pf", 64, 4096, 5, 5, 15, w)
    run(eng, "sf", 4096, 16, 3, 5, 10, w)
    run(eng, "sf", 1024, 16, 10, 4, 20, {"goal": 0.5, "pos": 1.0})
    if args.big:
        run(eng, "sf", 256, 16, 64, 8, 50, w, iters=2)
        run(eng, "sf", 1024, 64, 32, 4, 30, w, iters=2)
