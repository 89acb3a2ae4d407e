"""BASELINE config 5: horizon / agent / particle sweep of the fused rollout + objective + adjoint kernel on one GPU.

    python tools/sweep.py [--evals 1048576] [--out profiles/sweep.csv] [--quick]

Grid: t_planning in {10,20,30,40,50} x ados in {2,4,8,16,32,64} x particles ("modes") in {1,2,4,8}, social
forces, goal + interaction-acceleration objective with gradient, S*C = `--evals` evaluations per launch (C = 256 candidates
per scene). Device-side timing with CUDA events on the launching stream, 1 warm-up + 3 timed launches per shape; the
inputs of a launch (controls, gradients: 16 T bytes per evaluation, >= 168 MB at the default size) exceed the L2.
P = 1 rows are throughput-only (the reference raises at P = 1: torch.var of one particle is NaN).
Development / measurement tool, not the bench contract (bench.py)."""
import argparse
import csv
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantrap_b200 import engine, scenarios  # noqa: E402

FLOP_PAIR, FLOP_STEP, FLOP_TEST = 150.0, 150.0, 5.0      # DESIGN.md §6 (executed-work model)


def measure(eng, t_hor, n, p, evals, iters=3):
    c = 256
    s = max(1, evals // c)
    scene = scenarios.synthetic_scenes(s, n, p, env_kind="sf").to("cuda:0")
    controls = scenarios.candidate_controls(s, c, t_hor).to("cuda:0")
    mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
    spec = engine.ModuleSpec(weights={"goal": 0.5, "acc": 1.0}, mus_wo=mus_wo)
    stats = torch.zeros(4, dtype=torch.int64, device="cuda:0")
    out = eng.evaluate(scene, controls, spec, want=("total", "grad"), stats=stats)
    torch.cuda.synchronize()
    pairs = (int(stats[0]) + int(stats[1])) / (s * c)
    saturated = int(stats[2]) / (s * c)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(iters):
        eng.evaluate(scene, controls, spec, want=("total", "grad"), out=out)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / iters
    flop = FLOP_PAIR * pairs + FLOP_STEP * (t_hor - 1) * n * p + FLOP_TEST * (t_hor - 1) * n * p * n
    info = eng.launch_info(scene, controls, True)
    return dict(T=t_hor, N=n, P=p, evals=s * c, ms=round(ms, 3), evals_per_s=round(s * c / ms * 1e3, 1),
                pairs_per_eval=round(pairs, 1), saturated=round(saturated, 3), flop_per_eval=round(flop, 1),
                executed_tflops=round(s * c / ms * 1e3 * flop / 1e12, 3), block=info["block"], group=info["group_size"],
                blocks_per_sm=info["blocks_per_sm"], tape_in_smem=info["tape_in_smem"], kernel_kind=info["kernel_kind"])


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--evals", type=int, default=1 << 20)
    ap.add_argument("--out", default="gpurun_out/sweep.csv")
    ap.add_argument("--quick", action="store_true", help="corner points only")
    ap.add_argument("--dev", action="store_true", help="use the -DMTB_DEV_KNOBS library (launch-plan overrides from the environment)")
    args = ap.parse_args()
    if args.dev:
        import ctypes
        from mantrap_b200 import _cabi, build
        eng = engine.Engine.from_library(_cabi.bind(ctypes.CDLL(build.build(dev_knobs=True))), "cuda:0")
    else:
        eng = engine.Engine("cuda:0")
    ts, ns, ps = (10, 20, 30, 40, 50), (2, 4, 8, 16, 32, 64), (1, 2, 4, 8)
    if args.quick:
        ts, ns, ps = (10, 50), (2, 64), (2, 8)
    rows = []
    for t_hor in ts:
        for n in ns:
            for p in ps:
                # keep the biggest shapes inside a few seconds per launch: fewer evaluations, still >> one wave
                evals = args.evals if t_hor * n * n * p <= 50 * 32 * 32 * 4 else args.evals // 4
                row = measure(eng, t_hor, n, p, evals)
                rows.append(row)
                print(row, flush=True)
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    print("wrote", args.out)
