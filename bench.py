"""Benchmark of the mantrap_b200 hot path: rollout + objective + control-gradient evaluations per second.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (N>1: launched by torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the reference algorithm on the host CPU cores

Workload (BASELINE.json config 4, "synthetic ETH-shaped scenes"): social-forces environment, 20 pedestrians x 3
particles, t_planning = 12, dt = 0.4 s, double-integrator ego, goal + interaction-acceleration objectives, 256 random
control candidates per scene, 8192 scenes PER GPU (65536 scenes x 256 candidates on 8 GPUs; weak scaling). One step =
one pass of the hot path over the rank's whole batch: fused rollout/objective/adjoint kernel + per-scene argmin of the
best feasible candidate (+ one small NCCL all-gather of the winners when N > 1).

The JSON line follows the driver contract; see DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ADOS, N_PARTICLES, T_HORIZON, N_CANDIDATES = 20, 3, 12, 256
WEIGHTS = {"goal": 0.5, "acc": 1.0}
METRIC = "rollout+Jacobian evals/sec (scenes x candidates)"        # BASELINE.json's metric name
UNIT = "evals/s"


def flops_per_eval(t_hor, n, p):
    """Algorithmic work model of SURVEY §8(d): 150 (T-1) N P (N+1) FLOP per social-forces evaluation
    (forward 50 FLOP per pair term incl. 7 SFU ops, reverse sweep counted 2x)."""
    return 150.0 * (t_hor - 1) * n * p * (n + 1)


def sfu_per_eval(t_hor, n, p):
    return 3.0 * (t_hor - 1) * n * p * (7 * n + 5)


def bytes_per_eval(t_hor, n_mod):
    return 16.0 * t_hor + 4.0 * n_mod


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference algorithm on the host cores (cpu_baseline and --impl reference)
# ---------------------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """One process = one core: evaluates batches of the bench workload with the CPU oracle until `budget` seconds of
    timed work are spent (after one untimed warm-up batch). Returns (evaluations, seconds)."""
    worker, budget, batch, seed = args
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    import kernel_cases as kc
    scene = scenarios.synthetic_scenes(batch, N_ADOS, N_PARTICLES, seed=seed + worker, param_seed=seed + 1000 + worker)
    controls = scenarios.candidate_controls(batch, 1, T_HORIZON, seed=seed + 2000 + worker)[:, 0]
    osc = kc.oracle_scene(scene, 1)
    mus_wo, _ = to.rollout("sf", osc.ado0, osc.goal, osc.v0, osc.sigma, osc.weights, None, T_HORIZON, scene.dt)
    to.evaluate(osc, controls, mus_wo, WEIGHTS)                                    # warm-up (allocator, autograd)
    done, t0 = 0, time.perf_counter()
    while True:
        to.evaluate(osc, controls, mus_wo, WEIGHTS)
        done += batch
        elapsed = time.perf_counter() - t0
        if elapsed >= budget:
            return done, elapsed


def cpu_arm(seconds_target=12.0, batch=32, workers=None):
    """Times the CPU oracle (vectorised float32 torch port of the reference algorithm: forward rollout + objectives +
    torch.autograd gradient; one process per core, one thread each) on a bounded sample of the bench workload.
    Returns (evals/s, cores, description)."""
    import multiprocessing as mp
    workers = workers or os.cpu_count() or 1
    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        t0 = time.perf_counter()
        results = pool.map(_cpu_worker, [(w, seconds_target, batch, 7) for w in range(workers)])
        wall = time.perf_counter() - t0
    evals = sum(r[0] for r in results)
    timed = max(r[1] for r in results)
    sample = (f"{evals} evaluations of the bench workload (SF N={N_ADOS} P={N_PARTICLES} T={T_HORIZON}, goal+acc objective "
              f"+ autograd gradient) in {timed:.1f} s of timed work on {workers} single-threaded processes (batches of {batch}; "
              f"wall incl. start-up {wall:.1f} s)")
    return evals / timed, workers, sample


# BASELINE config 5 (sweep): a 12-shape subset (t_planning, ados, particles) timed on every bench run, GPU and CPU port
C5_SUBSET = [(10, 2, 2), (10, 8, 4), (10, 32, 2), (20, 4, 8), (20, 16, 4), (20, 64, 2), (30, 8, 8), (30, 32, 4), (40, 16, 2),
             (50, 2, 8), (50, 8, 4), (50, 64, 2)]


def _cpu_shape_worker(args):
    """CPU oracle (one core) on one sweep shape: evaluations per second of rollout + objectives + autograd gradient."""
    t_hor, n, p, budget = args
    import torch
    torch.set_num_threads(1)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import torch_oracle as to
    from mantrap_b200 import scenarios
    import kernel_cases as kc
    batch = 8
    scene = scenarios.synthetic_scenes(batch, n, p, seed=11, param_seed=12, regular_sigma=True)
    controls = scenarios.candidate_controls(batch, 1, t_hor, seed=13)[:, 0]
    osc = kc.oracle_scene(scene, 1)
    mus_wo, _ = to.rollout("sf", osc.ado0, osc.goal, osc.v0, osc.sigma, osc.weights, None, t_hor, scene.dt)
    to.evaluate(osc, controls, mus_wo, WEIGHTS)
    done, t0 = 0, time.perf_counter()
    while True:
        to.evaluate(osc, controls, mus_wo, WEIGHTS)
        done += batch
        elapsed = time.perf_counter() - t0
        if elapsed >= budget:
            return done / elapsed


def cpu_c5_subset(budget=2.0):
    """{shape: evals/s on ONE core} of the CPU oracle for the sweep subset (one process per shape, in parallel)."""
    import multiprocessing as mp
    ctx = mp.get_context("fork")
    with ctx.Pool(min(len(C5_SUBSET), os.cpu_count() or 1)) as pool:
        rates = pool.map(_cpu_shape_worker, [(t, n, p, budget) for t, n, p in C5_SUBSET])
    return dict(zip(C5_SUBSET, rates))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    values = []
    sample = ""
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    budget = max(3.0, min(30.0, 60.0 / max(1, args.steps)))        # each "step" = one bounded sample of the workload
    for _ in range(max(1, args.steps)):
        v, cores, sample = cpu_arm(seconds_target=budget)
        values.append(v)
    value = sum(values) / len(values)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": (time.perf_counter() - t0) * 1e3 / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(1, args.gpus)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference algorithm = oracle/torch_oracle.py (the Python reference cannot travel to the GPU box; "
                "its own per-call cost is ~0.3 evals/s/core, BASELINE.md §2)",
    }
    emit_line(line)


# ---------------------------------------------------------------------------------------------------------------
def workload_config(args, world):
    return {"workload": f"BASELINE config 4 (ETH-shaped): SocialForces N={N_ADOS} ados x P={N_PARTICLES} particles, "
                        f"t_planning={T_HORIZON}, dt=0.4, double-integrator ego, goal+interaction-acceleration objective, "
                        f"{args.scenes_per_gpu} scenes/GPU x {args.candidates} candidates; one eval = forward rollout + objectives + "
                        "control-limit / speed-limit rows + gradient w.r.t. the 2T controls (config 4 has no Jacobian rows: its "
                        "constraints have constant Jacobians; the Jacobian kernel is timed on config 3 under `secondary`)",
            "scenes_per_gpu": args.scenes_per_gpu, "candidates": args.candidates, "global_evals_per_step":
                args.scenes_per_gpu * args.candidates * world, "parallelism": f"scene-sharded x{world}",
            "cache": "inputs+outputs per step (controls 96 B + gradient 96 B + 12 B per eval) exceed the 126 MB L2; "
                     "no explicit flush"}


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi DURING the timed region. nvidia-smi needs a few hundred
    milliseconds to start (more on a busy 8-GPU box), which is as long as a short timed region: the process is therefore started
    before the warm-up steps and streams one row every 50 ms; `mark_start` / `mark_stop` bracket the timed region on the host
    clock and only rows that arrived inside the bracket are reported (`window`: "timed"). If none did (pathological), the rows
    closest to the bracket are used and `window` says "nearest"."""
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.t0 = self.t1 = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "50", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.monotonic(), [x.strip() for x in line.split(",")]))

    def mark_start(self):
        self.t0 = time.monotonic()

    def mark_stop(self):
        self.t1 = time.monotonic()

    def __exit__(self, *exc):
        if self.proc is not None:
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        rows = [(ts, r) for ts, r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        t0 = self.t0 if self.t0 is not None else float("-inf")
        t1 = self.t1 if self.t1 is not None else float("inf")
        inside = [r for ts, r in rows if t0 <= ts <= t1]
        window = "timed"
        if not inside and rows:
            window = "nearest"
            mid = 0.5 * (max(t0, rows[0][0]) + min(t1, rows[-1][0]))
            inside = [r for _, r in sorted(rows, key=lambda x: abs(x[0] - mid))[:3]]
        sm = sorted(float(r[0]) for r in inside)
        reasons = set()
        for r in inside:
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        smax = max((float(r[1]) for r in inside if r[1].replace(".", "").isdigit()), default=None)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm),
                "window": window}


def measure_peaks(eng, torch):
    """FP32 FMA and MUFU throughput of this device (dependent chains, 8-way ILP, 16 warps... per SM)."""
    lib, dev = eng.lib, eng.device
    grid, iters = 148 * 16, 20000
    out = torch.empty(grid * 256, dtype=torch.float32, device=dev)
    res = {}
    for kind, name, per_iter in ((0, "fp32_tflops", 16.0), (1, "mufu_tops", 8.0)):
        stream = torch.cuda.current_stream(dev).cuda_stream
        lib.mtb_probe_launch(kind, out.data_ptr(), grid, 100, stream)
        best = 1e30
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            lib.mtb_probe_launch(kind, out.data_ptr(), grid, iters, stream)
            b.record()
            torch.cuda.synchronize(dev)
            best = min(best, a.elapsed_time(b))
        res[name] = grid * 256 * iters * per_iter / (best * 1e-3) / 1e12
    return res


def _timed(torch, dev, fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(dev)
    return a.elapsed_time(b) / iters


def _eval_rate(eng, torch, dev, env_kind, s, c, n, p, t_hor, weights, seed=0, iters=3):
    """Device-timed gradient evaluations per second of one shape: (evals/s, ms per launch, saturated fraction, executed pair terms / eval)."""
    from mantrap_b200 import engine, scenarios
    scene = scenarios.synthetic_scenes(s, n, p, env_kind=env_kind, seed=1234 + seed, param_seed=5678 + seed).to(dev)
    controls = scenarios.candidate_controls(s, c, t_hor, seed=4321 + seed).to(dev)
    mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
    spec = engine.ModuleSpec(weights=weights, mus_wo=mus_wo)
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    out = eng.evaluate(scene, controls, spec, want=("total", "grad"), stats=stats)
    torch.cuda.synchronize(dev)
    st = stats.tolist()
    ms = _timed(torch, dev, lambda: eng.evaluate(scene, controls, spec, want=("total", "grad"), out=out), iters=iters, warm=1)
    return s * c / ms * 1e3, ms, st[2] / (s * c), (st[0] + st[1]) / (s * c)


def config1_measurements(torch):
    """BASELINE config 1 through the drop-in solver API (one process, one GPU): the four IPOPT callbacks, one IPOPT iteration
    (all four at one z), end-to-end solves with scipy's SLSQP over the same callbacks (labelled NOT IPOPT) and, where the
    third-party `cyipopt` is importable at run time, real IPOPT solves."""
    import numpy as np
    import mantrap_b200 as mantrap
    env = mantrap.environment.SocialForcesEnvironment(ego_position=torch.tensor([-5.0, 2.0]), ego_velocity=torch.zeros(2),
                                                      ego_type=mantrap.agents.DoubleIntegratorDTAgent, dt=0.4)
    for pos, goal in (((3.0, 2.0), (-4.0, 2.0)), ((0.0, -2.0), (0.0, 4.0)), ((-1.0, 4.5), (-1.0, -4.0))):
        env.add_ado(position=torch.tensor(pos), velocity=torch.zeros(2), goal=torch.tensor(goal))
    goal = torch.tensor([1.0, 1.0])
    modules = [mantrap.modules.GoalNormModule, mantrap.modules.InteractionAccelerationModule, mantrap.modules.ControlLimitModule,
               mantrap.modules.MinDistanceModule]
    solver = mantrap.solver.IPOPTSolver(env, goal=goal, t_planning=10, modules=modules)
    rng = np.random.RandomState(0)
    zs = [rng.uniform(-1, 1, 20) for _ in range(120)]
    out = {"workload": "SocialForces, double-integrator ego + 3 pedestrians x 5 particles, t_planning = 10, goal + interaction-"
                       "acceleration objective, control-limit + min-distance rows; wall clock of the Python call incl. the particle "
                       "draw, host<->device copies and the launch; a fresh z per call (no per-z cache hit)"}
    for name, fn in (("objective", solver.objective), ("gradient", solver.gradient), ("constraints", solver.constraints),
                     ("jacobian", solver.jacobian)):
        for z in zs[:20]:
            fn(z)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for z in zs[20:]:
            solver._cache.clear()
            fn(z)
        torch.cuda.synchronize()
        out[f"{name}_calls_per_s"] = 100.0 / (time.perf_counter() - t0)
    solver._cache.clear()
    t0 = time.perf_counter()
    for z in zs[20:]:
        solver.objective(z), solver.gradient(z), solver.constraints(z), solver.jacobian(z)
    torch.cuda.synchronize()
    out["ipopt_iterations_per_s"] = 100.0 / (time.perf_counter() - t0)
    out["ipopt_iteration_note"] = "objective + gradient + constraints + jacobian at one z = two launches (per-z cache)"
    # K independent callback sets per launch (K IPOPT replicas / line-search points of one scene): TrajOptSolver.evaluate_fused_batch
    zb = np.stack(zs[20:84])
    for _ in range(3):
        solver.evaluate_fused_batch(zb, want_jacobian=True)
    t0 = time.perf_counter()
    for _ in range(20):
        solver.evaluate_fused_batch(zb, want_jacobian=True)
    torch.cuda.synchronize()
    out["batched_ipopt_iterations_per_s"] = 20 * zb.shape[0] / (time.perf_counter() - t0)
    out["batched_note"] = "64 callback sets (values + gradient + rows + dense Jacobian) per launch pair through evaluate_fused_batch"
    slsqp = mantrap.solver.SLSQPSolver(env, goal=goal, t_planning=10, modules=modules)
    z0 = slsqp.warm_start()
    slsqp.optimize(z0, tag="bench", max_iter=30)
    n_solves = 10
    t0 = time.perf_counter()
    for k in range(n_solves):
        slsqp._cache.clear()                          # no per-z cache hits across solves: every callback evaluates
        slsqp.optimize(z0 + 0.05 * torch.as_tensor(rng.uniform(-1, 1, tuple(z0.shape)), dtype=z0.dtype), tag="bench", max_iter=30)
    out["slsqp_solves_per_s"] = n_solves / (time.perf_counter() - t0)
    out["slsqp_note"] = ("NOT IPOPT: scipy.optimize SLSQP (ftol 1e-4, <= 30 iterations; ~5 iterations from the hard warm start) over "
                         "the same four callbacks and bounds, per-z cache cleared between solves")
    try:
        try:
            import ipopt  # noqa: F401
        except ImportError:
            import cyipopt  # noqa: F401
        ips = mantrap.solver.IPOPTSolver(env, goal=goal, t_planning=10, modules=modules)
        ips.optimize(z0, tag="bench")
        t0 = time.perf_counter()
        for _ in range(3):
            ips._cache.clear()
            ips.optimize(z0, tag="bench")
        out["ipopt_solves_per_s"] = 3.0 / (time.perf_counter() - t0)
    except ImportError:
        out["ipopt_solves_per_s"] = None
        out["ipopt_note"] = "cyipopt (third-party, un-vendored by the reference) is not installed on this box: IPOPT solves/s unavailable"
    return out


def secondary_measurements(eng, torch, dev, cpu_c5=None):
    """Outside the timed region, N = 1: the other BASELINE configs and kernels of the path, device-timed (config 1: wall clock).
    Every gradient figure carries its `saturated_fraction` (a saturated clamp_max objective seeds no adjoint)."""
    from mantrap_b200 import engine, mpc, scenarios
    out = {}
    out["config1"] = config1_measurements(torch)
    # config 2: ONE potential-field scene x 4096 candidates (the sampling-baseline path), gradient of goal + interaction acceleration
    rate, ms, sat, _ = _eval_rate(eng, torch, dev, "pf", 1, 4096, 5, 5, 15, WEIGHTS, iters=10)
    out["config2"] = {"value": rate, "unit": "evals/s", "ms_per_launch": ms, "saturated_fraction": sat,
                      "workload": "PotentialField N=5 P=5 T=15, 1 scene x 4096 candidates (latency of one batched gradient evaluation)"}
    rate, ms, sat, _ = _eval_rate(eng, torch, dev, "pf", 64, 4096, 5, 5, 15, WEIGHTS)
    out["config2_throughput"] = {"value": rate, "unit": "evals/s", "ms_per_launch": ms, "saturated_fraction": sat,
                                 "workload": "PotentialField N=5 P=5 T=15, 64 scenes x 4096 candidates"}
    # config 3: gradient with the (unsaturated) acceleration objective and with the position objective (saturates by reference quirk
    # A.5-4: the zero-adjoint fast path), next to the dense constraint Jacobian
    for key, weights in (("config3_gradient_acc", {"goal": 0.5, "acc": 1.0}), ("config3_gradient_pos", {"goal": 0.5, "pos": 1.0})):
        rate, ms, sat, _ = _eval_rate(eng, torch, dev, "sf", 1024, 16, 10, 4, 20, weights)
        out[key] = {"value": rate, "unit": "evals/s", "ms_per_launch": ms, "saturated_fraction": sat,
                    "workload": f"SocialForces N=10 P=4 T=20, 1024 scenes x 16 candidates, objective {sorted(weights)}"}
    scene = scenarios.synthetic_scenes(1024, 10, 4).to(dev)
    controls = scenarios.candidate_controls(1024, 16, 20).to(dev)
    spec = engine.ModuleSpec(control_limits=True, speed_limits=True, min_distance=True)
    jac = eng.constraint_jacobian(scene, controls, spec)
    ms = _timed(torch, dev, lambda: eng.constraint_jacobian(scene, controls, spec))
    out["jacobian_config3"] = {"value": 1024 * 16 / ms * 1e3, "unit": "Jacobians/s", "rows": int(jac.shape[2]), "cols": int(jac.shape[3]),
                               "ms_per_launch": ms, "workload": "SocialForces N=10 P=4 T=20, 1024 scenes x 16 candidates, "
                               "control + speed + min-distance rows (forward-mode tangent kernel)"}
    # config 5: sweep subset, 2^18 evaluations per shape
    rows = []
    for t_hor, n, p in C5_SUBSET:
        c = 16
        s_c5 = (1 << 18) // c
        rate, ms, sat, pairs = _eval_rate(eng, torch, dev, "sf", s_c5, c, n, p, t_hor, WEIGHTS, iters=2)
        steps_pp = (t_hor - 1) * n * p
        tflops = rate * (150.0 * pairs + 150.0 * steps_pp + 5.0 * steps_pp * n) / 1e12
        row = {"t_planning": t_hor, "ados": n, "particles": p, "evals_per_s": rate, "ms_per_launch": ms, "saturated_fraction": sat,
               "executed_tflops": tflops}
        if cpu_c5 is not None:
            row["cpu_port_evals_per_s_per_core"] = cpu_c5[(t_hor, n, p)]
        rows.append(row)
    out["config5_subset"] = {"evals_per_shape": 1 << 18, "objective": "goal + interaction acceleration (gradient)", "shapes": rows}
    s = 2048
    loop = mpc.BatchedClosedLoop(eng, scenarios.synthetic_scenes(s, N_ADOS, N_PARTICLES, seed=41, param_seed=42), t_planning=T_HORIZON,
                                 num_candidates=N_CANDIDATES, attention="euclidean", seed=5)
    loop.solve(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    a.record()
    loop.solve(3)
    b.record()
    torch.cuda.synchronize(dev)
    ms = a.elapsed_time(b) / 3
    out["closed_loop"] = {"value": s / ms * 1e3, "unit": "scene-steps/s", "ms_per_step": ms,
                          "workload": f"{s} ETH-shaped scenes in lock step, {N_CANDIDATES} candidates per scene and step, euclidean "
                                      "attention, goal + interaction-acceleration objective, control + speed limits"}
    return out


def secondary_multi_gpu(eng, torch, dist, dev, rank, world):
    """N > 1: (config 2) ONE potential-field scene x 4096 candidates partitioned over the ranks + the order-preserving-key reduction
    of the winner; (config 5) the sweep subset with 2^18 evaluations PER RANK (weak scaling). Device-timed, max over ranks."""
    from mantrap_b200 import engine, parallel, scenarios
    out = {}
    scene = scenarios.synthetic_scenes(1, 5, 5, env_kind="pf", seed=77, param_seed=78).to(dev)       # the same scene on every rank
    controls_all = scenarios.candidate_controls(1, 4096, 15, seed=79)
    lo, hi = parallel.shard_range(4096, rank, world)
    controls = controls_all[:, lo:hi].to(dev).contiguous()
    mus_wo = eng.rollout_wo_ego(scene, 15)["mus"].contiguous()
    spec = engine.ModuleSpec(weights=WEIGHTS, mus_wo=mus_wo, control_limits=True, speed_limits=True)
    bufs = eng.evaluate(scene, controls, spec, want=("total", "grad", "violation"))

    def search():
        eng.evaluate(scene, controls, spec, want=("total", "grad", "violation"), out=bufs)
        idx, val, best = eng.argmin(bufs["total"], bufs["violation"], controls)
        return parallel.reduce_best_candidate_partitioned(val, idx, best, lo)

    ms = torch.tensor([_timed(torch, dev, search, iters=20, warm=3)], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    out["config2_candidate_partitioned"] = {"ms_per_search": float(ms.item()), "value": 4096 / float(ms.item()) * 1e3, "unit": "evals/s",
                                            "workload": f"PotentialField N=5 P=5 T=15, 1 scene x 4096 candidates split over {world} ranks, "
                                                        "gradient evaluation + argmin + 64-bit-key reduction of the winner"}
    rows = []
    for t_hor, n, p in C5_SUBSET:
        rate, ms1, sat, _ = _eval_rate(eng, torch, dev, "sf", (1 << 18) // 16, 16, n, p, t_hor, WEIGHTS, seed=rank, iters=2)
        t_ms = torch.tensor([ms1], dtype=torch.float64, device=dev)
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
        rows.append({"t_planning": t_hor, "ados": n, "particles": p, "evals_per_s": world * (1 << 18) / float(t_ms.item()) * 1e3,
                     "ms_per_launch": float(t_ms.item())})
    out["config5_subset"] = {"evals_per_shape_per_rank": 1 << 18, "shapes": rows}
    return out


_JSON_FD = None


def protect_stdout():
    """stdout carries the ONE JSON line and nothing else: native libraries (NCCL prints its version banner with NCCL_DEBUG set,
    whatever NCCL_DEBUG_FILE says), torch warnings and stray prints all write to file descriptor 1, so descriptor 1 is pointed
    at stderr for the whole run and the line is written to a private duplicate of the original stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scenes-per-gpu", type=int, default=8192)
    ap.add_argument("--candidates", type=int, default=N_CANDIDATES)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the measurements of the other BASELINE configs")
    args = ap.parse_args()
    protect_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        reference_arm(args)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:      # convenience: re-launch under torchrun as the driver would
        os.execvp(sys.executable, [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                   f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1", "--master-port", "29541",
                                   os.path.abspath(__file__), *sys.argv[1:]])
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    # CPU baseline first (fork-based pool, before CUDA is initialised in this process); rank 0 at N = 1 only.
    cpu_baseline, cpu_c5 = None, None
    if world == 1 and not args.no_cpu_baseline:
        v, cores, sample = cpu_arm(seconds_target=12.0)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        cpu_c5 = cpu_c5_subset()

    import torch
    import torch.distributed as dist
    from mantrap_b200 import api, build, engine, scenarios
    build.build()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL writes its banner ("NCCL version ...") and warnings to stdout by default: send them to stderr so that stdout
        # carries the ONE JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    eng = engine.Engine(dev)

    s_g, c, t_hor = args.scenes_per_gpu, args.candidates, T_HORIZON
    scene = scenarios.synthetic_scenes(s_g, N_ADOS, N_PARTICLES, seed=1234 + rank, param_seed=5678 + rank).to(dev)
    controls_host = scenarios.candidate_controls(s_g, c, t_hor, seed=4321 + rank, pinned=True)
    controls = controls_host.to(dev)
    mus_wo = eng.rollout_wo_ego(scene, t_hor)["mus"].contiguous()
    spec = engine.ModuleSpec(weights=WEIGHTS, mus_wo=mus_wo, control_limits=True, speed_limits=True)
    want = ("total", "grad", "violation")
    stats = torch.zeros(4, dtype=torch.int64, device=dev)
    bufs = eng.evaluate(scene, controls, spec, want=want, stats=stats)          # allocates outputs; counts pairs once
    torch.cuda.synchronize(dev)
    pair_stats = stats.tolist()
    from mantrap_b200 import parallel
    launches_per_step = 2

    def step(kernel_events=None):
        if kernel_events is not None:
            kernel_events[0].record()
        eng.evaluate(scene, controls, spec, want=want, out=bufs)
        if kernel_events is not None:
            kernel_events[1].record()
        idx, val, best = eng.argmin(bufs["total"], bufs["violation"], controls)
        if world > 1:   # the only exchange on the path: every rank learns every scene's best candidate (one all-gather)
            val, idx, best = parallel.gather_best_scene_partitioned(val, idx, best, s_g * world)
        return idx, val, best

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clocks:          # started before the warm-up: see the class docstring
        for _ in range(args.warmup):
            step()
        barrier()
        clocks.mark_start()
        start.record()
        for k in range(args.steps):
            step(ev[k])
        stop.record()
        barrier()
        clocks.mark_stop()
    ms_total = start.elapsed_time(stop)
    kernel_ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    t_ms = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(t_ms.item()) / args.steps
    evals_per_step = s_g * c * world
    value = evals_per_step / (ms_per_step * 1e-3)

    # ---- end to end through the public host-buffer API: pinned controls in, results (objective, violation,
    # gradient per evaluation + best candidate per scene) back to pinned host memory, every step.
    # Consecutive batches are software-pipelined (api.HostPipeline: copy-in of batch i+1 and copy-out of batch i-1 overlap the
    # kernels of batch i on three streams); every step still moves ALL of its inputs and results across PCIe inside the timed
    # region, and the region ends only when the last result is readable on the host.
    pipe = api.HostPipeline(eng, scene, spec, s_g, c, t_hor, depth=2)
    host_out = pipe.host[0]

    def e2e_step():
        slot = pipe.submit(controls_host)
        if world > 1:
            with torch.cuda.stream(pipe.s_run):
                idx, val, best = pipe.best_device(slot)
                parallel.gather_best_scene_partitioned(val, idx, best, s_g * world)
        return slot

    for _ in range(2):
        e2e_step()
    pipe.drain()
    barrier()
    e_start, e_stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e_start.record()
    pipe.s_h2d.wait_event(e_start)
    for _ in range(args.steps):
        slot = e2e_step()
    pipe.wait(slot)
    pipe.drain()
    e_stop.record()
    barrier()
    e_ms = torch.tensor([e_start.elapsed_time(e_stop)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    e2e_value = evals_per_step / (float(e_ms.item()) / args.steps * 1e-3)

    multi = secondary_multi_gpu(eng, torch, dist, dev, rank, world) if (world > 1 and not args.no_secondary) else None
    if rank == 0:
        peaks = measure_peaks(eng, torch)
        sm_clock = None
        try:
            measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            hbm_peak, peak_src = measured["hbm_gbs"], "MEASURED_PEAKS.json"
            sm_clock = measured.get("sm_max_mhz")
        except Exception:
            hbm_peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        spec_fp32 = 2 * 148 * 128 * (sm_clock or 1965.0) * 1e6 / 1e12
        fp32_peak = min(spec_fp32, peaks["fp32_tflops"])
        evals_per_launch = s_g * c
        # Algorithmic work: 150 FLOP per EXECUTED pair term (forward 50 incl. 7 MUFU + reverse 100; pairs
        # outside the 2 m interaction radius are skipped by the reference as well, social_forces.py:152, and are NOT
        # counted), 150 FLOP per particle-step (goal force, clamp, integration, mean/variance collapse, fwd + reverse)
        # and 5 FLOP per (particle, ado, step) for the interaction-radius test (forward only).
        steps_pp = (t_hor - 1) * N_ADOS * N_PARTICLES
        exec_pairs = (pair_stats[0] + pair_stats[1]) / evals_per_launch
        f_eval = 150.0 * exec_pairs + 150.0 * steps_pp + 5.0 * steps_pp * N_ADOS
        sfu_eval = 3.0 * (7.0 * exec_pairs + 5.0 * steps_pp)
        achieved = evals_per_launch * f_eval / (kernel_ms * 1e-3) / 1e12
        model_tflops = evals_per_launch * flops_per_eval(t_hor, N_ADOS, N_PARTICLES) / (kernel_ms * 1e-3) / 1e12
        hbm_achieved = evals_per_launch * bytes_per_eval(t_hor, 3) / (kernel_ms * 1e-3) / 1e9
        traffic, traffic_src = None, None
        try:        # DRAM bytes per evaluation from the committed ncu --set full capture of this kernel, scaled to this launch
            prof = json.load(open(os.path.join(ROOT, "profiles", "evaluate_pair_kernel_ncu.json")))
            traffic = prof["dram_bytes_per_eval"] * evals_per_launch
            traffic_src = prof["source"]
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": host_out.h2d_bytes(controls_host),
                    "d2h_bytes_per_step": host_out.d2h_bytes()},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks.summary(),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp32_peak, "traffic": traffic, "traffic_source": traffic_src,
                         "note": "compute-bound path (FP32 CUDA cores + MUFU; no contraction, HBM ~1e-3 of peak): the bound is "
                                 "the FP32 pipe, not hbm/tensor",
                         "kernel": "mtb::evaluate_pair_kernel<SF, GRAD> (two candidates per thread, packed FP32)", "kernel_ms": kernel_ms,
                         "kernel_share_of_step": kernel_ms / ms_per_step,
                         "work_model": "executed work: 150 FLOP x executed pair terms + 150 FLOP x particle-steps + 5 FLOP x "
                                       "radius tests (DESIGN.md §6)",
                         "flop_per_eval": f_eval, "executed_pairs_per_eval": exec_pairs,
                         "peak_source": f"min(2*148*128*f_max = {spec_fp32:.1f}, measured FMA probe = {peaks['fp32_tflops']:.1f}) TFLOP/s",
                         "mufu_frac": evals_per_launch * sfu_eval / (kernel_ms * 1e-3) / 1e12 / peaks["mufu_tops"],
                         "mufu_peak_tops": peaks["mufu_tops"],
                         "survey_model": {"flop_per_eval": flops_per_eval(t_hor, N_ADOS, N_PARTICLES), "tflops": model_tflops,
                                          "frac": model_tflops / fp32_peak,
                                          "note": "SURVEY 8(d) upper-bound model counting all N+1 pair terms per particle-step, "
                                                  "including the out-of-radius pairs nobody evaluates; reported for reference only"},
                         "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                                 "peak_source": peak_src}},
            "saturated_fraction": pair_stats[2] / evals_per_launch,
            "launch": eng.launch_info(scene, controls, True),
        }
        if world == 1 and not args.no_secondary:
            line["secondary"] = secondary_measurements(eng, torch, dev, cpu_c5)
        if multi is not None:
            line["secondary"] = multi
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        emit_line(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
