"""Multi-GPU execution of the batched hot path: one process per GPU, `torch.distributed` for the plumbing.

Every (scene, candidate) evaluation is independent: there is NO collective on the data path. The only
exchange is the tiny result of the per-scene best-candidate search (`solver/baselines/random_search.py:28-35` keeps the
best feasible candidate; the reference has no distributed code at all):

* scene-partitioned (BASELINE config 4): rank r owns the scenes `shard_range(S, r, world)` with all their candidates;
  the per-scene argmin is local; ONE all-gather of (value, index, controls) rows makes every rank know every scene's
  winner. Payload S * (2 + 2T) * 4 bytes (6.8 MB for S = 65536, T = 12): latency-bound over NVLink / NVSwitch.
* candidate-partitioned (BASELINE config 2: one scene, thousands of candidates): rank r owns a slice of the candidates
  of EVERY scene; the local winners are merged by an order-preserving 64-bit key (value bits << 32 | global candidate
  index) so that ties resolve to the lowest global index on every rank identically.

Works with any backend: NCCL on the GPUs, gloo in the CPU tests (world_size 2).
"""
import typing

import torch
import torch.distributed as dist


def shard_range(num_items: int, rank: int, world: int) -> typing.Tuple[int, int]:
    """Contiguous balanced shard [lo, hi) of `num_items` for `rank` (the first `num_items % world` ranks get one more)."""
    base, extra = divmod(num_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def _world(group=None) -> typing.Tuple[int, int]:
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def pack_best(best_val: torch.Tensor, best_idx: torch.Tensor, best_controls: torch.Tensor) -> torch.Tensor:
    """[S] value, [S] int32 index, [S,T,2] controls -> [S, 2 + 2T] float32 rows. The index column carries the int32 BITS
    (a bit-cast view, exact for every index; a float conversion would round global candidate indices >= 2^24)."""
    idx_bits = best_idx.to(torch.int32).contiguous().view(torch.float32)
    return torch.cat((best_val[:, None], idx_bits[:, None], best_controls.reshape(best_controls.shape[0], -1)), dim=1)


def unpack_best(packed: torch.Tensor, t_horizon: int):
    return packed[:, 0].contiguous(), packed[:, 1].contiguous().view(torch.int32), packed[:, 2:].reshape(-1, t_horizon, 2).contiguous()


def gather_best_scene_partitioned(best_val: torch.Tensor, best_idx: torch.Tensor, best_controls: torch.Tensor,
                                  num_scenes_total: int, group=None):
    """Scene-partitioned search: every rank contributes the winners of ITS scenes (shard_range order) and receives the
    winners of ALL scenes: (val [S], idx [S], controls [S,T,2]). One all-gather; ragged shards are padded to the largest."""
    rank, world = _world(group)
    t_horizon = best_controls.shape[1]
    packed = pack_best(best_val, best_idx, best_controls)
    if world == 1:
        return unpack_best(packed, t_horizon)
    sizes = [shard_range(num_scenes_total, r, world) for r in range(world)]
    width = max(hi - lo for lo, hi in sizes)
    assert packed.shape[0] == sizes[rank][1] - sizes[rank][0], "local winners do not match this rank's scene shard"
    if packed.shape[0] < width:
        packed = torch.cat((packed, packed.new_zeros((width - packed.shape[0], packed.shape[1]))))
    gathered = packed.new_empty((world * width, packed.shape[1]))
    dist.all_gather_into_tensor(gathered, packed.contiguous(), group=group)
    rows = torch.cat([gathered[r * width:r * width + (hi - lo)] for r, (lo, hi) in enumerate(sizes)])
    return unpack_best(rows, t_horizon)


def _order_key(val: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """Order-preserving float32 -> uint32 bits, packed with the candidate index into int64 (minimum = best value, then
    lowest index). Infeasible (idx < 0) -> int64 max."""
    bits = val.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    bits = torch.where(bits >= 0x80000000, (~bits) & 0xFFFFFFFF, bits | 0x80000000)
    key = ((bits - 0x80000000) << 32) | idx.to(torch.int64).clamp(min=0)       # signed-safe: shift the bits range
    return torch.where(idx < 0, torch.full_like(key, torch.iinfo(torch.int64).max), key)


def reduce_best_candidate_partitioned(best_val: torch.Tensor, best_idx_local: torch.Tensor, best_controls: torch.Tensor,
                                      candidate_offset: int, group=None):
    """Candidate-partitioned search: every rank holds, per scene, the winner among ITS candidates (local index, -1 when
    none is feasible). Returns the global winner per scene on every rank: (val [S], global idx [S], controls [S,T,2])."""
    rank, world = _world(group)
    t_horizon = best_controls.shape[1]
    global_idx = torch.where(best_idx_local >= 0, best_idx_local + candidate_offset, best_idx_local)
    if world == 1:
        return best_val, global_idx, best_controls
    key = _order_key(best_val, global_idx)
    keys = key.new_empty(world * key.shape[0])                                     # concatenated along dim 0
    dist.all_gather_into_tensor(keys, key.contiguous(), group=group)
    keys = keys.view(world, key.shape[0])
    packed = pack_best(best_val, global_idx, best_controls)
    gathered = packed.new_empty((world * packed.shape[0], packed.shape[1]))
    dist.all_gather_into_tensor(gathered, packed.contiguous(), group=group)
    gathered = gathered.view(world, *packed.shape)
    winner = keys.argmin(dim=0)                                                    # [S] rank holding the best candidate
    rows = gathered[winner, torch.arange(key.shape[0], device=key.device)]
    val, idx, controls = unpack_best(rows, t_horizon)
    none = keys.min(dim=0).values == torch.iinfo(torch.int64).max
    return (torch.where(none, torch.full_like(val, float("inf")), val), torch.where(none, torch.full_like(idx, -1), idx), controls)


def slice_scene(scene, lo: int, hi: int):
    """Scenes [lo, hi) of a SceneBatch (any device)."""
    import dataclasses
    fields = ("ego0", "ego_goal", "ado0", "ado_goal", "v0", "sigma", "weight", "noise")
    return dataclasses.replace(scene, **{k: (getattr(scene, k)[lo:hi].contiguous() if getattr(scene, k) is not None else None)
                                         for k in fields})


def sharded_search(engine, scene, controls: torch.Tensor, make_spec: typing.Callable, feas_tol: float, mode: str = "scenes",
                   group=None):
    """Best feasible candidate per scene over all ranks. `scene` / `controls` [S,C,T,2] describe the WHOLE problem (host
    or device); each rank evaluates only its shard. `make_spec(scene_shard) -> ModuleSpec` builds the module
    specification for the shard (e.g. runs the ego-free baseline rollout). Returns (val [S], idx [S], controls [S,T,2])
    identical on every rank, plus the local evaluation result dictionary."""
    rank, world = _world(group)
    s_total, c_total = controls.shape[0], controls.shape[1]
    if mode == "scenes":
        lo, hi = shard_range(s_total, rank, world)
        scene_r = slice_scene(scene, lo, hi).to(engine.device)
        controls_r = controls[lo:hi].to(engine.device).contiguous()
        offset = 0
    elif mode == "candidates":
        lo, hi = shard_range(c_total, rank, world)
        scene_r = scene.to(engine.device)
        controls_r = controls[:, lo:hi].to(engine.device).contiguous()
        offset = lo
    else:
        raise ValueError(mode)
    res = engine.evaluate(scene_r, controls_r, make_spec(scene_r), want=("total", "violation", "grad"))
    idx, val, best = engine.argmin(res["total"], res["violation"], controls_r, feas_tol)
    if mode == "scenes":
        out = gather_best_scene_partitioned(val, idx, best, s_total, group)
    else:
        out = reduce_best_candidate_partitioned(val, idx, best, offset, group)
    return (*out, res)


def sharded_closed_loop(engine, scene, time_steps: int, candidates: typing.Sequence[torch.Tensor] = None,
                        eps: typing.Sequence[torch.Tensor] = None, group=None, **loop_kwargs):
    """Scene-partitioned batched closed-loop MPC (`mpc.BatchedClosedLoop`, i.e. `TrajOptSolver.solve` for many scenes):
    rank r advances the scenes `shard_range(S, r, world)` for `time_steps` steps with NO exchange during the loop (scenes
    are independent), then ONE all-gather returns the trajectories of all scenes to every rank. `scene`, `candidates[k]`
    [S,C,T,2] and `eps[k]` [S,N,2] describe the whole problem; with `candidates` / `eps` omitted every rank draws its own
    (seeded by `seed` + rank). Returns a `ClosedLoopResult` over all S scenes."""
    from . import mpc
    rank, world = _world(group)
    s_total = scene.num_scenes
    lo, hi = shard_range(s_total, rank, world)
    kwargs = dict(loop_kwargs)
    kwargs["seed"] = kwargs.get("seed", 0) + rank
    loop = mpc.BatchedClosedLoop(engine, slice_scene(scene, lo, hi), **kwargs)
    out = loop.solve(time_steps, candidates=None if candidates is None else [c[lo:hi] for c in candidates],
                     eps=None if eps is None else [e[lo:hi] for e in eps])
    if world == 1:
        return out
    # one flat float32 row per scene: ego trajectory | ado trajectories | controls | objective | index | steps
    n_loc = hi - lo
    assert int(out.best_index.max()) < 2 ** 24 and time_steps < 2 ** 24, "float32 transport rows hold integers exactly below 2^24"
    rows = torch.cat([out.ego_trajectory.reshape(n_loc, -1), out.ado_trajectories.reshape(n_loc, -1), out.controls.reshape(n_loc, -1),
                      out.objective, out.best_index.to(torch.float32), out.steps.to(torch.float32)[:, None]], dim=1)
    sizes = [shard_range(s_total, r, world) for r in range(world)]
    width = max(b - a for a, b in sizes)
    if n_loc < width:
        rows = torch.cat((rows, rows.new_zeros((width - n_loc, rows.shape[1]))))
    gathered = rows.new_empty((world * width, rows.shape[1]))
    dist.all_gather_into_tensor(gathered, rows.contiguous(), group=group)
    rows = torch.cat([gathered[r * width:r * width + (b - a)] for r, (a, b) in enumerate(sizes)])
    k, n = time_steps, scene.num_ados
    cuts = [4 * (k + 1), 4 * n * (k + 1), 2 * k, k, k, 1]
    ego, ado, ctl, obj, idx, steps = torch.split(rows, cuts, dim=1)
    return mpc.ClosedLoopResult(ego.reshape(s_total, k + 1, 4).contiguous(), ado.reshape(s_total, n, k + 1, 4).contiguous(),
                                ctl.reshape(s_total, k, 2).contiguous(), obj.contiguous(), idx.to(torch.int32), steps[:, 0].to(torch.int32))
