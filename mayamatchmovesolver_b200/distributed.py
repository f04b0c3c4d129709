"""Frame-sharded solves across the GPUs of one box: one process per GPU.

What shards: a frame's observations touch only that
frame's animated parameters (adjust_relationships.cpp:565-617), so
  * per-frame mode (solve_v1's loop, adjust_base.cpp:1430-1484): every rank
    solves its own contiguous frame range, no collective on the data path;
    the per-frame results are gathered and merged with CommandResult::add /
    divide (adjust_results.h:89-117);
  * global mode: every rank runs the same LM control law; the reduced system of
    the static parameters is all-reduced over NCCL once per lmpar trial inside
    libmmsolver_b200.so (csrc/lm_global.cu).

The process group is only plumbing (exchange of the 128-byte communicator id
and the final gather of solved values): any object with `rank`, `world`,
`broadcast_bytes(b, src)` and `allgather(obj) -> list` works.  `TorchGroup`
adapts torch.distributed (gloo on CPU, nccl on GPUs); torch is imported only
there.
"""
import numpy as np

from . import _capi
from .problem import Comm, Problem, SolveResult, aggregate_results, comm_unique_id, shard_frames, shard_param_map  # noqa: F401


class TorchGroup:
    """torch.distributed as the plumbing group"""

    def __init__(self):
        import torch.distributed as dist
        self._dist = dist
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()

    def broadcast_bytes(self, b, src=0):
        objs = [b if self.rank == src else None]
        self._dist.broadcast_object_list(objs, src=src)
        return objs[0]

    def allgather(self, obj):
        out = [None] * self.world
        self._dist.all_gather_object(out, obj)
        return out


class SingleGroup:
    rank, world = 0, 1

    def broadcast_bytes(self, b, src=0):
        return b

    def allgather(self, obj):
        return [obj]


def _result_tuple(r):
    return tuple(getattr(r, name) if name != "reason_str" else None for name, _ in _capi.Result._fields_)


def _result_from_tuple(t):
    r = _capi.Result()
    for (name, _), v in zip(_capi.Result._fields_, t):
        if name != "reason_str":
            setattr(r, name, v)
    return r


def merge_shard_vectors(n_global, maps, values, fill=0.0):
    """scatter the shards' local parameter vectors into the unsharded order; static
    parameters (replicated) are taken from the last shard that holds them"""
    out = np.full(n_global, fill, dtype=np.float64)
    for m, v in zip(maps, values):
        out[np.asarray(m)] = np.asarray(v)
    return out


def solve_per_frame_sharded(scene, group, device=0, options=None, **opt_overrides):
    """frameSolveMode=1 over `group`: -> (aggregate SolveResult with unsharded x / attr values,
    [per-frame SolveResult] in frame order).  Identical on every rank."""
    opt_overrides.setdefault("frame_solve_mode", 1)
    with Problem(scene, options=options, shard=(group.rank, group.world), device=device, **opt_overrides) as p:
        agg, per = p.solve_per_frame()
        e2m, _ = p.index_maps()
        payload = dict(map=p.param_map, x=agg.x, av=agg.attr_values, f0=p.frame_begin,
                       per=[_result_tuple(r.raw) for r in per],
                       obs=(e2m[:, 0].copy(), e2m[:, 1] + p.frame_begin, agg.error_dist),
                       seconds=agg.timer_solve_s)
    parts = group.allgather(payload)
    parts.sort(key=lambda d: d["f0"])
    n_global = max(int(d["map"].max()) for d in parts) + 1
    x = merge_shard_vectors(n_global, [d["map"] for d in parts], [d["x"] for d in parts])
    av = merge_shard_vectors(n_global, [d["map"] for d in parts], [d["av"] for d in parts])
    raw = [_result_from_tuple(t) for d in parts for t in d["per"]]
    total = aggregate_results(raw)
    total.timer_solve_s = max(d["seconds"] for d in parts)  # ranks run concurrently
    # observations in the unsharded errorToMarkerList order: marker-major, frame-minor
    mk = np.concatenate([d["obs"][0] for d in parts])
    fr = np.concatenate([d["obs"][1] for d in parts])
    dist = np.concatenate([d["obs"][2] for d in parts])
    order = np.lexsort((fr, mk))
    result = SolveResult(total, x, av, dist[order])
    return result, [SolveResult(r, None, None, None) for r in raw]


def make_comm(group, device=0, local_backend=False):
    """collective: a long-lived communicator over `group` (None for a single rank)"""
    if group.world == 1:
        return None
    cid = comm_unique_id(local_backend) if group.rank == 0 else None
    cid = group.broadcast_bytes(cid, src=0)
    return Comm(cid, group.rank, group.world, device)


def solve_global_sharded(scene, group, device=0, options=None, local_backend=False, comm=None, local_slice=None,
                         gather=True, **opt_overrides):
    """frameSolveMode=0 over `group`: the sharded global solve.  -> SolveResult with unsharded
    x / attr values / per-observation distances, identical on every rank.  `comm`: a
    communicator from make_comm() to reuse; else one is made for this solve.
    `local_slice=(frame_begin, frames_global)`: `scene` holds only this rank's frames.
    `gather=False` skips the merge of the per-rank vectors (-> this rank's own SolveResult)."""
    cid = None
    if comm is None and group.world > 1:
        cid = comm_unique_id(local_backend) if group.rank == 0 else None
        cid = group.broadcast_bytes(cid, src=0)
    with Problem(scene, options=options, shard=(group.rank, group.world), local_slice=local_slice, device=device,
                 **opt_overrides) as p:
        if comm is not None:
            p.use_comm(comm)
        elif group.world > 1:
            p.attach_comm(cid, group.rank, group.world)
        r = p.solve()
        if not gather:
            return r
        e2m, _ = p.index_maps()
        payload = dict(map=p.param_map, x=r.x, av=r.attr_values, f0=p.frame_begin,
                       obs=(e2m[:, 0].copy(), e2m[:, 1] + p.frame_begin, r.error_dist))
    parts = group.allgather(payload)
    parts.sort(key=lambda d: d["f0"])
    n_global = max(int(d["map"].max()) for d in parts) + 1
    x = merge_shard_vectors(n_global, [d["map"] for d in parts], [d["x"] for d in parts])
    av = merge_shard_vectors(n_global, [d["map"] for d in parts], [d["av"] for d in parts])
    mk = np.concatenate([d["obs"][0] for d in parts])
    fr = np.concatenate([d["obs"][1] for d in parts])
    dist = np.concatenate([d["obs"][2] for d in parts])
    order = np.lexsort((fr, mk))
    return SolveResult(r.raw, x, av, dist[order])
