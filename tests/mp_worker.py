"""Worker of the multi-process tests (one process per rank, torch.distributed).

  cpu : gloo, world 2 -- host logic of the sharded path, no GPU
  gpu : nccl, one GPU per rank -- sharded global solve (NCCL all-reduce of the reduced
        system) and sharded per-frame solve against the unsharded solve on rank 0
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from mayamatchmovesolver_b200 import _capi, synth  # noqa: E402
from mayamatchmovesolver_b200.distributed import (TorchGroup, make_comm, merge_shard_vectors,  # noqa: E402
                                                  solve_global_sharded, solve_per_frame_sharded)
from mayamatchmovesolver_b200.problem import (Problem, aggregate_results, local_slice_param_map, shard_frames,  # noqa: E402
                                              shard_param_map)


def cpu_mode():
    dist.init_process_group("gloo")
    g = TorchGroup()
    F, B = 13, 5
    sb, _ = synth.make_scene(F, B, seed=11)
    n = 6 * F + 3 * B
    f0, f1 = shard_frames(F, g.world, g.rank)
    m = shard_param_map(sb, g.world, g.rank)
    # every rank "solves" x[global id] = 10 * id on its shard
    parts = g.allgather(dict(map=m, x=10.0 * m, f0=f0))
    parts.sort(key=lambda d: d["f0"])
    x = merge_shard_vectors(n, [d["map"] for d in parts], [d["x"] for d in parts])
    np.testing.assert_array_equal(x, 10.0 * np.arange(n))
    # a rank that generated only its own frames of a multi-camera scene numbers its
    # parameters like the library's cut of the full scene
    C_, F_, B_ = 3, 11, 4
    full, _ = synth.make_multicam_scene(C_, F_, B_, 7)
    g0, g1 = shard_frames(F_, g.world, g.rank)
    mine, _ = synth.make_multicam_scene(C_, F_, B_, 7, frame_range=(g0, g1))
    np.testing.assert_array_equal(local_slice_param_map(mine, g0, F_), shard_param_map(full, g.world, g.rank))
    a_full, a_mine = full.arrays(), mine.arrays()
    np.testing.assert_array_equal(a_full["mkr_x"].reshape(-1, F_)[:, g0:g1], a_mine["mkr_x"].reshape(-1, g1 - g0))
    np.testing.assert_array_equal(a_full["anim_values"].reshape(-1, F_)[:, g0:g1], a_mine["anim_values"].reshape(-1, g1 - g0))
    # per-frame records: frame f has error_final = f + 1
    mine = []
    for f in range(f0, f1):
        r = _capi.Result()
        r.success, r.error_final, r.error_avg, r.error_min, r.error_max, r.iteration_num = 1, f + 1.0, 2.0 * f, f, f, 1
        r.error_is_better = 1
        mine.append(tuple(getattr(r, k) for k, _ in _capi.Result._fields_ if k != "reason_str"))
    gathered = [t for part in g.allgather(mine) for t in part]
    recs = []
    for t in gathered:
        r = _capi.Result()
        for (k, v) in zip([k for k, _ in _capi.Result._fields_ if k != "reason_str"], t):
            setattr(r, k, v)
        recs.append(r)
    a = aggregate_results(recs)
    assert a.iteration_num == F and abs(a.error_final - (F + 1) / 2.0) < 1e-12 and a.error_max == F - 1
    # the communicator id travels as bytes
    cid = g.broadcast_bytes(bytes(range(128)) if g.rank == 0 else None)
    assert cid == bytes(range(128))
    print("rank %d ok" % g.rank)
    dist.destroy_process_group()


def gpu_mode():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = TorchGroup()
    # global solve, gauge-fixed so the solved values are determined
    sb, _ = synth.make_scene(F=24, B=20, seed=21, fixed_bundles=4, period=24)
    r = solve_global_sharded(sb, g, device=local)
    # the same solve twice more through one session communicator (mmb_comm_create)
    comm = make_comm(g, device=local)
    for _ in range(2):
        rs = solve_global_sharded(sb, g, device=local, comm=comm)
        assert (rs.reason_num, rs.iteration_num) == (r.reason_num, r.iteration_num)
        np.testing.assert_allclose(rs.attr_values, r.attr_values, rtol=1e-7, atol=1e-7)
    # peer-memory all-reduce (CUDA IPC arenas over NVLink): the sharded solve is ONE graph launch and
    # one synchronisation, like the unsharded one -- and gives the host-driven loop's answer bit for bit
    f0, f1 = shard_frames(24, g.world, g.rank)
    with Problem(sb, device=local, shard=(g.rank, g.world)) as p:
        p.use_comm(comm)
        rg = p.solve()
        stats = p.solve_stats()
    os.environ["MMB_LM_GRAPH"] = "0"
    with Problem(sb, device=local, shard=(g.rank, g.world)) as p:
        p.use_comm(comm)
        rh = p.solve()
        stats_h = p.solve_stats()
    del os.environ["MMB_LM_GRAPH"]
    if os.environ.get("MMB_PEER", "1") != "0":
        assert stats == (1, 1), stats
    assert stats_h[1] == 0 and stats_h[0] > 2
    assert rg.iteration_num == rh.iteration_num and np.array_equal(rg.x, rh.x) and np.array_equal(rg.error_dist, rh.error_dist)
    comm.close()
    sb2, _ = synth.make_scene(F=30, B=16, seed=22, solve_bundles=False, period=30)
    rp, per = solve_per_frame_sharded(sb2, g, device=local)
    # cfg5 shape: witness cameras, every rank generates and holds only its own frames
    C5, F5, B5 = 4, 18, 16
    h0, h1 = shard_frames(F5, g.world, g.rank)
    mine, _ = synth.make_multicam_scene(C5, F5, B5, 3, frame_range=(h0, h1), fixed_bundles=4)
    r5 = solve_global_sharded(mine, g, device=local, local_slice=(h0, F5))
    if g.rank == 0:
        full5, _ = synth.make_multicam_scene(C5, F5, B5, 3, fixed_bundles=4)
        with Problem(full5, device=local) as p:
            u5 = p.solve()
        assert (r5.reason_num, r5.iteration_num, r5.iteration_attempt_num) == (u5.reason_num, u5.iteration_num, u5.iteration_attempt_num), (r5, u5)
        assert np.max(np.abs(r5.attr_values - u5.attr_values) / np.maximum(np.abs(u5.attr_values), 1.0)) <= 1e-6
        assert np.max(np.abs(r5.error_dist - u5.error_dist)) <= 1e-5
    if g.rank == 0:
        with Problem(sb, device=local) as p:
            r1 = p.solve()
        assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (r1.reason_num, r1.iteration_num, r1.iteration_attempt_num), (r, r1)
        scale = np.maximum(np.abs(r1.attr_values), 1.0)
        assert np.max(np.abs(r.attr_values - r1.attr_values) / scale) <= 1e-6
        assert np.max(np.abs(r.error_dist - r1.error_dist)) <= 1e-5
        assert abs(r.error_final - r1.error_final) <= 1e-6 * r1.error_final
        with Problem(sb2, device=local, frame_solve_mode=1) as p:
            q1, per1 = p.solve_per_frame()
        np.testing.assert_array_equal(rp.attr_values, q1.attr_values)  # independent solves: identical
        np.testing.assert_array_equal(rp.error_dist, q1.error_dist)
        assert [x.iteration_num for x in per] == [x.iteration_num for x in per1]
        assert abs(rp.error_final - q1.error_final) <= 1e-12 * q1.error_final and rp.iteration_num == q1.iteration_num
    print("rank %d ok" % g.rank)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    cpu_mode() if sys.argv[1] == "cpu" else gpu_mode()
