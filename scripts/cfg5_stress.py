"""cfg5 (BASELINE.json configs[4]): 8 witness cameras, 10 000 frames x 5 000 bundles, one
global solve with the frames sharded over the GPUs of the box and the reduced (static)
system all-reduced over NCCL once per lmpar trial.  Each rank generates and holds only its
own frames (4e8 observations do not fit one host comfortably).

  torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/cfg5_stress.py
      [--cameras 8 --frames 10000 --bundles 5000 --iters 10 --out gpurun_out/cfg5.json]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.distributed import SingleGroup, TorchGroup, make_comm  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem, shard_frames  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cameras", type=int, default=8)
ap.add_argument("--frames", type=int, default=10000)
ap.add_argument("--bundles", type=int, default=5000)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--out", default="gpurun_out/cfg5.json")
a = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = TorchGroup()
else:
    g = SingleGroup()


def allmax(v):
    if world == 1:
        return float(v)
    t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allsum(v):
    if world == 1:
        return float(v)
    t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


f0, f1 = shard_frames(a.frames, g.world, g.rank)
t = time.perf_counter()
sb, _ = synth.make_multicam_scene(a.cameras, a.frames, a.bundles, 5, frame_range=(f0, f1))
sb.arrays()
t_scene = time.perf_counter() - t
comm = make_comm(g, device=local)
t = time.perf_counter()
with Problem(sb, shard=(g.rank, g.world), local_slice=(f0, a.frames), device=local, iter_max=a.iters) as p:
    if comm is not None:
        p.use_comm(comm)
    t_create = time.perf_counter() - t
    ms_eval, ms_jac, _ = p.bench_eval(reps=5)
    nobs = allsum(p.nobs)
    ms_eval = allmax(ms_eval)
    solves = []
    for rep in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t = time.perf_counter()
        r = p.solve()
        dt = allmax(time.perf_counter() - t)
        solves.append(dict(seconds=dt, reason_num=r.reason_num, nfev=r.iteration_num, njev=r.iteration_attempt_num,
                           error_final=r.error_final, error_avg_px=r.error_avg, error_max_px=r.error_max,
                           timer_solve_s=allmax(r.timer_solve_s), timer_jacobian_s=allmax(r.timer_jacobian_s),
                           timer_function_s=allmax(r.timer_function_s), timer_linear_s=allmax(r.timer_linear_s)))
    n_local = p.n
    stats = p.solve_stats()
t_scene, t_create = allmax(t_scene), allmax(t_create)
if g.rank == 0:
    n_static = 3 * a.bundles
    out = dict(workload="cfg5: %d witness cameras, %d frames x %d bundles, global solve, frames sharded over %d GPU(s)"
                        % (a.cameras, a.frames, a.bundles, world),
               n_gpus=world, observations=int(nobs), parameters=6 * a.cameras * a.frames + n_static,
               parameters_per_rank=n_local, static_parameters=n_static,
               eval_ms=ms_eval, evals_per_s=nobs / ms_eval * 1e3, scene_gen_s=t_scene, create_s=t_create,
               iter_max=a.iters, solves=solves, host_syncs_graph_launches=stats,
               reduced_system_allreduce_bytes=8 * sum(64 * 64 * (b + 1) for b in range((n_static + 1 + 63) // 64)))  # packed lower triangle
    print(json.dumps(out))
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    json.dump(out, open(a.out, "w"), indent=1)
if comm is not None:
    comm.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
