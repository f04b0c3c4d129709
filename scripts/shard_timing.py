"""torchrun helper: where a sharded cfg3-per-GPU solve spends its wall time on every rank
(create / use_comm / solve / destroy), to explain the weak-scaling e2e number."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.distributed import TorchGroup, make_comm
from mayamatchmovesolver_b200.problem import Problem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
g = TorchGroup()
comm = make_comm(g, device=local)
F1, B = 1000, 500
for rep in range(4):
    sb, _ = synth.make_scene(F1, B, 200 + rep, noise_px=0.5, frame_offset=rank * F1)
    sb.arrays()
    dist.barrier()
    t0 = time.perf_counter()
    p = Problem(sb, device=local, iter_max=20, local_slice=(rank * F1, F1 * world)) if world > 1 else Problem(sb, device=local, iter_max=20)
    t1 = time.perf_counter()
    if world > 1:
        p.use_comm(comm)
    r = p.solve()
    t2 = time.perf_counter()
    p.close()
    t3 = time.perf_counter()
    if rep:
        print("rank %d rep %d: create %.2f ms, solve %.2f ms (device %.2f: jac %.2f lin %.2f func %.2f), destroy %.2f ms, iters %d"
              % (rank, rep, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * r.timer_solve_s, 1e3 * r.timer_jacobian_s, 1e3 * r.timer_linear_s,
                 1e3 * r.timer_function_s, 1e3 * (t3 - t2), r.iteration_attempt_num), flush=True)
if comm is not None:
    comm.close()
dist.destroy_process_group()
