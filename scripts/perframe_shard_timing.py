"""torchrun helper: create / solve / destroy split of the per-frame workload (24 000 frames per rank)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
F = 24000
sb, _ = synth.make_scene(F, 50, 2, solve_bundles=False, frame_offset=rank * F)
sb.arrays()
for rep in range(4):
    dist.barrier()
    t0 = time.perf_counter()
    p = Problem(sb, device=local, frame_solve_mode=1, local_slice=(rank * F, F * world)) if world > 1 else Problem(sb, device=local, frame_solve_mode=1)
    t1 = time.perf_counter()
    r, per = p.solve_per_frame()
    t2 = time.perf_counter()
    p.close()
    t3 = time.perf_counter()
    if rep == 3:
        print("rank %d: create %.2f ms, solve %.2f ms (device %.2f), destroy %.2f ms -> total %.2f ms"
              % (rank, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * r.timer_solve_s, 1e3 * (t3 - t2), 1e3 * (t3 - t0)), flush=True)
dist.destroy_process_group()
