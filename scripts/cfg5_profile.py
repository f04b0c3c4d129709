"""cfg5-shaped solve on ONE GPU at a reduced size: per-phase times of the global LM
(MMB_LM_PROFILE host loop) to see how the kernels behave with 48-wide frame blocks and a
large static system.  usage: cfg5_profile.py C F B [iters [visibility]]"""
import os
import sys
import time

sys.path.insert(0, ".")
import numpy as np  # noqa: E402

from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

C, F, B = [int(v) for v in sys.argv[1:4]]
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 8
vis = float(sys.argv[5]) if len(sys.argv) > 5 else 1.0
t = time.perf_counter()
sb, _ = synth.make_multicam_scene(C, F, B, 5, visibility=vis)
sb.arrays()
print("scene s", time.perf_counter() - t, flush=True)
t = time.perf_counter()
with Problem(sb, iter_max=iters) as p:
    print("create s", time.perf_counter() - t, "n", p.n, "nobs", p.nobs, flush=True)
    ms = p.bench_eval(reps=5)
    print("eval (ms per step, ms Jacobian kernel)", ms, "evals/s", p.nobs / ms[0] * 1e3, flush=True)
    for rep in range(2):
        t = time.perf_counter()
        r = p.solve()
        print("solve s", time.perf_counter() - t, r, "timers", r.timer_solve_s, r.timer_jacobian_s, r.timer_function_s,
              r.timer_linear_s, "njev", r.iteration_attempt_num, "stats", p.solve_stats(), flush=True)
