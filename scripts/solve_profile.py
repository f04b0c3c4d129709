"""One cfg3-shaped global solve; MMB_LM_PROFILE=1 turns the LM phase timers on (host-driven loop)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 500
it = int(sys.argv[3]) if len(sys.argv) > 3 else 20
sb, _ = synth.make_scene(F, B, 3, noise_px=0.5)
for rep in range(2):
    t0 = time.perf_counter()
    with Problem(sb, iter_max=it) as p:
        t1 = time.perf_counter()
        r = p.solve()
        t2 = time.perf_counter()
        n = p.launch_count()
        stats = p.solve_stats()
    print("create %.4f s, solve %.4f s (jac %.4f, func %.4f, linear %.4f), launches %d: %r, LM iterations %d, host syncs %d, graph launches %d" %
          (t1 - t0, t2 - t1, r.timer_jacobian_s, r.timer_function_s, r.timer_linear_s, n, r, r.iteration_attempt_num, stats[0], stats[1]))
