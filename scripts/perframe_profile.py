"""cfg2 shape (per-frame mode): create / solve / destroy split at 240 frames and at a large batch."""
import sys
import time

sys.path.insert(0, ".")
from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

for F in (240, 24000):
    sb, _ = synth.make_scene(F, 50, 2, solve_bundles=False)
    sb.arrays()
    for rep in range(3):
        t0 = time.perf_counter()
        p = Problem(sb, frame_solve_mode=1)
        t1 = time.perf_counter()
        r, per = p.solve_per_frame()
        t2 = time.perf_counter()
        p.close()
        t3 = time.perf_counter()
    print("F=%d: create %.3f ms, solve %.3f ms (device timer %.3f ms), destroy %.3f ms -> %.0f frames/s; iterations/frame %.1f, avg err %.3f px"
          % (F, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * r.timer_solve_s * 1, 1e3 * (t3 - t2), F / (t3 - t0), r.iteration_num / F, r.error_avg))
