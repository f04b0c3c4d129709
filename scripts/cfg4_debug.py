import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem
LENS = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01, degree4_u=0.0, degree4_v=0.0)
for (F, B, kw) in [(30, 30, {}), (60, 40, dict(period=60)), (100, 100, {}), (300, 200, {}), (1000, 500, {})]:
    sb, _ = synth.make_scene(F, B, 4, noise_px=0.5, lens=LENS, solve_focal=True, solve_lens=tuple(LENS), in_frame=True, focal_init=float(os.environ.get("FOCAL_INIT", "33.0")), fixed_bundles=int(os.environ.get("FIXED", "0")), **kw)
    with Problem(sb, iter_max=100) as p:
        x0 = p.initial_parameters()
        f0, d0 = p.eval_residuals(x0)
        r = p.solve()
    print(F, B, kw, "init avg %.3f" % d0.mean(), r, "iters", r.iteration_attempt_num, "nan in dist:", np.isnan(r.error_dist).sum())
    print("  focal/lens:", r.attr_values[6 * F:6 * F + 7])
