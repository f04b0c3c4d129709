"""Small scenes through the newest kernels (sparse-visibility rows kernel, all cameras in one launch,
Cholesky beside the SYRK, two-half lmpar step) -- a quick end-to-end sanity run (also usable under a memory checker where one is available):
  compute-sanitizer --tool memcheck python scripts/sanitize_small.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

for vis in (1.0, 0.5):
    sb, _ = synth.make_multicam_scene(2, 64, 40, 5, visibility=vis)
    with Problem(sb, iter_max=4) as p:
        p.bench_eval(1, True)
        r = p.solve()
        print("multicam visibility", vis, r, p.solve_stats(), flush=True)
sb, _ = synth.make_scene(64, 30, 3, noise_px=0.5)
with Problem(sb, iter_max=6) as p:
    r = p.solve()
    print("single camera", r, p.solve_stats(), flush=True)
