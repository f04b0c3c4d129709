"""Workloads for the round-2 ncu captures (one evaluation step each, after a warm-up step):
  hbm   cfg3 x 16 (16 000 frames x 500 bundles, 1.54 GB per launch)    -> k_jacobian_rows
  mc    cfg5 shape, 8 cameras x 1250 frames x 1000 bundles (1.92 GB)   -> k_jacobian_rows, one launch, blockIdx.z = camera
  mcs   the same with Bernoulli(0.25) visibility and 4000 bundles: 10^7 observations in 4*10^7 slots -> k_jacobian_rows<SPARSE>
  cfg3  the bench workload (96 MB per launch)
usage: profile_r2.py hbm|mc|mcs|cfg3"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mayamatchmovesolver_b200 import synth  # noqa: E402
from mayamatchmovesolver_b200.problem import Problem  # noqa: E402

what = sys.argv[1]
if what == "hbm":
    sb, _ = synth.make_scene(16000, 500, 3)
elif what == "mc":
    sb, _ = synth.make_multicam_scene(8, 1250, 1000, 5)
elif what == "mcs":
    sb, _ = synth.make_multicam_scene(8, 1250, 4000, 5, visibility=0.25)
else:
    sb, _ = synth.make_scene(1000, 500, 3)
with Problem(sb) as p:
    for _ in range(3):
        ms = p.bench_eval(1, True)
    print(what, "nobs", p.nobs, "k", p.jacobian_kmax(), "ms (step, kernel)", ms[:2])
