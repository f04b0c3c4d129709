"""diagnostic: GPU vs block-sparse oracle on the cfg5 subsample (which attributes / distances differ)"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem
from oraclelib import OracleProblem

for vis, fixed in ((0.25, 6), (0.25, 20), (0.5, 6)):
    sb, _ = synth.make_multicam_scene(8, 64, 200, 5, visibility=vis, fixed_bundles=fixed)
    ro = OracleProblem(sb).solve_sparse(threads=8)
    with Problem(sb) as p:
        r = p.solve()
        r2 = p.solve()
        _, p2a = p.index_maps()
    sc = np.maximum(np.abs(ro.attr_values), 1.0)
    d = np.abs(r.attr_values - ro.attr_values) / sc
    k = np.argsort(d)[-5:]
    print("vis", vis, "fixed", fixed, ro, r)
    print("  attr rel max", d.max(), "worst", [(int(i), p2a[i].tolist(), float(d[i])) for i in k])
    print("  dist max", np.abs(r.error_dist - ro.error_dist).max(), "median attr", np.median(d))
    print("  run-to-run attr", np.abs(r.attr_values - r2.attr_values).max(), "dist", np.abs(r.error_dist - r2.error_dist).max())
