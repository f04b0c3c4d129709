"""Jacobian kernel time vs problem size (dev tool): separates fixed launch/prologue cost from per-observation cost."""
import sys
from mayamatchmovesolver_b200 import Problem
from mayamatchmovesolver_b200.synth import make_scene

sizes = [(1000, 64), (1000, 500), (1000, 2000)] if len(sys.argv) < 2 else [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for F, B in sizes:
    sb, _ = make_scene(F, B, 3)
    with Problem(sb) as p:
        p.bench_eval(5, True)
        ms, msj, _ = p.bench_eval(20, True)
        ms2, msj2, _ = p.bench_eval(20, False)
    nobs = F * B
    print("F=%d B=%d nobs=%d: jac %.2f us (%.0f GB/s), eval %.2f us; no-flush jac %.2f us" % (
        F, B, nobs, msj * 1e3, nobs * 192 / msj / 1e6, ms * 1e3, msj2 * 1e3), flush=True)
