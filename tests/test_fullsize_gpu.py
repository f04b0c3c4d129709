"""Parity at BASELINE.json's full sizes (cfg3: 1000 frames x 500 bundles, n = 7500; cfg4: the
same plus focal length and six 3DE radial-decentered deg-4 lens coefficients, n = 7507).

The oracle's own solve is dense (J = 60 GB at this size, the reference's cost structure), so
whole solves are compared with it at small sizes (test_solve_gpu.py) and here through
  * the full residual vector against the oracle's residual evaluation (500 000 observations),
  * Jacobian columns against forward differences of the ORACLE's residuals, built exactly as
    adjust_solveFunc.cpp:305-479 builds them, for parameters of every kind,
  * size-independent properties of the solve: noise-free markers are fitted to ~0 px and the
    gauge-invariant attributes (focal length, lens coefficients) return to their truth.
"""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.problem import Problem
from oraclelib import OracleProblem

pytestmark = pytest.mark.gpu

LENS = dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01,
            degree4_u=0.0, degree4_v=0.0)
F, B = 1000, 500


def cfg3(**kw):
    return synth.make_scene(F, B, 3, **kw)


def cfg4(**kw):
    return synth.make_scene(F, B, 4, lens=LENS, solve_focal=True, solve_lens=tuple(LENS), focal_init=33.0,
                            in_frame=True, **kw)


CONFIGS = {"cfg3": cfg3, "cfg4": cfg4}


@pytest.mark.parametrize("name", list(CONFIGS))
def test_full_size_residuals_match_oracle(name):
    sb, _ = CONFIGS[name]()
    o = OracleProblem(sb)
    with Problem(sb) as p:
        assert (p.n, p.nobs) == (o.n, o.nobs)
        e2m, p2a = p.index_maps()
        oe2m, op2a = o.index_maps()
        np.testing.assert_array_equal(e2m, oe2m)      # errorToMarkerList, bit-exact
        np.testing.assert_array_equal(p2a, op2a)      # paramToAttrList, bit-exact
        x = p.initial_parameters()
        np.testing.assert_array_equal(x, o.initial_parameters())
        f, d = p.eval_residuals(x)
    fo, do = o.eval_residuals(x)
    assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
    assert np.abs(d - do).max() <= 1e-9 * np.abs(do).max()


@pytest.mark.parametrize("name", list(CONFIGS))
def test_full_size_jacobian_columns_match_oracle_differences(name):
    sb, _ = CONFIGS[name]()
    o = OracleProblem(sb)
    delta = 1e-4
    with Problem(sb) as p:
        x = p.initial_parameters()
        vals, cols = p.eval_jacobian_blocks(x)       # [k, m], [nobs, k]
        n = p.n
    fo, _ = o.eval_residuals(x)
    # parameters of every kind: camera tx / rz at a few frames, focal + lens (cfg4), bundles
    picks = [0, 517, 999, 5 * F + 3, 5 * F + 999, 6 * F, 6 * F + 1, n - 1500 + 7, n - 1]
    if name == "cfg4":
        picks += [6 * F + 3, 6 * F + 6]
    scale = np.abs(vals).max()
    for pid in picks:
        xa = x.copy()
        xa[pid] += delta
        fa, _ = o.eval_residuals(xa)
        col_ref = (fa - fo) / delta                  # forward difference (adjust_solveFunc.cpp:455)
        col = np.zeros(2 * len(cols))
        for c in range(cols.shape[1]):
            rows = np.nonzero(cols[:, c] == pid)[0]
            col[2 * rows] = vals[c, 2 * rows]
            col[2 * rows + 1] = vals[c, 2 * rows + 1]
        touched = np.nonzero(col_ref)[0]
        assert len(touched) > 0, pid
        # rows the parameter cannot touch hold structural zeros on both sides
        assert np.abs(col - col_ref).max() <= 1e-9 * scale, (pid, np.abs(col - col_ref).max(), scale)


def test_cfg4_full_size_noise_free_recovers_focal_and_lens():
    """camera + bundles + focal + lens solved together from a 33 px start: with exact markers
    the fit goes to ~0 px and the gauge-invariant attributes head back to the truth."""
    sb, info = cfg4(noise_px=0.0)
    # The fit stops at ~0.04 px, not 0: the reference's residuals are |marker - point|
    # (adjust_measureErrors.cpp:481-499), so a forward difference of step delta = 1e-4 has the
    # wrong sign once a residual is smaller than delta * |J| ~ 0.1 px (kept quirk).
    with Problem(sb, iter_max=100) as p:
        x0 = p.initial_parameters()
        _, d0 = p.eval_residuals(x0)
        r = p.solve()
        assert p.solve_stats() == (1, 1)             # one graph launch, one synchronisation
    print(r, "initial avg", d0.mean(), "iterations", r.iteration_attempt_num, "solve s", r.timer_solve_s)
    focal = r.attr_values[6 * F]
    lens = r.attr_values[6 * F + 1:6 * F + 7]
    print("focal", focal, "lens", lens)
    assert r.success and r.error_avg < 0.1 and r.error_avg < 2e-3 * d0.mean()
    # focal length trades against depth at this parallax (weakly determined), the lens
    # coefficients are well determined
    assert abs(focal - 35.0) < 1.0                     # started at 33
    np.testing.assert_allclose(lens, [LENS[k] for k in LENS], atol=1e-2)  # started at 0; c2 trades against focal


GOLDEN_SCENES = {
    "cfg3": lambda: synth.make_scene(F, B, 3, fixed_bundles=4),
    "cfg4": lambda: synth.make_scene(F, B, 4, lens=LENS, solve_focal=True, solve_lens=tuple(LENS), focal_init=33.0,
                                     in_frame=True, fixed_bundles=4),
}


@pytest.mark.parametrize("name", list(GOLDEN_SCENES))
def test_full_size_solve_matches_cpu_golden(name):
    """Whole solves at full size against tests/golden/fullsize_<cfg>.npz: the block-sparse CPU
    oracle (oracle/oracle_sparse.h -- the reference's evaluation arithmetic and MINPACK control
    law, itself pinned to the dense faithful oracle at small sizes) on the same seeded scene,
    4 bundles held so that the solved values are determined.  north_star tolerances: solved
    attribute values 1e-6 relative, per-marker reprojection error 1e-6 px."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_%s.npz" % name))
    sb, _ = GOLDEN_SCENES[name]()
    with Problem(sb) as p:
        r = p.solve()
        assert p.solve_stats() == (1, 1)
    reason, nfev, njev, nfunc, njac, better = [int(v) for v in g["summary"]]
    print(name, r, "njev", r.iteration_attempt_num, "golden", reason, nfev, njev, "cpu s", float(g["cpu_seconds"]),
          "gpu solve s", r.timer_solve_s)
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (reason, nfev, njev)
    assert (r.iteration_function_num, r.iteration_jacobian_num, int(r.error_is_better)) == (nfunc, njac, better)
    scale = np.maximum(np.abs(g["attr_values"]), 1.0)
    assert np.max(np.abs(r.attr_values - g["attr_values"]) / scale) <= 1e-6
    stride = int(g["sample_stride"])
    ddiff = np.max(np.abs(r.error_dist[::stride] - g["dist_sample"]))
    print("max |dist - golden| px", ddiff)
    # north_star: 1e-6 px.  The solve is bit-reproducible (order-independent accumulation of the
    # normal equations), so this is one number, not a distribution
    assert ddiff <= 1e-6
    ef, ea, emin, emax = g["errors"]
    assert abs(r.error_final - ef) <= 1e-6 * ef and abs(r.error_avg - ea) <= 1e-6
    assert abs(r.error_min - emin) <= 1e-6 and abs(r.error_max - emax) <= 1e-6
