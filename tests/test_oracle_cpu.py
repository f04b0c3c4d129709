"""CPU tests (no GPU): pin the oracle against the reference's own known-answer
tests and against the reference's own lens code (oracle/_ref), and check the
host-side logic of the product (index maps / variant plan, via the test-only
host emulation of the CUDA thread bodies).

Pins:
  lib/rust/mmscenegraph/tests/reprojection.rs:36-99   single point KAT @1e-5
  lib/rust/mmscenegraph/tests/reprojection.rs:101-195 two bundles under a group
  tests/test/test_solver/test1.py:50-122              tx,ty -> -6.0, 3.6 @1e-4
  tests/test/test_solver/test_deviation_calculation.py:135-142 (pixel deviations)
  lib/cppbind/mmlens (compiled from /root/reference into oracle/_ref)
  MINPACK lmder via scipy.optimize.leastsq (same algorithm family as cminpack 1.3.8)
"""
import ctypes as C

import numpy as np
import pytest

from mayamatchmovesolver_b200 import synth
from mayamatchmovesolver_b200.scene import SceneBuilder
from oraclelib import EmuProblem, OracleProblem, _p, host_options, oracle, ref_lens


def test_rust_reprojection_single_point_kat():
    lib = oracle()
    cam = np.array([-2.0, 2.0, 5.0, 10.0, -10.0, -10.0, 1, 1, 1.0])
    proj = np.array([35.0, 36.0 / 25.4, 24.0 / 25.4, 0, 0, 2048, 1556, 1, 0.1, 10000.0, 1.0])
    pt = np.array([-0.5, 2.7, 0.0])
    out = np.zeros(2)
    lib.orc_reproject(_p(cam), 0, _p(proj), _p(pt), _p(out))
    assert abs(out[0] - 0.0865145148481126) < 1e-5   # EPSILON of the Rust test
    assert abs(out[1] - 0.0096299819122515) < 1e-5


def test_rust_two_bundles_under_group_kat():
    """reprojection.rs:101-195: group t=(0,0,-10) r=(0,15,0); bundles at x=-5/+5 under it;
    camera t=(0,5,10) r=(-10,0,0) ZXY -- through the whole scene evaluation (hierarchy)."""
    sb = SceneBuilder(frames=[1])
    grp = sb.add_transform(tx=0.0, ty=0.0, tz=-10.0, ry=15.0)
    ba = sb.add_transform(parent=grp, tx=-5.0)
    bb = sb.add_transform(parent=grp, tx=5.0)
    cam_t = sb.add_transform(tx=0.0, ty=5.0, tz=10.0, rx=-10.0, ry=0.0, rz=0.0, rotate_order="zxy")
    cam = sb.add_camera(cam_t, film_fit="horizontal", render_width=2048, render_height=1556,
                        focal_length_mm=35.0, sensor_width_mm=36.0, sensor_height_mm=24.0)
    for t in (ba, bb):
        sb.add_marker(cam, sb.add_bundle(t), 0.0, 0.0)
    sb.add_attr(ba["tx"])
    o = OracleProblem(sb, remove_unused_markers=0)   # the second marker is affected by no attribute
    f, d = o.eval_residuals(o.initial_parameters())
    # markers sit at (0,0): residuals are |point| * 2048
    pts = f.reshape(-1, 2) / 2048.0
    # residuals are absolute values: magnitudes @1e-5 (reprojection.rs:191-195)
    assert abs(pts[0, 0] - 0.243416) < 1e-5 and abs(pts[0, 1] - 0.111167) < 1e-5
    assert abs(pts[1, 0] - 0.2150060) < 1e-5 and abs(pts[1, 1] - 0.071858) < 1e-5


def _test1_scene():
    sb = SceneBuilder(frames=[1])
    cam_t = sb.add_transform(tx=-1.0, ty=1.0, tz=-5.0)
    # default Maya camera: 35 mm, 36x24 mm film back, Fill, 960x540 default resolution
    cam = sb.add_camera(cam_t, film_fit="fill", render_width=960, render_height=540)
    bt = sb.add_transform(tx=5.5, ty=6.4, tz=-25.0)
    sb.add_marker(cam, sb.add_bundle(bt), -0.243056042, 0.189583713)
    sb.add_attr(bt["tx"])
    sb.add_attr(bt["ty"])
    return sb


def test_reference_test1_kat():
    """tests/test/test_solver/test1.py:50-122 (cminpack_lmder, mmscenegraph)."""
    o = OracleProblem(_test1_scene(), iter_max=1000)
    r = o.solve()
    assert r.success
    assert abs(r.attr_values[0] - (-6.0)) < 1e-4   # baseutils.py assertApproxEqual eps
    assert abs(r.attr_values[1] - 3.6) < 1e-4


def test_reference_deviation_kat():
    """test_deviation_calculation.py:135-142: corner markers, imageWidth 2048; a marker at the
    frame edge is 0.5 * 2048 = 1024 px away from a point projected at the centre."""
    sb = SceneBuilder(frames=[1])
    cam_t = sb.add_transform()
    cam = sb.add_camera(cam_t, film_fit="horizontal", render_width=2048, render_height=1152)
    bt = sb.add_transform(tz=-10.0)  # on the optical axis -> projects to (0, 0)
    bnd = sb.add_bundle(bt)
    sb.add_marker(cam, bnd, 0.5, 0.0)
    sb.add_marker(cam, bnd, 0.0, 0.28125)   # 576 px / 2048
    sb.add_attr(bt["tx"])
    o = OracleProblem(sb)
    f, d = o.eval_residuals(o.initial_parameters())
    assert abs(d[0] - 1024.0) < 1e-9
    scale = (2048 / 1152) / (36.0 / 24.0)   # flat.rs:73-97 Horizontal: marker y scaled
    assert abs(d[1] - 576.0 * scale) < 1e-9


LENS_PARAMS = {
    1: [0.1, 1.0, 0.01, 0.02, 0.05],
    # lib/cppbind/mmlens/tests/test_once_3de_radial_std_deg4.cpp:48-63
    2: [0.1, 0.02, -0.01, 0.05, 0.01, -0.02, 10.0, 0.1],
    3: [0.02, 0.01, -0.015, 0.005, 0.03, -0.01, 0.002, 0.004, -0.003, 0.001, 5.0, 1.02, 0.98],
    4: [0.02, 0.01, -0.015, 0.005, 0.03, -0.01, 0.002, 0.004, -0.003, 0.001, 5.0, 1.02, 0.98, 1.1],
}


@pytest.mark.parametrize("model", [1, 2, 3, 4])
@pytest.mark.parametrize("direction", [0, 1])
def test_lens_models_match_reference_mmlens(model, direction):
    ref = ref_lens()
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    lib = oracle()
    rng = np.random.default_rng(model)
    pts = rng.uniform(-0.5, 0.5, (200, 2))
    co = np.zeros(14)
    co[:len(LENS_PARAMS[model])] = LENS_PARAMS[model]
    cam = np.array([3.5, 3.6, 2.4, 1.0, 0.05, -0.03])
    a = np.zeros_like(pts)
    b = np.zeros_like(pts)
    lib.orc_lens_apply(model, _p(co), _p(cam), direction, len(pts), _p(pts), _p(a))
    assert ref.ref_lens_apply(model, _p(co), _p(cam), direction, len(pts), _p(pts), _p(b)) == 0
    assert np.abs(a - b).max() < 1e-13, np.abs(a - b).max()


def test_lens_kat_from_survey_probe():
    """SURVEY.md 8c probe of the reference's own code: radial-deg4 with the coefficients of
    test_once_3de_radial_std_deg4.cpp:49-63."""
    ref = ref_lens()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    co = np.zeros(14)
    co[:8] = LENS_PARAMS[2]
    cam = np.array([3.0, 3.6, 2.4, 1.0, 0.0, 0.0])
    pt = np.array([[0.25, -0.2]])
    und = np.zeros_like(pt)
    red = np.zeros_like(pt)
    ref.ref_lens_apply(2, _p(co), _p(cam), 0, 1, _p(pt), _p(und))
    ref.ref_lens_apply(2, _p(co), _p(cam), 1, 1, _p(und), _p(red))
    assert np.abs(red - pt).max() < 1e-8   # undistort -> redistort round trip (test_both_*)


def test_bound_transforms_round_trip():
    lib = oracle()
    fm = float(np.finfo(np.float32).max)
    for v in (-3.0, 0.25, 7.5):
        assert lib.orc_bound_int_to_ext(lib.orc_bound_ext_to_int(v, -fm, fm, 0.0, 1.0), -fm, fm, 0.0, 1.0) == v
    # both bounds: asin / sin pair (adjust_base.cpp:194-258)
    x = lib.orc_bound_ext_to_int(2.0, 0.0, 5.0, 0.0, 1.0)
    assert abs(lib.orc_bound_int_to_ext(x, 0.0, 5.0, 0.0, 1.0) - 2.0) < 1e-12


def test_oracle_lmder_matches_scipy_minpack():
    """Iterate-level pin of the lmder restatement against MINPACK (scipy wraps the Fortran)."""
    from scipy.optimize import leastsq
    lib = oracle()
    rng = np.random.default_rng(0)
    m, n = 30, 5
    A = rng.normal(size=(m, n))
    t = rng.normal(size=m)

    def f(x):
        return np.tanh(A @ x) - t

    def jac(x):
        return (1 - np.tanh(A @ x) ** 2)[:, None] * A

    FCN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                      C.POINTER(C.c_double), C.c_int, C.c_int)

    def cb(ud, m_, n_, xp, fp, jp, ld, iflag):
        x = np.ctypeslib.as_array(xp, (n_,)).copy()
        if iflag == 1:
            np.ctypeslib.as_array(fp, (m_,))[:] = f(x)
        else:
            J = jac(x)
            out = np.ctypeslib.as_array(jp, (n_ * ld,))
            for j in range(n_):
                out[j * ld:j * ld + m_] = J[:, j]
        return 0
    lib.orc_lmder.argtypes = [FCN, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                              C.c_double, C.c_double, C.c_double, C.c_int, C.c_int, C.c_double,
                              C.POINTER(C.c_int), C.POINTER(C.c_int)]
    x0 = rng.normal(size=n) * 0.1
    for tol, maxfev in ((1e-6, 100), (1e-12, 200), (1e-6, 7)):
        x = x0.copy()
        fvec = np.zeros(m)
        nfev, njev = C.c_int(), C.c_int()
        info = lib.orc_lmder(FCN(cb), None, m, n, _p(x), _p(fvec), tol, tol, tol, maxfev, 1, 100.0,
                             C.byref(nfev), C.byref(njev))
        xs, cov, infodict, mesg, ier = leastsq(f, x0, Dfun=jac, full_output=1, ftol=tol, xtol=tol, gtol=tol,
                                               maxfev=maxfev, factor=100.0)
        assert info == ier
        assert nfev.value == infodict["nfev"]
        assert np.allclose(x, xs, rtol=1e-10, atol=1e-12)


def test_oracle_lmdif_matches_scipy_minpack():
    """lmdif (fdjac2 forward differences, nfev += n per Jacobian) against MINPACK's own lmdif:
    scipy.optimize.leastsq without Dfun."""
    from scipy.optimize import leastsq
    lib = oracle()
    rng = np.random.default_rng(1)
    m, n = 40, 6
    A = rng.normal(size=(m, n))
    t = rng.normal(size=m)

    def f(x):
        return np.tanh(A @ x) - t

    FCN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                      C.POINTER(C.c_double), C.c_int, C.c_int)

    def cb(ud, m_, n_, xp, fp, jp, ld, iflag):
        np.ctypeslib.as_array(fp, (m_,))[:] = f(np.ctypeslib.as_array(xp, (n_,)).copy())
        return 0
    lib.orc_lmdif.argtypes = [FCN, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                              C.c_double, C.c_double, C.c_double, C.c_int, C.c_double, C.c_int, C.c_double,
                              C.POINTER(C.c_int), C.POINTER(C.c_int)]
    x0 = rng.normal(size=n) * 0.1
    x0[2] = 0.0   # h = eps when x_j == 0
    for tol, maxfev, epsfcn in ((1e-6, 400, 1e-5), (1e-12, 400, 1e-5), (1e-6, 20, 1e-5), (1e-8, 400, 0.0)):
        x = x0.copy()
        fvec = np.zeros(m)
        nfev, njev = C.c_int(), C.c_int()
        info = lib.orc_lmdif(FCN(cb), None, m, n, _p(x), _p(fvec), tol, tol, tol, maxfev, epsfcn, 1, 100.0,
                             C.byref(nfev), C.byref(njev))
        xs, cov, infodict, mesg, ier = leastsq(f, x0, full_output=1, ftol=tol, xtol=tol, gtol=tol, maxfev=maxfev,
                                               epsfcn=epsfcn if epsfcn else None, factor=100.0)
        assert info == ier, (info, ier, mesg)
        assert nfev.value == infodict["nfev"]
        assert np.allclose(x, xs, rtol=1e-9, atol=1e-11)


def test_index_maps_reference_order():
    """adjust_relationships.cpp: errors marker-major/frame-minor over enabled & weight>0 pairs;
    parameters attr-major, animated attrs expand per frame; weights normalised per frame."""
    sb = SceneBuilder(frames=[10, 11, 12])
    cam_t = sb.add_transform(tx=[0.0, 0.1, 0.2], ty=0.0, tz=5.0)
    cam = sb.add_camera(cam_t)
    b0 = sb.add_transform(tx=1.0, tz=-10.0)
    b1 = sb.add_transform(tx=-1.0, tz=-12.0)
    sb.add_marker(cam, sb.add_bundle(b0), [0.1, 0.1, 0.1], 0.0, weight=[1.0, 2.0, 0.0], enable=[1, 1, 1])
    sb.add_marker(cam, sb.add_bundle(b1), [-0.1, -0.1, -0.1], 0.0, weight=[4.0, 1.0, 1.0], enable=[1, 0, 1])
    sb.add_attr(b0["tx"])
    sb.add_attr(cam_t["tx"])
    sb.add_attr(b1["tz"])
    o = OracleProblem(sb)
    e2m, p2a = o.index_maps()
    assert e2m.tolist() == [[0, 0], [0, 1], [1, 0], [1, 2]]
    assert p2a.tolist() == [[0, -1], [1, 0], [1, 1], [1, 2], [2, -1]]
    assert np.allclose(o.marker_weights(), [1.0 / 4.0, 2.0 / 2.0, 4.0 / 4.0, 1.0 / 1.0])


@pytest.mark.parametrize("adt", [0, 1])
@pytest.mark.parametrize("case", [
    dict(F=7, B=5, seed=11, drop_fraction=0.2, weights=True),
    dict(F=5, B=6, seed=4, lens=dict(degree2_distortion=0.05, degree2_u=0.002, degree4_distortion=0.01),
         solve_focal=True, solve_lens=("degree2_distortion", "degree4_distortion"), focal_init=33.0),
])
@pytest.mark.parametrize("bounded", [False, True])
def test_host_plan_logic_via_emulation(case, adt, bounded):
    """The product's host code (index maps, dependency analysis, variant plan) drives the same
    per-thread bodies the CUDA kernels run; executed on the host here they must reproduce the
    oracle's residuals and dense Jacobian.  `bounded`: box constraints, offset and scale on the
    solver attributes (every branch of the bound transform, sign-flipped FD steps)."""
    sb, info = synth.make_scene(**case)
    if bounded:
        from kat_scenes import apply_bounds
        apply_bounds(sb, info)
    o = OracleProblem(sb, auto_diff_type=adt)
    e = EmuProblem(sb, auto_diff_type=adt)
    x = o.initial_parameters() + np.random.default_rng(1).normal(0, 1e-2, o.n)
    fo, do = o.eval_residuals(x)
    fe, de = e.eval_residuals(x)
    assert np.abs(fo - fe).max() <= 1e-9 * np.abs(fo).max()
    Jo = o.eval_jacobian_dense(x)
    Je, _ = e.eval_jacobian_dense(x)
    assert np.abs(Jo - Je).max() <= 1e-9 * np.abs(Jo).max()
    assert np.array_equal(Jo != 0, Je != 0)


def test_oracle_structural_skip_is_exact():
    sb, _ = synth.make_scene(F=4, B=5, seed=2)
    o = OracleProblem(sb)
    x = o.initial_parameters()
    assert np.array_equal(o.eval_jacobian_dense(x, True), o.eval_jacobian_dense(x, False))


SPARSE_CASES = {
    "camera_bundles_12x16": dict(F=12, B=16, seed=9, fixed_bundles=3, period=12),
    "camera_bundles_40x30": dict(F=40, B=30, seed=5, fixed_bundles=4, period=40),
    "sparse_weighted": dict(F=20, B=14, seed=6, fixed_bundles=3, drop_fraction=0.15, weights=True, period=20),
    "bundles_only": dict(F=2, B=20, seed=1, solve_camera=False),
    "camera_only": dict(F=6, B=12, seed=2, solve_bundles=False),
    "lens_focal": dict(F=12, B=16, seed=4, fixed_bundles=16, solve_bundles=False,
                       lens=dict(degree2_distortion=0.05, degree2_u=0.002, degree2_v=-0.002, degree4_distortion=0.01),
                       solve_focal=True, solve_lens=("degree2_distortion", "degree2_u", "degree2_v", "degree4_distortion"),
                       focal_init=33.0, period=12),
}


@pytest.mark.parametrize("name", list(SPARSE_CASES))
def test_block_sparse_oracle_matches_dense_oracle(name):
    """oracle_sparse.h (normal equations + Schur complement; what makes full-size CPU solves and
    the tests/golden/fullsize_*.npz vectors possible) against the faithful dense path
    (one scene evaluation per column, pivoted-QR lmder): same LM decisions, same answers."""
    from mayamatchmovesolver_b200 import synth
    sb, _ = synth.make_scene(**SPARSE_CASES[name])
    rd = OracleProblem(sb).solve()
    rs = OracleProblem(sb).solve_sparse(threads=2)
    assert rs.status == 0
    assert (rs.reason_num, rs.iteration_num, rs.iteration_attempt_num) == (rd.reason_num, rd.iteration_num, rd.iteration_attempt_num)
    assert (rs.iteration_function_num, rs.iteration_jacobian_num, rs.error_is_better) == \
           (rd.iteration_function_num, rd.iteration_jacobian_num, rd.error_is_better)
    scale = np.maximum(np.abs(rd.attr_values), 1.0)
    assert np.max(np.abs(rs.attr_values - rd.attr_values) / scale) <= 1e-7
    assert np.max(np.abs(rs.error_dist - rd.error_dist)) <= 1e-6
    assert abs(rs.error_final - rd.error_final) <= 1e-7 * max(1.0, rd.error_final)
