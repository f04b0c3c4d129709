"""Transform hierarchies on the GPU path: a camera under an animated rig, bundles under a
group, non-default rotate orders -- the cases that do NOT take k_variants' block path or the
parentless-bundle shortcut (dag.rs:234-327 world = parent * local)."""
import numpy as np
import pytest

from mayamatchmovesolver_b200.problem import Problem
from mayamatchmovesolver_b200.scene import SceneBuilder
from oraclelib import OracleProblem

pytestmark = pytest.mark.gpu


def _scene(F=9, B=12, seed=5, solve_rig=True):
    rng = np.random.default_rng(seed)
    sb = SceneBuilder(frames=np.arange(1, F + 1))
    f = np.arange(F, dtype=np.float64)
    rig = sb.add_transform(tx=0.3 * np.sin(f / 4), ty=0.1 * f / F, tz=1.0 + 0.0 * f, ry=2.0 * np.sin(f / 5) + 0.0,
                           rotate_order="zxy")
    cam_tfm = sb.add_transform(parent=rig, tx=0.2 + 0.05 * f, ty=0.1 * np.cos(f / 3), tz=0.0 * f + 0.5,
                               rx=1.0 * np.sin(f / 6), ry=0.5 * f / F, rz=0.2 * np.cos(f / 2), rotate_order="yxz")
    cam = sb.add_camera(cam_tfm, film_fit="vertical", render_width=1920, render_height=1080, focal_length_mm=40.0)
    grp = sb.add_transform(tx=0.5, ty=-0.2, tz=-3.0, ry=10.0, sx=1.1, sy=0.9, sz=1.0)
    pts = np.stack([rng.uniform(-5, 5, B), rng.uniform(-3, 3, B), rng.uniform(-25, -12, B)], 1)
    tfms = []
    for b in range(B):
        t = sb.add_transform(parent=grp if b % 2 else None, tx=pts[b, 0], ty=pts[b, 1], tz=pts[b, 2])
        tfms.append(t)
        bn = sb.add_bundle(t)
        sb.add_marker(cam, bn, rng.uniform(-0.3, 0.3, F), rng.uniform(-0.2, 0.2, F))
    for k in ("tx", "ty", "rx", "ry", "rz"):
        sb.add_attr(cam_tfm[k])
    if solve_rig:
        sb.add_attr(rig["ry"])
        sb.add_attr(rig["tx"])
    sb.add_attr(grp["ry"])      # a static attribute shared by half of the bundles
    sb.add_attr(grp["sx"])
    for b in range(B):
        for k in ("tx", "ty", "tz"):
            sb.add_attr(tfms[b][k])
    return sb


@pytest.mark.parametrize("adt", [0, 1])
def test_hierarchy_residuals_and_jacobian_match_oracle(adt):
    sb = _scene()
    o = OracleProblem(sb, auto_diff_type=adt)
    with Problem(sb, auto_diff_type=adt) as p:
        assert (p.n, p.nobs) == (o.n, o.nobs)
        x = p.initial_parameters()
        xt = x + np.random.default_rng(1).normal(0, 1e-2, x.shape)
        f, d = p.eval_residuals(xt)
        fo, do = o.eval_residuals(xt)
        assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
        assert np.abs(d - do).max() <= 1e-9 * np.abs(do).max()
        J, Jo = p.eval_jacobian_dense(xt), o.eval_jacobian_dense(xt)
        assert np.array_equal(J != 0, Jo != 0) or np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
        assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()


def test_hierarchy_solve_matches_oracle():
    sb = _scene(solve_rig=False)
    ro = OracleProblem(sb, iter_max=15).solve()
    with Problem(sb, iter_max=15) as p:
        r = p.solve()
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (ro.reason_num, ro.iteration_num, ro.iteration_attempt_num)
    assert abs(r.error_avg - ro.error_avg) <= 1e-5 * max(1.0, ro.error_avg)
