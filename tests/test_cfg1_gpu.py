"""cfg1 (BASELINE.json configs[0]): one static camera, 20 bundles, ONE frame, markers read from
a marker file (.uv, the Load Marker format of the reference's tests/data/uvtrack): the bundle
positions are solved through the command surface and compared with the CPU oracle on the same
scene.  n = 60 (tx, ty, tz) > m = 40 is refused as in the reference (adjust_base.cpp:864-882,
test2.py:124); the solvable cfg1 problem is tx, ty per bundle (n = 40, m = 40)."""
import copy

import numpy as np
import pytest

from mayamatchmovesolver_b200 import loadmarker, mmSolver, solveresult, synth
from mayamatchmovesolver_b200._capi import MmbError
from mayamatchmovesolver_b200.problem import Problem
from mayamatchmovesolver_b200.scene import SceneBuilder
from oraclelib import OracleProblem

pytestmark = pytest.mark.gpu
B = 20


def _marker_file(tmp_path):
    """20 points seen by a camera at the origin looking down -z, written as a .uv (v1) file"""
    rng = np.random.default_rng(12)
    pts = np.stack([rng.uniform(-6, 6, B), rng.uniform(-3.5, 3.5, B), rng.uniform(-30, -14, B)], 1)
    x, y = synth.project_truth([0, 0, 0], [0, 0, 0], pts, 35.0, 36.0, 24.0, 2048, 1556)
    lines = [str(B)]
    for b in range(B):
        lines += ["point%02d" % b, "1", "1 %.17g %.17g 1.0" % (x[b] + 0.5, y[b] + 0.5)]
    path = tmp_path / "cfg1.uv"
    path.write_text("\n".join(lines) + "\n")
    return str(path), pts


def _scene(path, pts):
    _, tracks = loadmarker.read(path)
    assert len(tracks) == B
    sb = SceneBuilder(frames=[1])
    cam_tfm = sb.add_transform(name="cam_tfm")
    cam = sb.add_camera(cam_tfm, name="cam_shp", film_fit="horizontal", render_width=2048, render_height=1556,
                        sensor_width_mm=36.0, sensor_height_mm=24.0, focal_length_mm=35.0)
    nodes = loadmarker.add_tracks_to_scene(sb, cam, tracks)
    # start: the true depth, x / y off by up to half a unit
    rng = np.random.default_rng(13)
    for b, (_, _, tfm) in enumerate(nodes):
        sb.set_value(tfm["tx"], pts[b, 0] + rng.uniform(-0.5, 0.5))
        sb.set_value(tfm["ty"], pts[b, 1] + rng.uniform(-0.5, 0.5))
        sb.set_value(tfm["tz"], pts[b, 2])
    return sb, nodes


def test_cfg1_bundle_solve_from_marker_file(tmp_path):
    path, pts = _marker_file(tmp_path)
    sb, nodes = _scene(path, pts)
    attrs = []
    for b in range(B):
        attrs += [("point%02d_BND_tfm.tx" % b, "None", "None", "None", "None"),
                  ("point%02d_BND_tfm.ty" % b, "None", "None", "None", "None")]
    ref_scene = copy.deepcopy(sb)
    strings = mmSolver(sb, camera=[("cam_tfm", "cam_shp")],
                       marker=[("point%02d_MKR" % b, "cam_shp", "point%02d_BND" % b) for b in range(B)],
                       attr=attrs, frame=[1], iterations=100, solverType=2, sceneGraphMode=2)
    res = solveresult.SolveResultStrings(strings)
    assert res.get_success() and strings[0] == "success=1"
    # oracle on the same scene with the same attribute list
    for plug, *_ in attrs:
        ref_scene.add_attr(ref_scene.plug(plug))
    o = OracleProblem(ref_scene)
    ro = o.solve()
    assert (o.n, o.m) == (40, 40)
    assert res.solver_stats["stop_id"] == ro.reason_num
    got = np.array([[sb.get_value(n[2]["tx"]), sb.get_value(n[2]["ty"])] for n in nodes]).reshape(-1)
    np.testing.assert_allclose(got, ro.attr_values, rtol=1e-6, atol=1e-9)
    # the solved bundles sit on their markers' rays at the true depth: the true positions
    np.testing.assert_allclose(got.reshape(B, 2), pts[:, :2], atol=2e-3)
    assert res.error_stats["final_average"] < 0.05                     # px (|delta| residual floor)
    assert sorted(res.get_marker_error_list()) == sorted("point%02d_MKR" % b for b in range(B))


def test_cfg1_three_attributes_per_bundle_is_refused(tmp_path):
    path, pts = _marker_file(tmp_path)
    sb, nodes = _scene(path, pts)
    for _, _, tfm in nodes:
        for k in ("tx", "ty", "tz"):
            sb.add_attr(tfm[k])
    with Problem(sb) as p:
        assert (p.n, p.m) == (60, 40)
        with pytest.raises(MmbError):
            p.solve()


def test_cfg1_on_the_reference_marker_file():
    """cfg1 as BASELINE.json / SURVEY.md 8d specify it: the markers of
    tests/data/match_mover/loadmarker.rz2 (frame 41: the 20 points tracked there; fixture
    tests/golden/cfg1_loadmarker_frame41.rz2, read with the product's .rz2 reader), the camera and
    the bundle start positions of test_solve_operahouse.py; tx, ty per bundle solved through the
    command surface on the GPU and by the CPU oracle on the same scene."""
    from kat_scenes import cfg1_operahouse
    from oraclelib import oracle_mmsolver
    sb, flags, nodes = cfg1_operahouse()
    ro, oracle_values = oracle_mmsolver(sb, **flags)
    strings = mmSolver(sb, sceneGraphMode=2, **flags)
    res = solveresult.SolveResultStrings(strings)
    assert strings[0] == "success=1" and res.get_success()
    assert res.solver_stats["stop_id"] == ro.reason_num
    assert res.solver_stats["iteration_total_calls"] == ro.iteration_num
    got = np.array([[sb.get_value(n["tx"]), sb.get_value(n["ty"])] for n in nodes]).reshape(-1)
    want = np.array([oracle_values[a[0]] for a in flags["attr"]])
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9)
    assert abs(res.error_stats["final_average"] - ro.error_avg) <= 1e-6
    assert res.error_stats["final_average"] < 0.05      # px: every bundle moved onto its marker's ray
    assert len(res.get_marker_error_list()) == 20
