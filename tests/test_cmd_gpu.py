"""GPU: the reference's own solver test through the command surface -- test1.py:50-122
(`maya.cmds.mmSolver(camera=..., marker=..., attr=..., frame=..., iterations=..., verbose=True)`,
then `assertApproxEqual(tx, -6.0)`, `assertApproxEqual(ty, 3.6)`), and the result strings."""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import mmSolver, solveresult, synth
from mayamatchmovesolver_b200.scene import SceneBuilder

pytestmark = pytest.mark.gpu


def _test1_scene():
    sb = SceneBuilder(frames=[1])
    cam_t = sb.add_transform(name="cam_tfm", tx=-1.0, ty=1.0, tz=-5.0)
    cam = sb.add_camera(cam_t, name="cam_shp", film_fit="fill", render_width=960, render_height=540)
    bt = sb.add_transform(name="bundle_tfm", tx=5.5, ty=6.4, tz=-25.0)
    sb.add_marker(cam, sb.add_bundle(bt, name="bundle"), -0.243056042, 0.189583713, name="marker_tfm")
    return sb, bt


def test_reference_test1_through_the_command():
    sb, bt = _test1_scene()
    result = mmSolver(sb,
                      camera=[("cam_tfm", "cam_shp")],
                      marker=[("marker_tfm", "cam_shp", "bundle")],
                      attr=[("bundle_tfm.tx", "None", "None", "None", "None"),
                            ("bundle_tfm.ty", "None", "None", "None", "None")],
                      frame=[1], iterations=1000, solverType=2, sceneGraphMode=2, verbose=True)
    assert result[0] == "success=1"                       # test1.py:117
    # the command sets the solved values on the nodes (test1.py:121-122, eps 1e-4)
    assert abs(sb.get_value(bt["tx"]) - (-6.0)) < 1e-4
    assert abs(sb.get_value(bt["ty"]) - 3.6) < 1e-4
    r = solveresult.SolveResultStrings(result)
    assert r.get_success() and r.solver_stats["stop_id"] in (1, 2, 3, 4)
    assert r.error_stats["final_average"] < 1e-3
    assert set(r.per_marker_per_frame_error) == {"marker_tfm"} and list(r.per_frame_error) == [1.0]
    assert len(r.solve_parameter_list) == 2 and len(r.solve_error_list) == 2


def test_command_matches_problem_api_and_print_statistics():
    from mayamatchmovesolver_b200.problem import Problem
    sb, _ = synth.make_scene(F=12, B=16, seed=9, fixed_bundles=3, period=12)
    with Problem(sb) as p:
        r0 = p.solve()
    # printStatistics: the command reports and does NOT solve (constructPrintStats, adjust_base.cpp:618-646)
    stats = mmSolver(sb, write_back=False, printStatistics=["inputs"])
    assert stats[0] == "numberOfParameters=%d" % len(r0.x) and stats[1] == "numberOfErrors=%d" % (2 * len(r0.error_dist))
    rs = solveresult.SolveResultStrings(stats)
    assert rs.get_success() and rs.solver_stats["iteration_total_calls"] == 0
    with Problem(sb) as p:
        _, d0 = p.eval_residuals(None)
    assert abs(rs.error_stats["final_average"] - d0.mean()) <= 1e-5 * d0.mean()   # the initial deviation
    strings = mmSolver(sb, write_back=False, iterations=100, tauFactor=1.0, delta=1e-4)
    r = solveresult.SolveResultStrings(strings)
    assert r.solver_stats["stop_id"] == r0.reason_num and r.solver_stats["iteration_total_calls"] == r0.iteration_num
    assert abs(r.error_stats["final"] - r0.error_final) <= 1e-5 * r0.error_final     # 6 significant digits
    np.testing.assert_allclose(r.solve_parameter_list, r0.x, rtol=2e-5, atol=1e-9)
    assert len(r.per_frame_error) == 12 and len(r.per_marker_per_frame_error) == 16
    # per-frame error = mean deviation of that frame's markers (adjust_results.h:441-463)
    e2m_frame = np.tile(np.arange(12), 16)
    for fi in (0, 5, 11):
        assert abs(r.per_frame_error[float(fi + 1)] - r0.error_dist[e2m_frame == fi].mean()) < 1e-5


def test_per_frame_mode_through_the_command():
    sb, _ = synth.make_scene(F=20, B=12, seed=2, solve_bundles=False, period=20)
    strings = mmSolver(sb, frameSolveMode=1, sceneGraphMode=2, write_back=False)
    r = solveresult.SolveResultStrings(strings)
    assert r.get_success() and r.error_stats["final_average"] < 1.0 and len(r.per_frame_error) == 20
