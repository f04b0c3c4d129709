"""GPU: sceneGraphMode = 1 (Maya-DAG-convention residual arithmetic, behind-camera factor,
stiffness / smoothness rows) through the C-ABI against the oracle's literal restatement."""
import numpy as np
import pytest

from kat_scenes import KATS, smoothness
from mayamatchmovesolver_b200 import mmSolver, solveresult
from mayamatchmovesolver_b200._capi import MmbError
from mayamatchmovesolver_b200.problem import Problem
from mode1_scenes import FITS, eval_scene, solve_scene
from oraclelib import OracleProblem, oracle_mmsolver

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fit", list(FITS))
@pytest.mark.parametrize("adt", [0, 1])
@pytest.mark.parametrize("mode", [1, 2])
def test_residuals_and_jacobian_match_oracle(fit, adt, mode):
    sb = eval_scene(fit)
    o = OracleProblem(sb, scene_graph_mode=mode, auto_diff_type=adt)
    with Problem(sb, scene_graph_mode=mode, auto_diff_type=adt) as p:
        assert (p.n, p.nobs, p.m) == (o.n, o.nobs, o.m) and p.m == 2 * p.nobs + 2
        x = p.initial_parameters() + 0.01
        f, d = p.eval_residuals(x)
        fo, do = o.eval_residuals(x)
        assert (np.abs(f - fo) / np.maximum(np.abs(fo), 1.0)).max() <= 1e-9
        assert np.abs(d - do).max() <= 1e-9 * np.abs(do).max()
        J = p.eval_jacobian_dense(x)
        Jo = o.eval_jacobian_dense(x)
        row_scale = np.maximum(np.abs(Jo).max(axis=1, keepdims=True), 1.0)
        assert (np.abs(J - Jo) / row_scale).max() <= 1e-9
        assert np.array_equal(J[-2:] != 0, Jo[-2:] != 0)


@pytest.mark.parametrize("mode", [1, 2])
def test_solve_with_stiffness_rows_matches_oracle(mode):
    sb = solve_scene()
    ro = OracleProblem(sb, scene_graph_mode=mode).solve()
    with Problem(sb, scene_graph_mode=mode) as p:
        r = p.solve()
        fl = p.last_residuals()
    print(mode, "oracle", ro, "gpu", r)
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (ro.reason_num, ro.iteration_num, ro.iteration_attempt_num)
    scale = np.maximum(np.abs(ro.attr_values), 1.0)
    assert np.max(np.abs(r.attr_values - ro.attr_values) / scale) <= 1e-6
    assert np.max(np.abs(r.error_dist - ro.error_dist)) <= 1e-6
    assert abs(r.error_final - ro.error_final) <= 1e-6 * max(1.0, ro.error_final)
    assert len(fl) == 2 * len(r.error_dist) + 2 and (mode == 2) == bool(np.all(fl[-2:] == 0.0))


@pytest.mark.parametrize("name", list(KATS))
def test_reference_kat_in_the_commands_default_mode(name):
    """the reference's tests call `mmSolver(...)` without -sceneGraphMode: maya_dag"""
    sb, flags, expected = KATS[name]()
    ro, oracle_values = oracle_mmsolver(sb, scene_graph_mode=1, **flags)
    strings = mmSolver(sb, sceneGraphMode=1, **flags)
    assert strings[0] == "success=1"
    for plug, (want, eps) in expected.items():
        got = sb.get_value(sb.plug(plug))
        assert abs(got - want) < eps, (name, plug, got, want)
        assert abs(got - oracle_values[plug]) <= 1e-6 * max(1.0, abs(want))
    r = solveresult.SolveResultStrings(strings)
    assert r.solver_stats["stop_id"] == ro.reason_num and r.solver_stats["iteration_total_calls"] == ro.iteration_num


def test_reference_smoothness_test():
    """test_smoothness.py: `-attrSmoothness` through the command, 10 frames, iterations = 10: success=1"""
    sb, flags = smoothness()
    strings = mmSolver(sb, sceneGraphMode=1, **flags)
    assert strings[0] == "success=1"          # the reference's only assertion
    r = solveresult.SolveResultStrings(strings)
    assert r.solver_stats["iteration_total_calls"] <= 10
    # numberOfErrors counts the smoothness row
    stats = mmSolver(smoothness()[0], sceneGraphMode=1, printStatistics=["inputs"], **smoothness()[1])
    assert "numberOfMarkerErrors=20" in stats and "numberOfErrors=21" in stats and "numberOfAttrSmoothnessErrors=1" in stats


def test_mode1_refuses_per_frame_and_sharding():
    sb = solve_scene()
    with pytest.raises(MmbError):
        Problem(sb, scene_graph_mode=1, frame_solve_mode=1)
