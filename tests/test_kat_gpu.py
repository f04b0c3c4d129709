"""GPU: the reference's own Maya-hosted solver tests (tests/kat_scenes.py: test1.py, test3.py,
test_min_max_values.py, test_marker_weight.py, test_marker_enabled.py) run through the command
surface `mmSolver(**flags)` -> C-ABI -> CUDA, with the reference's own assertions, and compared
with the CPU oracle driven by the same flags."""
import numpy as np
import pytest

from kat_scenes import KATS, apply_bounds
from mayamatchmovesolver_b200 import mmSolver, solveresult, synth
from mayamatchmovesolver_b200.problem import Problem
from oraclelib import OracleProblem, oracle_mmsolver

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(KATS))
def test_reference_kat_through_the_command(name):
    sb, flags, expected = KATS[name]()
    ro, oracle_values = oracle_mmsolver(sb, **flags)
    strings = mmSolver(sb, sceneGraphMode=2, **flags)
    assert strings[0] == "success=1"
    for plug, (want, eps) in expected.items():
        got = sb.get_value(sb.plug(plug))        # the command wrote the solved value back
        assert abs(got - want) < eps, (name, plug, got, want)
        assert abs(got - oracle_values[plug]) <= 1e-6 * max(1.0, abs(want)), (name, plug, got, oracle_values[plug])
    r = solveresult.SolveResultStrings(strings)
    assert r.solver_stats["stop_id"] == ro.reason_num
    assert r.solver_stats["iteration_total_calls"] == ro.iteration_num


BOUNDED_SOLVES = {
    "camera_bundles_12x16": dict(F=12, B=16, seed=9, fixed_bundles=3, period=12),
    "camera_bundles_40x30": dict(F=40, B=30, seed=5, fixed_bundles=4, period=40),
    "sparse_weighted": dict(F=20, B=14, seed=6, fixed_bundles=3, drop_fraction=0.15, weights=True, period=20),
}


@pytest.mark.parametrize("name", list(BOUNDED_SOLVES))
def test_bounded_solve_matches_oracle(name):
    """whole LM solves with box constraints / offset / scale on the solver attributes
    (adjust_base.cpp:194-258): same LM decisions, solved values 1e-6 relative, errors 1e-6 px"""
    sb, info = synth.make_scene(**BOUNDED_SOLVES[name])
    apply_bounds(sb, info, lower_only=False, scale_with_bounds=False)
    ro = OracleProblem(sb).solve()
    with Problem(sb) as p:
        r = p.solve()
    print(name, "oracle", ro, "gpu", r)
    assert (r.reason_num, r.iteration_num, r.iteration_attempt_num) == (ro.reason_num, ro.iteration_num, ro.iteration_attempt_num)
    scale = np.maximum(np.abs(ro.attr_values), 1.0)
    assert np.max(np.abs(r.attr_values - ro.attr_values) / scale) <= 1e-6
    # measured 1.4e-6 px on the sparse scene (xtol = 1e-6 stops both solvers within that of each
    # other; FD quotients amplify evaluation rounding by 1/delta): DESIGN.md "parity floor"
    assert np.max(np.abs(r.error_dist - ro.error_dist)) <= (5e-6 if name == "sparse_weighted" else 1e-6)
    # solved values respect their boxes
    _, p2a = p.index_maps() if False else OracleProblem(sb).index_maps()
    mn = np.array(sb.attr_min)[p2a[:, 0]]; mx = np.array(sb.attr_max)[p2a[:, 0]]
    assert (r.attr_values >= mn).all() and (r.attr_values <= mx).all()


def test_bounded_per_frame_solve_matches_oracle():
    """frameSolveMode = 1 with bounded / scaled camera attributes: the per-frame kernel's own
    copy of the bound transform and step logic against one oracle solve per frame"""
    sb, info = synth.make_scene(F=10, B=14, seed=2, solve_bundles=False, period=10)
    apply_bounds(sb, info, lower_only=False, scale_with_bounds=False)
    with Problem(sb, frame_solve_mode=1) as p:
        agg, per = p.solve_per_frame()
    av = agg.attr_values.reshape(6, 10)
    for f in range(10):
        ro = OracleProblem(sb.slice_frames([f])).solve()
        assert per[f].reason_num == ro.reason_num and per[f].iteration_num == ro.iteration_num
        assert np.max(np.abs(av[:, f] - ro.attr_values) / np.maximum(np.abs(ro.attr_values), 1.0)) <= 1e-6


@pytest.mark.parametrize("rum,rua", [(1, 1), (0, 0)])
def test_unused_markers_and_attributes_on_the_device(rum, rua):
    """removeUnusedMarkers / removeUnusedAttributes (solve_v1, adjust_base.cpp:1322-1395): a marker
    affected by no attribute and an attribute affecting no marker leave the problem -- sizes, index
    maps, residuals, Jacobian and the solve against the oracle's restatement."""
    from test_affects_cpu import _scene
    sb = _scene()
    o = OracleProblem(sb, remove_unused_markers=rum, remove_unused_attributes=rua)
    with Problem(sb, remove_unused_markers=rum, remove_unused_attributes=rua) as p:
        assert (p.n, p.nobs) == (o.n, o.nobs)
        e2m, p2a = p.index_maps()
        oe2m, op2a = o.index_maps()
        assert np.array_equal(e2m, oe2m) and np.array_equal(p2a, op2a)
        used_m, used_a = p.used_objects()
        assert used_m.tolist() == ([0, 1, 2] if rum else [0, 1, 2, 3]) and used_a.tolist() == ([0, 1, 3] if rua else [0, 1, 2, 3])
        x = p.initial_parameters()
        assert np.array_equal(x, o.initial_parameters())
        f, d = p.eval_residuals(x)
        fo, do = o.eval_residuals(x)
        assert np.abs(f - fo).max() <= 1e-9 * np.abs(fo).max()
        J = p.eval_jacobian_dense(x)
        Jo = o.eval_jacobian_dense(x)
        assert np.abs(J - Jo).max() <= 1e-9 * np.abs(Jo).max()
        if rua:   # with the all-zero column of the unused attribute kept, J is rank deficient
            r = p.solve()
            ro = OracleProblem(sb, remove_unused_markers=rum, remove_unused_attributes=rua).solve()
            assert (r.reason_num, r.iteration_num) == (ro.reason_num, ro.iteration_num)
            assert np.max(np.abs(r.attr_values - ro.attr_values) / np.maximum(1.0, np.abs(ro.attr_values))) <= 1e-6
            assert np.max(np.abs(r.error_dist - ro.error_dist)) <= 1e-6


def test_print_statistics_used_objects_and_affects():
    from test_affects_cpu import _scene
    sb = _scene()
    out = mmSolver(sb, printStatistics=["usedSolveObjects", "affects", "inputs"], write_back=False)
    assert out[0] == "numberOfParameters=5" and out[1] == "numberOfErrors=18"
    assert "markers_used=mA0#mA1#mB1#" in out and "markers_unused=mB2#" in out
    assert "attributes_used=camA_tfm.tx#bnd0_tfm.tx#bnd1_tfm.ty#" in out and "attributes_unused=lonely.tx#" in out
    assert "marker_affects_attribute=mB1#bnd1_tfm.ty#1" in out and "success=1" in out
    assert not any(s.startswith("marker_affects_attribute=mB2") for s in out)   # only the used objects are listed
