"""CPU: sceneGraphMode = 1 -- the oracle's literal restatement of measureErrors_mayaDag
(adjust_measureErrors.cpp:116-387, maya_camera.cpp:75-414,863-894) pinned to the reference's own
Maya-hosted tests, which run in that (default) mode, and the product's device bodies executed on
the host against it."""
import numpy as np
import pytest

from kat_scenes import KATS, smoothness
from mode1_scenes import FITS, eval_scene, solve_scene
from oraclelib import EmuProblem, OracleProblem, oracle_mmsolver


@pytest.mark.parametrize("name", list(KATS))
def test_oracle_reproduces_reference_kat_in_maya_dag_mode(name):
    sb, flags, expected = KATS[name]()
    r, values = oracle_mmsolver(sb, scene_graph_mode=1, **flags)
    assert r.success == 1
    for plug, (want, eps) in expected.items():
        assert abs(values[plug] - want) < eps, (name, plug, values[plug], want)


def test_oracle_smoothness_kat():
    """test_smoothness.py:43-140 asserts `success=1` (the gaussian row reaches 1e121: exp(288))"""
    sb, flags = smoothness()
    r, _ = oracle_mmsolver(sb, scene_graph_mode=1, **flags)
    assert r.success == 1 and r.iteration_num <= 10


@pytest.mark.parametrize("fit", list(FITS))
@pytest.mark.parametrize("adt", [0, 1])
@pytest.mark.parametrize("mode", [1, 2])
def test_device_bodies_match_oracle(fit, adt, mode):
    sb = eval_scene(fit)
    o = OracleProblem(sb, scene_graph_mode=mode, auto_diff_type=adt)
    e = EmuProblem(sb, scene_graph_mode=mode, auto_diff_type=adt)
    assert (o.n, o.nobs, o.m) == (e.n, e.nobs, e.m) and o.m == 2 * o.nobs + 2   # two counted rows
    x = o.initial_parameters() + 0.01
    fo, do = o.eval_residuals(x)
    fe, de = e.eval_residuals(x)
    assert (np.abs(fo - fe) / np.maximum(np.abs(fo), 1.0)).max() <= 1e-9
    assert np.abs(do - de).max() <= 1e-9 * np.abs(do).max()
    if mode == 1:
        assert (np.abs(fo) > 1e6).sum() == 2 * sb.F          # the bundle behind the camera, every frame
        assert fo[-2] != 0.0 and fo[-1] != 0.0
    else:
        assert np.all(fo[-2:] == 0.0)                        # mmscenegraph mode leaves the rows at zero
    Jo = o.eval_jacobian_dense(x)
    Je, _ = e.eval_jacobian_dense(x)
    row_scale = np.maximum(np.abs(Jo).max(axis=1, keepdims=True), 1.0)
    assert (np.abs(Jo - Je) / row_scale).max() <= 1e-9
    assert np.array_equal(Jo != 0, Je != 0)


def test_mode1_differs_from_mode2_only_where_it_should():
    sb = eval_scene("horizontal", behind=False)
    x = OracleProblem(sb).initial_parameters()
    f1, _ = OracleProblem(sb, scene_graph_mode=1).eval_residuals(x)
    f2, _ = OracleProblem(sb, scene_graph_mode=2).eval_residuals(x)
    assert np.abs(f1[:-2] - f2[:-2]).max() > 10.0     # the film offsets move every point in mode 1
    sb0 = eval_scene("horizontal", behind=False)
    from mayamatchmovesolver_b200 import synth
    sbz, info = synth.make_scene(F=5, B=7, seed=21, fixed_bundles=2)
    x = OracleProblem(sbz).initial_parameters()
    g1, _ = OracleProblem(sbz, scene_graph_mode=1).eval_residuals(x)
    g2, _ = OracleProblem(sbz, scene_graph_mode=2).eval_residuals(x)
    assert np.abs(g1 - g2).max() <= 1e-9 * np.abs(g2).max()   # no offsets, nothing behind: same numbers
