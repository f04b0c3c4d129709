"""SolverStandard / SolverBasic executed on the GPU path: the staged solve (root frames ->
interpolated in-betweens -> per-frame refinement in ONE batched launch) reaches the error of
the direct global solve, each command starting from the values the previous one wrote."""
import copy

import numpy as np
import pytest

from mayamatchmovesolver_b200 import solver, synth
from mayamatchmovesolver_b200.problem import Problem
from mayamatchmovesolver_b200.solver import SolverBasic, SolverStandard

pytestmark = pytest.mark.gpu


def _avg_error(sb):
    with Problem(sb) as p:
        _, d = p.eval_residuals(p.initial_parameters())
    return float(np.mean(d))


def test_standard_solver_staged_solve():
    F, B = 41, 24
    sb, info = synth.make_scene(F, B, 7, init_noise_t=0.3, init_noise_r=1.0, init_noise_b=0.2, period=60.0)
    start = _avg_error(sb)
    direct = copy.deepcopy(sb)
    with Problem(direct) as p:
        rd = p.solve()
    s = SolverStandard()
    s.set_root_frame_list([1, 11, 21, 31, 41])
    s.set_frame_list(range(1, F + 1))
    staged = copy.deepcopy(sb)
    results = solver.execute(staged, s)
    assert len(results) == 2                                   # root command + per-frame command
    root, anim = results
    assert root.get_success() and anim.get_success()
    # the root command saw 5 frames, the per-frame command every frame (one record per frame)
    assert sorted(root.get_frame_error_list()) == [1, 11, 21, 31, 41]
    assert sorted(anim.get_frame_error_list()) == list(range(1, F + 1))
    after = _avg_error(staged)
    print("start %.3f px, staged %.4f px, direct global %.4f px" % (start, after, rd.error_avg))
    assert start > 10 * after
    # bundles come from 5 frames only, cameras per frame: within 15 % of the all-frames optimum
    assert after < 1.15 * rd.error_avg
    # global refinement from there lands on the direct solve's error
    s.set_global_solve(True)
    refined = copy.deepcopy(sb)
    solver.execute(refined, s)
    assert abs(_avg_error(refined) - rd.error_avg) < 2e-3


def test_only_root_frames_leaves_straight_lines():
    F, B = 21, 16
    sb, info = synth.make_scene(F, B, 8, period=60.0)
    s = SolverStandard()
    s.set_root_frame_list([1, 11, 21])
    s.set_frame_list(range(1, F + 1))
    s.set_only_root_frames(True)
    solver.execute(sb, s)
    tx = np.array([sb.get_value(info["cam_tfm"]["tx"], i) for i in range(F)])
    np.testing.assert_allclose(tx[:11], np.linspace(tx[0], tx[10], 11), atol=1e-12)
    np.testing.assert_allclose(tx[10:], np.linspace(tx[10], tx[20], 11), atol=1e-12)


def test_basic_solver_per_frame():
    F, B = 16, 20
    sb, info = synth.make_scene(F, B, 9, solve_bundles=False, init_noise_t=0.3, init_noise_r=1.0)
    start = _avg_error(sb)
    s = SolverBasic()
    s.set_frame_list(range(1, F + 1))
    (res,) = solver.execute(sb, s)
    assert res.get_success() and sorted(res.get_frame_error_list()) == list(range(1, F + 1))
    assert _avg_error(sb) < 0.7 and start > 5.0
