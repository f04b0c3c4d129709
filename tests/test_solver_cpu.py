"""Solver strategies (mayamatchmovesolver_b200/solver.py) against the reference's compile
logic: frame batching (doc examples of rootframe.py:62-77 and solverstandardutils.py:43-69),
attribute blocks (solverutils.py:43-127), and the Action sequence of compile_multi_frame
(solverstandardutils.py:433-735).  Host logic only: nothing is solved here."""
import numpy as np
import pytest

from mayamatchmovesolver_b200 import solver, synth
from mayamatchmovesolver_b200.solver import SolverBasic, SolverStandard, SolverStep


def test_generate_increment_frame_forward_doc_examples():
    g = solver.generate_increment_frame_forward
    assert g([1, 2]) == [[1, 2]]
    assert g([1, 2, 3, 4, 5]) == [[1, 2, 3], [1, 2, 3, 4, 5]]
    assert g(list(range(1, 10))) == [[1, 2, 3], [1, 2, 3, 4, 5, 6], [1, 2, 3, 4, 5, 6, 7, 8, 9]]


def test_two_frame_fwd():
    assert solver.gen_two_frame_fwd([7]) == [[7, 7]]
    assert solver.gen_two_frame_fwd([1, 5, 9]) == [[1, 5], [5, 9]]


def _scene():
    lens = dict(degree2_distortion=0.05, degree2_u=0.0, degree2_v=0.0, degree4_distortion=0.0)
    return synth.make_scene(12, 6, 1, lens=lens, solve_focal=True, solve_lens=("degree2_distortion",))[0]


def test_attribute_types_and_blocks():
    sb = _scene()
    attrs = solver.scene_attr_list(sb)
    types = [solver.get_attribute_solver_type(sb, a) for a in attrs]
    assert types[:3] == ["camera_translate"] * 3 and types[3:6] == ["camera_rotate"] * 3
    assert types[6] == "camera_intrinsic" and types[7] == "lens_distortion"
    assert set(types[8:]) == {"bundle_transform"} and len(types) == 8 + 18
    assert [solver.attribute_is_animated(sb, a) for a in attrs[:8]] == [True] * 6 + [False] * 2
    # no blocks: one stage with everything
    assert solver.create_attr_blocks(sb, False, attrs) == [(attrs, 1.0)]
    # bundles + camera + lens: the six-stage schedule, shares 0.2 0.1 0.1 0.1 0.1 0.4
    blocks = solver.create_attr_blocks(sb, True, attrs)
    assert [r for _, r in blocks] == [0.2, 0.1, 0.1, 0.1, 0.1, 0.4]
    assert len(blocks[0][0]) == 18 and len(blocks[1][0]) == 7 and len(blocks[2][0]) == 1 + 3 and blocks[5][0] == attrs
    # without lens: five stages
    no_lens = solver.filter_attr_list(sb, attrs, use_lens_distortion=False)
    assert len(no_lens) == len(attrs) - 1
    assert [r for _, r in solver.create_attr_blocks(sb, True, no_lens)] == [0.1, 0.1, 0.1, 0.1, 0.6]
    no_intr = solver.filter_attr_list(sb, attrs, use_camera_intrinsics=False)
    assert len(no_intr) == len(attrs) - 1


def test_solver_step_flags():
    sb = _scene()
    s = SolverStep()
    s.set_max_iterations(7)
    s.set_frame_list([1, 2, 3])
    s.set_attributes_use_static(False)
    s.set_tau_factor(0.5)
    s.set_frame_solve_mode(1)
    (action, vaction), = list(s.compile(sb, withtest=True))
    kw = action.kwargs
    assert action.func == "mmSolver" and kw["iterations"] == 7 and kw["frame"] == [1, 2, 3]
    assert kw["tauFactor"] == 0.5 and kw["frameSolveMode"] == 1 and kw["sceneGraphMode"] == 2 and kw["timeEvalMode"] == 0
    assert len(kw["attr"]) == 6 and all(a[0] < 0 for a in kw["attr"])   # animated only
    assert "delta" not in kw and "epsilon1" not in kw                  # unset values keep command defaults
    assert vaction.kwargs["printStatistics"] == ["inputs"]
    s.set_attributes_use_animated(False)
    assert list(s.compile(sb)) == []                                   # 'No Attributes found!'
    s2 = SolverStep()
    assert list(s2.compile(sb)) == []                                  # 'No Frames found!'
    # solverstep.py:642-657: 'auto' (0) reaches the command as maya_dag (1), which solves all frames
    # at once; a per-frame solve needs the mm scene graph
    s3 = SolverStep()
    s3.set_frame_list([1, 2])
    s3.set_scene_graph_mode(0)
    (a3, _), = list(s3.compile(sb))
    assert a3.kwargs["sceneGraphMode"] == 1
    s3.set_frame_solve_mode(1)
    with pytest.raises(AssertionError):
        list(s3.compile(sb))


def test_standard_action_sequence():
    sb = _scene()
    s = SolverStandard()
    s.set_root_frame_list([1, 6, 12])
    s.set_frame_list(range(1, 13))
    s.set_root_iteration_num(30)
    s.set_anim_iteration_num(10)
    acts = [a for a, _ in s.compile(sb)]
    assert [a.func for a in acts] == ["mmSolver", "interpolate_inbetween", "mmSolver"]
    root, interp, anim = acts
    assert root.kwargs["frame"] == [1, 6, 12] and root.kwargs["iterations"] == 30 and root.kwargs["frameSolveMode"] == 0
    assert len(root.kwargs["attr"]) == 26                       # static + animated
    assert interp.kwargs["frames"] == [2, 3, 4, 5, 7, 8, 9, 10, 11] and interp.kwargs["root_frames"] == [1, 6, 12]
    assert anim.kwargs["frame"] == list(range(1, 13)) and anim.kwargs["frameSolveMode"] == 1
    assert anim.kwargs["iterations"] == 10 and len(anim.kwargs["attr"]) == 6   # animated only, per frame
    # global solve instead of per-frame
    s.set_global_solve(True)
    acts = [a for a, _ in s.compile(sb)]
    assert acts[-1].kwargs["frameSolveMode"] == 0 and len(acts[-1].kwargs["attr"]) == 26
    # only root frames
    s.set_only_root_frames(True)
    assert [a.func for a, _ in s.compile(sb)] == ["mmSolver", "interpolate_inbetween"]
    # strategies
    s.set_only_root_frames(False)
    s.set_global_solve(False)
    for strategy, want in ((1, [[1, 6], [6, 12]]), (2, [[1, 6], [6, 12], [1, 6, 12]]), (3, [[1, 6, 12]])):
        s.set_root_frame_strategy(strategy)
        acts = [a for a, _ in s.compile(sb)]
        assert [a.kwargs["frame"] for a in acts[:-2]] == want
    # attribute blocks multiply the commands and split the iterations (min 3)
    s.set_root_frame_strategy(0)
    s.set_use_attr_blocks(True)
    acts = [a for a, _ in s.compile(sb)]
    roots = [a for a in acts if a.func == "mmSolver" and a.kwargs["frame"] == [1, 6, 12]]
    assert [a.kwargs["iterations"] for a in roots] == [6, 3, 3, 3, 3, 12]
    # single frame
    s.set_use_single_frame(True)
    s.set_single_frame(4)
    s.set_use_attr_blocks(False)
    (a,) = [a for a, _ in s.compile(sb)]
    assert a.kwargs["frame"] == [4] and a.kwargs["iterations"] == 100 and all(x[0] < 0 for x in a.kwargs["attr"])


def test_basic_action_sequence():
    sb = _scene()
    s = SolverBasic()
    s.set_frame_list([3, 4, 5])
    acts = [a for a, _ in s.compile(sb)]
    assert [a.func for a in acts] == ["mmSolver", "euler_filter"]
    assert acts[0].kwargs["frameSolveMode"] == 1 and acts[0].kwargs["frame"] == [3, 4, 5]
    assert len(acts[0].kwargs["attr"]) == 6


def test_interpolate_and_euler_filter():
    sb, info = synth.make_scene(9, 4, 2)
    tx = info["cam_tfm"]["tx"]
    for i in range(9):
        sb.set_value(tx, 100.0 if i not in (0, 4, 8) else float(i), i)
    attrs = solver.scene_attr_list(sb)
    solver.interpolate_inbetween(sb, attrs[:1], [1, 5, 9], [2, 3, 4, 6, 7, 8])
    np.testing.assert_allclose([sb.get_value(tx, i) for i in range(9)], np.arange(9.0))
    rx = info["cam_tfm"]["rx"]
    for i, v in enumerate([170.0, 179.0, -178.0, -170.0, 175.0, 171.0, 160.0, 150.0, 140.0]):
        sb.set_value(rx, v, i)
    solver.euler_filter(sb, [(rx, None, None, None, None)])
    got = np.array([sb.get_value(rx, i) for i in range(9)])
    np.testing.assert_allclose(got, [170, 179, 182, 190, 175, 171, 160, 150, 140])
    assert np.abs(np.diff(got)).max() < 180
