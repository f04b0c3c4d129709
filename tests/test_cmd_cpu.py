"""CPU: the mmSolver command surface -- flag names, the `key=value` result strings
(adjust_results.h:871-938) and the reader of those strings, pinned to what the REFERENCE's
own parser (python/mmSolver/_api/solveresult.py) returns for the same strings
(tests/golden/solveresult_golden.json, made by tests/golden/make_solveresult_golden.py)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from mayamatchmovesolver_b200 import _capi, build, cmd, solveresult
from mayamatchmovesolver_b200.scene import FLT_MAX, SceneBuilder

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(HERE, "golden", "solveresult_golden.json")) as fh:
        return json.load(fh)


def test_reader_matches_reference_parser(golden):
    data = solveresult.parse_command_result(golden["sample"])
    assert sorted(data.keys()) == golden["parsed_keys"]
    assert solveresult.solver_stats(data) == golden["solver_stats"]
    assert solveresult.error_stats(data) == golden["error_stats"]
    assert solveresult.timer_stats(data) == golden["timer_stats"]
    assert solveresult.print_stats(data) == golden["print_stats"]
    assert {repr(k): v for k, v in solveresult.error_per_frame(data).items()} == golden["error_per_frame"]
    got = {m: {repr(k): v for k, v in d.items()} for m, d in solveresult.error_per_marker_per_frame(data).items()}
    assert got == golden["error_per_marker_per_frame"]
    r = solveresult.SolveResultStrings(golden["sample"])
    assert r.get_success() and r.get_final_error() == 8.13955
    assert r.solve_parameter_list == [1.5, -2.25, 3e-07] and r.solve_error_list == [0.1, 0.2, 0.3, 0.4]


def _format(r, n, nobs, e2m, frames, names, x, fvec, dist, flags):
    build.build()
    lib = _capi.load()
    e2m = np.ascontiguousarray(e2m, np.int32)
    frames = np.ascontiguousarray(frames, np.int32)
    arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
    p = lambda a: np.ascontiguousarray(a, np.float64).ctypes.data_as(_capi.c_f64p)
    needed = C.c_int64()
    st = lib.mmb_format_results(C.byref(r), n, nobs, e2m.ctypes.data_as(_capi.c_i32p), frames.ctypes.data_as(_capi.c_i32p),
                                len(frames), arr, len(names), p(x), p(fvec), p(dist), flags, None, 0, C.byref(needed))
    assert st == 0
    buf = C.create_string_buffer(needed.value)
    st = lib.mmb_format_results(C.byref(r), n, nobs, e2m.ctypes.data_as(_capi.c_i32p), frames.ctypes.data_as(_capi.c_i32p),
                                len(frames), arr, len(names), p(x), p(fvec), p(dist), flags, buf, needed.value, C.byref(needed))
    assert st == 0
    # a buffer that is too small is refused, with the size it needs
    small = C.create_string_buffer(8)
    st = lib.mmb_format_results(C.byref(r), n, nobs, e2m.ctypes.data_as(_capi.c_i32p), frames.ctypes.data_as(_capi.c_i32p),
                                len(frames), arr, len(names), p(x), p(fvec), p(dist), flags, small, 8, C.byref(needed))
    assert st == _capi.MMB_ERR_INVALID_ARGUMENT and needed.value == len(buf.value) + 1
    return buf.value.decode().split("\n")[:-1]


def test_formatter_emits_the_commands_strings(golden):
    """the C-ABI formatter reproduces the golden sample (the strings the reference parser was
    fed), key order of CommandResult::appendToMStringArray included"""
    r = _capi.Result()
    r.success, r.reason_num, r.error_final, r.error_avg, r.error_max, r.error_min = 1, 2, 8.13955, 0.51381, 1.67977, 0.043739
    r.iteration_num, r.iteration_function_num, r.iteration_jacobian_num, r.user_interrupted = 17, 17, 672, 0
    r.timer_solve_s, r.timer_function_s, r.timer_jacobian_s = 0.0513, 0.0014, 0.0038
    e2m = [[0, 0], [0, 1], [1, 0], [1, 1]]          # marker-major, frame-minor
    dist = [0.25, 0.5, 0.75, 1e-05]
    lines = _format(r, 3, 2 * 30, e2m, [1, 2], ["mkr_A", "mkr_B"], [1.5, -2.25, 3e-07], np.zeros(120), dist, 8)
    sample = golden["sample"]
    # numberOf* counts come from the sizes handed in (n_obs = 60 -> 120 errors); patch the two
    # lines whose inputs differ from the sample, then everything up to the lists must be equal
    assert lines[0] == "numberOfParameters=3" and lines[1] == "numberOfErrors=120"
    assert lines[2:25] == sample[2:25]
    # observation arrays of 4 entries for the per-marker / per-frame block
    lines = _format(r, 3, 4, e2m, [1, 2], ["mkr_A", "mkr_B"], [1.5, -2.25, 3e-07], [0.1, 0.2, 0.3, 0.4, 9, 9, 9, 9], dist, 1 | 2 | 4)
    assert lines[20:24] == sample[25:29]           # error_per_marker_per_frame, marker-major
    assert lines[24:26] == sample[29:31]           # error_per_frame: mean deviation per frame
    assert lines[26] == sample[31]                 # solve_parameter_list=...#
    assert lines[27] == "solve_error_list=0.1#0.2#0.3#0.4#9#9#9#9#"
    # and the reference-parser mirror reads them back
    back = solveresult.SolveResultStrings(lines)
    assert back.per_marker_per_frame_error["mkr_B"][2.0] == 1e-05 and back.per_frame_error[1.0] == 0.5


def _named_scene():
    sb = SceneBuilder(frames=[10, 11, 12, 13])
    ct = sb.add_transform(name="cam_tfm", tx=[0, 0.1, 0.2, 0.3], ty=0.0, tz=5.0)
    cam = sb.add_camera(ct, name="cam_shp")
    for i in range(3):
        bt = sb.add_transform(name="bnd%d" % i, tx=float(i), ty=1.0, tz=-10.0)
        bnd = sb.add_bundle(bt, name="bundle%d" % i)
        sb.add_marker(cam, bnd, 0.01 * i, 0.02, name="mkr%d" % i)
    return sb


def test_object_flags_select_markers_attrs_frames():
    sb = _named_scene()
    out = cmd.select_scene(sb, camera=[("cam_tfm", "cam_shp")],
                           marker=[("mkr0", "cam_shp", "bundle0"), ("mkr2", "cam_shp", "bundle2")],
                           attr=[("bnd0.tx", "None", "None", "None", "None"), ("cam_tfm.tx", "-1.5", 2, "0.5", "None"),
                                 ("bad", "flag")],
                           frame=[12, 10])
    assert list(out.frames) == [12, 10] and out.F == 2
    # only the listed markers are part of the solve (markerList), in flag order
    assert len(out.mkr_cam) == 2 and out.names["mkr0"].index == 0 and out.names["mkr2"].index == 1
    assert "mkr1" not in out.names and out.mkr_bnd == [sb.names["bundle0"].index, sb.names["bundle2"].index]
    assert out.attr_ref == [sb.names["bnd0"]["tx"], sb.names["cam_tfm"]["tx"]]
    # non-numeric strings keep the defaults of maya_attr.cpp:88-91
    assert out.attr_min == [-FLT_MAX, -1.5] and out.attr_max == [FLT_MAX, 2.0]
    assert out.attr_off == [0.0, 0.5] and out.attr_scl == [1.0, 1.0]
    np.testing.assert_array_equal(out._anim[0], [0.2, 0.0])   # cam_tfm.tx at frames 12, 10
    with pytest.raises(ValueError):
        cmd.select_scene(sb, frame=[99])
    with pytest.raises(ValueError):
        cmd.select_scene(sb, marker=[("mkr0", "cam_shp", "bundle1")])


def test_flag_names_long_and_short():
    opts, rest = cmd._options_from_kwargs(dict(iterations=7, t=3.0, e1=1e-3, epsilon2=1e-4, adt=1, frameSolveMode=1,
                                               sgm=2, imw=1920.0, aob=False, verbose=True))
    assert opts == dict(iter_max=7, tau=1.0, eps1=1e-3, eps2=1e-4, auto_diff_type=1, frame_solve_mode=1,
                        scene_graph_mode=2, image_width=1920.0, accept_only_better=0)   # tauFactor clamped to [0, 1]
    assert rest == dict(verbose=True)
    with pytest.raises(TypeError):
        cmd.mmSolver(_named_scene(), notAFlag=1)


def test_mmSolverType_mirrors_the_reference_test(monkeypatch):
    """tests/test/test_solver/test_solver_type.py:49-121 (src/mmSolver/cmd/MMSolverTypeCmd.cpp,
    getSolverTypes / getSolverTypeDefault adjust_base.cpp:80-130), with the reference's assertions"""
    monkeypatch.delenv("MMSOLVER_DEFAULT_SOLVER", raising=False)
    a_list = cmd.mmSolverType(query=True, list=True)
    assert len(a_list) > 0
    for solver_type in a_list:
        assert "=" in solver_type and isinstance(solver_type, str)
    assert a_list == ["1=cminpack_lmdif", "2=cminpack_lmder"]
    b_list = cmd.mmSolverType(query=True, list=True, name=True, index=False)
    assert len(b_list) == len(a_list) and all("=" not in n for n in b_list)
    c_list = cmd.mmSolverType(query=True, list=True, name=False, index=True)
    assert len(c_list) == len(a_list) and all(isinstance(i, int) for i in c_list)
    with pytest.raises(RuntimeError):
        cmd.mmSolverType(query=True, list=True, name=False, index=False)
    with pytest.raises(RuntimeError):
        cmd.mmSolverType(list=True)                       # not a query
    with pytest.raises(RuntimeError):
        cmd.mmSolverType(query=True, list=True, default=True)
    with pytest.raises(RuntimeError):
        cmd.mmSolverType(query=True)
    solver_type = cmd.mmSolverType(query=True, default=True)
    solver_name = cmd.mmSolverType(query=True, default=True, name=True, index=False)
    solver_index = cmd.mmSolverType(query=True, default=True, name=False, index=True)
    assert "=" in solver_type and "=" not in solver_name and len(solver_name) < len(solver_type)
    assert solver_index == cmd.SOLVER_TYPE_CMINPACK_LMDER == 2
    # the index the command reports is the one the C-ABI option takes
    from mayamatchmovesolver_b200.problem import default_options
    assert default_options().solver_type == solver_index
    # MMSOLVER_DEFAULT_SOLVER picks the default (and is where the default's name comes from)
    monkeypatch.setenv("MMSOLVER_DEFAULT_SOLVER", "cminpack_lmdif")
    assert cmd.mmSolverType(query=True, default=True) == "1=cminpack_lmdif"
    monkeypatch.setenv("MMSOLVER_DEFAULT_SOLVER", "ceres")   # not in this build: the default stays
    assert cmd.mmSolverType(q=True, idx=True, n=False, **{"def": True}) == 2


def test_mmSolverSceneGraph_answers_whether_the_scene_can_be_built():
    """src/mmSolver/cmd/MMSolverSceneGraphCmd.cpp:140-208: maya_dag always True; mm scene graph:
    True / False (never an exception) from the scene checks; other modes are errors"""
    from mayamatchmovesolver_b200 import synth
    sb, _ = synth.make_scene(F=4, B=5, seed=1)
    assert cmd.mmSolverSceneGraph(sb) is True                                        # default mode: maya_dag
    assert cmd.mmSolverSceneGraph(sb, sceneGraphMode=cmd.SCENE_GRAPH_MODE_MM_SCENE_GRAPH) is True
    assert cmd.mmSolverSceneGraph(sb, sgm=2, frame=list(sb.frames)[:2]) is True        # short flag, object flags
    import copy
    bad = copy.deepcopy(sb)
    bad._arrays_cache = None
    bad.mkr_bnd[0] = 10 ** 6          # a marker whose bundle does not exist
    assert cmd.mmSolverSceneGraph(bad, sceneGraphMode=2) is False
    assert cmd.mmSolverSceneGraph(bad, sceneGraphMode=1) is True
    assert cmd.mmSolverSceneGraph(sb, sceneGraphMode=2, frame=[10 ** 6]) is False      # a frame the scene does not have
    with pytest.raises(RuntimeError):
        cmd.mmSolverSceneGraph(sb, mode="construct")
    with pytest.raises(RuntimeError):
        cmd.mmSolverSceneGraph(sb, sceneGraphMode=3)
