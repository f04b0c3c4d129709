"""`mmSolver(...)`: the reference's Maya command surface over plain arrays.

The reference's Python API reaches the solver through `maya.cmds.mmSolver(**kwargs)`
(python/mmSolver/_api/solverstep.py:549-712; flags: src/mmSolver/cmd/arg_flags_solve_info.h,
arg_flags_solve_object.h, arg_flags_solve_log.h).  There is no Maya here, so node names refer
to the named nodes of a SceneBuilder; everything else -- flag names (long and short), values,
defaults, error behaviour, the returned list of `key=value` strings
(adjust_results.h:871-938) -- is the command's.

    strings = mmSolver(scene,
                       camera=[("cam_tfm", "cam_shp")],
                       marker=[("mkr1", "cam_shp", "bnd1"), ...],
                       attr=[("cam_tfm.tx", "None", "None", "None", "None"), ...],
                       frame=[1, 2, 3], iterations=100, frameSolveMode=0, sceneGraphMode=2)
    result = solveresult.SolveResultStrings(strings)

Without camera / marker / attr / frame flags the scene is solved as it stands.  Solved values
are written back into the scene's attribute store when the command would set them
(set_maya_attribute_values, adjust_base.cpp:297-342, under acceptOnlyBetter).
"""
import copy
import ctypes as C

import numpy as np

from . import _capi
from ._capi import c_f64p, c_i32p, check
from .problem import Problem, default_options
from .scene import FLT_MAX

# long flag -> (short flag, option field, converter); arg_flags_solve_info.h:40-172
_FLAGS = {
    "solverType": ("st", "solver_type", int),
    "sceneGraphMode": ("sgm", "scene_graph_mode", int),
    "frameSolveMode": ("fsm", "frame_solve_mode", int),
    "timeEvalMode": ("tem", "time_eval_mode", int),
    "iterations": ("it", "iter_max", int),
    "tauFactor": ("t", "tau", float),
    "epsilon1": ("e1", "eps1", float),
    "epsilon2": ("e2", "eps2", float),
    "epsilon3": ("e3", "eps3", float),
    "delta": ("dt", "delta", float),
    "autoDiffType": ("adt", "auto_diff_type", int),
    "autoParamScaling": ("aps", "auto_param_scale", int),
    "robustLossType": ("rlt", "robust_loss_type", int),
    "robustLossScale": ("rls", "robust_loss_scale", float),
    "acceptOnlyBetter": ("aob", "accept_only_better", int),
    "removeUnusedMarkers": ("rum", "remove_unused_markers", int),
    "removeUnusedAttributes": ("rua", "remove_unused_attributes", int),
    "imageWidth": ("imw", "image_width", float),
}
_SHORT = {v[0]: k for k, v in _FLAGS.items()}
_OBJECT_FLAGS = {"camera": "c", "marker": "m", "attr": "a", "frame": "f"}
_IGNORED = {"verbose": "v", "logLevel": "lgl", "resultsNode": "rsn", "setMarkerDeviationAttrs": "smd",
            "attrStiffness": "asf", "attrSmoothness": "asm"}
_PRINT_STATS = ("inputs", "affects", "usedSolveObjects", "deviation")

# sceneGraphMode / solverType / frameSolveMode values (adjust_defines.h:92-152, _api/constant.py)
SCENE_GRAPH_MODE_MAYA_DAG = 1
SCENE_GRAPH_MODE_MM_SCENE_GRAPH = 2
SOLVER_TYPE_CMINPACK_LMDIF = 1
SOLVER_TYPE_CMINPACK_LMDER = 2
FRAME_SOLVE_MODE_ALL_FRAMES_AT_ONCE = 0
FRAME_SOLVE_MODE_PER_FRAME = 1


def _is_double(v):
    """MString::isDouble: 'None' and other non-numeric strings keep the attribute default
    (arg_flags_solve_object.cpp:288-318)"""
    if isinstance(v, (int, float)) and not isinstance(v, bool):
        return True
    try:
        float(v)
        return True
    except (TypeError, ValueError):
        return False


def _options_from_kwargs(kwargs):
    opts = {}
    rest = {}
    for key, value in kwargs.items():
        name = _SHORT.get(key, key)
        if name in _FLAGS:
            _, field, conv = _FLAGS[name]
            if name == "tauFactor":
                value = min(max(float(value), 0.0), 1.0)  # arg_flags_solve_info.cpp: clamped to [0, 1]
            opts[field] = conv(value)
        else:
            rest[name] = value
    return opts, rest


def select_scene(scene, camera=None, marker=None, attr=None, frame=None):
    """Apply the command's object flags to a copy of `scene`.
    marker: only the listed markers take part; attr: the listed plugs become the solver
    attributes (in flag order); frame: the listed frame numbers (in flag order)."""
    sb = copy.deepcopy(scene)
    sb._arrays_cache = None
    if frame is not None:
        numbers = [int(f) for f in frame]
        index = {int(n): i for i, n in enumerate(scene.frames)}
        missing = [n for n in numbers if n not in index]
        if missing:
            raise ValueError("frame numbers not in the scene: %r" % missing)
        sb = sb.slice_frames([index[n] for n in numbers])
    if camera is not None:
        for tfm_name, shp_name in camera:
            if tfm_name not in sb.names or shp_name not in sb.names:
                raise ValueError("unknown camera %r / %r" % (tfm_name, shp_name))
    if marker is not None:
        keep = set()
        for mkr_name, cam_name, bnd_name in marker:
            node = sb.names[mkr_name]
            if sb.mkr_cam[node.index] != sb.names[cam_name].index or sb.mkr_bnd[node.index] != sb.names[bnd_name].index:
                raise ValueError("marker %r is not linked to camera %r / bundle %r" % (mkr_name, cam_name, bnd_name))
            keep.add(node.index)
        for i in range(len(sb.mkr_en)):
            if i not in keep:
                sb.mkr_en[i] = np.zeros_like(sb.mkr_en[i])
        sb._arrays_cache = None
    if attr is not None:
        sb.attr_ref, sb.attr_min, sb.attr_max, sb.attr_off, sb.attr_scl = [], [], [], [], []
        sb._arrays_cache = None
        for entry in attr:
            if len(entry) != 5:
                # the command logs an error and skips the flag (arg_flags_solve_object.cpp:262-273)
                continue
            plug, mn, mx, off, scl = entry
            sb.add_attr(sb.plug(plug) if isinstance(plug, str) else int(plug), min=float(mn) if _is_double(mn) else None, max=float(mx) if _is_double(mx) else None,
                        offset=float(off) if _is_double(off) else None, scale=float(scl) if _is_double(scl) else None)
    return sb


def format_results(problem, result, scene, flags):
    """CommandResult::appendToMStringArray through the C-ABI -> list of str"""
    lib = _capi.load()
    e2m, _ = problem.index_maps()
    fvec = np.zeros(problem.m)
    check(lib.mmb_get_last_residuals(problem._h, fvec.ctypes.data_as(c_f64p)), "mmb_get_last_residuals")
    frames = np.ascontiguousarray(scene.frames, dtype=np.int32)
    by_index = {n.index: n.name for n in scene.names.values() if n.kind == "marker"}
    M = len(scene.mkr_cam)
    names = (C.c_char_p * M)(*[(by_index.get(i) or ("marker%d" % i)).encode() for i in range(M)])
    e2m = np.ascontiguousarray(e2m, dtype=np.int32)
    x = np.ascontiguousarray(result.x)
    dist = np.ascontiguousarray(result.error_dist)
    needed = C.c_int64()
    raw = C.byref(result.raw)
    lib.mmb_format_results(raw, problem.n, problem.nobs, e2m.ctypes.data_as(c_i32p), frames.ctypes.data_as(c_i32p),
                           len(frames), names, M, x.ctypes.data_as(c_f64p), fvec.ctypes.data_as(c_f64p),
                           dist.ctypes.data_as(c_f64p), flags, None, 0, C.byref(needed))
    buf = C.create_string_buffer(needed.value)
    check(lib.mmb_format_results(raw, problem.n, problem.nobs, e2m.ctypes.data_as(c_i32p),
                                 frames.ctypes.data_as(c_i32p), len(frames), names, M, x.ctypes.data_as(c_f64p),
                                 fvec.ctypes.data_as(c_f64p), dist.ctypes.data_as(c_f64p), flags, buf, needed.value,
                                 C.byref(needed)), "mmb_format_results")
    return [s for s in buf.value.decode().split("\n") if s]


def mmSolver(scene, write_back=True, device=0, **kwargs):
    """-> list of `key=value` strings, as `maya.cmds.mmSolver` returns them."""
    opts, rest = _options_from_kwargs(kwargs)
    objects = {}
    print_stats_flags = []
    for key, value in rest.items():
        long_name = {v: k for k, v in _OBJECT_FLAGS.items()}.get(key, key)
        if long_name in _OBJECT_FLAGS:
            objects[long_name] = value
        elif key in ("printStatistics", "pst"):
            print_stats_flags = [value] if isinstance(value, str) else list(value)
        elif key in _IGNORED or key in _IGNORED.values():
            continue  # accepted for compatibility; Maya-side features (log level, result nodes, v1 stiffness)
        else:
            raise TypeError("mmSolver: unknown flag %r" % key)
    for name in print_stats_flags:
        if name not in _PRINT_STATS:
            raise ValueError("printStatistics: unknown value %r" % name)
    # solverstep.py:647-657: per-frame needs mmscenegraph, the Maya DAG mode solves all frames at once
    sgm = opts.get("scene_graph_mode", SCENE_GRAPH_MODE_MM_SCENE_GRAPH)
    if "scene_graph_mode" not in opts:
        opts["scene_graph_mode"] = SCENE_GRAPH_MODE_MM_SCENE_GRAPH  # the only mode without Maya
    if opts.get("frame_solve_mode", 0) == FRAME_SOLVE_MODE_PER_FRAME:
        assert sgm == SCENE_GRAPH_MODE_MM_SCENE_GRAPH, "per-frame solving needs sceneGraphMode = mmscenegraph"
    sb = select_scene(scene, **objects) if objects else scene
    options = default_options(device=device, **opts)
    flags = _capi.MMB_FMT_ERROR_PER_MARKER_PER_FRAME | _capi.MMB_FMT_ERROR_PER_FRAME | _capi.MMB_FMT_SOLVE_VALUES
    if "inputs" in print_stats_flags:
        flags |= _capi.MMB_FMT_INPUT_STATS
    with Problem(sb, options=options) as p:
        r = p.solve_per_frame()[0] if options.frame_solve_mode == FRAME_SOLVE_MODE_PER_FRAME else p.solve()
        strings = format_results(p, r, sb, flags)
        _, p2a = p.index_maps()
    if write_back and r.success:
        # attr_values already holds the initial values when acceptOnlyBetter rejected the solve
        frame_index = {int(n): i for i, n in enumerate(scene.frames)}
        for i in range(len(r.attr_values)):
            a, f = int(p2a[i, 0]), int(p2a[i, 1])
            ref = sb.attr_ref[a]
            if f < 0:
                scene.set_value(ref, r.attr_values[i])
            else:
                scene.set_value(ref, r.attr_values[i], frame_index[int(sb.frames[f])])
    return strings
