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
_IGNORED = {"verbose": "v", "logLevel": "lgl", "resultsNode": "rsn", "setMarkerDeviationAttrs": "smd"}
_STIFF_FLAGS = {"attrStiffness": "asf", "attrSmoothness": "asm"}
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


def select_scene(scene, camera=None, marker=None, attr=None, frame=None, attrStiffness=None, attrSmoothness=None):
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
        keep = []
        for mkr_name, cam_name, bnd_name in marker:
            node = sb.names[mkr_name]
            if sb.mkr_cam[node.index] != sb.names[cam_name].index or sb.mkr_bnd[node.index] != sb.names[bnd_name].index:
                raise ValueError("marker %r is not linked to camera %r / bundle %r" % (mkr_name, cam_name, bnd_name))
            keep.append(node.index)
        sb.keep_markers(keep)
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
    # -attrStiffness / -attrSmoothness "node.attr" "weightPlug" "variancePlug" "valuePlug"
    # (arg_flags_attr_details.cpp:60-200): the attribute is looked up among the -attr flags
    for flags, smooth in ((attrStiffness, False), (attrSmoothness, True)):
        if flags is None:
            continue
        if smooth:
            sb.smoothness = []
        else:
            sb.stiffness = []
        for entry in flags:
            if len(entry) != 4:
                continue   # the command logs an error and skips the flag
            as_ref = lambda v: int(sb.plug(v)) if isinstance(v, str) else int(v)   # plug name or attribute-store ref
            ref, wref, vref, valref = (as_ref(v) for v in entry)
            if ref not in sb.attr_ref:
                continue   # "is not a declared attribute"
            sb.add_stiffness(sb.attr_ref.index(ref), wref, vref, valref, smoothness=smooth)
    return sb


def format_results(problem, result, scene, flags, solved=True):
    """CommandResult::appendToMStringArray through the C-ABI -> list of str.
    `scene`: frames + the names of the markers the handle numbers (its used list)."""
    lib = _capi.load()
    e2m, _ = problem.index_maps()
    fvec = np.zeros(problem.m)
    if solved:
        check(lib.mmb_get_last_residuals(problem._h, fvec.ctypes.data_as(c_f64p)), "mmb_get_last_residuals")
    frames = np.ascontiguousarray(scene.frames, dtype=np.int32)
    marker_names = getattr(scene, "marker_names", None)
    if marker_names is None:
        marker_names = _names(scene, "marker", len(scene.mkr_cam))
    M = len(marker_names)
    names = (C.c_char_p * M)(*[n.encode() for n in marker_names])
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


def _names(scene, kind, count):
    by_index = {n.index: n.name for n in scene.names.values() if n.kind == kind}
    return [by_index.get(i) or ("%s%d" % (kind, i)) for i in range(count)]


def _attr_names(scene):
    """'node.attr' of every solver attribute of the scene (its plug name when the node is named)"""
    by_ref = {}
    for node in scene.names.values():
        for attr, ref in node.items():
            by_ref.setdefault(ref, "%s.%s" % (node.name, attr))
    return [by_ref.get(ref, "attr%d" % i) for i, ref in enumerate(scene.attr_ref)]


def affects_strings(scene, used_only=None):
    """AffectsResult::appendToMStringArray (adjust_results.h:664-680):
    `marker_affects_attribute=<marker>#<node.attr>#<0|1>` for every (marker, attribute) pair"""
    from .problem import affects
    table, mu, au = affects(scene)
    mk_names = _names(scene, "marker", len(scene.mkr_cam))
    at_names = _attr_names(scene)
    out = []
    for i in range(table.shape[0]):
        if used_only is not None and not used_only[0][i]:
            continue
        for j in range(table.shape[1]):
            if used_only is not None and not used_only[1][j]:
                continue
            out.append("marker_affects_attribute=%s#%s#%d" % (mk_names[i], at_names[j], int(table[i, j])))
    return out


def solver_types():
    """getSolverTypes (adjust_base.cpp:80-93): the (index, name) pairs, in the reference's order"""
    return [(SOLVER_TYPE_CMINPACK_LMDIF, "cminpack_lmdif"), (SOLVER_TYPE_CMINPACK_LMDER, "cminpack_lmder")]


def solver_type_default():
    """getSolverTypeDefault (adjust_base.cpp:96-130): index cminpack_lmder unless the environment
    variable MMSOLVER_DEFAULT_SOLVER names another type; as in the reference the NAME is only
    filled in from that variable (it is empty without it), and an unknown value keeps the default"""
    import os
    index, name = SOLVER_TYPE_CMINPACK_LMDER, ""
    wanted = os.environ.get("MMSOLVER_DEFAULT_SOLVER")
    if wanted is not None:
        for i, n in solver_types():
            if wanted == n:
                index, name = i, n
    return index, name


def mmSolverType(query=False, list=False, default=False, name=None, index=None, **short):
    """The `mmSolverType` command (src/mmSolver/cmd/MMSolverTypeCmd.cpp:96-258): the solver types
    this build has, or the default one, as 'index=name' strings, names or indices.  The short flag
    names of MMSolverTypeCmd.h:63-76 (q, ls, def, n, idx) are accepted too (`def` only through
    `**{"def": True}`: it is a Python keyword)."""
    query = short.pop("q", query)
    list = short.pop("ls", list)
    default = short.pop("def", default)
    name = short.pop("n", name)
    index = short.pop("idx", index)
    if short:
        raise TypeError("mmSolverType: unknown flag(s) %s" % sorted(short))
    if not query:
        raise RuntimeError("mmSolverType must query using the 'query' flag")
    if bool(list) == bool(default):
        raise RuntimeError("mmSolverType; must give 'list' or 'default' flag, not both flags at once")
    want_name = True if name is None else bool(name)
    want_index = True if index is None else bool(index)
    if not want_name and not want_index:
        raise RuntimeError("mmSolverType must have 'name' or 'index' flags")

    def item(i, n):
        if want_name:
            return ("%d=" % i if want_index else "") + n
        return i

    if list:
        return [item(i, n) for i, n in solver_types()]
    return item(*solver_type_default())


def mmSolverAffects(scene, mode="returnString", **kwargs):
    """The `mmSolverAffects` command (src/mmSolver/cmd/MMSolverAffectsCmd.cpp): which solver
    attribute can affect which marker.  mode 'returnString' -> the `marker_affects_attribute=`
    strings; mode 'addAttrsToMarkers' stores the table on the scene (`scene.affects`, the
    stand-in for the per-marker Maya attributes the command creates) and returns None."""
    _, rest = _options_from_kwargs(kwargs)
    objects = {}
    for key, value in rest.items():
        long_name = {v: k for k, v in _OBJECT_FLAGS.items()}.get(key, key)
        if long_name in _OBJECT_FLAGS:
            objects[long_name] = value
    objects.pop("frame", None)
    sb = select_scene(scene, **objects) if objects else scene
    if mode == "returnString":
        return affects_strings(sb)
    if mode == "addAttrsToMarkers":
        from .problem import affects
        table, _, _ = affects(sb)
        mk_names = _names(sb, "marker", len(sb.mkr_cam))
        at_names = _attr_names(sb)
        scene.affects = {(mk_names[i], at_names[j]): bool(table[i, j]) for i in range(table.shape[0]) for j in range(table.shape[1])}
        return None
    raise ValueError("mmSolverAffects: mode value is invalid: mode=%r" % mode)


def mmSolverSceneGraph(scene, mode="debugConstruct", **kwargs):
    """The `mmSolverSceneGraph` command (src/mmSolver/cmd/MMSolverSceneGraphCmd.cpp:140-208): can
    the scene of these cameras / markers / attributes / frames be built for the given scene graph
    mode?  The reference's Python API asks this before it picks sceneGraphMode = 2
    (`construct_scene_graph`); maya_dag (1) always answers True.  Here the scene already is a flat
    scene: the answer is whether the library's own scene checks (the host half of
    mmb_problem_create, through the host-only mmb_affects) accept it.  A scene that cannot be
    built answers False, it does not raise -- as the command does."""
    if mode != "debugConstruct":
        raise RuntimeError("mmSolverSceneGraph: Mode value is invalid: mode=%r" % (mode,))
    options, rest = _options_from_kwargs(kwargs)
    sgm = int(options.get("scene_graph_mode", SCENE_GRAPH_MODE_MAYA_DAG))   # the command's default
    if sgm == SCENE_GRAPH_MODE_MAYA_DAG:
        return True
    if sgm != SCENE_GRAPH_MODE_MM_SCENE_GRAPH:
        raise RuntimeError("mmSolverSceneGraph: Scene Graph Mode value is invalid: value=%d" % sgm)
    objects = {}
    for key, value in rest.items():
        long_name = {v: k for k, v in _OBJECT_FLAGS.items()}.get(key, key)
        if long_name in _OBJECT_FLAGS:
            objects[long_name] = value
    from .problem import affects
    from ._capi import MmbError
    try:
        sb = select_scene(scene, **objects) if objects else scene
        affects(sb)
    except (MmbError, ValueError, KeyError, IndexError):
        return False
    return True


def mmSolver(scene, write_back=True, device=0, **kwargs):
    """-> list of `key=value` strings, as `maya.cmds.mmSolver` returns them."""
    opts, rest = _options_from_kwargs(kwargs)
    objects = {}
    print_stats_flags = []
    for key, value in rest.items():
        long_name = {v: k for k, v in _OBJECT_FLAGS.items()}.get(key, key)
        if long_name in _OBJECT_FLAGS:
            objects[long_name] = value
        elif key in _STIFF_FLAGS or key in _STIFF_FLAGS.values():
            objects[{v: k for k, v in _STIFF_FLAGS.items()}.get(key, key)] = value
        elif key in ("printStatistics", "pst"):
            print_stats_flags = [value] if isinstance(value, str) else list(value)
        elif key in _IGNORED or key in _IGNORED.values():
            continue  # accepted for compatibility; Maya-side features (log level, result nodes)
        else:
            raise TypeError("mmSolver: unknown flag %r" % key)
    for name in print_stats_flags:
        if name not in _PRINT_STATS:
            raise ValueError("printStatistics: unknown value %r" % name)
    # solverstep.py:647-657: per-frame needs mmscenegraph, the Maya DAG mode solves all frames at once
    sgm = opts.get("scene_graph_mode", SCENE_GRAPH_MODE_MM_SCENE_GRAPH)
    if "scene_graph_mode" not in opts:
        # the reference command defaults to maya_dag (1); without Maya the default here is the
        # mmscenegraph arithmetic.  sceneGraphMode=1 selects the Maya-DAG-convention arithmetic of
        # the residual (row-vector projection, behind-camera factor, stiffness / smoothness rows).
        opts["scene_graph_mode"] = SCENE_GRAPH_MODE_MM_SCENE_GRAPH
    if opts.get("frame_solve_mode", 0) == FRAME_SOLVE_MODE_PER_FRAME:
        assert sgm == SCENE_GRAPH_MODE_MM_SCENE_GRAPH, "per-frame solving needs sceneGraphMode = mmscenegraph"
    sb = select_scene(scene, **objects) if objects else scene
    options = default_options(device=device, **opts)
    do_not_solve = bool(print_stats_flags)  # constructPrintStats, adjust_base.cpp:618-646
    flags = _capi.MMB_FMT_ERROR_PER_MARKER_PER_FRAME | _capi.MMB_FMT_ERROR_PER_FRAME | _capi.MMB_FMT_SOLVE_VALUES
    if "inputs" in print_stats_flags:
        flags |= _capi.MMB_FMT_INPUT_STATS
    # used / unused solve objects (solve_v1, adjust_base.cpp:1322-1395)
    from .problem import affects
    _, marker_used, attr_used = affects(sb)
    if not (options.remove_unused_markers or options.remove_unused_attributes):
        marker_used[:] = True
        attr_used[:] = True
    mk_names = _names(sb, "marker", len(sb.mkr_cam))
    at_names = _attr_names(sb)
    extra = []
    if "usedSolveObjects" in print_stats_flags:
        # SolverObjectUsageResult::appendToMStringArray (adjust_results.h:755-790)
        groups = (("markers_used=", [n for n, u in zip(mk_names, marker_used) if u]),
                  ("markers_unused=", [n for n, u in zip(mk_names, marker_used) if not u]),
                  ("attributes_used=", [n for n, u in zip(at_names, attr_used) if u]),
                  ("attributes_unused=", [n for n, u in zip(at_names, attr_used) if not u]))
        extra += [key + "".join(n + "#" for n in names) for key, names in groups if names]
    if "affects" in print_stats_flags:
        use_m = marker_used if options.remove_unused_markers else np.ones_like(marker_used)
        use_a = attr_used if options.remove_unused_attributes else np.ones_like(attr_used)
        extra += affects_strings(sb, used_only=(use_m, use_a))
    n_used_m = int(marker_used.sum()) if options.remove_unused_markers else len(mk_names)
    n_used_a = int(attr_used.sum()) if options.remove_unused_attributes else len(at_names)
    if n_used_m == 0 or n_used_a == 0:
        # adjust_base.cpp:844-862: not enough used markers / attributes
        return extra + (["success=1"] if do_not_solve else ["success=0"])

    def splice(strings):
        k = 5 if "inputs" in print_stats_flags else 0   # the numberOf* block comes first
        return strings[:k] + extra + strings[k:]

    with Problem(sb, options=options) as p:
        n_rows = p.m - 2 * p.nobs       # stiffness + smoothness rows (MMB_FMT_STIFFNESS / _SMOOTHNESS)
        n_smooth = _count_rows(sb, sb.smoothness)
        flags |= ((n_rows - n_smooth) & 0x7ff) << 8 | (n_smooth & 0x7ff) << 19
        used_m, used_a = p.used_objects()
        names_scene = _UsedView(sb, [mk_names[i] for i in used_m])
        if do_not_solve:
            # adjust_base.cpp:1105-1135: the initial measurement only
            r = _initial_measure_result(p, options)
            return splice(format_results(p, r, names_scene, flags, solved=False))
        if p.n > p.m:
            # adjust_base.cpp:864-882: more parameters than errors -> success=0, nothing solved
            r = _initial_measure_result(p, options)
            r.raw.success = 0
            return format_results(p, r, names_scene, flags, solved=False)
        r = p.solve_per_frame()[0] if options.frame_solve_mode == FRAME_SOLVE_MODE_PER_FRAME else p.solve()
        strings = format_results(p, r, names_scene, flags)
        _, p2a = p.index_maps()
    if write_back and r.success:
        # attr_values already holds the initial values when acceptOnlyBetter rejected the solve
        frame_index = {int(n): i for i, n in enumerate(scene.frames)}
        for i in range(len(r.attr_values)):
            a, f = int(used_a[p2a[i, 0]]), int(p2a[i, 1])
            ref = sb.attr_ref[a]
            if f < 0:
                scene.set_value(ref, r.attr_values[i])
            else:
                scene.set_value(ref, r.attr_values[i], frame_index[int(sb.frames[f])])
    return strings


def _count_rows(scene, entries):
    """entries whose weight, at the current frame, is > 0 (adjust_relationships.cpp:183-219)"""
    return sum(1 for e in entries if scene.get_value(e[1], scene.current_frame) > 0.0)


class _UsedView:
    """what format_results needs of a scene: the frames and the names of the USED markers"""

    def __init__(self, scene, marker_names):
        self.frames = scene.frames
        self.marker_names = marker_names


def _initial_measure_result(p, options):
    from .problem import SolveResult
    raw = _capi.Result()
    raw.success = 1
    dist = np.zeros(p.nobs)
    if options.accept_only_better:
        _, dist = p.eval_residuals(None)
        finite = dist[np.isfinite(dist)]
        if len(finite):
            raw.error_avg, raw.error_min, raw.error_max = float(finite.mean()), float(finite.min()), float(finite.max())
    return SolveResult(raw, np.zeros(p.n), np.zeros(p.n), dist)
