"""Solver strategies above the hot path: the reference's
`SolverStep`, `SolverBasic` and `SolverStandard` (python/mmSolver/_api/solverstep.py,
solverbasic.py, solverstandard.py, solverstandardutils.py, solverutils.py) compiled to the
same list of Actions -- `mmSolver(**flags)` calls plus the key-frame edits between them --
and executed against a SceneBuilder through `cmd.mmSolver` (the GPU path).

Without Maya an attribute is an `-attr` flag tuple `(plug, min, max, offset, scale)` whose
plug is a 'node.attr' name or an attribute-store ref of the scene; markers and cameras are the
scene's own.  The Maya-only Actions are restated on the attribute store:

  cutKey(non-root frames) + keyTangent(linear)  ->  linear interpolation of the animated
      solver attributes between the root frames (`interpolate_inbetween`,
      solverstandardutils.py:212-260)
  euler filter                                  ->  per-channel 360-degree unwrap of the
      rotate attributes (`euler_filter`, solverutils.py compile_euler_filter)

Scene-graph mode is always mmscenegraph (there is no Maya DAG here), so the in-between frames
are ONE per-frame-mode command (solverstandardutils.py:392-417) -- one batched launch on the
GPU -- not one command per frame.
"""
import numpy as np

from . import cmd
from .solveresult import SolveResultStrings

# _api/constant.py:355-401
ROOT_FRAME_STRATEGY_GLOBAL_VALUE = 0
ROOT_FRAME_STRATEGY_FWD_PAIR_VALUE = 1
ROOT_FRAME_STRATEGY_FWD_PAIR_AND_GLOBAL_VALUE = 2
ROOT_FRAME_STRATEGY_FWD_INCREMENT_VALUE = 3
ROOT_FRAME_STRATEGY_VALUE_LIST = [0, 1, 2, 3]
AUTO_DIFF_TYPE_FORWARD = 0
MIN_NUM_ITERATIONS = 3   # solverutils.py:40

ATTR_SOLVER_TYPE_REGULAR = "regular"
ATTR_SOLVER_TYPE_BUNDLE_TRANSFORM = "bundle_transform"
ATTR_SOLVER_TYPE_CAMERA_TRANSLATE = "camera_translate"
ATTR_SOLVER_TYPE_CAMERA_ROTATE = "camera_rotate"
ATTR_SOLVER_TYPE_CAMERA_INTRINSIC = "camera_intrinsic"
ATTR_SOLVER_TYPE_LENS_DISTORTION = "lens_distortion"


class Action:
    """_api/action.py: a function name with its arguments, run later"""

    def __init__(self, func, args=(), kwargs=None):
        self.func, self.args, self.kwargs = func, list(args), dict(kwargs or {})

    def __repr__(self):
        keys = {k: (v if not isinstance(v, (list, tuple)) or len(v) < 4 else "<%d>" % len(v)) for k, v in self.kwargs.items()}
        return "Action(%s, %r)" % (self.func, keys)


# ---- attributes ----------------------------------------------------------------------------
def _ref_of(scene, plug):
    return scene.plug(plug) if isinstance(plug, str) else int(plug)


def scene_attr_list(scene):
    """the scene's own solver attributes as `-attr` flag tuples (refs as plugs)"""
    return [(scene.attr_ref[i], scene.attr_min[i], scene.attr_max[i], scene.attr_off[i], scene.attr_scl[i])
            for i in range(len(scene.attr_ref))]


def attribute_is_animated(scene, attr):
    return _ref_of(scene, attr[0]) < 0


def get_attribute_solver_type(scene, attr):
    """compile.py:223-255 on the scene's object tables"""
    ref = _ref_of(scene, attr[0])
    cam_tfms, bnd_tfms = set(scene.cam_tfm), set(scene.bnd_tfm)
    for t, refs in enumerate(scene.tfm_attrs):
        if ref in refs:
            k = refs.index(ref)
            if t in bnd_tfms:
                return ATTR_SOLVER_TYPE_BUNDLE_TRANSFORM if k < 3 else ATTR_SOLVER_TYPE_REGULAR
            if t in cam_tfms:
                if k < 3:
                    return ATTR_SOLVER_TYPE_CAMERA_TRANSLATE
                if k < 6:
                    return ATTR_SOLVER_TYPE_CAMERA_ROTATE
            return ATTR_SOLVER_TYPE_REGULAR
    for refs in scene.cam_attrs:
        if ref in refs:
            return ATTR_SOLVER_TYPE_CAMERA_INTRINSIC
    for refs in scene.lens_attrs:
        if ref in list(refs):
            return ATTR_SOLVER_TYPE_LENS_DISTORTION
    return ATTR_SOLVER_TYPE_REGULAR


def filter_attr_list(scene, attr_list, use_camera_intrinsics=True, use_lens_distortion=True):
    """solverutils.py:153-171"""
    out = []
    for attr in attr_list:
        t = get_attribute_solver_type(scene, attr)
        if t == ATTR_SOLVER_TYPE_CAMERA_INTRINSIC and not use_camera_intrinsics:
            continue
        if t == ATTR_SOLVER_TYPE_LENS_DISTORTION and not use_lens_distortion:
            continue
        out.append(attr)
    return out


def create_attr_blocks(scene, use_attr_blocks, attr_list):
    """solverutils.py:43-127 -> [(attrs, share of the iterations)]"""
    if not use_attr_blocks:
        return [(list(attr_list), 1.0)]
    a, b, c, d = [], [], [], []
    for attr in attr_list:
        t = get_attribute_solver_type(scene, attr)
        if t in (ATTR_SOLVER_TYPE_REGULAR, ATTR_SOLVER_TYPE_BUNDLE_TRANSFORM):
            a.append(attr)
        elif t in (ATTR_SOLVER_TYPE_CAMERA_TRANSLATE, ATTR_SOLVER_TYPE_CAMERA_INTRINSIC):
            b.append(attr)
        elif t == ATTR_SOLVER_TYPE_CAMERA_ROTATE:
            c.append(attr)
        else:
            d.append(attr)
    full = list(attr_list)
    if a and (b or c) and not d:
        return [(a, 0.1), (b + c, 0.1), (a, 0.1), (b + c, 0.1), (full, 0.6)]
    if a and not (b or c) and d:
        return [(a, 0.1), (d, 0.1), (a, 0.1), (d, 0.1), (full, 0.6)]
    if a and (b or c) and d:
        return [(a, 0.2), (b + c, 0.1), (d + c, 0.1), (b + c, 0.1), (d + c, 0.1), (full, 0.4)]
    return [(full, 1.0)]


# ---- frames --------------------------------------------------------------------------------
def gen_two_frame_fwd(int_list):
    """solverstandardutils.py:43-69"""
    if len(int_list) == 1:
        return [[int_list[0], int_list[0]]]
    return [list(int_list[i:i + 2]) for i in range(len(int_list) - 1)]


def generate_increment_frame_forward(int_list):
    """rootframe.py:62-91"""
    if len(int_list) < 3:
        return [list(int_list)]
    out = []
    for i in range(1, len(int_list) - 1):
        frm = list(int_list[0:i + pow(2, i)])
        out.append(frm)
        if len(frm) >= len(int_list):
            break
    return out


# ---- SolverStep ----------------------------------------------------------------------------
class SolverStep:
    """_api/solverstep.py: one mmSolver command.  Unset values (None) leave the command's
    defaults, as in the reference (compile(): solverstep.py:549-712)."""

    _FIELDS = dict(max_iterations="iterations", solver_type="solverType", scene_graph_mode="sceneGraphMode",
                   frame_solve_mode="frameSolveMode", delta_factor="delta", auto_diff_type="autoDiffType",
                   auto_parameter_scaling="autoParamScaling", tau_factor="tauFactor", gradient_error_factor="epsilon1",
                   parameter_error_factor="epsilon2", error_factor="epsilon3", robust_loss_type="robustLossType",
                   robust_loss_scale="robustLossScale", time_eval_mode="timeEvalMode",
                   remove_unused_markers="removeUnusedMarkers", remove_unused_attributes="removeUnusedAttributes")

    def __init__(self):
        self._data = {k: None for k in self._FIELDS}
        self._data["time_eval_mode"] = 0
        self._data["scene_graph_mode"] = cmd.SCENE_GRAPH_MODE_MM_SCENE_GRAPH
        self._frames = []
        self._use_animated, self._use_static = True, True
        self._use_stiffness = self._use_smoothness = False

    def __getattr__(self, name):
        # get_<field> / set_<field> for every command option (solverstep.py:100-540)
        if name.startswith("get_") and name[4:] in self._FIELDS:
            return lambda: self._data[name[4:]]
        if name.startswith("set_") and name[4:] in self._FIELDS:
            def setter(value, key=name[4:]):
                self._data[key] = value
            return setter
        raise AttributeError(name)

    def get_frame_list(self):
        return list(self._frames)

    def set_frame_list(self, frames):
        self._frames = [int(f) for f in frames]

    def get_attributes_use_animated(self):
        return self._use_animated

    def set_attributes_use_animated(self, v):
        self._use_animated = bool(v)

    def get_attributes_use_static(self):
        return self._use_static

    def set_attributes_use_static(self, v):
        self._use_static = bool(v)

    def set_use_stiffness(self, v):
        """pass the scene's `-attrStiffness` entries to the command (solverstep.py:596-606); the rows are
        evaluated in sceneGraphMode 1 and stay zero in mode 2, as in the reference"""
        self._use_stiffness = bool(v)

    def get_use_stiffness(self):
        return self._use_stiffness

    def get_use_smoothness(self):
        return self._use_smoothness

    def set_use_smoothness(self, v):
        self._use_smoothness = bool(v)

    def set_precomputed_data(self, v):   # a Maya-query cache in the reference; nothing to cache here
        pass

    def compile(self, scene, mkr_list=None, attr_list=None, withtest=False):
        """yields (Action, validation Action or None)"""
        attr_list = scene_attr_list(scene) if attr_list is None else attr_list
        attrs = [a for a in attr_list
                 if (self._use_animated and attribute_is_animated(scene, a)) or
                    (self._use_static and not attribute_is_animated(scene, a))]
        if not attrs or not self._frames:
            return   # 'No Attributes found!' / 'No Frames found!'
        kwargs = dict(attr=[tuple(a) for a in attrs], frame=list(self._frames))
        if mkr_list is not None:
            if len(mkr_list) == 0:
                return   # 'No Markers found!'
            kwargs["marker"] = list(mkr_list)
        # solverstep.py:642-657: the command has no 'auto' scene graph (0 is passed as maya_dag);
        # maya_dag solves all frames at once, per-frame solves need the mm scene graph
        sgm, fsm = self._data["scene_graph_mode"], self._data["frame_solve_mode"]
        if sgm is not None:
            sgm = max(cmd.SCENE_GRAPH_MODE_MAYA_DAG, int(sgm))
            if sgm == cmd.SCENE_GRAPH_MODE_MAYA_DAG:
                assert fsm in (None, cmd.FRAME_SOLVE_MODE_ALL_FRAMES_AT_ONCE)
        if fsm == cmd.FRAME_SOLVE_MODE_PER_FRAME:
            assert sgm == cmd.SCENE_GRAPH_MODE_MM_SCENE_GRAPH
        for key, flag in self._FIELDS.items():
            if self._data[key] is not None:
                v = self._data[key]
                kwargs[flag] = int(v) if isinstance(v, bool) else v
        if sgm is not None:
            kwargs["sceneGraphMode"] = sgm
        # stiffness / smoothness flags of the attributes that take part (solverstep.py:596-636)
        refs = {int(a[0]) if not isinstance(a[0], str) else scene.plug(a[0]) for a in attrs}
        for use, entries, flag in ((self._use_stiffness, scene.stiffness, "attrStiffness"),
                                   (self._use_smoothness, scene.smoothness, "attrSmoothness")):
            rows = [(scene.attr_ref[e[0]], e[1], e[2], e[3]) for e in entries if scene.attr_ref[e[0]] in refs] if use else []
            if rows:
                kwargs[flag] = rows
        action = Action("mmSolver", kwargs=kwargs)
        vaction = None
        if withtest:
            vkw = dict(kwargs)
            vkw["printStatistics"] = ["inputs"]
            vaction = Action("mmSolver", kwargs=vkw)
        yield action, vaction


Solver = SolverStep


def compile_solver_step_blocks(sol, scene, mkr_list, attr_blocks, withtest):
    """solverutils.py:130-150"""
    total = sol.get_max_iterations()
    for attrs, ratio in attr_blocks:
        if len(attrs) == 0:
            continue
        sol.set_max_iterations(max(MIN_NUM_ITERATIONS, int(total * ratio)))
        for pair in sol.compile(scene, mkr_list, attrs, withtest):
            yield pair
    sol.set_max_iterations(total)


def _step(frames, iterations, use_static, frame_solve_mode, solver_type, remove_unused):
    sol = SolverStep()
    sol.set_max_iterations(iterations)
    sol.set_frame_list(frames)
    sol.set_attributes_use_animated(True)
    sol.set_attributes_use_static(use_static)
    sol.set_auto_diff_type(AUTO_DIFF_TYPE_FORWARD)
    sol.set_use_smoothness(False)
    sol.set_use_stiffness(False)
    sol.set_solver_type(solver_type)
    sol.set_scene_graph_mode(cmd.SCENE_GRAPH_MODE_MM_SCENE_GRAPH)
    sol.set_frame_solve_mode(frame_solve_mode)
    sol.set_remove_unused_markers(remove_unused)
    sol.set_remove_unused_attributes(remove_unused)
    return sol


class _SolverBase:
    def __init__(self):
        self._frames = []
        self._use_single_frame, self._single_frame = False, None
        self._use_attr_blocks = False
        self._anim_iter_num, self._lineup_iter_num = 100, 100
        self._eval_object_relationships = False
        self._solve_focal_length, self._solve_lens_distortion = True, True
        self._solver_type = None

    def get_frame_list(self):
        return list(self._frames)

    def set_frame_list(self, frames):
        self._frames = [int(f) for f in frames]

    def set_use_single_frame(self, v):
        self._use_single_frame = bool(v)

    def set_single_frame(self, f):
        self._single_frame = None if f is None else int(f)

    def set_use_attr_blocks(self, v):
        self._use_attr_blocks = bool(v)

    def set_anim_iteration_num(self, v):
        self._anim_iter_num = int(v)

    def set_lineup_iteration_num(self, v):
        self._lineup_iter_num = int(v)

    def set_eval_object_relationships(self, v):
        self._eval_object_relationships = bool(v)

    def set_solve_focal_length(self, v):
        self._solve_focal_length = bool(v)

    def set_solve_lens_distortion(self, v):
        self._solve_lens_distortion = bool(v)

    def set_solver_type(self, v):
        self._solver_type = v

    def _prepare(self, scene, attr_list):
        attr_list = scene_attr_list(scene) if attr_list is None else list(attr_list)
        return filter_attr_list(scene, attr_list, self._solve_focal_length, self._solve_lens_distortion)

    def _single_frame_actions(self, scene, mkr_list, attr_blocks, withtest):
        # solverbasic.py:463-486 / solverstandardutils.py compile_single_frame
        sol = _step([self._single_frame], self._lineup_iter_num, False, cmd.FRAME_SOLVE_MODE_ALL_FRAMES_AT_ONCE,
                    self._solver_type, self._eval_object_relationships)
        return compile_solver_step_blocks(sol, scene, mkr_list, attr_blocks, withtest)


class SolverBasic(_SolverBase):
    """solverbasic.py:397-551: animated attributes only, every frame on its own."""

    def compile(self, scene, mkr_list=None, attr_list=None, withtest=False):
        attr_list = self._prepare(scene, attr_list)
        attr_blocks = create_attr_blocks(scene, self._use_attr_blocks, attr_list)
        if self._use_single_frame:
            for pair in self._single_frame_actions(scene, mkr_list, attr_blocks, withtest):
                yield pair
            return
        sol = _step(self._frames, self._anim_iter_num, False, cmd.FRAME_SOLVE_MODE_PER_FRAME, self._solver_type,
                    self._eval_object_relationships)
        for pair in compile_solver_step_blocks(sol, scene, mkr_list, attr_blocks, withtest):
            yield pair
        rotate = [a for a in attr_list if get_attribute_solver_type(scene, a) == ATTR_SOLVER_TYPE_CAMERA_ROTATE]
        if rotate:   # use_euler_filter = True (solverbasic.py:406)
            yield Action("euler_filter", kwargs=dict(attr=rotate)), None


class SolverStandard(_SolverBase):
    """solverstandard.py / solverstandardutils.py compile_multi_frame: root frames (static +
    animated attributes together) -> in-between keys removed -> in-between frames (animated
    attributes per frame, or everything at once with global_solve)."""

    def __init__(self):
        super().__init__()
        self._root_frames = []
        self._only_root_frames = False
        self._global_solve = False
        self._root_frame_strategy = ROOT_FRAME_STRATEGY_GLOBAL_VALUE
        self._block_iter_num, self._root_iter_num = 5, 100
        self._use_euler_filter = False

    def get_root_frame_list(self):
        return list(self._root_frames)

    def set_root_frame_list(self, frames):
        self._root_frames = [int(f) for f in frames]

    def set_only_root_frames(self, v):
        self._only_root_frames = bool(v)

    def set_global_solve(self, v):
        self._global_solve = bool(v)

    def set_root_frame_strategy(self, v):
        assert v in ROOT_FRAME_STRATEGY_VALUE_LIST
        self._root_frame_strategy = v

    def set_block_iteration_num(self, v):
        self._block_iter_num = int(v)

    def set_root_iteration_num(self, v):
        self._root_iter_num = int(v)

    def set_use_euler_filter(self, v):
        self._use_euler_filter = bool(v)

    def compile(self, scene, mkr_list=None, attr_list=None, withtest=False):
        attr_list = self._prepare(scene, attr_list)
        attr_blocks = create_attr_blocks(scene, self._use_attr_blocks, attr_list)
        if self._use_single_frame:
            for pair in self._single_frame_actions(scene, mkr_list, attr_blocks, withtest):
                yield pair
            return
        roots = list(self._root_frames)
        non_root = sorted(set(self._frames) - set(roots))
        all_frames = sorted(set(self._frames) | set(roots))
        if not roots:
            return   # 'Not enough Markers given for root frames.'
        st, ru = self._solver_type, self._eval_object_relationships
        all_at_once = cmd.FRAME_SOLVE_MODE_ALL_FRAMES_AT_ONCE
        # root frames (solverstandardutils.py:616-680)
        if self._root_frame_strategy == ROOT_FRAME_STRATEGY_GLOBAL_VALUE:
            batches = [roots]
        elif self._root_frame_strategy == ROOT_FRAME_STRATEGY_FWD_PAIR_VALUE:
            batches = gen_two_frame_fwd(roots)
        elif self._root_frame_strategy == ROOT_FRAME_STRATEGY_FWD_PAIR_AND_GLOBAL_VALUE:
            batches = gen_two_frame_fwd(roots) + [roots]
        else:
            batches = generate_increment_frame_forward(roots)
        for frames in batches:
            sol = _step(frames, self._root_iter_num, True, all_at_once, st, ru)
            for pair in compile_solver_step_blocks(sol, scene, mkr_list, attr_blocks, withtest):
                yield pair
        # keys of the in-between frames go, tangents become linear (:212-260)
        animated = [a for a in attr_list if attribute_is_animated(scene, a)]
        if non_root and animated:
            yield Action("interpolate_inbetween", kwargs=dict(attr=animated, root_frames=sorted(roots),
                                                              frames=non_root)), None
        if self._only_root_frames:
            return
        if self._use_euler_filter:
            rotate = [a for a in attr_list if get_attribute_solver_type(scene, a) == ATTR_SOLVER_TYPE_CAMERA_ROTATE]
            if rotate:
                yield Action("euler_filter", kwargs=dict(attr=rotate)), None
        # in-between frames (:263-430)
        if self._global_solve:
            sol = _step(all_frames, self._anim_iter_num, True, all_at_once, st, ru)
        else:
            sol = _step(all_frames, self._anim_iter_num, False, cmd.FRAME_SOLVE_MODE_PER_FRAME, st, ru)
        for pair in compile_solver_step_blocks(sol, scene, mkr_list, attr_blocks, withtest):
            yield pair


# ---- the Maya-side Actions on the attribute store -------------------------------------------
def interpolate_inbetween(scene, attr, root_frames, frames):
    """cutKey on `frames` + linear tangents: the animated attributes follow straight lines
    between their values on the root frames (held constant outside them)."""
    index = {int(n): i for i, n in enumerate(scene.frames)}
    roots = [f for f in root_frames if f in index]
    xs = np.asarray(roots, dtype=np.float64)
    for a in attr:
        ref = _ref_of(scene, a[0])
        if ref >= 0:
            continue
        ys = np.asarray([scene.get_value(ref, index[f]) for f in roots])
        for f in frames:
            if f in index:
                scene.set_value(ref, float(np.interp(float(f), xs, ys)), index[f])


def euler_filter(scene, attr):
    """neighbouring keys never more than 180 degrees apart (per channel)"""
    for a in attr:
        ref = _ref_of(scene, a[0])
        if ref >= 0:
            continue
        vals = np.asarray([scene.get_value(ref, i) for i in range(scene.F)])
        vals = np.rad2deg(np.unwrap(np.deg2rad(vals)))
        for i, v in enumerate(vals):
            scene.set_value(ref, float(v), i)


def execute_actions(scene, actions, device=0):
    """_api/execute.py, without the Maya UI around it: run the Actions in order; the solved
    values of one command are the start values of the next.  -> [SolveResultStrings]"""
    results = []
    for action in actions:
        if action.func == "mmSolver":
            results.append(SolveResultStrings(cmd.mmSolver(scene, device=device, **action.kwargs)))
        elif action.func == "interpolate_inbetween":
            interpolate_inbetween(scene, **action.kwargs)
        elif action.func == "euler_filter":
            euler_filter(scene, **action.kwargs)
        else:
            raise ValueError("unknown action %r" % action.func)
    return results


def execute(scene, solver, mkr_list=None, attr_list=None, device=0):
    """compile `solver` against `scene` and run it"""
    actions = [a for a, _ in solver.compile(scene, mkr_list, attr_list)]
    return execute_actions(scene, actions, device=device)
