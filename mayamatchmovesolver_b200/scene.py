"""Flat scene description (mmb_scene_desc) built from Python.

Mirrors what the reference hands to its solve callback in mmSceneGraph mode:
the FlatScene + AttrDataBlock built by construct_scene_graph
(src/mmSolver/mayahelper/maya_scene_graph.cpp) plus the `-attr` list of the
mmSolver command (arg_flags_solve_object.cpp:288-318).  Without Maya there are
no node names, so attributes are named by attribute-store references.
"""
import ctypes as C

import numpy as np

from . import _capi

FLT_MAX = float(np.finfo(np.float32).max)
ROTATE_ORDERS = {"xyz": 0, "yzx": 1, "zxy": 2, "xzy": 3, "yxz": 4, "zyx": 5}
FILM_FITS = {"fill": 0, "horizontal": 1, "vertical": 2, "overscan": 3}
LENS_MODELS = {"passthrough": 0, "3de_classic": 1, "3de_radial_std_deg4": 2,
               "3de_anamorphic_std_deg4": 3, "3de_anamorphic_std_deg4_rescaled": 4}
TFM_ATTR_NAMES = ("tx", "ty", "tz", "rx", "ry", "rz", "sx", "sy", "sz")
CAM_ATTR_NAMES = ("sensor_width_mm", "sensor_height_mm", "focal_length_mm", "lens_offset_x_mm",
                  "lens_offset_y_mm", "near_clip", "far_clip", "camera_scale")
LENS_COEFF_NAMES = {
    1: ("distortion", "anamorphic_squeeze", "curvature_x", "curvature_y", "quartic_distortion"),
    2: ("degree2_distortion", "degree2_u", "degree2_v", "degree4_distortion", "degree4_u", "degree4_v",
        "cylindric_direction", "cylindric_bending"),
    3: ("degree2_cx02", "degree2_cy02", "degree2_cx22", "degree2_cy22", "degree4_cx04", "degree4_cy04",
        "degree4_cx24", "degree4_cy24", "degree4_cx44", "degree4_cy44", "lens_rotation", "squeeze_x",
        "squeeze_y"),
    4: ("degree2_cx02", "degree2_cy02", "degree2_cx22", "degree2_cy22", "degree4_cx04", "degree4_cy04",
        "degree4_cx24", "degree4_cy24", "degree4_cx44", "degree4_cy44", "lens_rotation", "squeeze_x",
        "squeeze_y", "rescale"),
}


class Node(dict):
    """index + {attr name: attribute-store ref}; `name` is the Maya node name the
    mmSolver command flags would use (optional)"""

    def __init__(self, index, refs, kind="", name=None):
        super().__init__(refs)
        self.index = index
        self.kind = kind
        self.name = name


class SceneBuilder:
    def __init__(self, frames):
        self.frames = np.asarray(frames, dtype=np.int32)
        self.F = len(self.frames)
        self._static = []
        self._anim = []
        self.tfm_parent, self.tfm_roo, self.tfm_attrs = [], [], []
        self.cam_tfm, self.cam_attrs, self.cam_fit, self.cam_rw, self.cam_rh, self.cam_lens = [], [], [], [], [], []
        self.bnd_tfm = []
        self.mkr_cam, self.mkr_bnd, self.mkr_x, self.mkr_y, self.mkr_w, self.mkr_en = [], [], [], [], [], []
        self.lens_model, self.lens_attrs, self.lens_cam = [], [], []
        self.attr_ref, self.attr_min, self.attr_max, self.attr_off, self.attr_scl = [], [], [], [], []
        # -attrStiffness / -attrSmoothness entries: [attr index, weight ref, variance ref, value ref]
        self.stiffness, self.smoothness = [], []
        self.current_frame = 0   # index into frames of "Maya's current time"
        self.names = {}  # node name -> Node (what the command's -camera / -marker / -attr flags name)
        self._arrays_cache = None  # flattened arrays of the current state (dropped by every mutator)

    def _register(self, node):
        if node.name is not None:
            self.names[node.name] = node
        return node

    def plug(self, plug_name):
        """'node.attr' -> attribute-store ref (the MPlug of an `-attr` flag)"""
        node_name, _, attr = plug_name.rpartition(".")
        node = self.names[node_name]
        return node[attr]

    # ---- attribute store (datablock.rs) --------------------------------
    def value(self, v):
        """scalar -> static slot, length-F sequence -> animated slot; returns the ref"""
        self._arrays_cache = None
        if np.ndim(v) == 0:
            self._static.append(float(v))
            return len(self._static) - 1
        arr = np.asarray(v, dtype=np.float64)
        assert arr.shape == (self.F,), "animated value needs one entry per frame"
        self._anim.append(arr.copy())
        return -(len(self._anim) - 1) - 1

    def get_value(self, ref, frame_index=0):
        return self._static[ref] if ref >= 0 else float(self._anim[-ref - 1][frame_index])

    def set_value(self, ref, v, frame_index=None):
        self._arrays_cache = None
        if ref >= 0:
            self._static[ref] = float(v)
        elif frame_index is None:
            self._anim[-ref - 1][:] = v
        else:
            self._anim[-ref - 1][frame_index] = v

    # ---- nodes ---------------------------------------------------------
    def add_transform(self, parent=None, rotate_order="xyz", name=None, **vals):
        defaults = dict(tx=0.0, ty=0.0, tz=0.0, rx=0.0, ry=0.0, rz=0.0, sx=1.0, sy=1.0, sz=1.0)
        defaults.update(vals)
        refs = {k: self.value(defaults[k]) for k in TFM_ATTR_NAMES}
        self.tfm_parent.append(-1 if parent is None else int(getattr(parent, "index", parent)))
        self.tfm_roo.append(ROTATE_ORDERS[rotate_order] if isinstance(rotate_order, str) else int(rotate_order))
        self.tfm_attrs.append([refs[k] for k in TFM_ATTR_NAMES])
        return self._register(Node(len(self.tfm_parent) - 1, refs, "transform", name))

    def add_camera(self, tfm, film_fit="horizontal", render_width=2048, render_height=1556, lens=None, name=None,
                   **vals):
        # defaults of a Maya camera in mm (dag.rs:120-172 reads these 8)
        defaults = dict(sensor_width_mm=36.0, sensor_height_mm=24.0, focal_length_mm=35.0,
                        lens_offset_x_mm=0.0, lens_offset_y_mm=0.0, near_clip=0.1, far_clip=10000.0,
                        camera_scale=1.0)
        defaults.update(vals)
        refs = {k: self.value(defaults[k]) for k in CAM_ATTR_NAMES}
        self.cam_tfm.append(int(getattr(tfm, "index", tfm)))
        self.cam_attrs.append([refs[k] for k in CAM_ATTR_NAMES])
        self.cam_fit.append(FILM_FITS[film_fit] if isinstance(film_fit, str) else int(film_fit))
        self.cam_rw.append(int(render_width))
        self.cam_rh.append(int(render_height))
        self.cam_lens.append(-1 if lens is None else int(getattr(lens, "index", lens)))
        return self._register(Node(len(self.cam_tfm) - 1, refs, "camera", name))

    def add_bundle(self, tfm, name=None):
        self._arrays_cache = None
        self.bnd_tfm.append(int(getattr(tfm, "index", tfm)))
        return self._register(Node(len(self.bnd_tfm) - 1, {}, "bundle", name))

    def add_marker(self, cam, bnd, x, y, weight=1.0, enable=True, name=None):
        F = self.F
        self._arrays_cache = None
        self.mkr_cam.append(int(getattr(cam, "index", cam)))
        self.mkr_bnd.append(int(getattr(bnd, "index", bnd)))
        self.mkr_x.append(np.broadcast_to(np.asarray(x, dtype=np.float64), (F,)).copy())
        self.mkr_y.append(np.broadcast_to(np.asarray(y, dtype=np.float64), (F,)).copy())
        self.mkr_w.append(np.broadcast_to(np.asarray(weight, dtype=np.float64), (F,)).copy())
        self.mkr_en.append(np.broadcast_to(np.asarray(enable, dtype=np.uint8), (F,)).copy())
        return self._register(Node(len(self.mkr_cam) - 1, {}, "marker", name))

    def add_lens(self, model, camera_params=(3.0, 3.6, 2.4, 1.0, 0.0, 0.0), name=None, **coeffs):
        mid = LENS_MODELS[model] if isinstance(model, str) else int(model)
        names = LENS_COEFF_NAMES.get(mid, ())
        defaults = {k: 0.0 for k in names}
        if mid == 1:
            defaults["anamorphic_squeeze"] = 1.0
        if mid in (3, 4):
            defaults["squeeze_x"] = 1.0
            defaults["squeeze_y"] = 1.0
        if mid == 4:
            defaults["rescale"] = 1.0
        for k in coeffs:
            assert k in defaults, "unknown lens coefficient %s" % k
        defaults.update(coeffs)
        refs = {k: self.value(defaults[k]) for k in names}
        row = [refs[k] for k in names] + [_capi.MMB_ATTR_NONE] * (_capi.MMB_LENS_MAX_COEFFS - len(names))
        self.lens_model.append(mid)
        self.lens_attrs.append(row)
        self.lens_cam.append([float(v) for v in camera_params])
        return self._register(Node(len(self.lens_model) - 1, refs, "lens", name))

    def add_attr(self, ref, min=None, max=None, offset=None, scale=None):
        """One `-attr node.attr min max offset scale` flag; None keeps the
        reference default (maya_attr.cpp:88-91: -/+FLT_MAX, 0, 1)."""
        self._arrays_cache = None
        self.attr_ref.append(int(ref))
        self.attr_min.append(-FLT_MAX if min is None else float(min))
        self.attr_max.append(FLT_MAX if max is None else float(max))
        self.attr_off.append(0.0 if offset is None else float(offset))
        self.attr_scl.append(1.0 if scale is None else float(scale))
        return len(self.attr_ref) - 1

    def add_custom_attrs(self, node, **vals):
        """dynamic attributes on a named node (`maya.cmds.addAttr`): values for `plug('node.attr')`"""
        for k, v in vals.items():
            node[k] = self.value(v)
        return node

    def add_stiffness(self, attr_index, weight, variance, value, smoothness=False):
        """One `-attrStiffness` (or `-attrSmoothness`) flag: solver attribute index + the three
        attribute-store refs (scalars are stored as new static values)."""
        self._arrays_cache = None
        refs = [v if isinstance(v, (int, np.integer)) and not isinstance(v, bool) else self.value(v)
                for v in (weight, variance, value)]
        (self.smoothness if smoothness else self.stiffness).append([int(attr_index)] + [int(r) for r in refs])

    def keep_markers(self, indices):
        """Drop every marker that is not listed (the command's `-marker` flags name the markers
        of the solve; the others are not part of markerList at all)."""
        self._arrays_cache = None
        keep = list(dict.fromkeys(int(i) for i in indices))   # the flag order, as markerList has it
        new_index = {old: new for new, old in enumerate(keep)}
        for name in ("mkr_cam", "mkr_bnd", "mkr_x", "mkr_y", "mkr_w", "mkr_en"):
            lst = getattr(self, name)
            setattr(self, name, [lst[i] for i in keep])
        for key in [k for k, n in self.names.items() if n.kind == "marker"]:
            node = self.names[key]
            if node.index in new_index:
                node.index = new_index[node.index]
            else:
                del self.names[key]

    def slice_frames(self, frame_indices):
        """A copy of the scene restricted to some frames (what solve_v1 hands to
        solveFrames for one entry of its per-frame loop, adjust_base.cpp:1447)."""
        import copy
        idx = np.asarray(frame_indices, dtype=np.int64)
        o = copy.deepcopy(self)
        o._arrays_cache = None
        o.frames = self.frames[idx].copy()
        o.F = len(idx)
        o._anim = [a[idx].copy() for a in self._anim]
        for name in ("mkr_x", "mkr_y", "mkr_w", "mkr_en"):
            setattr(o, name, [a[idx].copy() for a in getattr(self, name)])
        return o

    # ---- flatten -------------------------------------------------------
    def arrays(self):
        """flat numpy arrays of the scene (cached until the scene is changed through its
        methods; code that edits the lists directly must set `_arrays_cache = None`)"""
        if self._arrays_cache is not None:
            return self._arrays_cache
        F = self.F
        i32 = lambda v, shape=None: np.ascontiguousarray(np.asarray(v, dtype=np.int32).reshape(shape) if shape else np.asarray(v, dtype=np.int32))
        f64 = lambda v: np.ascontiguousarray(np.asarray(v, dtype=np.float64))
        M = len(self.mkr_cam)
        a = dict(
            frames=i32(self.frames),
            static_values=f64(self._static),
            anim_values=f64(self._anim).reshape(-1) if self._anim else np.zeros(0),
            tfm_parent=i32(self.tfm_parent), tfm_rotate_order=i32(self.tfm_roo),
            tfm_attrs=i32(self.tfm_attrs).reshape(-1),
            cam_tfm=i32(self.cam_tfm), cam_attrs=i32(self.cam_attrs).reshape(-1),
            cam_film_fit=i32(self.cam_fit), cam_render_width=i32(self.cam_rw),
            cam_render_height=i32(self.cam_rh), cam_lens=i32(self.cam_lens),
            bnd_tfm=i32(self.bnd_tfm), mkr_cam=i32(self.mkr_cam), mkr_bnd=i32(self.mkr_bnd),
            mkr_x=f64(self.mkr_x).reshape(-1) if M else np.zeros(0),
            mkr_y=f64(self.mkr_y).reshape(-1) if M else np.zeros(0),
            mkr_weight=f64(self.mkr_w).reshape(-1) if M else np.zeros(0),
            mkr_enable=np.ascontiguousarray(np.asarray(self.mkr_en, dtype=np.uint8).reshape(-1)) if M else np.zeros(0, np.uint8),
            lens_model=i32(self.lens_model), lens_attrs=i32(self.lens_attrs).reshape(-1),
            lens_camera_params=f64(self.lens_cam).reshape(-1),
            attr_ref=i32(self.attr_ref), attr_min=f64(self.attr_min), attr_max=f64(self.attr_max),
            attr_offset=f64(self.attr_off), attr_scale=f64(self.attr_scl),
        )
        for key, rows in (("stiff", self.stiffness), ("smooth", self.smoothness)):
            tab = np.asarray(rows, dtype=np.int32).reshape(-1, 4)
            for j, name in enumerate(("attr", "weight", "variance", "value")):
                a["%s_%s" % (key, name)] = np.ascontiguousarray(tab[:, j])
        self._arrays_cache = a
        return a

    def desc(self):
        """-> (SceneDesc, keepalive dict of numpy arrays)"""
        a = self.arrays()
        d = _capi.SceneDesc()
        d.abi_version = _capi.MMB_ABI_VERSION
        d.num_frames = self.F
        d.num_static = len(self._static)
        d.num_anim = len(self._anim)
        d.num_transforms = len(self.tfm_parent)
        d.num_cameras = len(self.cam_tfm)
        d.num_bundles = len(self.bnd_tfm)
        d.num_markers = len(self.mkr_cam)
        d.num_lenses = len(self.lens_model)
        d.num_attrs = len(self.attr_ref)
        d.num_stiffness = len(self.stiffness)
        d.num_smoothness = len(self.smoothness)
        d.current_frame = int(self.current_frame)
        for name, ctype in _capi.SceneDesc._fields_:
            if name in a:
                setattr(d, name, a[name].ctypes.data_as(ctype))
        return d, a


def ptr(arr, ctype):
    return arr.ctypes.data_as(ctype) if arr is not None else None
