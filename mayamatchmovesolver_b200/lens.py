"""Bulk lens (un)distortion of point buffers through the C-ABI (mmb_lens_apply): the
reference's LensModel.applyModelUndistort / applyModelDistort loops
(lib/cppbind/mmlens/src/lens_model_3de_*.cpp) on the GPU."""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import c_f64p, c_i32p, check
from .scene import LENS_COEFF_NAMES, LENS_MODELS

UNDISTORT, DISTORT = 0, 1
DEFAULT_CAMERA = (3.0, 3.6, 2.4, 1.0, 0.0, 0.0)   # focal_cm fbw_cm fbh_cm pixel_aspect lco_x lco_y (lens_model.h:42)


def _coeff_array(model, coeffs):
    mid = LENS_MODELS[model] if isinstance(model, str) else int(model)
    arr = np.zeros(_capi.MMB_LENS_MAX_COEFFS)
    if isinstance(coeffs, dict):
        names = LENS_COEFF_NAMES.get(mid, ())
        defaults = {k: 0.0 for k in names}
        if mid == 1:
            defaults["anamorphic_squeeze"] = 1.0
        if mid in (3, 4):
            defaults["squeeze_x"] = defaults["squeeze_y"] = 1.0
        if mid == 4:
            defaults["rescale"] = 1.0
        for k in coeffs:
            if k not in defaults:
                raise KeyError("unknown lens coefficient %r" % k)
        defaults.update(coeffs)
        arr[:len(names)] = [defaults[k] for k in names]
    elif coeffs is not None:
        v = np.asarray(coeffs, dtype=np.float64)
        arr[:len(v)] = v
    return mid, arr


def apply(model, coeffs, points, direction=DISTORT, camera_params=DEFAULT_CAMERA, device=0):
    """points [..., 2] in -0.5..0.5 units -> same shape"""
    lib = _capi.load()
    mid, arr = _coeff_array(model, coeffs)
    pts = np.ascontiguousarray(points, dtype=np.float64)
    assert pts.shape[-1] == 2
    out = np.empty_like(pts)
    cam = np.ascontiguousarray(camera_params, dtype=np.float64)
    check(lib.mmb_lens_apply(int(device), mid, arr.ctypes.data_as(c_f64p), cam.ctypes.data_as(c_f64p), int(direction),
                             pts.size // 2, pts.ctypes.data_as(c_f64p), out.ctypes.data_as(c_f64p)), "mmb_lens_apply")
    return out


NO_ALLOC_MAX_POINTS = 262144   # MMB_LENS_NO_ALLOC_MAX_POINTS


def apply_layered(layers, points, direction=DISTORT, device=0, no_alloc=False):
    """layers: [(model, coeffs, camera_params), ...] innermost first (a lens-node chain)"""
    lib = _capi.load()
    models = np.zeros(len(layers), np.int32)
    co = np.zeros((len(layers), _capi.MMB_LENS_MAX_COEFFS))
    cams = np.zeros((len(layers), 6))
    for k, (model, coeffs, cam) in enumerate(layers):
        models[k], co[k] = _coeff_array(model, coeffs)
        cams[k] = cam
    pts = np.ascontiguousarray(points, dtype=np.float64)
    out = np.empty_like(pts)
    # no_alloc: the entry point that uses the library's static device scratch (at most
    # NO_ALLOC_MAX_POINTS points, calls must not overlap on one GPU)
    fn = lib.mmb_lens_apply_layered_no_alloc if no_alloc else lib.mmb_lens_apply_layered
    check(fn(int(device), len(layers), models.ctypes.data_as(c_i32p), co.ctypes.data_as(c_f64p),
             cams.ctypes.data_as(c_f64p), int(direction), pts.size // 2,
             pts.ctypes.data_as(c_f64p), out.ctypes.data_as(c_f64p)),
          "mmb_lens_apply_layered_no_alloc" if no_alloc else "mmb_lens_apply_layered")
    return out


def bench(model, coeffs, direction, n_points, reps=10, camera_params=DEFAULT_CAMERA, device=0):
    """-> ms per pass over a device-resident grid of n_points"""
    lib = _capi.load()
    mid, arr = _coeff_array(model, coeffs)
    cam = np.ascontiguousarray(camera_params, dtype=np.float64)
    ms = C.c_double()
    check(lib.mmb_bench_lens(int(device), mid, arr.ctypes.data_as(c_f64p), cam.ctypes.data_as(c_f64p), int(direction),
                             int(n_points), int(reps), C.byref(ms)), "mmb_bench_lens")
    return ms.value
