"""The reference's Maya-hosted solver tests restated as array scenes:
every number below is in the reference's test source; `create_camera` there makes a default
Maya camera (35 mm, 36 x 24 mm film back, film fit Fill, render globals 960 x 540) and the
marker group scales marker tx / ty to -0.5 .. 0.5 of the film back.

  tests/test/test_solver/test1.py:50-122
  tests/test/test_solver/test3.py:53-124
  tests/test/test_solver/test_min_max_values.py:42-220
  tests/test/test_solver/test_marker_weight.py:42-398
  tests/test/test_solver/test_marker_enabled.py:42-112

Each entry: name -> (builder, mmSolver kwargs, {plug: (expected, eps)}).
`assertApproxEqual` (tests/test/baseutils.py) defaults to eps = 0.0001.
TEST INFRASTRUCTURE ONLY.
"""
from mayamatchmovesolver_b200.scene import SceneBuilder

NONE4 = ("None", "None", "None", "None")


def _camera(sb, tx, ty, tz):
    cam_t = sb.add_transform(name="cam_tfm", tx=tx, ty=ty, tz=tz)
    sb.add_camera(cam_t, name="cam_shp", film_fit="fill", render_width=960, render_height=540)
    return cam_t


def _one_bundle_scene(cam_tz, bundle_pos, marker_xy):
    sb = SceneBuilder(frames=[1])
    _camera(sb, -1.0, 1.0, cam_tz)
    bt = sb.add_transform(name="bundle_tfm", tx=bundle_pos[0], ty=bundle_pos[1], tz=bundle_pos[2])
    sb.add_marker(sb.names["cam_shp"], sb.add_bundle(bt, name="bundle"), marker_xy[0], marker_xy[1], name="marker_tfm")
    return sb


def _one_bundle_flags(attrs, **extra):
    kw = dict(camera=[("cam_tfm", "cam_shp")], marker=[("marker_tfm", "cam_shp", "bundle")], attr=attrs, frame=[1],
              verbose=True)
    kw.update(extra)
    return kw


def test1():
    sb = _one_bundle_scene(-5.0, (5.5, 6.4, -25.0), (-0.243056042, 0.189583713))
    flags = _one_bundle_flags([("bundle_tfm.tx",) + NONE4, ("bundle_tfm.ty",) + NONE4], iterations=1000, solverType=2)
    return sb, flags, {"bundle_tfm.tx": (-6.0, 1e-4), "bundle_tfm.ty": (3.6, 1e-4)}


def test3():
    """nodal camera on one frame: rx, ry solved, delta = 1e-5"""
    sb = _one_bundle_scene(-5.0, (5.5, 6.4, -25.0), (-0.243056042, 0.189583713))
    flags = _one_bundle_flags([("cam_tfm.rx",) + NONE4, ("cam_tfm.ry",) + NONE4], delta=0.00001, solverType=2)
    return sb, flags, {"cam_tfm.rx": (7.44014, 1e-3), "cam_tfm.ry": (-32.3891, 1e-3)}


def _min_max(mn, mx, expect_tx):
    sb = _one_bundle_scene(10.0, (0.0, 0.0, 0.0), (-0.486112083, 0.189583713))
    flags = _one_bundle_flags([("bundle_tfm.tx", mn, mx, "None", "None"), ("bundle_tfm.ty",) + NONE4])
    return sb, flags, {"bundle_tfm.tx": (expect_tx, 1e-4), "bundle_tfm.ty": (2.3, 1e-4)}


def min_max_both():
    return _min_max("-5.0", "5.0", -5.0)


def min_max_lower_only():
    return _min_max("-5.0", "None", -5.0)


def min_max_upper_only():
    return _min_max("None", "5.0", -6.0)


def _two_markers(w1, w2, cam=(-1.0, 1.0, -5.0), mx=0.243056042, enable2=True):
    sb = SceneBuilder(frames=[1])
    _camera(sb, *cam)
    grp = sb.add_transform(name="group1", tz=-10.0)
    b1 = sb.add_bundle(sb.add_transform(parent=grp, name="bundle_01_tfm"), name="bundle_01")
    b2 = sb.add_bundle(sb.add_transform(parent=grp, name="bundle_02_tfm"), name="bundle_02")
    cam_shp = sb.names["cam_shp"]
    sb.add_marker(cam_shp, b1, -mx, 0.189583713, weight=w1, name="marker_01")
    sb.add_marker(cam_shp, b2, mx, 0.189583713, weight=w2, enable=enable2, name="marker_02")
    flags = dict(camera=[("cam_tfm", "cam_shp")],
                 marker=[("marker_01", "cam_shp", "bundle_01"), ("marker_02", "cam_shp", "bundle_02")],
                 attr=[("group1.tx",) + NONE4, ("group1.ty",) + NONE4], frame=[1], verbose=True)
    return sb, flags


def weight_high():
    sb, flags = _two_markers(100.0, 1.0)
    return sb, flags, {"group1.tx": (-2.2252424, 1e-4), "group1.ty": (1.65, 1e-4)}


def weight_low():
    sb, flags = _two_markers(1.0, 0.01)
    return sb, flags, {"group1.tx": (-2.2252424, 1e-4), "group1.ty": (1.65, 1e-4)}


def weight_ratio():
    sb, flags = _two_markers(100.0, 50.0, cam=(0.0, 0.0, 0.0), mx=0.097222417)
    return sb, flags, {"group1.tx": (-0.333333333333, 1e-4), "group1.ty": (1.3, 1e-4)}


def weight_same():
    sb, flags = _two_markers(0.5, 0.5)
    return sb, flags, {"group1.tx": (-1.00000134, 1e-4), "group1.ty": (1.65000055, 1e-4)}


def weight_none():
    sb, flags = _two_markers(1.0, 0.0)
    return sb, flags, {"group1.tx": (-2.25, 1e-3), "group1.ty": (1.65, 1e-3)}


def marker_enabled():
    sb, flags = _two_markers(1.0, 1.0, enable2=False)
    return sb, flags, {"group1.tx": (-2.24999755, 1e-4), "group1.ty": (1.65000644, 1e-4)}


KATS = {
    "test1": test1, "test3": test3,
    "min_max_both": min_max_both, "min_max_lower_only": min_max_lower_only, "min_max_upper_only": min_max_upper_only,
    "weight_high": weight_high, "weight_low": weight_low, "weight_ratio": weight_ratio, "weight_same": weight_same,
    "weight_none": weight_none, "marker_enabled": marker_enabled,
}


def apply_bounds(sb, info, near=True, lower_only=True, scale_with_bounds=True):
    """Box constraints / offset / scale on the solver attributes of a synth.make_scene scene
    (`-attr node.attr min max offset scale`, arg_flags_solve_object.cpp:288-318): every
    branch of parameterBoundFromInternalToExternal (adjust_base.cpp:194-258) and, with
    `near`, start values within delta of a bound so that calculateParameterDelta
    (adjust_solveFunc.cpp:148-180) flips the sign of the step."""
    import numpy as np
    sb._arrays_cache = None
    cam = info["cam_tfm"]
    by_ref = {ref: i for i, ref in enumerate(sb.attr_ref)}

    def vals(ref):
        return np.array([sb.get_value(ref, f) for f in range(sb.F)]) if ref < 0 else np.array([sb.get_value(ref)])

    def setb(ref, mn=None, mx=None, off=None, scl=None):
        i = by_ref.get(ref)
        if i is None:
            return
        if mn is not None: sb.attr_min[i] = float(mn)
        if mx is not None: sb.attr_max[i] = float(mx)
        if off is not None: sb.attr_off[i] = float(off)
        if scl is not None: sb.attr_scl[i] = float(scl)

    v = vals(cam["tx"]); setb(cam["tx"], mn=v.min() - 3.0, mx=v.max() + 3.0)                 # both bounds
    v = vals(cam["ty"]); setb(cam["ty"], mx=v.max() + 2.0)                                   # upper only
    if lower_only:  # the external value is pinned at min (formula as written): a zero Jacobian column
        v = vals(cam["tz"]); setb(cam["tz"], mn=v.min() - 2.0)
    setb(cam["rx"], off=0.5, scl=2.0)                                                        # offset / scale, no bounds
    # bounds AND offset / scale: the reference's two transforms are not inverses of each other
    # then (external e comes back as e / scale - offset), so solves keep them apart
    sws = scale_with_bounds
    v = vals(cam["ry"]); setb(cam["ry"], mn=v.min() - 10.0, mx=v.max() + 10.0, off=-1.0 if sws else None, scl=0.5 if sws else None)
    for b, t in enumerate(info["bnd_tfms"]):
        if b % 3 == 0:
            v = vals(t["tx"]); setb(t["tx"], mn=v[0] - 1.0, mx=v[0] + 1.0)
        elif b % 3 == 1 and near:
            # the INTERNAL value asin(..) = 0.96 is compared with the EXTERNAL bounds (as the
            # reference does): wherever it exceeds max - delta the forward step flips to -delta
            v = vals(t["ty"]); setb(t["ty"], mn=v[0] - 5.0, mx=v[0] + 0.5)
        else:
            v = vals(t["tz"]); setb(t["tz"], mn=v[0] - 4.0, mx=v[0] + 4.0, scl=1.5 if sws else None)
    return sb


def cfg1_operahouse(noise_seed=1):
    """BASELINE.json configs[0]: one static camera + 20 bundles on ONE frame, markers read from
    the reference's own marker file (tests/golden/cfg1_loadmarker_frame41.rz2 = frame 41 of
    tests/data/match_mover/loadmarker.rz2), camera pose / intrinsics and bundle start positions of
    tests/test/test_api/test_solve_operahouse.py:62-100,166-214 (+ N(0, 0.05) noise, seed 1).
    n > m is refused by the reference (adjust_base.cpp:864), so tx, ty per bundle are solved
    (n = 40, m = 40), as in test1.py.  -> (scene, mmSolver flags, [bundle transform nodes])"""
    import json
    import os
    import numpy as np
    from mayamatchmovesolver_b200 import loadmarker
    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    data = json.load(open(os.path.join(here, "cfg1_operahouse.json")))
    _, tracks = loadmarker.read(os.path.join(here, "cfg1_loadmarker_frame41.rz2"))
    assert [t.name for t in tracks] == data["tracks"] and len(tracks) == 20
    frame = data["frame"]
    sb = SceneBuilder(frames=[frame])
    tx, ty, tz, rx, ry, rz = data["camera_pose"]
    cam_tfm = sb.add_transform(name="cam_tfm", tx=tx, ty=ty, tz=tz, rx=rx, ry=ry, rz=rz, rotate_order=data["rotate_order"])
    w, h = data["image_size"]
    cam = sb.add_camera(cam_tfm, name="cam_shp", film_fit="horizontal", render_width=w, render_height=h,
                        sensor_width_mm=data["film_back_mm"][0], sensor_height_mm=data["film_back_mm"][1],
                        focal_length_mm=data["focal_length_mm"])
    nodes = loadmarker.add_tracks_to_scene(sb, cam, tracks)
    rng = np.random.default_rng(noise_seed)
    for t, (_, _, tfm) in zip(tracks, nodes):
        pos = np.asarray(data["bundle_positions"][t.name]) + rng.normal(0, 0.05, 3)
        for k, v in zip(("tx", "ty", "tz"), pos):
            sb.set_value(tfm[k], v)
    attrs = []
    for t in tracks:
        attrs += [(t.name + "_BND_tfm.tx",) + NONE4, (t.name + "_BND_tfm.ty",) + NONE4]
    flags = dict(camera=[("cam_tfm", "cam_shp")], marker=[(t.name + "_MKR", "cam_shp", t.name + "_BND") for t in tracks],
                 attr=attrs, frame=[frame], iterations=100, solverType=2)
    return sb, flags, [n[2] for n in nodes]


def smoothness():
    """tests/test/test_solver/test_smoothness.py:43-140: animated nodal camera over 10 frames, one
    marker, rx / ry solved, `-attrSmoothness cam.rx cam.weightValue cam.varianceValue cam.smoothValue`
    with weight 1, variance 0.5, value 10; iterations = 10.  Runs in the command's default
    sceneGraphMode (maya_dag).  The reference only asserts `success=1`."""
    import numpy as np
    frames = np.arange(1, 11)
    t = (frames - 1) / 9.0
    sb = SceneBuilder(frames=frames)
    cam_t = sb.add_transform(name="cam_tfm", tx=-1.0, ty=1.0, tz=-5.0, rx=-2.0 + 4.0 * t, ry=-2.5 + 5.0 * t)
    sb.add_camera(cam_t, name="cam_shp", film_fit="fill", render_width=960, render_height=540)
    sb.add_custom_attrs(cam_t, weightValue=1.0, varianceValue=0.5, smoothValue=10.0)
    bt = sb.add_transform(name="bundle_tfm", tx=-2.5, ty=2.4, tz=-15.0)
    # marker keyed in Maya units under the camera at tz = -10: tx -2.5 -> 3.0, ty 1.5 -> 1.3; as a
    # fraction of the film back (35 mm, 36 x 24 mm): x / (10 * 36 / 35), y / (10 * 24 / 35)
    mx = (-2.5 + 5.5 * t) / (10.0 * 36.0 / 35.0)
    my = (1.5 - 0.2 * t) / (10.0 * 24.0 / 35.0)
    sb.add_marker(sb.names["cam_shp"], sb.add_bundle(bt, name="bundle"), mx, my, name="marker_tfm")
    flags = dict(camera=[("cam_tfm", "cam_shp")], marker=[("marker_tfm", "cam_shp", "bundle")],
                 attr=[("cam_tfm.rx",) + NONE4, ("cam_tfm.ry",) + NONE4], frame=[int(f) for f in frames],
                 attrSmoothness=[("cam_tfm.rx", "cam_tfm.weightValue", "cam_tfm.varianceValue", "cam_tfm.smoothValue")],
                 iterations=10, verbose=True)
    return sb, flags
