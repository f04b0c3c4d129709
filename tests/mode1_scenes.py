"""Scenes that exercise sceneGraphMode = 1 (the Maya-DAG-convention arithmetic of the residual,
maya_camera.cpp:75-414,863-894 + adjust_measureErrors.cpp:116-387): film offsets (which only
Maya's row-vector layout lets reach x / y), every film fit (Horizontal scales Y there, X in
camera.rs), a bundle behind the camera (x 1e6) and stiffness / smoothness rows.  TEST ONLY."""
from mayamatchmovesolver_b200 import synth

FITS = {"fill": 0, "horizontal": 1, "vertical": 2, "overscan": 3}


def eval_scene(fit="horizontal", behind=True, F=5, B=7, seed=21):
    sb, info = synth.make_scene(F=F, B=B, seed=seed, fixed_bundles=2)
    cam = info["cam"]
    sb.set_value(cam["lens_offset_x_mm"], 1.3)
    sb.set_value(cam["lens_offset_y_mm"], -0.7)
    sb.cam_fit[0] = FITS[fit]
    if behind:
        sb.set_value(info["bnd_tfms"][3]["tz"], 40.0)
    sb.current_frame = 2
    # solver attrs: camera tx ty tz rx ry rz (0..5), then bundle attrs
    sb.add_stiffness(1, 0.7, 0.5, float(sb.get_value(info["cam_tfm"]["ty"], 2)) + 0.3)
    sb.add_stiffness(2, 0.0, 1.0, 0.0)                       # weight 0: not counted
    sb.add_stiffness(4, 1.0, 2.0, 1.0, smoothness=True)
    sb._arrays_cache = None
    return sb


def solve_scene():
    """a well-conditioned solve with mild stiffness on camera tx and smoothness on ry at frame 3"""
    sb, info = synth.make_scene(F=10, B=14, seed=23, fixed_bundles=3, period=10)
    sb.set_value(info["cam"]["lens_offset_x_mm"], 0.4)
    sb.current_frame = 3
    sb.add_stiffness(0, 2.0, 3.0, float(sb.get_value(info["cam_tfm"]["tx"], 3)) + 0.2)
    sb.add_stiffness(4, 1.5, 4.0, float(sb.get_value(info["cam_tfm"]["ry"], 3)) - 0.5, smoothness=True)
    sb._arrays_cache = None
    return sb
