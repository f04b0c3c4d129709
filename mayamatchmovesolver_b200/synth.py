"""Deterministic synthetic scenes of the BASELINE.json config shapes
(SURVEY.md section 8d).  numpy only; used by bench.py and tests/.

The scene mathematics used to *generate* ground-truth markers is a small numpy
restatement of SURVEY appendix B (projection of the truth bundles through the
truth camera); it is data generation, not the solve path.
"""
import numpy as np

from .scene import SceneBuilder

DEG = np.pi / 180.0


def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    X = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Y = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Z = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Z @ Y @ X  # rotate order XYZ (transform.rs:433)


def project_truth(cam_t, cam_r_deg, pts, focal_mm, sensor_w_mm, sensor_h_mm, render_w, render_h):
    """Horizontal film fit (camera.rs:183-242).  Returns marker x,y in the
    -0.5..0.5 units the markers are stored in (before flat.rs film-fit scale)."""
    R = _rot(*(np.asarray(cam_r_deg) * DEG))
    pc = (pts - np.asarray(cam_t)) @ R  # inv(R) = R^T applied to column vectors
    size_x = 0.1 / focal_mm * sensor_w_mm
    image_aspect = render_w / render_h
    size_y = size_x / image_aspect
    p00 = 1.0 / (size_x * 0.5) * 0.1
    p11 = 1.0 / (size_y * 0.5) * 0.1
    x = p00 * pc[:, 0] / (-pc[:, 2]) * 0.5
    y = p11 * pc[:, 1] / (-pc[:, 2]) * 0.5
    # measureErrors compares against marker * film-fit scale; invert it here
    film_aspect = sensor_w_mm / sensor_h_mm
    y = y / (image_aspect / film_aspect)
    return x, y


def camera_path(F, rng, amp_t=2.0, amp_r=5.0, period=120.0, frame_offset=0):
    f = np.arange(F) + frame_offset
    ph = rng.uniform(0, 2 * np.pi, 5)
    t = np.stack([amp_t * np.sin(2 * np.pi * f / period + ph[0]),
                  amp_t * np.sin(2 * np.pi * f / period + ph[1]),
                  amp_t * np.sin(2 * np.pi * f / period + ph[2])], 1)
    r = np.stack([amp_r * np.sin(2 * np.pi * f / period + ph[3]),
                  amp_r * np.sin(2 * np.pi * f / period + ph[4]),
                  np.zeros(F)], 1)
    return t, r


def make_scene(F, B, seed, solve_camera=True, solve_bundles=True, noise_px=0.5, init_noise_t=0.1,
               init_noise_r=0.5, init_noise_b=0.05, fixed_bundles=0, lens=None, solve_focal=False,
               solve_lens=(), focal_truth=35.0, focal_init=None, drop_fraction=0.0, weights=False, period=120.0,
               in_frame=False, lens_model="3de_radial_std_deg4", frame_offset=0):
    """One animated camera, B static bundles, full (or thinned) visibility.

    cfg2: make_scene(240, 50, 2, solve_bundles=False)
    cfg3: make_scene(1000, 500, 3)
    cfg4: make_scene(1000, 500, 4, lens={...}, solve_focal=True, solve_lens=(...), in_frame=True)
    `fixed_bundles` bundles are left out of the attr list (removes the gauge
    freedom so solved values are comparable between solvers).
    `in_frame` keeps every bundle inside the image on every frame (a plate only has
    markers inside the frame; outside it the degree-4 lens polynomial is not
    invertible and the fixed-point inverse of LDPK diverges) -- used by cfg4.
    `lens_model` other than the radial one takes its coefficients from `lens` too, but
    the markers are left undistorted (residual / Jacobian parity scenes, not fits).
    `frame_offset`: generate frames offset .. offset+F-1 of a longer shot (same bundles, the camera
    path continues): a rank's own slice of a sharded per-frame workload.
    Returns (SceneBuilder, info dict).
    """
    rng = np.random.default_rng(seed)
    sensor_w, sensor_h, rw, rh = 36.0, 24.0, 2048, 1556
    if in_frame:
        pts = np.stack([rng.uniform(-4, 4, B), rng.uniform(-2.2, 2.2, B), rng.uniform(-30, -14, B)], 1)
    else:
        pts = np.stack([rng.uniform(-10, 10, B), rng.uniform(-10, 10, B), rng.uniform(-30, -10, B)], 1)
    cam_t, cam_r = camera_path(F, rng, period=period, frame_offset=frame_offset)
    sb = SceneBuilder(frames=np.arange(1, F + 1) + frame_offset)
    lens_node = None
    if lens is not None:
        # truth coefficients are only used to make the markers; the scene
        # starts from zero coefficients when they are solved
        init = {k: (0.0 if k in solve_lens else v) for k, v in lens.items()}
        lens_node = sb.add_lens(lens_model, camera_params=(focal_truth * 0.1, sensor_w * 0.1, sensor_h * 0.1, 1.0, 0.0, 0.0), **init)
    t0 = cam_t + rng.normal(0, init_noise_t, cam_t.shape) if solve_camera else cam_t
    r0 = cam_r + rng.normal(0, init_noise_r, cam_r.shape) if solve_camera else cam_r
    r0[:, 2] = cam_r[:, 2] + (rng.normal(0, init_noise_r, F) if solve_camera else 0.0)
    cam_tfm = sb.add_transform(tx=t0[:, 0], ty=t0[:, 1], tz=t0[:, 2], rx=r0[:, 0], ry=r0[:, 1], rz=r0[:, 2])
    f_init = focal_truth if focal_init is None else focal_init
    cam = sb.add_camera(cam_tfm, film_fit="horizontal", render_width=rw, render_height=rh, lens=lens_node,
                        sensor_width_mm=sensor_w, sensor_height_mm=sensor_h, focal_length_mm=f_init)
    b0 = pts + (rng.normal(0, init_noise_b, pts.shape) if solve_bundles else 0.0)
    if fixed_bundles:
        b0[:fixed_bundles] = pts[:fixed_bundles]
    bnd_tfms = []
    mx = np.zeros((B, F))
    my = np.zeros((B, F))
    for f in range(F):
        x, y = project_truth(cam_t[f], cam_r[f], pts, focal_truth, sensor_w, sensor_h, rw, rh)
        mx[:, f] = x
        my[:, f] = y
    if lens is not None and lens_model == "3de_radial_std_deg4" and lens.get("cylindric_bending", 0.0) == 0.0:
        mx, my = _distort_markers(mx, my, lens, (focal_truth * 0.1, sensor_w * 0.1, sensor_h * 0.1, 1.0, 0.0, 0.0), rw / rh, sensor_w / sensor_h)
    mx = mx + rng.normal(0, noise_px / 2048.0, mx.shape)
    my = my + rng.normal(0, noise_px / 2048.0, my.shape)
    enable = np.ones((B, F), np.uint8)
    if drop_fraction > 0:
        enable = (rng.uniform(size=(B, F)) >= drop_fraction).astype(np.uint8)
    w = rng.uniform(0.5, 1.0, (B, F)) if weights else np.ones((B, F))
    for b in range(B):
        t = sb.add_transform(tx=b0[b, 0], ty=b0[b, 1], tz=b0[b, 2])
        bnd_tfms.append(t)
        bn = sb.add_bundle(t)
        sb.add_marker(cam, bn, mx[b], my[b], weight=w[b], enable=enable[b])
    if solve_camera:
        for k in ("tx", "ty", "tz", "rx", "ry", "rz"):
            sb.add_attr(cam_tfm[k])
    if solve_focal:
        sb.add_attr(cam["focal_length_mm"])
    for k in solve_lens:
        sb.add_attr(lens_node[k])
    if solve_bundles:
        for b in range(fixed_bundles, B):
            for k in ("tx", "ty", "tz"):
                sb.add_attr(bnd_tfms[b][k])
    info = dict(F=F, B=B, seed=seed, truth_points=pts, truth_cam_t=cam_t, truth_cam_r=cam_r,
                cam_tfm=cam_tfm, cam=cam, bnd_tfms=bnd_tfms, lens=lens_node)
    return sb, info


def _lens_radial_core(x, y, c):
    c2, u2, v2, c4, u4, v4 = (c.get(k, 0.0) for k in ("degree2_distortion", "degree2_u", "degree2_v",
                                                      "degree4_distortion", "degree4_u", "degree4_v"))
    r2 = x * x + y * y
    rad = 1 + c2 * r2 + c4 * r2 * r2
    return (x * rad + (r2 + 2 * x * x) * (u2 + u4 * r2) + 2 * x * y * (v2 + v4 * r2),
            y * rad + (r2 + 2 * y * y) * (v2 + v4 * r2) + 2 * x * y * (u2 + u4 * r2))


def _distort_markers(mx, my, lens, cam6, render_aspect, sensor_aspect):
    """Markers of a distorted plate: apply the lens *distort* direction (the
    fixed-point inverse the solver applies to projected points) to the ideal
    marker positions.  Data generation only (cylindric terms must be 0)."""
    assert lens.get("cylindric_bending", 0.0) == 0.0
    scale_y = render_aspect / sensor_aspect  # flat.rs Horizontal: residual space is (mx, my*scale)
    w, h = cam6[1], cam6[2]
    radius = np.sqrt(w * w + h * h) / 2
    qx = (mx * w) / radius
    qy = (my * scale_y * h) / radius
    px, py = qx.copy(), qy.copy()
    for _ in range(60):
        fx, fy = _lens_radial_core(px, py, lens)
        px = px + qx - fx
        py = py + qy - fy
    return px * radius / w, py * radius / h / scale_y


def witness_camera_paths(C, frames, focus=(0.0, 0.0, -22.0), radius=22.0, arc_deg=80.0, period=120.0):
    """C cameras on an arc around `focus`, each looking at it, with a small deterministic
    wobble that is a function of the absolute frame number (so that a rank can generate its
    own frame range of a sharded scene).  -> t [C, F, 3], r [C, F, 3] degrees (rotate order xyz)"""
    f = np.asarray(frames, dtype=np.float64)
    t = np.zeros((C, len(f), 3))
    r = np.zeros((C, len(f), 3))
    for c in range(C):
        th = (-0.5 + (c + 0.5) / C) * arc_deg
        w = 2 * np.pi * f / period + 0.7 * c
        t[c, :, 0] = focus[0] + radius * np.sin(th * DEG) + 0.5 * np.sin(w)
        t[c, :, 1] = focus[1] + 0.3 * np.sin(w + 1.0)
        t[c, :, 2] = focus[2] + radius * np.cos(th * DEG) + 0.5 * np.sin(w + 2.0)
        r[c, :, 0] = 1.0 * np.sin(w + 3.0)
        r[c, :, 1] = th + 1.5 * np.sin(w + 4.0)
        r[c, :, 2] = 0.5 * np.sin(w + 5.0)
    return t, r


def make_multicam_scene(C, F, B, seed, frame_range=None, noise_px=0.5, init_noise_t=0.05, init_noise_r=0.2,
                        init_noise_b=0.05, fixed_bundles=0, solve_camera=True, solve_bundles=True, visibility=1.0):
    """cfg5 shape: C witness cameras (animated, 6 attributes each solved per frame) all seeing
    the same B static bundles on every frame: M = C*B markers, Nobs = C*B*F.

    `visibility` < 1: a Bernoulli(visibility) mask per (camera, bundle, frame) decides which
    markers are enabled (cfg5 as SURVEY.md 8d words it: p = 0.25), seeded per absolute frame.

    `frame_range=(f0, f1)` generates only frames f0..f1-1 of the F-frame scene (this rank's
    shard, for Problem(..., local_slice=(f0, F))): truth bundles, start values of the static
    attributes and the camera paths do not depend on the range; marker noise and the start
    noise of the animated attributes are seeded per absolute frame.
    Returns (SceneBuilder, info dict).
    """
    f0, f1 = (0, F) if frame_range is None else frame_range
    frames = np.arange(f0, f1)
    Fl = len(frames)
    rng = np.random.default_rng(seed)
    sensor_w, sensor_h, rw, rh, focal = 36.0, 24.0, 2048, 1556, 35.0
    pts = np.stack([rng.uniform(-4, 4, B), rng.uniform(-2.2, 2.2, B), rng.uniform(-26, -18, B)], 1)
    b0 = pts + (rng.normal(0, init_noise_b, pts.shape) if solve_bundles else 0.0)
    if fixed_bundles:
        b0[:fixed_bundles] = pts[:fixed_bundles]
    cam_t, cam_r = witness_camera_paths(C, frames)
    # per-frame random streams: the same numbers whatever range the frame is generated in
    def per_frame_normal(tag, shape_tail, sigma):
        out = np.zeros((Fl,) + shape_tail)
        if sigma == 0.0:
            return out
        for i, fr in enumerate(frames):
            out[i] = np.random.default_rng([seed, tag, int(fr)]).normal(0, sigma, shape_tail)
        return out
    nt = per_frame_normal(1, (C, 3), init_noise_t if solve_camera else 0.0)
    nr = per_frame_normal(2, (C, 3), init_noise_r if solve_camera else 0.0)
    nm = per_frame_normal(3, (C, B, 2), noise_px / 2048.0)
    sb = SceneBuilder(frames=frames + 1)
    cams, cam_tfms = [], []
    for c in range(C):
        t0 = cam_t[c] + nt[:, c]
        r0 = cam_r[c] + nr[:, c]
        tf = sb.add_transform(tx=t0[:, 0], ty=t0[:, 1], tz=t0[:, 2], rx=r0[:, 0], ry=r0[:, 1], rz=r0[:, 2])
        cam_tfms.append(tf)
        cams.append(sb.add_camera(tf, film_fit="horizontal", render_width=rw, render_height=rh,
                                  sensor_width_mm=sensor_w, sensor_height_mm=sensor_h, focal_length_mm=focal))
    bnds, bnd_tfms = [], []
    for b in range(B):
        tf = sb.add_transform(tx=b0[b, 0], ty=b0[b, 1], tz=b0[b, 2])
        bnd_tfms.append(tf)
        bnds.append(sb.add_bundle(tf))
    ones = np.ones(Fl)
    en = np.ones(Fl, np.uint8)
    vis = None
    if visibility < 1.0:
        vis = np.zeros((Fl, C, B), np.uint8)
        for i, fr in enumerate(frames):
            vis[i] = np.random.default_rng([seed, 4, int(fr)]).uniform(size=(C, B)) < visibility
    for c in range(C):
        mx = np.zeros((B, Fl))
        my = np.zeros((B, Fl))
        for i in range(Fl):
            x, y = project_truth(cam_t[c, i], cam_r[c, i], pts, focal, sensor_w, sensor_h, rw, rh)
            mx[:, i] = x + nm[i, c, :, 0]
            my[:, i] = y + nm[i, c, :, 1]
        for b in range(B):
            sb.add_marker(cams[c], bnds[b], mx[b], my[b], weight=ones, enable=en if vis is None else vis[:, c, b])
    if solve_camera:
        for tf in cam_tfms:
            for k in ("tx", "ty", "tz", "rx", "ry", "rz"):
                sb.add_attr(tf[k])
    if solve_bundles:
        for b in range(fixed_bundles, B):
            for k in ("tx", "ty", "tz"):
                sb.add_attr(bnd_tfms[b][k])
    info = dict(C=C, F=F, B=B, seed=seed, frame_range=(f0, f1), truth_points=pts, truth_cam_t=cam_t, truth_cam_r=cam_r,
                cam_tfms=cam_tfms, cams=cams, bnd_tfms=bnd_tfms)
    return sb, info
