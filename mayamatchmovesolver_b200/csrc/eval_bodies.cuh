// Per-thread bodies of the scene / residual / FD-Jacobian kernels.
//
// Path replaced (reference):
//   setParameters_mmSceneGraph          adjust_setParameters.cpp:174-250
//   FlatScene::evaluate                 lib/rust/mmscenegraph/src/scene/flat.rs:172-358
//   measureErrors_mmSceneGraph          adjust_measureErrors.cpp:392-521
//   solveFunc_calculateJacobianMatrix*  adjust_solveFunc.cpp:305-525
//
// The reference re-evaluates the whole scene once per Jacobian column.  Here
// every (entity, frame, perturbed parameter) "variant" is evaluated once in a
// tiny launch, and a single streaming launch over observations combines them.
#pragma once
#include "device_math.cuh"

namespace mmb {

struct PlanEntryD {
    int32_t attr;
    int16_t cam_var, bnd_var, lens_var, pad;
};

// Everything the kernels read; plain device pointers.
struct DevScene {
    int F, T, C, B, M, L, A, n, nobs, kmax;
    int central;        // autoDiffType == 1
    int dense_grid;     // obs i == marker*F + frame
    int any_lens;
    int lens_literal;   // reference's [markerIndex + frameIndex] lens read
    double delta;       // SolverOptions.delta
    int lmdif;          // solverType == cminpack_lmdif: fdjac2 relative forward step
    double fd_eps;      // sqrt(max(|delta|, epsmch))
    double image_width;
    // attribute store (AttrDataBlock): live values
    const double *stat; // [num_static]  read-only: solved values come from x
    const double *anim; // [num_anim * F]
    // scene
    const int32_t *tfm_parent, *tfm_roo, *tfm_attrs;
    const int32_t *cam_tfm, *cam_attrs, *cam_fit, *cam_rw, *cam_rh, *cam_lens;
    const int32_t *bnd_tfm, *mkr_cam, *mkr_bnd;
    const int32_t *lens_model, *lens_attrs;
    const double *lens_cam;
    // solver attrs / params
    const int32_t *attr_ref, *attr_param_base;
    const int32_t *stat_attr, *anim_attr;  // attribute-store slot -> solver attr (or -1)
    const uint8_t *attr_anim;
    const uint8_t *attr_plain;  // unbounded, scale 1, offset 0
    const double *attr_min, *attr_max, *attr_off, *attr_scl;
    const int32_t *par_attr, *par_frame;
    // observations (errorToMarkerList order)
    const int32_t *err_marker, *err_frame;
    const double *obs_mx, *obs_my, *obs_w;
    // variant tables
    double *cam_tab, *bnd_tab, *lens_tab;
    const int64_t *cam_off, *bnd_off, *lens_off;
    const int32_t *cam_fs, *bnd_fs, *lens_fs;
    const int32_t *cam_ndep, *bnd_ndep, *lens_ndep;  // number of deps
    // variant jobs
    const int32_t *cam_job_ent, *cam_job_var, *cam_job_attr;
    const int64_t *cam_job_start;
    const int32_t *cam_thread_job, *bnd_thread_job, *lens_thread_job;  // thread -> job
    int cam_njobs; long long cam_jobs_total;
    // block path of k_variants: one CTA per (animated, parentless camera, 32 frames)
    const int32_t *cam_job0;             // [C] job of variant 0
    const uint8_t *cam_blk;              // [C] camera is evaluated by the block path
    const int32_t *vblk_cam, *vblk_fb;   // [vblk_count] camera and frame block of CTA i
    int vblk_count, vblk_all;            // vblk_all: every camera is on the block path
    const int32_t *bnd_job_ent, *bnd_job_var, *bnd_job_attr;
    const int64_t *bnd_job_start;
    int bnd_njobs; long long bnd_jobs_total;
    const int32_t *lens_job_ent, *lens_job_var, *lens_job_attr;
    const int64_t *lens_job_start;
    int lens_njobs; long long lens_jobs_total;
    // plan
    const PlanEntryD *plan;  // [M*kmax]
    const int32_t *mkr_k;    // [M]
    // per-parameter FD step actually used (deltaA, deltaB) for x
    double *par_delta_a, *par_delta_b;  // [n]
    // sceneGraphMode = 1 (Maya-DAG-convention arithmetic) and the stiffness / smoothness rows
    int maya_dag;
    int cur_frame;                       // "Maya's current time" as a frame index
    int n_extra;                         // stiffness rows first, then smoothness rows
    const int32_t *ex_attr, *ex_w, *ex_var, *ex_val;   // [n_extra] solver attr, weight / variance / value refs
    double *cam_fwd;                     // [camera][variant][4][frames]: forward axis xyz, forward . position
};

// ---- attribute values --------------------------------------------------------
// setParameters (adjust_setParameters.cpp:174-250) never materialises here: a
// solved attribute is read straight from the parameter vector through
// parameterBoundFromInternalToExternal, an unsolved one from the (read-only)
// attribute store.  One attribute may be overridden by x[p] + delta (the
// perturbed parameter of a Jacobian column).
struct AttrView {
    const DevScene *s;
    const double *x;  // nullptr: the scene's own values (before any setParameters)
    int f;
    int ovr_attr;
    double ovr_delta;
    MMB_HD double get(int32_t ref) const {
        const int a = x ? (ref >= 0 ? s->stat_attr[ref] : s->anim_attr[-ref - 1]) : -1;
        if (a < 0) {
            if (ref >= 0) return s->stat[ref];
            return s->anim[(size_t)(-ref - 1) * s->F + f];
        }
        const int p = s->attr_param_base[a] + (s->attr_anim[a] ? f : 0);
        double xv = x[p];
        if (a == ovr_attr) xv += ovr_delta;
        // no bounds, scale 1, offset 0: (x / 1) - 0 clamped to +-FLT_MAX is x itself
        if (s->attr_plain[a]) return fmin(fmax(xv, -3.40282346638528859811704183484516925e+38), 3.40282346638528859811704183484516925e+38);
        return bound_int_to_ext(xv, s->attr_min[a], s->attr_max[a], s->attr_off[a], s->attr_scl[a]);
    }
};

// Which job / frame does flat thread index t belong to? (binary search)
MMB_HD int find_job(const int64_t *start, int njobs, long long t) {
    int lo = 0, hi = njobs;  // start[lo] <= t < start[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (start[mid] <= t) lo = mid; else hi = mid;
    }
    return lo;
}

constexpr int kCamRec = 16;  // doubles per camera variant record
// Camera table index of (variant v, row j, frame f) for a camera with nvar variants.
// Animated cameras are stored in blocks of 32 frames, [frame block][variant][row pair][frame][2]:
// rows 2k and 2k+1 of one frame are adjacent (one 16-byte access), consecutive frames are
// consecutive (coalesced), and everything a 32-frame tile needs -- all variants -- is ONE
// contiguous run that a single TMA bulk copy brings into shared memory.
constexpr int kCamBlock = 32;
MMB_HD size_t cam_idx(int v, int j, int fs, int f, int nvar) {
    if (fs == 1) return (size_t)v * kCamRec + (size_t)j;
    const size_t fb = (size_t)(f / kCamBlock);
    const int fl = f % kCamBlock;
    return ((((fb * nvar + v) * (kCamRec / 2) + (size_t)(j >> 1)) * kCamBlock + fl) << 1) + (size_t)(j & 1);
}
MMB_HD int cam_nvar_of(int ndep, int central) { return 1 + (central ? 2 : 1) * ndep; }
constexpr int kBndRec = 4;   // doubles per bundle variant record

// View for variant `var` of an entity with `ndep` dependencies at frame f.
// calculateParameterDelta (adjust_solveFunc.cpp:148-180) is evaluated here; the
// steps are also published per parameter for the generic Jacobian kernel.
MMB_HD AttrView make_view(const DevScene &s, const double *x, int f, int attr, int var, int ndep,
                          double &inv_delta_a) {
    AttrView v;
    v.s = &s; v.x = x; v.f = f;
    v.ovr_attr = -1;
    v.ovr_delta = 0.0;
    inv_delta_a = 0.0;
    if (var > 0) {
        const int p = s.attr_param_base[attr] + (s.attr_anim[attr] ? f : 0);
        double da, db;
        if (s.lmdif) {  // fdjac2: h = eps * |x_j|, eps when x_j == 0; no bound logic
            da = s.fd_eps * fabs(x[p]);
            if (da == 0.0) da = s.fd_eps;
            db = -da;
        } else {
            da = parameter_delta(x[p], s.delta, 1.0, s.attr_min[attr], s.attr_max[attr]);
            db = parameter_delta(x[p], s.delta, -1.0, s.attr_min[attr], s.attr_max[attr]);
        }
        s.par_delta_a[p] = da;  // same value from every entity that depends on p
        s.par_delta_b[p] = db;
        inv_delta_a = 1.0 / da;
        v.ovr_attr = attr;
        v.ovr_delta = var > ndep ? db : da;  // central: variants ndep+1.. use deltaB
    }
    return v;
}

MMB_HD Aff world_matrix(const DevScene &s, const AttrView &av, int t) {
    // dag.rs:234-327: world = parent_world * local, accumulated leaf -> root
    Aff w;
    bool first = true;
    while (t >= 0) {
        const int32_t *a = &s.tfm_attrs[t * 9];
        const Aff l = local_matrix(av.get(a[0]), av.get(a[1]), av.get(a[2]), av.get(a[3]), av.get(a[4]),
                                   av.get(a[5]), av.get(a[6]), av.get(a[7]), av.get(a[8]), s.tfm_roo[t]);
        w = first ? l : aff_mul(l, w);
        first = false;
        t = s.tfm_parent[t];
    }
    return w;
}

// One thread per (camera, variant, frame): kCamRec doubles
//   [0..11] rows 0,1,3 of P*inv(W);  [12],[13] marker film-fit scale;
//   [14] 1/deltaA of the perturbed parameter (variants only)
// table layout: cam_off[c] + cam_idx(v, j, fs, f, nvar)
// the record of camera c, variant v, frame f from its world matrix and the camera attribute
// values cav[0..4] = sensor w, sensor h, focal, offset x, offset y; cav[5] = camera scale
MMB_HD void camera_variant_store(const DevScene &s, int c, int v, int f, const Aff &w, const double *cav, double inv_da) {
    const int fs = s.cam_fs[c];
    const double sw = cav[0], sh = cav[1];
    const Proj p = projection_entries(cav[2], sw, sh, cav[3], cav[4], (double)s.cam_rw[c], (double)s.cam_rh[c],
                                      s.cam_fit[c], 0.1 /* dag.rs:152 */, cav[5], s.maya_dag);
    double out[14];
    camera_rows(w, p, out, s.maya_dag);
    if (s.maya_dag) {
        // Camera::getForwardDirection / getWorldPosition (maya_camera.cpp:814-856): (0, 0, -1) * worldMat
        // is minus the third column of the column-convention world matrix; only the sign of
        // forward . (bundle - position) is used (adjust_measureErrors.cpp:262-270)
        const double fx = -w.m[2], fy = -w.m[6], fz = -w.m[10];
        const long long fpad = fs == 1 ? 1 : ((long long)fs + kCamBlock - 1) / kCamBlock * kCamBlock;
        double *fw = s.cam_fwd + s.cam_off[c] / 4 + (size_t)v * 4 * fpad + (fs == 1 ? 0 : f);
        fw[0] = fx; fw[fpad] = fy; fw[2 * fpad] = fz;
        fw[3 * fpad] = fx * w.m[3] + fy * w.m[7] + fz * w.m[11];
    }
    marker_film_fit_scale(s.cam_fit[c], sw / sh, (double)s.cam_rw[c] / (double)s.cam_rh[c], out[12], out[13]);
    double *dst = s.cam_tab + s.cam_off[c];
    const int nvar = cam_nvar_of(s.cam_ndep[c], s.central);
#pragma unroll
    for (int j = 0; j < 14; ++j) dst[cam_idx(v, j, fs, f, nvar)] = out[j];
    dst[cam_idx(v, 14, fs, f, nvar)] = inv_da;
}

// Variant v of camera c at frame f when the frame's 17 attribute values (9 transform, 8 camera;
// vals[j * stride]) have already been read through the attribute view: only the attribute the
// variant perturbs is looked up again.  refs / attr_of: attribute-store ref and solver attribute
// of the 17 slots.  Cameras whose transform has no parent only.
MMB_HD void camera_variant_gathered(const DevScene &s, const double *x, int c, int v, int f, const double *vals,
                                    int stride, const int *refs, const int *attr_of) {
    const int attr = s.cam_job_attr[s.cam_job0[c] + v];
    double inv_da;
    const AttrView av = make_view(s, x, f, attr, v, s.cam_ndep[c], inv_da);
    double a[17];
#pragma unroll
    for (int j = 0; j < 17; ++j) a[j] = (v > 0 && attr_of[j] == attr) ? av.get(refs[j]) : vals[j * stride];
    const Aff w = local_matrix(a[0], a[1], a[2], a[3], a[4], a[5], a[6], a[7], a[8], s.tfm_roo[s.cam_tfm[c]]);
    const double cav[6] = {a[9], a[10], a[11], a[12], a[13], a[16]};
    camera_variant_store(s, c, v, f, w, cav, inv_da);
}

MMB_HD void camera_variant_thread(const DevScene &s, const double *x, long long t) {
    const int job = s.cam_thread_job[t];
    const int c = s.cam_job_ent[job], v = s.cam_job_var[job], attr = s.cam_job_attr[job];
    const int f = (int)(t - s.cam_job_start[job]);
    double inv_da;
    const AttrView av = make_view(s, x, f, attr, v, s.cam_ndep[c], inv_da);
    const Aff w = world_matrix(s, av, s.cam_tfm[c]);
    const int32_t *ca = &s.cam_attrs[c * 8];
    const double cav[6] = {av.get(ca[0]), av.get(ca[1]), av.get(ca[2]), av.get(ca[3]), av.get(ca[4]), av.get(ca[7])};
    camera_variant_store(s, c, v, f, w, cav, inv_da);
}


// One thread per (bundle, variant, frame): world position + 1/deltaA (kBndRec doubles)
MMB_HD void bundle_variant_thread(const DevScene &s, const double *x, long long t) {
    const int job = s.bnd_thread_job[t];
    const int b = s.bnd_job_ent[job], v = s.bnd_job_var[job], attr = s.bnd_job_attr[job];
    const int f = (int)(t - s.bnd_job_start[job]);
    const int fs = s.bnd_fs[b];
    double inv_da;
    const AttrView av = make_view(s, x, f, attr, v, s.bnd_ndep[b], inv_da);
    const int tfm = s.bnd_tfm[b];
    double px, py, pz;
    if (s.tfm_parent[tfm] < 0) {
        // the world position of a parentless transform is its translation (T * R * S puts
        // tx, ty, tz in the last column untouched): no rotation, no scale, three lookups
        const int32_t *a = &s.tfm_attrs[tfm * 9];
        px = av.get(a[0]); py = av.get(a[1]); pz = av.get(a[2]);
    } else {
        const Aff w = world_matrix(s, av, tfm);
        px = w.m[3]; py = w.m[7]; pz = w.m[11];
    }
    double *dst = s.bnd_tab + s.bnd_off[b];
    dst[((size_t)v * kBndRec + 0) * fs + f] = px;
    dst[((size_t)v * kBndRec + 1) * fs + f] = py;
    dst[((size_t)v * kBndRec + 2) * fs + f] = pz;
    dst[((size_t)v * kBndRec + 3) * fs + f] = inv_da;
}

// One thread per (lens, variant, frame): prepared lens record (24 doubles)
MMB_HD void lens_variant_thread(const DevScene &s, const double *x, long long t) {
    const int job = s.lens_thread_job[t];
    const int l = s.lens_job_ent[job], v = s.lens_job_var[job], attr = s.lens_job_attr[job];
    const int f = (int)(t - s.lens_job_start[job]);
    const int fs = s.lens_fs[l];
    double inv_da;
    const AttrView av = make_view(s, x, f, attr, v, s.lens_ndep[l], inv_da);
    double c[14], rec[kLensRec];
    for (int k = 0; k < 14; ++k) {
        const int32_t r = s.lens_attrs[l * 14 + k];
        c[k] = (r == INT32_MIN) ? 0.0 : av.get(r);
    }
    lens_prepare(s.lens_model[l], c, &s.lens_cam[l * 6], rec);
    double *dst = s.lens_tab + s.lens_off[l];
    for (int j = 0; j < kLensRec; ++j) dst[((size_t)v * kLensRec + j) * fs + f] = rec[j];
}

// ---- observation evaluation ------------------------------------------------
struct ObsCtx {
    int marker, frame, cam, bnd, lens, lens_frame;
    double mx, my, w;       // raw marker, sqrt(weight)
};

constexpr int kCamVals = 18;  // 14 + the forward record of sceneGraphMode 1
MMB_HD void load_cam(const DevScene &s, int c, int v, int f, double *cm /*kCamVals*/) {
    const int fs = s.cam_fs[c];
    const int ff = fs == 1 ? 0 : f;
    const double *src = s.cam_tab + s.cam_off[c];
    const int nvar = cam_nvar_of(s.cam_ndep[c], s.central);
#pragma unroll
    for (int j = 0; j < 14; ++j) cm[j] = src[cam_idx(v, j, fs, ff, nvar)];
    if (s.maya_dag) {
        const long long fpad = fs == 1 ? 1 : ((long long)fs + kCamBlock - 1) / kCamBlock * kCamBlock;
        const double *fw = s.cam_fwd + s.cam_off[c] / 4 + (size_t)v * 4 * fpad + ff;
        cm[14] = fw[0]; cm[15] = fw[fpad]; cm[16] = fw[2 * fpad]; cm[17] = fw[3 * fpad];
    }
}
MMB_HD void load_bnd(const DevScene &s, int b, int v, int f, double *bp /*3*/) {
    const int fs = s.bnd_fs[b];
    const int ff = fs == 1 ? 0 : f;
    const double *src = s.bnd_tab + s.bnd_off[b] + (size_t)v * kBndRec * fs + ff;
    bp[0] = src[0]; bp[1] = src[fs]; bp[2] = src[2 * (size_t)fs];
}
MMB_HD void load_lens(const DevScene &s, int l, int v, int f, double *rec /*24*/) {
    const int fs = s.lens_fs[l];
    const int ff = fs == 1 ? 0 : f;
    const double *src = s.lens_tab + s.lens_off[l] + (size_t)v * kLensRec * fs + ff;
    for (int j = 0; j < kLensRec; ++j) rec[j] = src[(size_t)j * fs];
}

// reprojection.rs:27-62 on the reduced rows; then the residual pair
// (adjust_measureErrors.cpp:481-499).  ex, ey are the weighted residuals.
MMB_HD void project_point(const double *cm, const double *bp, double &px, double &py) {
    const double sx = cm[0] * bp[0] + cm[1] * bp[1] + cm[2] * bp[2] + cm[3];
    const double sy = cm[4] * bp[0] + cm[5] * bp[1] + cm[6] * bp[2] + cm[7];
    const double sw = cm[8] * bp[0] + cm[9] * bp[1] + cm[10] * bp[2] + cm[11];
    px = sx / sw * 0.5;
    py = sy / sw * 0.5;
}

MMB_HD ObsCtx make_obs(const DevScene &s, int i) {
    ObsCtx o;
    if (s.dense_grid) { o.marker = i / s.F; o.frame = i - o.marker * s.F; }
    else { o.marker = s.err_marker[i]; o.frame = s.err_frame[i]; }
    o.cam = s.mkr_cam[o.marker]; o.bnd = s.mkr_bnd[o.marker];
    o.mx = s.obs_mx[i]; o.my = s.obs_my[i]; o.w = sqrt(s.obs_w[i]);
    o.lens = -1; o.lens_frame = o.frame;
    if (s.any_lens) {
        int lm = o.marker;
        if (s.lens_literal) {  // adjust_measureErrors.cpp:463 literal index
            const int idx = o.marker + o.frame;
            lm = idx / s.F; o.lens_frame = idx - lm * s.F;
            if (lm >= s.M) lm = -1;
        }
        if (lm >= 0) {
            const int l = s.cam_lens[s.mkr_cam[lm]];
            if (l >= 0 && s.lens_model[l] != 0) o.lens = l;
        }
    }
    return o;
}

// fwd: the camera's forward record (sceneGraphMode 1, generic kernels only) or nullptr
MMB_HD void residual_from(const DevScene &s, const ObsCtx &o, const double *cm, const double *bp,
                          const double *lens_rec, double &ex, double &ey, double &dist, const double *fwd = nullptr) {
    double px, py;
    project_point(cm, bp, px, py);
    if (lens_rec) {
        double ox, oy;
        lens_distort(lens_rec, px, py, ox, oy);
        px = ox; py = oy;
    }
    const double mkx = o.mx * cm[12], mky = o.my * cm[13];
    const double dx = fabs(mkx - px), dy = fabs(mky - py);
    const double dxp = dx * s.image_width, dyp = dy * s.image_width;
    ex = dxp * o.w;
    ey = dyp * o.w;
    if (fwd) {
        // behind the camera: the residual pair (not the reported distance) x 1e6
        // (adjust_measureErrors.cpp:262-282)
        const double dot = fwd[0] * bp[0] + fwd[1] * bp[1] + fwd[2] * bp[2] - fwd[3];
        if (dot < 0.0) { ex *= 1e+6; ey *= 1e+6; }
    }
    dist = sqrt((dx * dx) + (dy * dy)) * s.image_width;
}

// ---- stiffness / smoothness rows (adjust_measureErrors.cpp:311-387) ---------------------------
// error = (1 / gaussian(attr value, value, variance) - 1) * weight, every value read at the
// current time.  `av` decides which parameter is perturbed (the FD quotient of the row).
MMB_HD double extra_row_value(const DevScene &s, const AttrView &av, int e) {
    if (!s.maya_dag) return 0.0;   // measureErrors_mmSceneGraph leaves the rows untouched: zero
    const double weight = av.get(s.ex_w[e]), variance = av.get(s.ex_var[e]), value = av.get(s.ex_val[e]);
    const double attr_value = av.get(s.attr_ref[s.ex_attr[e]]);
    const double g = exp(-(pow((attr_value - value), 2.0) / (2.0 * (pow(variance, 2.0)))));
    return ((1.0 / g) - 1.0) * weight;
}
// row value at x (fvec entry) and its one non-zero Jacobian entry: the parameter of the row's
// attribute at the current frame, forward or central quotient exactly as the marker rows
MMB_HD void extra_row_thread(const DevScene &s, const double *x, int e, double *row, double *jext, int32_t *pid_out) {
    double inv;
    const AttrView base = make_view(s, x, s.cur_frame, -1, 0, 0, inv);
    const double r0 = extra_row_value(s, base, e);
    if (row) row[e] = r0;
    if (!jext) return;
    const int a = s.ex_attr[e];
    const int pid = s.attr_param_base[a] + (s.attr_anim[a] ? s.cur_frame : 0);
    double j = 0.0;
    if (s.maya_dag) {
        const AttrView va = make_view(s, x, s.cur_frame, a, 1, 1, inv);
        const double ra = extra_row_value(s, va, e);
        const double da = s.par_delta_a[pid];
        bool forward = !s.central;
        double db = 0.0;
        if (!forward) { db = s.par_delta_b[pid]; if (da == db) forward = true; }
        if (forward) j = (ra - r0) * (1.0 / da);
        else {
            const AttrView vb = make_view(s, x, s.cur_frame, a, 2, 1, inv);
            const double rb = extra_row_value(s, vb, e);
            j = (ra - rb) * (0.5 / (fabs(da) + fabs(db)));
        }
    }
    jext[e] = j;
    pid_out[e] = pid;
}

// Residuals only (solveFunc iflag=1).  fvec[2i], fvec[2i+1], dist[i].
MMB_HD void residual_thread(const DevScene &s, int i, double *fvec, double *dist) {
    const ObsCtx o = make_obs(s, i);
    double cm[kCamVals], bp[3], rec[kLensRec];
    load_cam(s, o.cam, 0, o.frame, cm);
    load_bnd(s, o.bnd, 0, o.frame, bp);
    if (o.lens >= 0) load_lens(s, o.lens, 0, o.lens_frame, rec);
    double ex, ey, d;
    residual_from(s, o, cm, bp, o.lens >= 0 ? rec : nullptr, ex, ey, d, s.maya_dag ? cm + 14 : nullptr);
    fvec[2 * (size_t)i] = ex;
    fvec[2 * (size_t)i + 1] = ey;
    if (dist) dist[i] = d;
}

// Residuals + block Jacobian (solveFunc iflag=1 and iflag=2 fused).
//   jac[c * m + 2i + {0,1}], col_ids[i * kmax + c] = parameter index or -1
// dist_last (optional): the distance the reference leaves in errorDistanceList
// after its Jacobian loop -- the evaluation with the LAST solver attribute
// perturbed (adjust_measureErrors.cpp:497 overwrites it on every call).
MMB_HD void jacobian_thread(const DevScene &s, int i, double *fvec, double *dist, double *jac,
                            int32_t *col_ids, double *dist_last = nullptr) {
    const ObsCtx o = make_obs(s, i);
    const size_t m = 2 * (size_t)s.nobs;
    double cm[kCamVals], bp[3], rec[kLensRec];
    load_cam(s, o.cam, 0, o.frame, cm);
    load_bnd(s, o.bnd, 0, o.frame, bp);
    const bool has_lens = o.lens >= 0;
    if (has_lens) load_lens(s, o.lens, 0, o.lens_frame, rec);
    double ex, ey, d;
    residual_from(s, o, cm, bp, has_lens ? rec : nullptr, ex, ey, d, s.maya_dag ? cm + 14 : nullptr);
    if (fvec) { fvec[2 * (size_t)i] = ex; fvec[2 * (size_t)i + 1] = ey; }
    if (dist) dist[i] = d;
    double dlast = d;
    const int k = s.mkr_k[o.marker];
    const PlanEntryD *pl = s.plan + (size_t)o.marker * s.kmax;
    const int cnd = s.cam_ndep[o.cam], bnd_nd = s.bnd_ndep[o.bnd];
    const int lnd = has_lens ? s.lens_ndep[o.lens] : 0;
    for (int c = 0; c < s.kmax; ++c) {
        double jx = 0.0, jy = 0.0;
        int32_t pid = -1;
        if (c < k) {
            const PlanEntryD e = pl[c];
            pid = s.attr_param_base[e.attr] + (s.attr_anim[e.attr] ? o.frame : 0);
            double cmv[kCamVals], bpv[3], recv[kLensRec];
            const double *pcm = cm, *pbp = bp, *prec = has_lens ? rec : nullptr;
            if (e.cam_var) { load_cam(s, o.cam, e.cam_var, o.frame, cmv); pcm = cmv; }
            if (e.bnd_var) { load_bnd(s, o.bnd, e.bnd_var, o.frame, bpv); pbp = bpv; }
            if (e.lens_var && has_lens) { load_lens(s, o.lens, e.lens_var, o.lens_frame, recv); prec = recv; }
            double ax, ay, ad;
            residual_from(s, o, pcm, pbp, prec, ax, ay, ad, s.maya_dag ? pcm + 14 : nullptr);
            const double da = s.par_delta_a[pid];
            bool forward = !s.central;
            double db = 0.0;
            if (!forward) { db = s.par_delta_b[pid]; if (da == db) forward = true; }
            if (forward) {
                const double inv_delta = 1.0 / da;
                jx = (ax - ex) * inv_delta;
                jy = (ay - ey) * inv_delta;
                if (e.attr == s.A - 1) dlast = ad;
            } else {
                pcm = cm; pbp = bp; prec = has_lens ? rec : nullptr;
                if (e.cam_var) { load_cam(s, o.cam, e.cam_var + cnd, o.frame, cmv); pcm = cmv; }
                if (e.bnd_var) { load_bnd(s, o.bnd, e.bnd_var + bnd_nd, o.frame, bpv); pbp = bpv; }
                if (e.lens_var && has_lens) { load_lens(s, o.lens, e.lens_var + lnd, o.lens_frame, recv); prec = recv; }
                double bx, by, bd;
                residual_from(s, o, pcm, pbp, prec, bx, by, bd, s.maya_dag ? pcm + 14 : nullptr);
                // adjust_solveFunc.cpp:468 (0.5/(|dA|+|dB|), as written)
                const double inv_delta = 0.5 / (fabs(da) + fabs(db));
                jx = (ax - bx) * inv_delta;
                jy = (ay - by) * inv_delta;
                if (e.attr == s.A - 1) dlast = bd;
            }
        }
        jac[(size_t)c * m + 2 * (size_t)i] = jx;
        jac[(size_t)c * m + 2 * (size_t)i + 1] = jy;
        if (col_ids) col_ids[(size_t)i * s.kmax + c] = pid;
    }
    if (dist_last) dist_last[i] = dlast;
}

}  // namespace mmb
