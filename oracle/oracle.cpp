// TEST INFRASTRUCTURE ONLY -- see oracle_core.h.  extern "C" surface of the
// CPU oracle (loaded with ctypes by tests/ and bench.py's CPU legs).
#include "oracle_core.h"

// adjust_solveFunc.cpp:148-180 (internal value vs external bounds: kept)
static double calculateParameterDelta(double value, double delta, double sign,
                                      double xmin, double xmax) {
    double new_sign = sign;
    if ((value + delta) > xmax) new_sign = -1;
    if ((value - delta) < xmin) new_sign = 1;
    return delta * new_sign;
}

// adjust_solveFunc.cpp:187-226
static void determineMarkersToBeEvaluated(orc_problem *p, const double *x,
                                          std::vector<bool> &evalMeasurements) {
    const double approxDelta = std::fabs(p->opt.delta) * 0.5;
    bool noneChanged = true;
    std::vector<bool> changed(p->n, false);
    for (int i = 0; i < p->n; ++i) {
        // number::isApproxEqual: fabs(a-b) < eps
        const bool c = !(std::fabs(x[i] - p->previousParam[i]) < approxDelta);
        changed[i] = c; if (c) noneChanged = false;
    }
    evalMeasurements.assign(p->nobs, false);
    // errorToParamList (adjust_relationships.cpp:565-617) evaluated lazily.
    for (int j = 0; j < p->nobs; ++j) {
        const int mk = p->errorToMarker[j].first, mf = p->errorToMarker[j].second;
        bool any = false;
        for (int i = 0; i < p->n && !any; ++i) {
            if (!(changed[i] || noneChanged)) continue;
            const int ai = p->paramToAttr[i].first, af = p->paramToAttr[i].second;
            if (!p->markerToAttr[mk][ai]) continue;
            if (af >= 0 && af != mf) continue;
            any = true;
        }
        evalMeasurements[j] = any;
    }
}

// solveFunc iflag=1: adjust_solveFunc.cpp:262-303
static void func_residuals(orc_problem *p, const double *x, double *fvec, bool jacobian_call = false) {
    // incrementNormalIteration / incrementJacobianIteration (adjust_solveFunc.cpp:227-259):
    // an lmdif fdjac2 call (iflag = 2, doCalcJacobian = false) counts as a Jacobian iteration
    ++p->funcEvalNum;
    if (jacobian_call) ++p->jacIterNum; else ++p->iterNum;
    std::vector<bool> evalM(p->nobs, true), frameEn(p->F, true);
    setParameters(p, x);
    double a, b, c;
    measureErrors(p, frameEn, evalM, fvec, a, b, c);
}

// solveFunc iflag=2: adjust_solveFunc.cpp:305-525.  `structural_skip` only
// skips observations that cannot depend on the parameter (their FD quotient
// is exactly 0 in the reference too); it changes cost, not values: the
// reference's measureErrors would leave the UNPERTURBED distance of such an
// observation in errorDistanceList, which is written here without evaluating.
static void func_jacobian(orc_problem *p, const double *x, const double *fvec,
                          double *fjac, int ldfjac, bool structural_skip) {
    const int n = p->n, m = p->m;
    std::vector<bool> evalMeasurements;
    determineMarkersToBeEvaluated(p, x, evalMeasurements);
    std::vector<double> paramA(n), errA(m), paramB, errB;
    const double delta = p->opt.delta;
    // lmder calls iflag=2 right after an accepted iflag=1 at the same x: errorDist = dist(x)
    const std::vector<double> baseDist = p->errorDist;
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) paramA[j] = x[j];
        for (int j = 0; j < m; ++j) errA[j] = fvec[j];
        const int ai = p->paramToAttr[i].first;
        const double value = x[i];
        const double deltaA = calculateParameterDelta(value, delta, 1, p->attr_min[ai], p->attr_max[ai]);
        std::vector<bool> evalM = evalMeasurements;
        if (structural_skip)
            for (int j = 0; j < p->nobs; ++j)
                if (!p->markerToAttr[p->errorToMarker[j].first][ai]) evalM[j] = false;
        ++p->funcEvalNum; ++p->jacIterNum;
        paramA[i] += deltaA;
        setParameters(p, paramA.data());
        double a, b, c;
        measureErrors(p, p->paramFrame[i], evalM, errA.data(), a, b, c);
        if (structural_skip)
            for (int j = 0; j < p->nobs; ++j)
                if (evalMeasurements[j] && !evalM[j] && p->paramFrame[i][p->errorToMarker[j].second]) p->errorDist[j] = baseDist[j];
        bool forward = p->opt.auto_diff_type == 0;
        double deltaB = 0.0;
        if (!forward) {
            deltaB = calculateParameterDelta(value, delta, -1, p->attr_min[ai], p->attr_max[ai]);
            if (deltaA == deltaB) forward = true;
        }
        if (forward) {
            const double inv_delta = 1.0 / deltaA;
            for (int j = 0; j < m; ++j) fjac[(size_t)i * ldfjac + j] = (errA[j] - fvec[j]) * inv_delta;
        } else {
            ++p->funcEvalNum; ++p->jacIterNum;
            paramB.assign(x, x + n);
            // The reference initialises errorListB with zeros (:412), not with
            // a copy of `errors`, so rows it does not re-evaluate become
            // errors[j]*inv_delta instead of 0.  central_literal reproduces
            // that; the default is the intended behaviour (copy of errors).
            if (p->central_literal) errB.assign(m, 0.0); else errB.assign(fvec, fvec + m);
            paramB[i] += deltaB;
            setParameters(p, paramB.data());
            measureErrors(p, p->paramFrame[i], evalM, errB.data(), a, b, c);
            const double inv_delta = 0.5 / (std::fabs(deltaA) + std::fabs(deltaB));
            for (int j = 0; j < m; ++j) fjac[(size_t)i * ldfjac + j] = (errA[j] - errB[j]) * inv_delta;
        }
    }
}

struct CbData { orc_problem *p; bool skip; bool dif = false; };
static int lm_callback(void *ud, int m, int n, const double *x, double *fvec,
                       double *fjac, int ldfjac, int iflag) {
    CbData *d = (CbData *)ud; (void)m; (void)n;
    if (iflag == 1) func_residuals(d->p, x, fvec);
    else if (iflag == 2 && d->dif) func_residuals(d->p, x, fvec, true);  // solveFunc_cminpack_lmdif
    else if (iflag == 2) func_jacobian(d->p, x, fvec, fjac, ldfjac, d->skip);
    return 0;
}

static const char *k_reasons[9] = {
    "Improper input parameters",
    "Both actual and predicted relative reductions in the sum of squares are at most ftol.",
    "Relative error between two consecutive iterates is at most xtol.",
    "Conditions for info = 1 and info = 2 both hold.",
    "The cosine of the angle between fvec and any column of the Jacobian is at most gtol in absolute value.",
    "Number of calls to fcn has reached or exceeded maxfev.",
    "'ftol' is too small. No further reduction in the sum of squares is possible.",
    "'xtol' is too small. No further improvement in the approximate solution x is possible.",
    "'gtol' is too small. fvec is orthogonal to the columns of the Jacobian to machine precision."};

// compute_error_stats adjust_base.cpp:346-372
static void error_stats(const orc_problem *p, double &avg, double &mn, double &mx) {
    avg = 0; mn = std::numeric_limits<double>::max(); mx = -0.0;
    for (int i = 0; i < p->nobs; ++i) {
        const double e = p->errorDist[i];
        if (!std::isfinite(e)) continue;
        avg += e; if (e < mn) mn = e; if (e > mx) mx = e;
    }
    avg /= p->nobs;
}

#ifdef _OPENMP
#include <omp.h>
#endif
extern "C" int orc_initial_parameters(orc_problem *p, double *x);
#include "oracle_sparse.h"

extern "C" {

int orc_create(const mmb_scene_desc *s, const mmb_options *opt, orc_problem **out) {
    if (!s || !opt || !out) return MMB_ERR_INVALID_ARGUMENT;
    orc_problem *p = new orc_problem();
    p->opt = *opt;
    p->F = s->num_frames; p->T = s->num_transforms; p->C = s->num_cameras;
    p->B = s->num_bundles; p->M = s->num_markers; p->L = s->num_lenses; p->A = s->num_attrs;
    const size_t F = p->F;
    p->frames.assign(s->frames, s->frames + F);
    p->stat.assign(s->static_values, s->static_values + s->num_static);
    p->anim.assign(s->anim_values, s->anim_values + (size_t)s->num_anim * F);
    p->stat0 = p->stat; p->anim0 = p->anim;
    p->tfm_parent.assign(s->tfm_parent, s->tfm_parent + p->T);
    p->tfm_roo.assign(s->tfm_rotate_order, s->tfm_rotate_order + p->T);
    p->tfm_attrs.assign(s->tfm_attrs, s->tfm_attrs + p->T * 9);
    p->cam_tfm.assign(s->cam_tfm, s->cam_tfm + p->C);
    p->cam_attrs.assign(s->cam_attrs, s->cam_attrs + p->C * 8);
    p->cam_fit.assign(s->cam_film_fit, s->cam_film_fit + p->C);
    p->cam_rw.assign(s->cam_render_width, s->cam_render_width + p->C);
    p->cam_rh.assign(s->cam_render_height, s->cam_render_height + p->C);
    p->cam_lens.assign(p->C, -1);
    if (s->cam_lens) p->cam_lens.assign(s->cam_lens, s->cam_lens + p->C);
    p->bnd_tfm.assign(s->bnd_tfm, s->bnd_tfm + p->B);
    p->mkr_cam.assign(s->mkr_cam, s->mkr_cam + p->M);
    p->mkr_bnd.assign(s->mkr_bnd, s->mkr_bnd + p->M);
    p->mkr_x.assign(s->mkr_x, s->mkr_x + p->M * F);
    p->mkr_y.assign(s->mkr_y, s->mkr_y + p->M * F);
    p->mkr_w.assign(s->mkr_weight, s->mkr_weight + p->M * F);
    p->mkr_en.assign(s->mkr_enable, s->mkr_enable + p->M * F);
    if (p->L > 0) {
        p->lens_model.assign(s->lens_model, s->lens_model + p->L);
        p->lens_attrs.assign(s->lens_attrs, s->lens_attrs + p->L * MMB_LENS_MAX_COEFFS);
        p->lens_cam.assign(s->lens_camera_params, s->lens_camera_params + p->L * 6);
    }
    p->attr_ref.assign(s->attr_ref, s->attr_ref + p->A);
    p->attr_min.assign(s->attr_min, s->attr_min + p->A);
    p->attr_max.assign(s->attr_max, s->attr_max + p->A);
    p->attr_off.assign(s->attr_offset, s->attr_offset + p->A);
    p->attr_scl.assign(s->attr_scale, s->attr_scale + p->A);
    for (int i = 0; i < s->num_stiffness; ++i)
        p->stiffAttrs.push_back({s->stiff_attr[i], s->stiff_weight[i], s->stiff_variance[i], s->stiff_value[i]});
    for (int i = 0; i < s->num_smoothness; ++i)
        p->smoothAttrs.push_back({s->smooth_attr[i], s->smooth_weight[i], s->smooth_variance[i], s->smooth_value[i]});
    p->currentFrame = s->current_frame;
    split_used_objects(p);
    build_index_maps(p);
    *out = p;
    return MMB_OK;
}

// which of the caller's markers / attributes take part (solve_v1's used / unused split)
int orc_used_objects(const orc_problem *p, uint8_t *marker_used, uint8_t *attr_used) {
    if (marker_used) for (size_t i = 0; i < p->marker_used_in.size(); ++i) marker_used[i] = p->marker_used_in[i];
    if (attr_used) for (size_t i = 0; i < p->attr_used_in.size(); ++i) attr_used[i] = p->attr_used_in[i];
    return MMB_OK;
}

void orc_destroy(orc_problem *p) { delete p; }

int orc_sizes(const orc_problem *p, int32_t *n, int32_t *nobs, int32_t *m) {
    if (n) *n = p->n; if (nobs) *nobs = p->nobs; if (m) *m = p->m;
    return MMB_OK;
}

int orc_index_maps(const orc_problem *p, int32_t *e2m, int32_t *p2a) {
    if (e2m) for (int i = 0; i < p->nobs; ++i) { e2m[2 * i] = p->errorToMarker[i].first; e2m[2 * i + 1] = p->errorToMarker[i].second; }
    if (p2a) for (int i = 0; i < p->n; ++i) { p2a[2 * i] = p->paramToAttr[i].first; p2a[2 * i + 1] = p->paramToAttr[i].second; }
    return MMB_OK;
}

int orc_weights(const orc_problem *p, double *w) {
    for (int i = 0; i < p->nobs; ++i) w[i] = p->markerWeight[i];
    return MMB_OK;
}

void orc_set_faithful(orc_problem *p, int faithful) { p->faithful_eval = faithful != 0; }
void orc_set_central_literal(orc_problem *p, int v) { p->central_literal = v != 0; }

// get_initial_parameters adjust_base.cpp:260-295
int orc_initial_parameters(orc_problem *p, double *x) {
    for (int i = 0; i < p->n; ++i) {
        const int ai = p->paramToAttr[i].first, fi = p->paramToAttr[i].second;
        const int32_t ref = p->attr_ref[ai];
        const double v = ref >= 0 ? p->stat0[ref] : p->anim0[(size_t)(-ref - 1) * p->F + (fi < 0 ? 0 : fi)];
        x[i] = boundExternalToInternal(v, p->attr_min[ai], p->attr_max[ai], p->attr_off[ai], p->attr_scl[ai]);
    }
    return MMB_OK;
}

int orc_eval_residuals(orc_problem *p, const double *x, double *fvec, double *dist) {
    func_residuals(p, x, fvec);
    if (dist) for (int i = 0; i < p->nobs; ++i) dist[i] = p->errorDist[i];
    return MMB_OK;
}

// fvec must hold the residuals at x (as cminpack passes them to iflag=2).
int orc_eval_jacobian_dense(orc_problem *p, const double *x, const double *fvec,
                            double *fjac, int32_t ldfjac, int32_t structural_skip) {
    setParameters(p, x);  // state cminpack leaves: last f-eval was at x
    func_jacobian(p, x, fvec, fjac, ldfjac, structural_skip != 0);
    return MMB_OK;
}

// Whole solve: solveFrames (adjust_base.cpp:1076-1287) for all frames at once.
int orc_solve(orc_problem *p, mmb_result *res, double *x_out, double *attr_out,
              double *dist_out, int32_t structural_skip) {
    std::memset(res, 0, sizeof(*res));
    if (p->nobs == 0 || p->n == 0) { res->success = 0; return MMB_ERR_NOT_ENOUGH_OBJECTS; }
    if (p->n > p->m) { res->success = 0; return MMB_ERR_MORE_PARAMS_THAN_ERRORS; }
    p->stat = p->stat0; p->anim = p->anim0;
    p->iterNum = p->jacIterNum = p->funcEvalNum = 0;
    const int n = p->n, m = p->m;
    std::vector<double> errors(m, 0.0), x(n), prev(n);
    double iavg = 0, imin = 0, imax = 0;
    if (p->opt.accept_only_better) {
        std::vector<bool> fe(p->F, true), em(p->nobs, true);
        double a, b, c;
        measureErrors(p, fe, em, errors.data(), a, b, c);
        error_stats(p, iavg, imin, imax);
    }
    orc_initial_parameters(p, x.data());
    prev = x;
    // solve_3d_cminpack_lmder adjust_cminpack_lmder.cpp:64-198
    const int ldfjac = std::max(m, n);
    std::vector<double> fjac((size_t)n * ldfjac, 0.0), qtf(n), wa1(n), wa2(n), wa3(n), wa4(m), diag(n, 1.0);
    std::vector<int> ipvt(n);
    const int mode = p->opt.auto_param_scale == 1 ? 1 : 2;
    const double factor = p->opt.tau * 100.0;
    int nfev = 0, njev = 0;
    // solverType 1 = cminpack_lmdif (adjust_cminpack_lmdif.cpp:60-198: epsfcn = |delta|)
    const bool dif = p->opt.solver_type == 1;
    CbData cb{p, structural_skip != 0, dif};
    int info = lmder(lm_callback, &cb, m, n, x.data(), errors.data(), fjac.data(), ldfjac,
                     p->opt.eps1, p->opt.eps2, p->opt.eps3, p->opt.iter_max, diag.data(), mode,
                     factor, &nfev, &njev, ipvt.data(), qtf.data(), wa1.data(), wa2.data(),
                     wa3.data(), wa4.data(), nullptr, dif, std::fabs(p->opt.delta));
    res->error_final = enorm(m, errors.data());
    res->success = p->iterNum > 0;
    res->reason_num = info;
    res->reason_str = (info >= 0 && info < 9) ? k_reasons[info] : "";
    res->iteration_num = nfev;
    res->iteration_function_num = p->iterNum;
    res->iteration_jacobian_num = p->jacIterNum;
    res->iteration_attempt_num = njev;
    error_stats(p, res->error_avg, res->error_min, res->error_max);
    bool better = true;
    if (p->opt.accept_only_better) better = res->error_avg <= iavg;
    res->error_is_better = better;
    const std::vector<double> &fin = better ? x : prev;
    for (int i = 0; i < n; ++i) {
        const int ai = p->paramToAttr[i].first;
        if (x_out) x_out[i] = x[i];
        if (attr_out) attr_out[i] = boundInternalToExternal(fin[i], p->attr_min[ai], p->attr_max[ai], p->attr_off[ai], p->attr_scl[ai]);
    }
    if (dist_out) for (int i = 0; i < p->nobs; ++i) dist_out[i] = p->errorDist[i];
    return MMB_OK;
}

// Block-sparse solve (oracle_sparse.h): same evaluation arithmetic and control law, normal
// equations + Schur complement instead of the dense QR.  threads <= 0: all cores.
// max_outer > 0 stops after that many Jacobian evaluations (bounded timing samples).
int orc_solve_sparse(orc_problem *p, mmb_result *res, double *x_out, double *attr_out, double *dist_out,
                     int32_t threads, int32_t max_outer) {
    std::memset(res, 0, sizeof(*res));
    if (p->nobs == 0 || p->n == 0) { res->success = 0; return MMB_ERR_NOT_ENOUGH_OBJECTS; }
    if (p->n > p->m) { res->success = 0; return MMB_ERR_MORE_PARAMS_THAN_ERRORS; }
    if (p->opt.auto_diff_type != 0) return MMB_ERR_UNSUPPORTED;
    if (p->opt.scene_graph_mode == 1 || p->m != 2 * p->nobs) return MMB_ERR_UNSUPPORTED;  // marker rows, mode 2 only
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#else
    (void)threads;
#endif
    p->stat = p->stat0; p->anim = p->anim0;
    p->iterNum = p->jacIterNum = p->funcEvalNum = 0;
    orcs::Sparse s;
    orcs::sparse_init(s, p);
    const int n = p->n, m = p->m;
    std::vector<double> fvec(m, 0.0), x(n), prev(n), dist(p->nobs, 0.0);
    double iavg = 0, imin = 0, imax = 0;
    if (p->opt.accept_only_better) {
        // measureErrors at the scene's own values (adjust_base.cpp:1080-1103)
        orcs::Getter G{p, {0, 0, 0.0, false}};
        for (int i = 0; i < p->nobs; ++i) orcs::obs_eval(G, i, fvec[2 * i], fvec[2 * i + 1], p->errorDist[i]);
        error_stats(p, iavg, imin, imax);
    }
    orc_initial_parameters(p, x.data());
    prev = x;
    int nfev = 0, njev = 0;
    const int info = orcs::sparse_lmder(s, x, fvec, dist, &nfev, &njev, max_outer);
    p->errorDist = dist;
    res->error_final = enorm(m, fvec.data());
    res->success = p->iterNum > 0;
    res->reason_num = info;
    res->reason_str = (info >= 0 && info < 9) ? k_reasons[info] : "";
    res->iteration_num = nfev;
    res->iteration_function_num = p->iterNum;
    res->iteration_jacobian_num = p->jacIterNum;
    res->iteration_attempt_num = njev;
    error_stats(p, res->error_avg, res->error_min, res->error_max);
    bool better = true;
    if (p->opt.accept_only_better) better = res->error_avg <= iavg;
    res->error_is_better = better;
    const std::vector<double> &fin = better ? x : prev;
    for (int i = 0; i < n; ++i) {
        const int ai = p->paramToAttr[i].first;
        if (x_out) x_out[i] = x[i];
        if (attr_out) attr_out[i] = boundInternalToExternal(fin[i], p->attr_min[ai], p->attr_max[ai], p->attr_off[ai], p->attr_scl[ai]);
    }
    if (dist_out) for (int i = 0; i < p->nobs; ++i) dist_out[i] = dist[i];
    return MMB_OK;
}

// ---- low-level hooks used to pin the restatement ------------------------
void orc_reproject(const double *cam_trs9, int cam_roo, const double *proj11,
                   const double *pt3, double *out2) {
    // proj11: focal_mm fbw_in fbh_in offx_in offy_in img_w img_h film_fit near far scale
    M4 cw = local_matrix(cam_trs9[0], cam_trs9[1], cam_trs9[2], cam_trs9[3], cam_trs9[4], cam_trs9[5],
                         cam_trs9[6], cam_trs9[7], cam_trs9[8], cam_roo);
    M4 pr = projection_matrix(proj11[0], proj11[1], proj11[2], proj11[3], proj11[4], proj11[5],
                              proj11[6], (int)proj11[7], proj11[8], proj11[9], proj11[10]);
    M4 b = m4_identity(); b.m[3] = pt3[0]; b.m[7] = pt3[1]; b.m[11] = pt3[2];
    reproject_normalised(cw, pr, b, out2[0], out2[1]);
}
void orc_projection_matrix(const double *proj11, double *out16) {
    M4 pr = projection_matrix(proj11[0], proj11[1], proj11[2], proj11[3], proj11[4], proj11[5],
                              proj11[6], (int)proj11[7], proj11[8], proj11[9], proj11[10]);
    std::memcpy(out16, pr.m, sizeof(pr.m));
}
void orc_local_matrix(const double *trs9, int roo, double *out16) {
    M4 mm = local_matrix(trs9[0], trs9[1], trs9[2], trs9[3], trs9[4], trs9[5], trs9[6], trs9[7], trs9[8], roo);
    std::memcpy(out16, mm.m, sizeof(mm.m));
}
// direction 0 = undistort, 1 = distort (applyModelDistort)
void orc_lens_apply(int model, const double *coeffs, const double *cam6, int direction,
                    int npts, const double *xy_in, double *xy_out) {
    LensEval L; L.model = model;
    for (int k = 0; k < 6; ++k) L.cam[k] = cam6[k];
    for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k) L.c[k] = coeffs[k];
    lens_prepare(L);
    for (int i = 0; i < npts; ++i) {
        if (direction == 0) lens_undistort(L, xy_in[2 * i], xy_in[2 * i + 1], xy_out[2 * i], xy_out[2 * i + 1]);
        else lens_distort(L, xy_in[2 * i], xy_in[2 * i + 1], xy_out[2 * i], xy_out[2 * i + 1]);
    }
}
double orc_bound_int_to_ext(double v, double mn, double mx, double off, double scl) { return boundInternalToExternal(v, mn, mx, off, scl); }
double orc_bound_ext_to_int(double v, double mn, double mx, double off, double scl) { return boundExternalToInternal(v, mn, mx, off, scl); }

// Generic lmder entry for the scipy cross-check (callback from Python).
int orc_lmder(lmder_fcn fcn, void *ud, int m, int n, double *x, double *fvec, double ftol,
              double xtol, double gtol, int maxfev, int mode, double factor, int *nfev, int *njev) {
    const int ld = std::max(m, n);
    std::vector<double> fjac((size_t)n * ld), qtf(n), wa1(n), wa2(n), wa3(n), wa4(m), diag(n, 1.0);
    std::vector<int> ipvt(n);
    return lmder(fcn, ud, m, n, x, fvec, fjac.data(), ld, ftol, xtol, gtol, maxfev, diag.data(), mode,
                 factor, nfev, njev, ipvt.data(), qtf.data(), wa1.data(), wa2.data(), wa3.data(), wa4.data());
}

// lmdif through the same entry (scipy cross-check): fcn fills fvec for iflag 1 and 2.
int orc_lmdif(lmder_fcn fcn, void *ud, int m, int n, double *x, double *fvec, double ftol,
              double xtol, double gtol, int maxfev, double epsfcn, int mode, double factor, int *nfev, int *njev) {
    const int ld = std::max(m, n);
    std::vector<double> fjac((size_t)n * ld), qtf(n), wa1(n), wa2(n), wa3(n), wa4(m), diag(n, 1.0);
    std::vector<int> ipvt(n);
    return lmder(fcn, ud, m, n, x, fvec, fjac.data(), ld, ftol, xtol, gtol, maxfev, diag.data(), mode,
                 factor, nfev, njev, ipvt.data(), qtf.data(), wa1.data(), wa2.data(), wa3.data(), wa4.data(),
                 nullptr, true, epsfcn);
}

}  // extern "C"
