// Per-frame solve mode (frameSolveMode = 1): batched independent LM solves.
//
// Reference: solve_v1 loops the frames and calls solveFrames() with a one-frame
// list each time (adjust_base.cpp:1430-1484), summing the results with
// CommandResult::add / divide (adjust_results.h:89-117).  When every solved
// attribute is animated the frames are independent problems, and here ONE
// kernel launch runs them all: one CTA per frame carries the whole MINPACK
// lmder loop (residuals, FD Jacobian, J^T J, damped Cholesky, lmpar, ratio test,
// accept/reject, termination) with its normal equations in shared memory and no
// host involvement at all.  Frames shard across GPUs with no collective.
#include <algorithm>
#include <chrono>
#include <cstring>

#include "lm.h"
#include "shard.h"
#include "reduce.cuh"

namespace mmb {

constexpr int kPfThreads = 64;   // 234 registers per thread: 64 threads keep four frames per SM in flight (128: two)

struct PfFrameResult {
    int info, nfev, njev, func_evals;
    double fnorm, err_avg, err_min, err_max, init_avg;
    int better, pad;
};

struct PfArgs {
    int na, F, kmax, maxfev, mode, accept_only_better;
    int nfev_per_jac;  // lmdif: na function calls per Jacobian count towards maxfev
    double ftol, xtol, gtol, factor;
    const int32_t *frame_obs_start;  // [F+1]
    const int32_t *frame_obs;        // [nobs] observation ids grouped by frame
    const int32_t *slot_attr;        // [na] solver attr of eliminated slot j
    const int8_t *mkr_slot_col;      // [M*na] Jacobian column of slot j for marker m, or -1
    double *x, *x2;                  // [n] parameter vectors (parameter order)
    double *fvec, *fvec2, *dist, *dist2, *jac;
    PfFrameResult *out;              // [F]
};

struct PfShared {
    double *H, *L;  // [na*na] each, dynamic shared memory
    double g[kMaxNA], diag[kMaxNA], acn[kMaxNA], p[kMaxNA], w[kMaxNA], xs[kMaxNA];
    double red[32];
    double jtj_part[2][27];      // per-warp partial sums of pf_jtj<6> (64 threads)
    double cvals[17];            // attribute values of the frame's camera (pf_variants)
    int crefs[17], cattr[17];
    double fnorm, fnorm1, xnorm, delta, par, gnorm, pnorm, jp2, gscaled, dxnorm, qn2;
    double parl, paru, fp;
    int iter, nfev, njev, info, fail, lmpar_iter, lmpar_done, accepted, done, func_evals;
};

// entity variants of frame f (per-frame entities only; static ones are prepared once)
__device__ void pf_variants(const DevScene &s, PfShared &sh, const double *x, int f, bool base_only) {
    // animated parentless cameras: the frame's 17 attribute values are read once (17 threads),
    // the variants are evaluated from shared memory (k_variants' block path, one frame wide)
    for (int c = 0; c < s.C; ++c) {
        if (!s.cam_blk[c]) continue;
        if (threadIdx.x < 17) {
            const int j = threadIdx.x, tfm = s.cam_tfm[c];
            const int32_t ref = j < 9 ? s.tfm_attrs[tfm * 9 + j] : s.cam_attrs[c * 8 + j - 9];
            AttrView b;
            b.s = &s; b.x = x; b.f = f; b.ovr_attr = -1; b.ovr_delta = 0.0;
            sh.crefs[j] = ref;
            sh.cattr[j] = x ? (ref >= 0 ? s.stat_attr[ref] : s.anim_attr[-ref - 1]) : -1;
            sh.cvals[j] = b.get(ref);
        }
        __syncthreads();
        const int nvar = base_only ? 1 : cam_nvar_of(s.cam_ndep[c], s.central);
        for (int v = threadIdx.x; v < nvar; v += blockDim.x)
            camera_variant_gathered(s, x, c, v, f, sh.cvals, 1, sh.crefs, sh.cattr);
        __syncthreads();
    }
    for (int j = threadIdx.x; j < s.cam_njobs; j += blockDim.x) {
        const int e = s.cam_job_ent[j];
        if (s.cam_blk[e] || s.cam_fs[e] == 1 || (base_only && s.cam_job_var[j] != 0)) continue;
        camera_variant_thread(s, x, s.cam_job_start[j] + f);
    }
    for (int j = threadIdx.x; j < s.bnd_njobs; j += blockDim.x) {
        const int e = s.bnd_job_ent[j];
        if (s.bnd_fs[e] == 1 || (base_only && s.bnd_job_var[j] != 0)) continue;
        bundle_variant_thread(s, x, s.bnd_job_start[j] + f);
    }
    for (int j = threadIdx.x; j < s.lens_njobs; j += blockDim.x) {
        const int e = s.lens_job_ent[j];
        if (s.lens_fs[e] == 1 || (base_only && s.lens_job_var[j] != 0)) continue;
        lens_variant_thread(s, x, s.lens_job_start[j] + f);
    }
    __syncthreads();
}

// residuals of frame f at x -> returns sum of squares (all threads)
__device__ double pf_residuals(const DevScene &s, const PfArgs &a, PfShared &sh, const double *x, int f,
                               double *fvec, double *dist) {
    pf_variants(s, sh, x, f, true);
    const int q0 = a.frame_obs_start[f], q1 = a.frame_obs_start[f + 1];
    double acc = 0.0;
    for (int q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        const int i = a.frame_obs[q];
        residual_thread(s, i, fvec, dist);
        const double ex = fvec[2 * (size_t)i], ey = fvec[2 * (size_t)i + 1];
        acc += ex * ex + ey * ey;
    }
    return block_sum(acc, sh.red);
}

// compute_error_stats over the frame's observations (adjust_base.cpp:346-372)
__device__ void pf_error_stats(const PfArgs &a, PfShared &sh, int f, const double *dist, double &avg, double &mn,
                               double &mx) {
    const int q0 = a.frame_obs_start[f], q1 = a.frame_obs_start[f + 1];
    double s = 0.0, lo = 1.7976931348623157e308, hi = -0.0;
    for (int q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        const double e = dist[a.frame_obs[q]];
        if (!isfinite(e)) continue;
        s += e; lo = fmin(lo, e); hi = fmax(hi, e);
    }
    s = block_sum(s, sh.red);
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { sh.red[threadIdx.x >> 5] = lo; sh.red[8 + (threadIdx.x >> 5)] = hi; }
    __syncthreads();
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) { lo = fmin(lo, sh.red[w2]); hi = fmax(hi, sh.red[8 + w2]); }
    __syncthreads();
    avg = s / (double)(q1 - q0); mn = lo; mx = hi;
}

// Cholesky of H + par*D^2 into L, p = solve; optionally w = L^-1 (D^2 p / dxnorm).
// Single thread: the blocks are tiny (na = 6 for a camera).
template <int NA> __device__ void pf_factor_solve_reg(PfShared &sh, double par);
__device__ void pf_factor_solve(PfShared &sh, int na, double par) {
    if (na == 6) { pf_factor_solve_reg<6>(sh, par); return; }
    if (na == 3) { pf_factor_solve_reg<3>(sh, par); return; }
    sh.fail = 0;
    for (int r = 0; r < na; ++r)
        for (int c = 0; c <= r; ++c) sh.L[r * na + c] = sh.H[r * na + c] + (r == c ? par * sh.diag[r] * sh.diag[r] : 0.0);
    for (int j = 0; j < na; ++j) {
        double pv = sh.L[j * na + j];
        for (int k = 0; k < j; ++k) pv -= sh.L[j * na + k] * sh.L[j * na + k];
        if (!(pv > 0.0)) { sh.fail = 1; pv = 1.0; }
        pv = sqrt(pv);
        sh.L[j * na + j] = pv;
        for (int r = j + 1; r < na; ++r) {
            double v = sh.L[r * na + j];
            for (int k = 0; k < j; ++k) v -= sh.L[r * na + k] * sh.L[j * na + k];
            sh.L[r * na + j] = v / pv;
        }
    }
    // L y = g ; L^T p = y
    for (int r = 0; r < na; ++r) {
        double v = sh.g[r];
        for (int c = 0; c < r; ++c) v -= sh.L[r * na + c] * sh.p[c];
        sh.p[r] = v / sh.L[r * na + r];
    }
    for (int r = na - 1; r >= 0; --r) {
        double v = sh.p[r];
        for (int c = r + 1; c < na; ++c) v -= sh.L[c * na + r] * sh.p[c];
        sh.p[r] = v / sh.L[r * na + r];
    }
    double dx = 0.0;
    for (int r = 0; r < na; ++r) { const double t = sh.diag[r] * sh.p[r]; dx += t * t; }
    sh.dxnorm = sh.fail ? INFINITY : sqrt(dx);
    double qn = 0.0;
    for (int r = 0; r < na; ++r) {
        double v = sh.diag[r] * sh.diag[r] * sh.p[r] / sh.dxnorm;
        for (int c = 0; c < r; ++c) v -= sh.L[r * na + c] * sh.w[c];
        sh.w[r] = v / sh.L[r * na + r];
        qn += sh.w[r] * sh.w[r];
    }
    sh.qn2 = qn;
}

// The same with everything in registers (NA known at compile time): the shared-memory version
// is a chain of dependent shared loads, divisions and square roots by one thread (34 us per LM
// iteration at na = 6, a third of the per-frame solve); pivots are inverted once per column.
template <int NA>
__device__ void pf_factor_solve_reg(PfShared &sh, double par) {
    double L[NA][NA], g[NA], dg[NA], pv_inv[NA], p[NA], w[NA];
#pragma unroll
    for (int r = 0; r < NA; ++r) {
        g[r] = sh.g[r]; dg[r] = sh.diag[r];
#pragma unroll
        for (int c = 0; c <= r; ++c) L[r][c] = sh.H[r * NA + c] + (r == c ? par * dg[r] * dg[r] : 0.0);
    }
    int fail = 0;
#pragma unroll
    for (int j = 0; j < NA; ++j) {
        double pv = L[j][j];
#pragma unroll
        for (int k = 0; k < j; ++k) pv -= L[j][k] * L[j][k];
        if (!(pv > 0.0)) { fail = 1; pv = 1.0; }
        pv = sqrt(pv);
        L[j][j] = pv;
        pv_inv[j] = 1.0 / pv;
#pragma unroll
        for (int r = j + 1; r < NA; ++r) {
            double v = L[r][j];
#pragma unroll
            for (int k = 0; k < j; ++k) v -= L[r][k] * L[j][k];
            L[r][j] = v * pv_inv[j];
        }
    }
#pragma unroll
    for (int r = 0; r < NA; ++r) {
        double v = g[r];
#pragma unroll
        for (int c = 0; c < r; ++c) v -= L[r][c] * p[c];
        p[r] = v * pv_inv[r];
    }
#pragma unroll
    for (int r = NA - 1; r >= 0; --r) {
        double v = p[r];
#pragma unroll
        for (int c = r + 1; c < NA; ++c) v -= L[c][r] * p[c];
        p[r] = v * pv_inv[r];
    }
    double dx = 0.0;
#pragma unroll
    for (int r = 0; r < NA; ++r) { const double t = dg[r] * p[r]; dx += t * t; }
    const double dxnorm = fail ? INFINITY : sqrt(dx);
    const double dxi = 1.0 / dxnorm;
    double qn = 0.0;
#pragma unroll
    for (int r = 0; r < NA; ++r) {
        double v = dg[r] * dg[r] * p[r] * dxi;
#pragma unroll
        for (int c = 0; c < r; ++c) v -= L[r][c] * w[c];
        w[r] = v * pv_inv[r];
        qn += w[r] * w[r];
    }
#pragma unroll
    for (int r = 0; r < NA; ++r) { sh.p[r] = p[r]; sh.w[r] = w[r]; }
    sh.fail = fail; sh.dxnorm = dxnorm; sh.qn2 = qn;
}

// MINPACK lmpar on the normal equations (thread 0)
__device__ void pf_lmpar(PfShared &sh, int na) {
    const double p1 = 0.1, p001 = 0.001, dwarf = 2.2250738585072014e-308;
    const double delta = sh.delta;
    pf_factor_solve(sh, na, 0.0);
    bool fail = sh.fail != 0;
    double fp = sh.dxnorm - delta;
    if (!fail && fp <= p1 * delta) { sh.par = 0.0; return; }
    double parl = 0.0;
    if (!fail && sh.qn2 > 0.0) parl = ((fp / delta) / sh.qn2);
    double paru = sh.gscaled / delta;
    if (paru == 0.0) paru = dwarf / fmin(delta, p1);
    double par = fmin(fmax(sh.par, parl), paru);
    if (par == 0.0) par = fail ? 0.0 : sh.gscaled / sh.dxnorm;
    for (int iter = 1;; ++iter) {
        if (par == 0.0) par = fmax(dwarf, p001 * paru);
        pf_factor_solve(sh, na, par);
        if (sh.fail) {  // numerically indefinite: raise the damping
            parl = fmax(parl, par);
            par = fmax(par * 10.0, p001 * paru);
            if (iter > 30) break;
            continue;
        }
        const double temp = fp;
        fp = sh.dxnorm - delta;
        if (fabs(fp) <= p1 * delta || (parl == 0.0 && fp <= temp && temp < 0.0) || iter >= 10) break;
        const double parc = ((fp / delta) / sh.qn2);
        if (fp > 0.0) parl = fmax(parl, par);
        if (fp < 0.0) paru = fmin(paru, par);
        par = fmax(parl, par + parc);
    }
    sh.par = par;
}

// J^T J (lower) and J^T f with the OBSERVATIONS spread over the threads: every thread forms the
// NA (NA + 1) / 2 + NA products of its observations, the sums go through warp shuffles and one
// shared-memory stage.  (One thread per product walking all observations was 21 us per LM
// iteration at 50 observations: 50 dependent strided loads per thread.)
template <int NA>
__device__ void pf_jtj(const DevScene &s, const PfArgs &a, PfShared &sh, int q0, int q1, size_t m) {
    constexpr int NP = NA * (NA + 1) / 2 + NA;
    static_assert(NP <= 27, "PfShared::jtj_part");
    double acc[NP];
#pragma unroll
    for (int t = 0; t < NP; ++t) acc[t] = 0.0;
    for (int q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
        const int i = a.frame_obs[q];
        const int mk = s.dense_grid ? i / s.F : s.err_marker[i];
        double jx[NA], jy[NA];
#pragma unroll
        for (int r = 0; r < NA; ++r) {
            const int cr = a.mkr_slot_col[(size_t)mk * NA + r];
            jx[r] = cr < 0 ? 0.0 : a.jac[(size_t)cr * m + 2 * (size_t)i];
            jy[r] = cr < 0 ? 0.0 : a.jac[(size_t)cr * m + 2 * (size_t)i + 1];
        }
        const double fx = a.fvec[2 * (size_t)i], fy = a.fvec[2 * (size_t)i + 1];
        int t = 0;
#pragma unroll
        for (int r = 0; r < NA; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c, ++t) acc[t] += jx[r] * jx[c] + jy[r] * jy[c];
#pragma unroll
        for (int r = 0; r < NA; ++r, ++t) acc[t] += jx[r] * fx + jy[r] * fy;
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int t = 0; t < NP; ++t) {
        const double v = warp_sum(acc[t]);
        if (lane == 0) sh.jtj_part[w][t] = v;
    }
    __syncthreads();
    if (threadIdx.x < NP) {
        double v = 0.0;
        for (int ww = 0; ww < (int)(blockDim.x >> 5); ++ww) v += sh.jtj_part[ww][threadIdx.x];
        int t = 0;
#pragma unroll
        for (int r = 0; r < NA; ++r)
#pragma unroll
            for (int c = 0; c <= r; ++c, ++t)
                if (t == (int)threadIdx.x) sh.H[r * NA + c] = v;
#pragma unroll
        for (int r = 0; r < NA; ++r, ++t)
            if (t == (int)threadIdx.x) sh.g[r] = v;
    }
}

// MINB = 1: no register cap (255 registers, 64 threads: four frames per SM in flight), for batches
// below one wave; MINB = 8: 128 registers, launched with 32 threads per frame: sixteen frames per
// SM -- a frame's solve is a latency chain, so frames in flight are throughput (24 000 frames:
// 33.1 ms uncapped, 30.0 ms capped at 64 threads, 26.9 ms capped at 32 threads)
template <int MINB>
__global__ void __launch_bounds__(kPfThreads, MINB) k_perframe_lm(DevScene s, PfArgs a, int frame0) {
    __shared__ PfShared sh;
    extern __shared__ double pf_dyn[];
    const int f = frame0 + blockIdx.x;
    if (f >= a.F) return;
    const int na = a.na, tid = threadIdx.x;
    if (tid == 0) { sh.H = pf_dyn; sh.L = pf_dyn + (size_t)na * na; }
    __syncthreads();
    const int q0 = a.frame_obs_start[f], q1 = a.frame_obs_start[f + 1];
    const size_t m = 2 * (size_t)s.nobs;
    PfFrameResult res;
    memset(&res, 0, sizeof(res));
    // initial measureErrors at the scene's own values (acceptOnlyBetter, adjust_base.cpp:1080)
    double init_avg = 0.0, t1, t2;
    if (a.accept_only_better) {
        pf_residuals(s, a, sh, nullptr, f, a.fvec, a.dist2);
        pf_error_stats(a, sh, f, a.dist2, init_avg, t1, t2);
    }
    for (int j = tid; j < na; j += blockDim.x) {
        const int attr = a.slot_attr[j];
        sh.xs[j] = a.x[s.attr_param_base[attr] + f];
        sh.diag[j] = 1.0;
    }
    __syncthreads();
    const int nrows = 2 * (q1 - q0);
    double ss = pf_residuals(s, a, sh, a.x, f, a.fvec, a.dist);
    if (tid == 0) {
        sh.fnorm = sqrt(ss); sh.par = 0.0; sh.iter = 1; sh.nfev = 1; sh.njev = 0; sh.info = 0; sh.done = 0;
        sh.func_evals = 1; sh.delta = 0.0; sh.xnorm = 0.0; sh.gnorm = 0.0;
        if (na <= 0 || nrows < na || a.ftol < 0.0 || a.xtol < 0.0 || a.gtol < 0.0 || a.maxfev <= 0 || a.factor <= 0.0) sh.done = 1;
    }
    __syncthreads();
    const bool refused = (q1 == q0) || nrows < na;  // adjust_base.cpp:844-882 pre-checks
    while (!sh.done && !refused) {
        // ---- Jacobian at x (solveFunc iflag = 2) -------------------------------
        pf_variants(s, sh, a.x, f, false);
        for (int q = q0 + tid; q < q1; q += blockDim.x)
            jacobian_thread(s, a.frame_obs[q], nullptr, nullptr, a.jac, nullptr, a.dist2);
        __syncthreads();
        // J^T J (lower) and J^T f: deterministic order
        if (na == 6 && blockDim.x <= 64) pf_jtj<6>(s, a, sh, q0, q1, m);
        else
        for (int t = tid; t < na * (na + 1) / 2 + na; t += blockDim.x) {
            int r, c;
            const bool grad = t >= na * (na + 1) / 2;
            if (grad) { r = t - na * (na + 1) / 2; c = 0; }
            else { r = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5); while ((r + 1) * (r + 2) / 2 <= t) ++r; while (r * (r + 1) / 2 > t) --r; c = t - r * (r + 1) / 2; }
            double acc = 0.0;
            for (int q = q0; q < q1; ++q) {
                const int i = a.frame_obs[q];
                const int mk = s.dense_grid ? i / s.F : s.err_marker[i];
                const int cr = a.mkr_slot_col[(size_t)mk * na + r];
                if (cr < 0) continue;
                const double jx = a.jac[(size_t)cr * m + 2 * (size_t)i], jy = a.jac[(size_t)cr * m + 2 * (size_t)i + 1];
                if (grad) acc += jx * a.fvec[2 * (size_t)i] + jy * a.fvec[2 * (size_t)i + 1];
                else {
                    const int cc = a.mkr_slot_col[(size_t)mk * na + c];
                    if (cc < 0) continue;
                    acc += jx * a.jac[(size_t)cc * m + 2 * (size_t)i] + jy * a.jac[(size_t)cc * m + 2 * (size_t)i + 1];
                }
            }
            if (grad) sh.g[r] = acc; else sh.H[r * na + c] = acc;
        }
        __syncthreads();
        if (tid == 0) {
            sh.njev += 1;
            sh.nfev += a.nfev_per_jac;
            const bool first = sh.iter == 1;
            double xs2 = 0.0, gmax = 0.0, gs = 0.0;
            for (int j = 0; j < na; ++j) {
                const double ac = sqrt(fmax(sh.H[j * na + j], 0.0));
                sh.acn[j] = ac;
                double dg = sh.diag[j];
                if (a.mode != 2 && first) dg = (ac == 0.0) ? 1.0 : ac;
                if (first) { const double t = dg * sh.xs[j]; xs2 += t * t; }
                if (sh.fnorm != 0.0 && ac != 0.0) gmax = fmax(gmax, fabs(sh.g[j] / sh.fnorm) / ac);
                if (a.mode != 2) dg = fmax(dg, ac);
                sh.diag[j] = dg;
                const double t2 = sh.g[j] / dg;
                gs += t2 * t2;
            }
            if (first) {
                sh.xnorm = sqrt(xs2);
                sh.delta = a.factor * sh.xnorm;
                if (sh.delta == 0.0) sh.delta = a.factor;
            }
            sh.gnorm = gmax; sh.gscaled = sqrt(gs);
            if (gmax <= a.gtol) { sh.info = 4; sh.done = 1; }
        }
        __syncthreads();
        // ---- inner loop: lmpar + trial until the step is accepted -------------
        while (!sh.done) {
            if (tid == 0) {
                pf_lmpar(sh, na);
                double pn = 0.0;
                for (int j = 0; j < na; ++j) {
                    sh.p[j] = -sh.p[j];
                    const double t = sh.diag[j] * sh.p[j];
                    pn += t * t;
                }
                sh.pnorm = sqrt(pn);
                if (sh.iter == 1) sh.delta = fmin(sh.delta, sh.pnorm);
            }
            __syncthreads();
            for (int j = tid; j < na; j += blockDim.x)
                a.x2[s.attr_param_base[a.slot_attr[j]] + f] = sh.xs[j] + sh.p[j];
            __syncthreads();
            const double ss1 = pf_residuals(s, a, sh, a.x2, f, a.fvec2, a.dist2);
            // ||J p||^2 from the block Jacobian
            double jp = 0.0;
            for (int q = q0 + tid; q < q1; q += blockDim.x) {
                const int i = a.frame_obs[q];
                const int mk = s.dense_grid ? i / s.F : s.err_marker[i];
                double sx = 0.0, sy = 0.0;
                for (int j = 0; j < na; ++j) {
                    const int cj = a.mkr_slot_col[(size_t)mk * na + j];
                    if (cj < 0) continue;
                    sx += a.jac[(size_t)cj * m + 2 * (size_t)i] * sh.p[j];
                    sy += a.jac[(size_t)cj * m + 2 * (size_t)i + 1] * sh.p[j];
                }
                jp += sx * sx + sy * sy;
            }
            jp = block_sum(jp, sh.red);
            if (tid == 0) {
                const double p1 = 0.1, p5 = 0.5, p25 = 0.25, p75 = 0.75, p0001 = 1e-4, epsmch = 2.220446049250313e-16;
                const double fnorm = sh.fnorm, fnorm1 = sqrt(ss1), pnorm = sh.pnorm;
                sh.nfev += 1; sh.func_evals += 1;
                double actred = -1.0;
                if (p1 * fnorm1 < fnorm) { const double t = fnorm1 / fnorm; actred = 1.0 - t * t; }
                const double temp1 = sqrt(jp) / fnorm, temp2 = (sqrt(sh.par) * pnorm) / fnorm;
                const double prered = temp1 * temp1 + temp2 * temp2 / p5;
                const double dirder = -(temp1 * temp1 + temp2 * temp2);
                double ratio = 0.0;
                if (prered != 0.0) ratio = actred / prered;
                if (ratio <= p25) {
                    double temp = actred >= 0.0 ? p5 : p5 * dirder / (dirder + p5 * actred);
                    if (p1 * fnorm1 >= fnorm || temp < p1) temp = p1;
                    sh.delta = temp * fmin(sh.delta, pnorm / p1);
                    sh.par /= temp;
                } else if (sh.par == 0.0 || ratio >= p75) {
                    sh.delta = pnorm / p5;
                    sh.par = p5 * sh.par;
                }
                sh.accepted = 0;
                if (ratio >= p0001) {
                    sh.accepted = 1;
                    double xs2 = 0.0;
                    for (int j = 0; j < na; ++j) { sh.xs[j] += sh.p[j]; const double t = sh.diag[j] * sh.xs[j]; xs2 += t * t; }
                    sh.xnorm = sqrt(xs2);
                    sh.fnorm = fnorm1;
                    sh.iter += 1;
                }
                int info = 0;
                const bool small_red = fabs(actred) <= a.ftol && prered <= a.ftol && p5 * ratio <= 1.0;
                if (small_red) info = 1;
                if (sh.delta <= a.xtol * sh.xnorm) info = 2;
                if (small_red && info == 2) info = 3;
                if (info == 0) {
                    if (sh.nfev >= a.maxfev) info = 5;
                    if (fabs(actred) <= epsmch && prered <= epsmch && p5 * ratio <= 1.0) info = 6;
                    if (sh.delta <= epsmch * sh.xnorm) info = 7;
                    if (sh.gnorm <= epsmch) info = 8;
                }
                if (info != 0) { sh.info = info; sh.done = 1; }
            }
            __syncthreads();
            if (sh.accepted) {
                for (int j = tid; j < na; j += blockDim.x) a.x[s.attr_param_base[a.slot_attr[j]] + f] = sh.xs[j];
                for (int q = q0 + tid; q < q1; q += blockDim.x) {
                    const int i = a.frame_obs[q];
                    a.fvec[2 * (size_t)i] = a.fvec2[2 * (size_t)i];
                    a.fvec[2 * (size_t)i + 1] = a.fvec2[2 * (size_t)i + 1];
                }
            }
            __syncthreads();
            if (sh.accepted) break;
        }
    }
    // results: enorm(fvec), error stats of the LAST measured distances
    double ssf = 0.0;
    for (int q = q0 + tid; q < q1; q += blockDim.x) {
        const int i = a.frame_obs[q];
        ssf += a.fvec[2 * (size_t)i] * a.fvec[2 * (size_t)i] + a.fvec[2 * (size_t)i + 1] * a.fvec[2 * (size_t)i + 1];
    }
    ssf = block_sum(ssf, sh.red);
    double eavg, emin, emax;
    pf_error_stats(a, sh, f, a.dist2, eavg, emin, emax);
    if (tid == 0) {
        res.info = refused ? -100 : sh.info;
        res.nfev = sh.nfev; res.njev = sh.njev; res.func_evals = sh.func_evals;
        res.fnorm = sqrt(ssf); res.err_avg = eavg; res.err_min = emin; res.err_max = emax;
        res.init_avg = init_avg;
        res.better = a.accept_only_better ? (eavg <= init_avg) : 1;
        a.out[f] = res;
    }
}

// ---- host side ------------------------------------------------------------------------------
mmb_status lm_solve_per_frame(DeviceProblem *p, mmb_result *res, double *x_out, double *attr_out, double *dist_out,
                              mmb_result *per_frame_out) {
    std::memset(res, 0, sizeof(*res));
    const HostProblem &hp = p->hp;
    if (hp.nobs == 0 || hp.n == 0) { set_last_error("no used markers or attributes"); return MMB_ERR_NOT_ENOUGH_OBJECTS; }
    if (hp.ns != 0) {
        set_last_error("per-frame solve with static attributes carries state from frame to frame; only all-animated attribute sets run batched");
        return MMB_ERR_UNSUPPORTED;
    }
    if (hp.na > kMaxNA) { set_last_error("more than 64 animated attributes"); return MMB_ERR_UNSUPPORTED; }
    MMB_CUDA_CHECK(cudaSetDevice(p->device));
    mmb_status st;
    if ((st = ensure_jacobian_buffers(p)) != MMB_OK) return st;
    const auto t0 = std::chrono::steady_clock::now();
    DeviceArena A;
    A.bind(p->stream);
    // observations grouped by frame
    std::vector<int32_t> start(hp.F + 1, 0), ids(hp.nobs);
    for (int i = 0; i < hp.nobs; ++i) start[hp.err_frame[i] + 1]++;
    for (int f = 0; f < hp.F; ++f) start[f + 1] += start[f];
    { std::vector<int32_t> fill(start.begin(), start.end() - 1);
      for (int i = 0; i < hp.nobs; ++i) ids[fill[hp.err_frame[i]]++] = i; }
    std::vector<int32_t> slot_attr(std::max(hp.na, 1), 0);
    for (int a = 0; a < hp.A; ++a) if (hp.attr_anim[a]) slot_attr[hp.attr_slot[a]] = a;
    std::vector<int8_t> slot_col((size_t)hp.M * std::max(hp.na, 1), -1);
    for (int mk = 0; mk < hp.M; ++mk)
        for (int c = 0; c < hp.mkr_k[mk]; ++c) {
            const int a = hp.plan[(size_t)mk * hp.kmax + c].attr;
            slot_col[(size_t)mk * hp.na + hp.attr_slot[a]] = (int8_t)c;
        }
    std::vector<double> x0(hp.n);
    for (int i = 0; i < hp.n; ++i) {
        const int a = hp.par_attr[i], f = hp.par_frame[i];
        const int32_t ref = hp.attr_ref[a];
        const double v = ref >= 0 ? hp.stat0[ref] : hp.anim0[(size_t)(-ref - 1) * hp.F + f];
        x0[i] = bound_external_to_internal(v, hp.attr_min[a], hp.attr_max[a], hp.attr_off[a], hp.attr_scl[a]);
    }
    PfArgs a{};
    a.na = hp.na; a.F = hp.F; a.kmax = p->ds.kmax; a.maxfev = hp.opt.iter_max;
    a.nfev_per_jac = hp.opt.solver_type == 1 ? hp.na : 0;
    a.mode = hp.opt.auto_param_scale == 1 ? 1 : 2; a.accept_only_better = hp.opt.accept_only_better;
    a.ftol = hp.opt.eps1; a.xtol = hp.opt.eps2; a.gtol = hp.opt.eps3; a.factor = hp.opt.tau * 100.0;
    int32_t *d_i = nullptr; int8_t *d_8 = nullptr; double *d_x2 = nullptr, *d_f2 = nullptr, *d_d2 = nullptr;
    MMB_CUDA_CHECK(A.upload(&d_i, start)); a.frame_obs_start = d_i;
    MMB_CUDA_CHECK(A.upload(&d_i, ids)); a.frame_obs = d_i;
    MMB_CUDA_CHECK(A.upload(&d_i, slot_attr)); a.slot_attr = d_i;
    MMB_CUDA_CHECK(A.upload(&d_8, slot_col)); a.mkr_slot_col = d_8;
    MMB_CUDA_CHECK(A.alloc(&d_x2, (size_t)hp.n));
    MMB_CUDA_CHECK(A.alloc(&d_f2, (size_t)hp.m));
    MMB_CUDA_CHECK(A.alloc(&d_d2, (size_t)hp.nobs));
    MMB_CUDA_CHECK(A.alloc(&a.out, (size_t)hp.F));
    MMB_CUDA_CHECK(cudaMemcpyAsync(p->d_x, x0.data(), sizeof(double) * hp.n, cudaMemcpyHostToDevice, p->stream));
    MMB_CUDA_CHECK(cudaMemcpyAsync(d_x2, x0.data(), sizeof(double) * hp.n, cudaMemcpyHostToDevice, p->stream));
    a.x = p->d_x; a.x2 = d_x2; a.fvec = p->d_fvec; a.fvec2 = d_f2; a.dist = p->d_dist; a.dist2 = d_d2; a.jac = p->d_jac;
    // static entities never depend on solved attributes here: evaluate their base once
    if ((st = launch_variants(p, nullptr, false)) != MMB_OK) return st;
    const size_t dyn = 2 * sizeof(double) * (size_t)hp.na * hp.na;
    static int pf_minb = -1;
    if (pf_minb < 0) { const char *e = getenv("MMB_PF_MINB"); pf_minb = e ? atoi(e) : 8; }
    // few frames (less than one wave of the capped instance): the uncapped instance is faster per frame
    const bool dense = pf_minb >= 8 && hp.F > 148 * 4;
    static int pf_threads = -1;   // threads per frame (a multiple of 32, <= kPfThreads)
    if (pf_threads < 0) { const char *e = getenv("MMB_PF_THREADS"); pf_threads = e ? atoi(e) : 32; }
    const int threads = (dense && (pf_threads == 32 || pf_threads == 64)) ? pf_threads : kPfThreads;
    if (dense) {
        MMB_CUDA_CHECK(cudaFuncSetAttribute(k_perframe_lm<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        k_perframe_lm<8><<<hp.F, threads, dyn, p->stream>>>(p->ds, a, 0);
    } else {
        MMB_CUDA_CHECK(cudaFuncSetAttribute(k_perframe_lm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        k_perframe_lm<1><<<hp.F, kPfThreads, dyn, p->stream>>>(p->ds, a, 0);
    }
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    std::vector<PfFrameResult> fr(hp.F);
    std::vector<double> x(hp.n);
    MMB_CUDA_CHECK(cudaMemcpyAsync(fr.data(), a.out, sizeof(PfFrameResult) * hp.F, cudaMemcpyDeviceToHost, p->stream));
    MMB_CUDA_CHECK(cudaMemcpyAsync(x.data(), p->d_x, sizeof(double) * hp.n, cudaMemcpyDeviceToHost, p->stream));
    if (dist_out) MMB_CUDA_CHECK(cudaMemcpyAsync(dist_out, d_d2, sizeof(double) * hp.nobs, cudaMemcpyDeviceToHost, p->stream));
    MMB_CUDA_CHECK(cudaStreamSynchronize(p->stream));
    // per-frame results, then CommandResult::add / divide (adjust_results.h:89-117)
    std::vector<mmb_result> per(hp.F);
    for (int f = 0; f < hp.F; ++f) {
        const PfFrameResult &r = fr[f];
        mmb_result &o = per[f];
        std::memset(&o, 0, sizeof(o));
        o.success = (r.info != -100 && r.func_evals > 0) ? 1 : 0;
        o.reason_num = r.info; o.error_final = r.fnorm; o.error_avg = r.err_avg;
        o.error_min = r.err_min; o.error_max = r.err_max; o.iteration_num = r.nfev;
        o.iteration_function_num = r.func_evals;
        o.iteration_jacobian_num = r.njev * hp.na * (hp.opt.auto_diff_type == 1 ? 2 : 1);
        o.iteration_attempt_num = r.njev; o.error_is_better = r.better;
    }
    aggregate_results(per.data(), hp.F, res);
    if (per_frame_out) std::memcpy(per_frame_out, per.data(), sizeof(mmb_result) * hp.F);
    for (int i = 0; i < hp.n; ++i) {
        const int at = hp.par_attr[i], f = hp.par_frame[i];
        if (x_out) x_out[i] = x[i];
        if (attr_out) {
            const bool better = fr[f < 0 ? 0 : f].better != 0;
            attr_out[i] = bound_internal_to_external(better ? x[i] : x0[i], hp.attr_min[at], hp.attr_max[at], hp.attr_off[at], hp.attr_scl[at]);
        }
    }
    res->timer_solve_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return MMB_OK;
}

}  // namespace mmb
