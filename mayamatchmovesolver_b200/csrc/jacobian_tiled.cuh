// Register-tiled residual + FD-Jacobian kernel for the common scene shape:
// one camera, full visibility (observation i = marker*F + frame), forward
// differences, no lens, static bundles, and every Jacobian column perturbs
// either the camera or the bundle (never both).  Same mathematics as
// jacobian_thread (eval_bodies.cuh); what changes is the schedule:
//   * a thread owns one frame and MPC consecutive markers, so the camera rows of
//     that frame (base + one variant at a time) live in registers and each
//     variant is fetched once per MPC observations instead of once per
//     observation (L2 -> SM traffic / MPC),
//   * the MPC bundles of the CTA (base + perturbed positions + 1/delta) are
//     staged once in shared memory and read as broadcasts,
//   * 1/delta of every perturbed parameter rides inside the variant records, so
//     the inner loops contain no dependent index chains,
//   * marker loads and residual / Jacobian stores stay fully coalesced
//     (consecutive threads = consecutive frames = consecutive observations).
// HBM traffic is the algorithmic 48 + 16k bytes per observation.
#pragma once
#include "eval_bodies.cuh"

namespace mmb {

constexpr int kTileFrames = 128;  // threads per CTA (frames)
constexpr int kTileMaxVC = 16;    // camera deps supported
constexpr int kTileMaxVB = 7;     // bundle deps supported

struct TiledPlan {
    int cam;                     // the camera
    int VC, VB;                  // camera deps / max bundle deps
    int uniform;                 // every marker has the same column layout
    int8_t cam_col_u[kTileMaxVC];  // uniform layout: column of camera dep v
    int8_t bnd_col_u[kTileMaxVB + 1];
    long long cam_coloff_u[kTileMaxVC];      // col * nobs (double2 units), uniform layout
    long long bnd_coloff_u[kTileMaxVB + 1];
    const int32_t *cam_dep_attr; // [VC] solver attr of camera dep v
    const int8_t *mkr_cam_col;   // [M*VC] Jacobian column of camera dep v for marker m (-1: none)
    const int8_t *mkr_bnd_col;   // [M*VB] Jacobian column of bundle dep v
    const int32_t *mkr_bnd_attr; // [M*VB] solver attr of bundle dep v (-1: none)
    int last_attr_cam_var;       // 0-based camera dep index of attr A-1 (the variant the
                                 // reference's errorDistanceList ends on), or -1
};

#ifdef __CUDACC__
// 1/x to (nearly) full precision without the slow-path branch of the IEEE
// division: hardware seed (20 bits) + two Newton steps.
__device__ __forceinline__ double rcp_nr(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
    return r;
}
// a / b given r ~ 1/b, with one residual correction (error <= 1 ulp)
__device__ __forceinline__ double div_nr(double a, double b, double r) {
    const double q = a * r;
    return fma(fma(-b, q, a), r, q);
}

// |marker - projected point| for one (camera rows, bundle position) pair
__device__ __forceinline__ void eval_pair(const double *cm, double b0, double b1, double b2, double mkx,
                                          double mky, double &dx, double &dy) {
    const double sx = cm[0] * b0 + cm[1] * b1 + cm[2] * b2 + cm[3];
    const double sy = cm[4] * b0 + cm[5] * b1 + cm[6] * b2 + cm[7];
    const double sw = cm[8] * b0 + cm[9] * b1 + cm[10] * b2 + cm[11];
    const double r = rcp_nr(sw);
    const double px = div_nr(sx, sw, r) * 0.5;  // reprojection.rs:55-62
    const double py = div_nr(sy, sw, r) * 0.5;
    dx = fabs(mkx - px);
    dy = fabs(mky - py);
}

template <int MPC, int MINB>
__global__ void __launch_bounds__(kTileFrames, MINB) k_jacobian_tiled(DevScene s, TiledPlan tp, double *__restrict__ fvec,
                                                                   double *__restrict__ dist, double *__restrict__ jac,
                                                                   int32_t *__restrict__ cols,
                                                                   double *__restrict__ dist_last, int m_begin) {
    // bundle records of the CTA's markers: [marker j][variant 0..VB][x y z 1/delta]
    __shared__ double sb[MPC][kTileMaxVB + 1][kBndRec];
    const int f = blockIdx.x * kTileFrames + threadIdx.x;
    const int m0 = m_begin + blockIdx.y * MPC;
    const int nrec = (tp.VB + 1) * kBndRec;
    for (int t = threadIdx.x; t < MPC * nrec; t += kTileFrames) {
        const int j = t / nrec, r = t - j * nrec;
        const int m = m0 + j;
        double v = 0.0;
        if (m < s.M) {
            const int b = s.mkr_bnd[m];
            const int nv = (1 + s.bnd_ndep[b]) * kBndRec;
            if (r < nv) v = s.bnd_tab[s.bnd_off[b] + r];  // static bundle: frame stride 1
        }
        sb[j][r / kBndRec][r % kBndRec] = v;
    }
    __syncthreads();
    if (f >= s.F) return;
    const size_t mtot = 2 * (size_t)s.nobs;
    const double imw = s.image_width;
    const int cfs = s.cam_fs[tp.cam];
    const double *ctab = s.cam_tab + s.cam_off[tp.cam] + (cfs == 1 ? 0 : f);
    double cm[14];
#pragma unroll
    for (int j = 0; j < 14; ++j) cm[j] = ctab[(size_t)j * cfs];

    // ---- base evaluation of the MPC markers of this frame ------------------
    double mkx[MPC], mky[MPC], wim[MPC], ex[MPC], ey[MPC];
    const size_t i0 = (size_t)m0 * s.F + f;  // observation of marker m0; next marker: + F
#pragma unroll
    for (int j = 0; j < MPC; ++j) {
        mkx[j] = mky[j] = wim[j] = 0.0;
        if (m0 + j < s.M) {
            const size_t i = i0 + (size_t)j * s.F;
            mkx[j] = s.obs_mx[i] * cm[12];
            mky[j] = s.obs_my[i] * cm[13];
            wim[j] = sqrt(s.obs_w[i]);
        }
    }
#pragma unroll
    for (int j = 0; j < MPC; ++j) {
        ex[j] = ey[j] = 0.0;
        if (m0 + j >= s.M) continue;
        const size_t i = i0 + (size_t)j * s.F;
        double dx, dy;
        eval_pair(cm, sb[j][0][0], sb[j][0][1], sb[j][0][2], mkx[j], mky[j], dx, dy);
        // adjust_measureErrors.cpp:481-499 (same operation order as residual_from)
        ex[j] = dx * imw * wim[j];
        ey[j] = dy * imw * wim[j];
        if (fvec) *reinterpret_cast<double2 *>(&fvec[2 * i]) = make_double2(ex[j], ey[j]);
        if (dist || dist_last) {
            const double d = sqrt((dx * dx) + (dy * dy)) * imw;
            if (dist) dist[i] = d;
            if (dist_last) dist_last[i] = d;  // overwritten below by the last attr's variant
        }
    }

    // ---- camera columns: one variant in registers at a time ------------------
    for (int v = 0; v < tp.VC; ++v) {
        double cv[12];
        const double *vt = ctab + (size_t)(v + 1) * kCamRec * cfs;
#pragma unroll
        for (int j = 0; j < 12; ++j) cv[j] = vt[(size_t)j * cfs];
        const double inv_delta = vt[(size_t)14 * cfs];
        const bool is_last = dist_last && tp.last_attr_cam_var == v;
        const int ucol = tp.cam_col_u[v];
#pragma unroll
        for (int j = 0; j < MPC; ++j) {
            const int m = m0 + j;
            if (m >= s.M) continue;
            const int col = tp.uniform ? ucol : tp.mkr_cam_col[(size_t)m * tp.VC + v];
            if (col < 0) continue;
            const size_t i = i0 + (size_t)j * s.F;
            double dx, dy;
            eval_pair(cv, sb[j][0][0], sb[j][0][1], sb[j][0][2], mkx[j], mky[j], dx, dy);
            const double ax = dx * imw * wim[j], ay = dy * imw * wim[j];
            *reinterpret_cast<double2 *>(&jac[(size_t)col * mtot + 2 * i]) =
                make_double2((ax - ex[j]) * inv_delta, (ay - ey[j]) * inv_delta);
            if (cols) {
                const int attr = tp.cam_dep_attr[v];
                cols[i * s.kmax + col] = s.attr_param_base[attr] + (s.attr_anim[attr] ? f : 0);
            }
            if (is_last) dist_last[i] = sqrt((dx * dx) + (dy * dy)) * imw;
        }
    }

    // ---- bundle columns: base camera rows, perturbed bundle position -----------
#pragma unroll
    for (int j = 0; j < MPC; ++j) {
        const int m = m0 + j;
        if (m >= s.M) continue;
        const size_t i = i0 + (size_t)j * s.F;
        for (int v = 0; v < tp.VB; ++v) {
            const int col = tp.uniform ? tp.bnd_col_u[v] : tp.mkr_bnd_col[(size_t)m * tp.VB + v];
            if (col < 0) continue;
            const double inv_delta = sb[j][v + 1][3];
            double dx, dy;
            eval_pair(cm, sb[j][v + 1][0], sb[j][v + 1][1], sb[j][v + 1][2], mkx[j], mky[j], dx, dy);
            const double ax = dx * imw * wim[j], ay = dy * imw * wim[j];
            *reinterpret_cast<double2 *>(&jac[(size_t)col * mtot + 2 * i]) =
                make_double2((ax - ex[j]) * inv_delta, (ay - ey[j]) * inv_delta);
            if (cols || dist_last) {
                const int attr = tp.mkr_bnd_attr[(size_t)m * tp.VB + v];
                if (cols) cols[i * s.kmax + col] = s.attr_param_base[attr] + (s.attr_anim[attr] ? f : 0);
                if (dist_last && attr == s.A - 1) dist_last[i] = sqrt((dx * dx) + (dy * dy)) * imw;
            }
        }
    }
}

// Fast instance: uniform column layout, full tile of markers (no tail), no
// column-id output.  Identical arithmetic to k_jacobian_tiled; the per-store
// index arithmetic is reduced to one 64-bit add.
template <int MPC, int MINB, bool DIST_LAST>
__global__ void __launch_bounds__(kTileFrames, MINB) k_jacobian_tiled_fast(DevScene s, TiledPlan tp,
                                                                        double *__restrict__ fvec,
                                                                        double *__restrict__ dist,
                                                                        double *__restrict__ jac,
                                                                        double *__restrict__ dist_last) {
    __shared__ double sb[MPC][kTileMaxVB + 1][kBndRec];
    const int f = blockIdx.x * kTileFrames + threadIdx.x;
    const int m0 = blockIdx.y * MPC;
    const int nrec = (tp.VB + 1) * kBndRec;
    for (int t = threadIdx.x; t < MPC * nrec; t += kTileFrames) {
        const int j = t / nrec, r = t - j * nrec;
        const int b = s.mkr_bnd[m0 + j];
        sb[j][r / kBndRec][r % kBndRec] = s.bnd_tab[s.bnd_off[b] + r];
    }
    __syncthreads();
    if (f >= s.F) return;
    const double imw = s.image_width;
    const int F = s.F;
    const int cfs = s.cam_fs[tp.cam];
    const double *ctab = s.cam_tab + s.cam_off[tp.cam] + (cfs == 1 ? 0 : f);
    double cm[14];
#pragma unroll
    for (int j = 0; j < 14; ++j) cm[j] = ctab[(size_t)j * cfs];
    const size_t i0 = (size_t)m0 * F + f;
    const double *pmx = s.obs_mx + i0, *pmy = s.obs_my + i0, *pw = s.obs_w + i0;
    double2 *jac2 = reinterpret_cast<double2 *>(jac) + i0;
    double mkx[MPC], mky[MPC], wim[MPC], ex[MPC], ey[MPC];
#pragma unroll
    for (int j = 0; j < MPC; ++j) {
        mkx[j] = pmx[(size_t)j * F] * cm[12];
        mky[j] = pmy[(size_t)j * F] * cm[13];
        wim[j] = sqrt(pw[(size_t)j * F]) ;
    }
#pragma unroll
    for (int j = 0; j < MPC; ++j) {
        double dx, dy;
        eval_pair(cm, sb[j][0][0], sb[j][0][1], sb[j][0][2], mkx[j], mky[j], dx, dy);
        ex[j] = dx * imw * wim[j];
        ey[j] = dy * imw * wim[j];
        if (fvec) reinterpret_cast<double2 *>(fvec)[i0 + (size_t)j * F] = make_double2(ex[j], ey[j]);
        if (dist || DIST_LAST) {
            const double d = sqrt((dx * dx) + (dy * dy)) * imw;
            if (dist) dist[i0 + (size_t)j * F] = d;
            if (DIST_LAST) dist_last[i0 + (size_t)j * F] = d;
        }
    }
    for (int v = 0; v < tp.VC; ++v) {
        double cv[12];
        const double *vt = ctab + (size_t)(v + 1) * kCamRec * cfs;
#pragma unroll
        for (int j = 0; j < 12; ++j) cv[j] = vt[(size_t)j * cfs];
        const double inv_delta = vt[(size_t)14 * cfs];
        double2 *jp = jac2 + tp.cam_coloff_u[v];
        const bool is_last = DIST_LAST && tp.last_attr_cam_var == v;
#pragma unroll
        for (int j = 0; j < MPC; ++j) {
            double dx, dy;
            eval_pair(cv, sb[j][0][0], sb[j][0][1], sb[j][0][2], mkx[j], mky[j], dx, dy);
            const double ax = dx * imw * wim[j], ay = dy * imw * wim[j];
            jp[(size_t)j * F] = make_double2((ax - ex[j]) * inv_delta, (ay - ey[j]) * inv_delta);
            if (is_last) dist_last[i0 + (size_t)j * F] = sqrt((dx * dx) + (dy * dy)) * imw;
        }
    }
    for (int v = 0; v < tp.VB; ++v) {
        double2 *jp = jac2 + tp.bnd_coloff_u[v];
#pragma unroll
        for (int j = 0; j < MPC; ++j) {
            const double inv_delta = sb[j][v + 1][3];
            double dx, dy;
            eval_pair(cm, sb[j][v + 1][0], sb[j][v + 1][1], sb[j][v + 1][2], mkx[j], mky[j], dx, dy);
            const double ax = dx * imw * wim[j], ay = dy * imw * wim[j];
            jp[(size_t)j * F] = make_double2((ax - ex[j]) * inv_delta, (ay - ey[j]) * inv_delta);
            if (DIST_LAST) {
                if (tp.mkr_bnd_attr[(size_t)(m0 + j) * tp.VB + v] == s.A - 1)
                    dist_last[i0 + (size_t)j * F] = sqrt((dx * dx) + (dy * dy)) * imw;
            }
        }
    }
}
#endif

}  // namespace mmb
