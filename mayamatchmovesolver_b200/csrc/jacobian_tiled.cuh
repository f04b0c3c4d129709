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
    int m_base, m_count;         // its markers: m_base .. m_base + m_count - 1 (camera-major marker order)
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
    int bnd_contig;              // bundle records of markers 0..M-1 are consecutive in bnd_tab
    long long bnd_rec0;          // ... starting at this offset (doubles)
    const double *mk_blk;        // marker x, y, weight re-blocked [frame block of 32][marker][3][32]
    int last_attr_cam_var;       // 0-based camera dep index of attr A-1 (the variant the
                                 // reference's errorDistanceList ends on), or -1
    // sparse visibility (disabled / zero-weight marker frames leave no observation): for every
    // (32-frame block, marker) the index of its first observation and which frames have one;
    // the observation of frame lane is obs_base + popc(obs_mask & lanes below).  nullptr: dense grid.
    const int32_t *obs_base;     // [frame blocks][M]
    const uint32_t *obs_mask;    // [frame blocks][M]
};

// Several cameras in one launch of k_jacobian_rows: blockIdx.z picks the camera, whose part of
// the plan (the fields that differ between cameras) comes from this record.
struct CamSlice {
    int cam, m_base, m_count, bnd_contig, last_attr_cam_var;
    long long bnd_rec0;
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
    // row pairs of (variant 0, this frame); next pair: + cps, next variant: + cvs (cam_idx)
    const int cps = cfs == 1 ? 1 : kCamBlock;
    const size_t cvs = (size_t)(kCamRec / 2) * cps;
    const double2 *ctab = reinterpret_cast<const double2 *>(s.cam_tab + s.cam_off[tp.cam]) +
                          (cfs == 1 ? (size_t)0 : (size_t)(f / kCamBlock) * (tp.VC + 1) * cvs + (size_t)(f % kCamBlock));
    double cm[14];
#pragma unroll
    for (int j = 0; j < 7; ++j) { const double2 t = ctab[(size_t)j * cps]; cm[2 * j] = t.x; cm[2 * j + 1] = t.y; }

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
        const double2 *vt = ctab + (size_t)(v + 1) * cvs;
#pragma unroll
        for (int j = 0; j < 6; ++j) { const double2 t = vt[(size_t)j * cps]; cv[2 * j] = t.x; cv[2 * j + 1] = t.y; }
        const double inv_delta = vt[(size_t)7 * cps].x;
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
    // row pairs of (variant 0, this frame); next pair: + cps, next variant: + cvs (cam_idx)
    const int cps = cfs == 1 ? 1 : kCamBlock;
    const size_t cvs = (size_t)(kCamRec / 2) * cps;
    const double2 *ctab = reinterpret_cast<const double2 *>(s.cam_tab + s.cam_off[tp.cam]) +
                          (cfs == 1 ? (size_t)0 : (size_t)(f / kCamBlock) * (tp.VC + 1) * cvs + (size_t)(f % kCamBlock));
    double cm[14];
#pragma unroll
    for (int j = 0; j < 7; ++j) { const double2 t = ctab[(size_t)j * cps]; cm[2 * j] = t.x; cm[2 * j + 1] = t.y; }
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
        const double2 *vt = ctab + (size_t)(v + 1) * cvs;
#pragma unroll
        for (int j = 0; j < 6; ++j) { const double2 t = vt[(size_t)j * cps]; cv[2 * j] = t.x; cv[2 * j + 1] = t.y; }
        const double inv_delta = vt[(size_t)7 * cps].x;
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

// ---- shared-memory camera rows: the one-wave instance -----------------------------------
// A CTA owns 32 consecutive frames and `mt` consecutive markers.  The camera rows of
// its frames (base + every perturbed variant, (VC+1) x 14 rows x 32 frames) are brought
// into shared memory once with 1-D TMA bulk copies (cp.async.bulk, one mbarrier per
// variant, so the first evaluations start while later variants are still in flight)
// and are then shared by the CTA's W warps, each of which walks its own marker pairs.
// A thread keeps only the base rows and two markers in registers; the grid is sized by
// the host to exactly one resident wave (kernels.cu).
constexpr int kRowFrames = 32;
constexpr int kRowSlots = 14;  // rows 0..11, then base: film-fit scale x,y / variant: 1/delta

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MMB_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MMB_DONE_%=;\n"
        "bra MMB_WAIT_%=;\n"
        "MMB_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// eval_pair with the three dot products as nested FMAs and the quotient taken as
// s * (1/w) without the final residual correction (<= 2 ulp; both sides of every finite
// difference use the same formula)
__device__ __forceinline__ void eval_pair_fma(const double *cm, double b0, double b1, double b2, double mkx,
                                              double mky, double &dx, double &dy) {
    const double sx = fma(cm[0], b0, fma(cm[1], b1, fma(cm[2], b2, cm[3])));
    const double sy = fma(cm[4], b0, fma(cm[5], b1, fma(cm[6], b2, cm[7])));
    const double sw = fma(cm[8], b0, fma(cm[9], b1, fma(cm[10], b2, cm[11])));
    const double r = rcp_nr(sw) * -0.5;  // reprojection.rs:55-62
    dx = fabs(fma(sx, r, mkx));
    dy = fabs(fma(sy, r, mky));
}

constexpr int kRowPairs = kCamRec / 2;  // 16-byte row pairs per camera variant

// VCT / VBT: compile-time camera / bundle variant counts (0 = read them from the plan): with
// them the variant loops unroll completely and the FP64 chains (~20 cycles per dependent
// operation) of different variants interleave.
template <int W, int MINB, bool DIST_LAST, int VCT = 0, int VBT = 0, bool SPARSE = false>
__global__ void __launch_bounds__(32 * W, MINB) k_jacobian_rows(DevScene s, TiledPlan tp, int mt,
                                                                double *__restrict__ fvec, double *__restrict__ dist,
                                                                double *__restrict__ jac, double *__restrict__ dist_last,
                                                                const CamSlice *__restrict__ slices) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int nv = tp.VC + 1;
    const int nrec = (tp.VB + 1) * kBndRec;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw);                        // [nv + 1]
    double2 *scam = reinterpret_cast<double2 *>(smem_raw + 256);                    // [nv][8 row pairs][32 frames]
    double *smk = reinterpret_cast<double *>(scam + (size_t)nv * kRowPairs * kRowFrames);  // [mt][3][32] marker x, y, weight
    double *sbnd = SPARSE ? smk : smk + (size_t)mt * 3 * kRowFrames;                // [mt][VB+1][4]  (sparse: no marker block)
    int cam_ = tp.cam, m_base_ = tp.m_base, m_count_ = tp.m_count, bnd_contig_ = tp.bnd_contig, last_cam_var_ = tp.last_attr_cam_var;
    long long bnd_rec0_ = tp.bnd_rec0;
    if (slices) {
        const CamSlice cs = slices[blockIdx.z];
        cam_ = cs.cam; m_base_ = cs.m_base; m_count_ = cs.m_count; bnd_contig_ = cs.bnd_contig; last_cam_var_ = cs.last_attr_cam_var;
        bnd_rec0_ = cs.bnd_rec0;
        if ((int)blockIdx.y * mt >= m_count_) return;   // the grid is sized for the camera with the most markers
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int f0 = blockIdx.x * kRowFrames, f = f0 + lane;
    const int F = s.F;
    const int m0 = m_base_ + blockIdx.y * mt;
    const int mcount = min(mt, m_base_ + m_count_ - m0);
    uint64_t *bar_mk = &bars[nv];
    if (threadIdx.x == 0) {
        for (int v = 0; v <= nv; ++v) mbar_init(&bars[v], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // Every input of the tile is one contiguous run (cam_idx block layout, frame-blocked
    // marker copy, marker-ordered bundle records): nv + 2 TMA bulk copies per CTA.
    if (threadIdx.x == 0) {
        const uint32_t var_bytes = (uint32_t)(kRowPairs * kRowFrames * sizeof(double2));
        const uint32_t mk_bytes = (uint32_t)mcount * 3u * kRowFrames * 8u;
        const uint32_t bnd_bytes = bnd_contig_ ? (uint32_t)(mcount * nrec) * 8u : 0u;
        const double2 *ctab = reinterpret_cast<const double2 *>(s.cam_tab + s.cam_off[cam_]) +
                              (size_t)blockIdx.x * nv * kRowPairs * kRowFrames;
        // the marker tracks are problem constants: fetch them (HBM latency) before waiting on
        // the variants kernel this launch may overlap with (programmatic dependent launch)
        if (!SPARSE) {
            mbar_expect_tx(bar_mk, mk_bytes);
            bulk_g2s(smk, tp.mk_blk + ((size_t)blockIdx.x * s.M + m0) * 3 * kRowFrames, mk_bytes, bar_mk);
        }
        asm volatile("griddepcontrol.wait;" ::: "memory");
        mbar_expect_tx(&bars[0], var_bytes + bnd_bytes);
        bulk_g2s(scam, ctab, var_bytes, &bars[0]);
        if (bnd_contig_) bulk_g2s(sbnd, s.bnd_tab + bnd_rec0_ + (size_t)(m0 - m_base_) * nrec, bnd_bytes, &bars[0]);
        for (int v = 1; v < nv; ++v) {
            mbar_expect_tx(&bars[v], var_bytes);
            bulk_g2s(scam + (size_t)v * kRowPairs * kRowFrames, ctab + (size_t)v * kRowPairs * kRowFrames, var_bytes, &bars[v]);
        }
    }
    if (!bnd_contig_) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        for (int t = threadIdx.x; t < mcount * nrec; t += 32 * W) {
            const int j = t / nrec, r = t - j * nrec;
            sbnd[t] = s.bnd_tab[s.bnd_off[s.mkr_bnd[m0 + j]] + r];
        }
        __syncthreads();
    }
    if (SPARSE) {
        // ---- sparse visibility: compacted lanes ------------------------------------------------
        // With lane = frame a marker seen on one frame in four leaves three lanes in four idle.
        // Here a warp takes KB markers at a time and deals their live (marker, frame) pairs to its
        // lanes: pair t of the batch is the r-th set bit of marker u's frame mask (__fns).  The
        // camera rows of the lane's frame come from the staged block by frame index, the bundle
        // record by marker; the observation index is obs_base + r, so the lanes of one marker
        // still store to consecutive addresses.
        constexpr int KB = 16;
        const double imw_s = s.image_width;
        const int VBs = VBT ? VBT : tp.VB, VCs = VCT ? VCT : tp.VC;
        const int32_t *const ob_ = tp.obs_base + (size_t)blockIdx.x * s.M;
        const uint32_t *const om_ = tp.obs_mask + (size_t)blockIdx.x * s.M;
        double2 *const jac2s = reinterpret_cast<double2 *>(jac);
        mbar_wait(&bars[0], 0);
        for (int v = 1; v < nv; ++v) mbar_wait(&bars[v], 0);
        // The lanes of a pass want different frames: 16-byte reads of [row pair][frame] collide
        // (frames 8 apart share their banks).  Re-lay every row pair in place as four rows of
        // 32-bit words, [word][frame]: bank = frame, so any set of frames reads without conflict
        // and equal frames broadcast.
        uint32_t *const swd = reinterpret_cast<uint32_t *>(scam);
        for (int rp = w; rp < nv * kRowPairs; rp += W) {
            const double2 q2 = scam[rp * kRowFrames + lane];
            __syncwarp();
            uint32_t *o = swd + (size_t)rp * 4 * kRowFrames + lane;
            o[0] = (uint32_t)__double2loint(q2.x); o[kRowFrames] = (uint32_t)__double2hiint(q2.x);
            o[2 * kRowFrames] = (uint32_t)__double2loint(q2.y); o[3 * kRowFrames] = (uint32_t)__double2hiint(q2.y);
        }
        __syncthreads();
        auto ldw = [&](const uint32_t *q) { return __hiloint2double((int)q[kRowFrames], (int)q[0]); };
        for (int j0 = w * KB; j0 < mcount; j0 += W * KB) {
            // lane u (and u + 16) holds marker j0 + u: its frame mask, first observation, and the
            // number of live pairs of the markers before it in the batch
            const int ju = j0 + (lane & (KB - 1));
            const uint32_t km_l = ju < mcount ? om_[m0 + ju] : 0u;
            const int ob_l = ju < mcount ? ob_[m0 + ju] : 0;
            int incl = __popc(km_l);
#pragma unroll
            for (int d = 1; d < KB; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, d, KB); if ((lane & (KB - 1)) >= d) incl += y; }
            const int total = __shfl_sync(0xffffffffu, incl, KB - 1);
            const int excl = incl - __popc(km_l);
            for (int tb = 0; tb < total; tb += 32) {
                const int t = tb + lane;
                int u = 0;    // the last marker whose pairs start at or before t (it has pairs: the next one starts after t)
#pragma unroll
                for (int st = KB / 2; st > 0; st >>= 1) { const int e = __shfl_sync(0xffffffffu, excl, u + st); if (e <= t) u += st; }
                const uint32_t kmu = __shfl_sync(0xffffffffu, km_l, u);
                const int r = t - __shfl_sync(0xffffffffu, excl, u);
                const int obu = __shfl_sync(0xffffffffu, ob_l, u);
                if (t >= total) continue;
                const int fl = __fns(kmu, 0, r + 1);          // the frame (lane of the block) of this pair
                const int j = j0 + u;
                const size_t i = (size_t)obu + r;
                const uint32_t *cw = swd + fl;
                double cmb[14];
#pragma unroll
                for (int q = 0; q < 14; ++q) cmb[q] = ldw(cw + 2 * q * kRowFrames);
                const double2 *bq = reinterpret_cast<const double2 *>(sbnd + (size_t)j * nrec);
                // the marker track straight from the observation arrays (consecutive lanes, consecutive
                // observations): a frame-blocked copy would be read for every slot, live or not
                const double mkx = __ldg(s.obs_mx + i) * cmb[12], mky = __ldg(s.obs_my + i) * cmb[13], wq = sqrt(__ldg(s.obs_w + i));
                const double2 b01 = bq[0], b23 = bq[1];
                const double px = b01.x, py = b01.y, pz = b23.x;
                double dx, dy;
                eval_pair_fma(cmb, px, py, pz, mkx, mky, dx, dy);
                const double c0 = imw_s * wq;
                if (fvec) reinterpret_cast<double2 *>(fvec)[i] = make_double2(dx * c0, dy * c0);
                if (dist || DIST_LAST) {
                    const double d0 = sqrt((dx * dx) + (dy * dy)) * imw_s;
                    if (dist) dist[i] = d0;
                    if (DIST_LAST) dist_last[i] = d0;
                }
                double2 *const jp = jac2s + i;
#pragma unroll(VBT ? VBT : 1)
                for (int v = 0; v < VBs; ++v) {
                    const double2 q01 = bq[2 * (v + 1)], q23 = bq[2 * (v + 1) + 1];
                    double ax, ay;
                    eval_pair_fma(cmb, q01.x, q01.y, q23.x, mkx, mky, ax, ay);
                    const double g0 = c0 * q23.y;
                    jp[tp.bnd_coloff_u[v]] = make_double2((ax - dx) * g0, (ay - dy) * g0);
                    if (DIST_LAST && tp.mkr_bnd_attr[(size_t)(m0 + j) * VBs + v] == s.A - 1) dist_last[i] = sqrt((ax * ax) + (ay * ay)) * imw_s;
                }
                const uint32_t *sv = cw + (size_t)kRowPairs * 4 * kRowFrames;
#pragma unroll(VCT ? VCT : 1)
                for (int v = 0; v < VCs; ++v, sv += kRowPairs * 4 * kRowFrames) {
                    double cv[12];
#pragma unroll
                    for (int q = 0; q < 12; ++q) cv[q] = ldw(sv + 2 * q * kRowFrames);
                    const double inv_delta = ldw(sv + 2 * 14 * kRowFrames);
                    double ax, ay;
                    eval_pair_fma(cv, px, py, pz, mkx, mky, ax, ay);
                    const double g0 = c0 * inv_delta;
                    jp[tp.cam_coloff_u[v]] = make_double2((ax - dx) * g0, (ay - dy) * g0);
                    if (DIST_LAST && last_cam_var_ == v) dist_last[i] = sqrt((ax * ax) + (ay * ay)) * imw_s;
                }
            }
        }
        return;
    }
    if (f >= F) return;
    const double imw = s.image_width;
    mbar_wait(&bars[0], 0);
    double cm[14];
#pragma unroll
    for (int j = 0; j < 7; ++j) { const double2 t = scam[j * kRowFrames + lane]; cm[2 * j] = t.x; cm[2 * j + 1] = t.y; }
    mbar_wait(bar_mk, 0);
    bool variants_ready = false;
    double2 *const jac2 = reinterpret_cast<double2 *>(jac);
    const int VB = VBT ? VBT : tp.VB, VC = VCT ? VCT : tp.VC;
    for (int jj = 2 * w; jj < mcount; jj += 2 * W) {
        // an odd marker count evaluates its last marker twice (same values to the same addresses)
        const int j1 = min(jj + 1, mcount - 1);
        // i0, i1: the two observations of this lane's frame (dense grid)
        const size_t i0 = (size_t)(m0 + jj) * F + f, i1 = (size_t)(m0 + j1) * F + f;
        const double *b0 = sbnd + (size_t)jj * nrec;
        const double *b1 = sbnd + (size_t)j1 * nrec;
        const double *k0 = smk + (size_t)jj * 3 * kRowFrames + lane, *k1 = smk + (size_t)j1 * 3 * kRowFrames + lane;
        const double mkx0 = k0[0] * cm[12], mky0 = k0[kRowFrames] * cm[13], w0 = sqrt(k0[2 * kRowFrames]);
        const double mkx1 = k1[0] * cm[12], mky1 = k1[kRowFrames] * cm[13], w1 = sqrt(k1[2 * kRowFrames]);
        const double p0x = b0[0], p0y = b0[1], p0z = b0[2], p1x = b1[0], p1y = b1[1], p1z = b1[2];
        double dx0, dy0, dx1, dy1;
        eval_pair_fma(cm, p0x, p0y, p0z, mkx0, mky0, dx0, dy0);
        eval_pair_fma(cm, p1x, p1y, p1z, mkx1, mky1, dx1, dy1);
        // (|a| - |e|) * imageWidth * sqrt(weight) / delta: subtracting the two unweighted
        // deviations first loses nothing to rounding
        const double c0 = imw * w0, c1 = imw * w1;
        if (fvec) {
            double2 *fv = reinterpret_cast<double2 *>(fvec);
            fv[i0] = make_double2(dx0 * c0, dy0 * c0);
            fv[i1] = make_double2(dx1 * c1, dy1 * c1);
        }
        if (dist || DIST_LAST) {
            const double d0 = sqrt((dx0 * dx0) + (dy0 * dy0)) * imw, d1 = sqrt((dx1 * dx1) + (dy1 * dy1)) * imw;
            if (dist) { dist[i0] = d0; dist[i1] = d1; }
            if (DIST_LAST) { dist_last[i0] = d0; dist_last[i1] = d1; }
        }
        double2 *const jp0 = jac2 + i0, *const jp1 = jac2 + i1;
        // bundle columns: base camera rows, perturbed bundle
#pragma unroll(VBT ? VBT : 1)
        for (int v = 0; v < VB; ++v) {
            const double *q0 = b0 + (v + 1) * kBndRec, *q1 = b1 + (v + 1) * kBndRec;
            double ax, ay, bx, by;
            eval_pair_fma(cm, q0[0], q0[1], q0[2], mkx0, mky0, ax, ay);
            eval_pair_fma(cm, q1[0], q1[1], q1[2], mkx1, mky1, bx, by);
            const long long co = tp.bnd_coloff_u[v];
            const double g0 = c0 * q0[3], g1 = c1 * q1[3];
            jp0[co] = make_double2((ax - dx0) * g0, (ay - dy0) * g0);
            jp1[co] = make_double2((bx - dx1) * g1, (by - dy1) * g1);
            if (DIST_LAST) {
                if (tp.mkr_bnd_attr[(size_t)(m0 + jj) * VB + v] == s.A - 1) dist_last[i0] = sqrt((ax * ax) + (ay * ay)) * imw;
                if (tp.mkr_bnd_attr[(size_t)(m0 + j1) * VB + v] == s.A - 1) dist_last[i1] = sqrt((bx * bx) + (by * by)) * imw;
            }
        }
        // camera columns: perturbed camera rows from shared memory, base bundle
        if (!variants_ready) {
            for (int v = 1; v < nv; ++v) mbar_wait(&bars[v], 0);
            variants_ready = true;
        }
        const double2 *sv = scam + (size_t)kRowPairs * kRowFrames + lane;
#pragma unroll(VCT ? VCT : 1)
        for (int v = 0; v < VC; ++v, sv += kRowPairs * kRowFrames) {
            double cv[12];
#pragma unroll
            for (int j = 0; j < 6; ++j) { const double2 t = sv[j * kRowFrames]; cv[2 * j] = t.x; cv[2 * j + 1] = t.y; }
            const double inv_delta = sv[7 * kRowFrames].x;
            double ax, ay, bx, by;
            eval_pair_fma(cv, p0x, p0y, p0z, mkx0, mky0, ax, ay);
            eval_pair_fma(cv, p1x, p1y, p1z, mkx1, mky1, bx, by);
            const long long co = tp.cam_coloff_u[v];
            const double g0 = c0 * inv_delta, g1 = c1 * inv_delta;
            jp0[co] = make_double2((ax - dx0) * g0, (ay - dy0) * g0);
            jp1[co] = make_double2((bx - dx1) * g1, (by - dy1) * g1);
            if (DIST_LAST && last_cam_var_ == v) {
                dist_last[i0] = sqrt((ax * ax) + (ay * ay)) * imw;
                dist_last[i1] = sqrt((bx * bx) + (by * by)) * imw;
            }
        }
    }
}
#endif

}  // namespace mmb
