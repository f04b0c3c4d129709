// CUDA kernels (sm_100a) of the residual / FD-Jacobian path and the handle
// that owns their HBM buffers.  Per-thread bodies live in eval_bodies.cuh.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

#include "comm.h"
#include "device_problem.h"

namespace mmb {

static thread_local std::string g_last_error;
void set_last_error(const std::string &msg) { g_last_error = msg; }
const std::string &get_last_error() { return g_last_error; }

// ---- kernels ---------------------------------------------------------------
// kind 0 camera, 1 bundle, 2 lens; base_only skips perturbed variants
// Camera variants of 32 frames of one camera by one CTA: the 17 attribute values of the frames
// (9 transform + 8 camera attributes) go through the attribute view ONCE into shared memory;
// each (variant, frame) thread then takes them from there and looks up only the attribute its
// variant perturbs.  The flat one-thread-per-job path did the 17 lookups (100 dependent, cold
// loads and most of its 1 500 instructions) in every variant thread.
constexpr int kVarThreads = 256;
__device__ void camera_block_variants(const DevScene &s, const double *x, int base_only) {
    __shared__ double vals[17][32];
    __shared__ int refs[17], attr_of[17];
    const int tid = threadIdx.x, lane = tid & 31;
    const int c = s.vblk_cam[blockIdx.x], f0 = s.vblk_fb[blockIdx.x] * 32;
    const int tfm = s.cam_tfm[c];
    if (tid < 17) {
        const int32_t ref = tid < 9 ? s.tfm_attrs[tfm * 9 + tid] : s.cam_attrs[c * 8 + tid - 9];
        refs[tid] = ref;
        attr_of[tid] = x ? (ref >= 0 ? s.stat_attr[ref] : s.anim_attr[-ref - 1]) : -1;
    }
    __syncthreads();
    for (int e = tid; e < 17 * 32; e += kVarThreads) {
        const int j = e >> 5, ff = f0 + (e & 31);
        if (ff < s.F) {
            AttrView b;
            b.s = &s; b.x = x; b.f = ff; b.ovr_attr = -1; b.ovr_delta = 0.0;
            vals[j][e & 31] = b.get(refs[j]);
        }
    }
    __syncthreads();
    const int f = f0 + lane;
    if (f >= s.F) return;
    const int nvar = base_only ? 1 : cam_nvar_of(s.cam_ndep[c], s.central);
    for (int v = tid >> 5; v < nvar; v += kVarThreads / 32)
        camera_variant_gathered(s, x, c, v, f, &vals[0][lane], 32, refs, attr_of);
}

__global__ void __launch_bounds__(kVarThreads) k_variants(DevScene s, const double *x, int base_only) {
    // a dependent launch (k_jacobian_rows) may start its prologue now; it waits for this
    // grid's results with griddepcontrol.wait
    asm volatile("griddepcontrol.launch_dependents;");
    if ((int)blockIdx.x < s.vblk_count) { camera_block_variants(s, x, base_only); return; }
    const long long nc = s.cam_jobs_total, nb = s.bnd_jobs_total, nl = s.lens_jobs_total;
    // flat one-thread-per-job path (bundles, lenses, cameras the block path does not take): ONE
    // warp per CTA -- these are long dependent chains with fewer warps than the device has
    // schedulers; spread over all SMs they do not share a front end
    if (threadIdx.x >= 32) return;
    const long long t = (s.vblk_all ? nc : 0) + (blockIdx.x - s.vblk_count) * 32LL + threadIdx.x;
    if (t < nc) {
        const int job = s.cam_thread_job[t];
        if (s.cam_blk[s.cam_job_ent[job]]) return;   // done by the block path
        if (base_only && s.cam_job_var[job] != 0) return;
        camera_variant_thread(s, x, t);
    } else if (t < nc + nb) {
        const long long u = t - nc;
        if (base_only && s.bnd_job_var[s.bnd_thread_job[u]] != 0) return;
        bundle_variant_thread(s, x, u);
    } else if (t < nc + nb + nl) {
        const long long u = t - nc - nb;
        if (base_only && s.lens_job_var[s.lens_thread_job[u]] != 0) return;
        lens_variant_thread(s, x, u);
    }
}

__global__ void __launch_bounds__(256) k_residuals(DevScene s, double *fvec, double *dist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < s.nobs) residual_thread(s, i, fvec, dist);
}

__global__ void __launch_bounds__(256) k_jacobian(DevScene s, double *fvec, double *dist, double *jac,
                                                  int32_t *cols, double *dist_last) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < s.nobs) jacobian_thread(s, i, fvec, dist, jac, cols, dist_last);
}

// Lens in the loop (cfg4): forward-difference Jacobian with ONE evaluation per thread.
// CTA = (32 consecutive frames, a group of markers); warp e evaluates evaluation e of the current
// marker (e = 0: base, e = c + 1: column c) for the 32 frames, lanes = frames.  The camera
// variant records of the 32 frames (one contiguous block of the cam_idx layout) and the lens
// variant records are staged in shared memory once per CTA; residual pairs go through shared
// memory and the quotients are formed by the same threads afterwards, so every column store is a
// 512-byte coalesced run.  Same per-evaluation code as k_jacobian (residual_from): identical
// values.  With the LDPK fixed-point inverse an evaluation is a ~400-operation dependent FP64
// chain; the thread-per-observation kernel runs k + 1 of them back to back at 25 % occupancy,
// here they run side by side.
// Needs: dense grid, one animated camera (block layout), static lens and bundles, forward
// differences.
constexpr int kLensTileMaxEval = 24;
constexpr int kLensTileMB = 4;   // markers per barrier round
// GENERAL = false: one camera (index 0) with a lens, staged once; true: the camera (and lens, if
// any) of the current marker, restaged when it changes (markers of one camera are normally
// consecutive).
template <int MAXE, int MINB, bool GENERAL>
__global__ void __launch_bounds__(32 * MAXE, MINB) k_jacobian_lens(DevScene s, int mt, double *fvec, double *dist,
                                                                   double *jac, int32_t *cols, double *dist_last,
                                                                   int nv_max, int nvl_max) {
    extern __shared__ __align__(16) double lsm[];
    const int NE = blockDim.x >> 5, tid = threadIdx.x, lane = tid & 31, e = tid >> 5;
    double *scam = lsm;                                            // [nv][8 row pairs][32 frames][2]
    double *slens = scam + (size_t)nv_max * 512;                   // [nvl][24]
    double2 *sres = reinterpret_cast<double2 *>(slens + ((nvl_max * kLensRec + 1) & ~1));   // [MB][NE][32]
    double *sd = reinterpret_cast<double *>(sres + kLensTileMB * NE * 32);                  // [MB][NE][32]
    int cam = -1, lens = -1;
    const int f = blockIdx.x * 32 + lane;
    const bool live = f < s.F;
    const int m0 = blockIdx.y * mt, m1 = min(s.M, m0 + mt);
    const size_t m = 2 * (size_t)s.nobs;
    const double2 *scam2 = reinterpret_cast<const double2 *>(scam);
    for (int mk0 = m0; mk0 < m1;) {
        const int want = GENERAL ? s.mkr_cam[mk0] : 0;
        if (want != cam) {
            __syncthreads();
            cam = want;
            const int l = s.cam_lens[cam];
            lens = (l >= 0 && s.lens_model[l] != 0) ? l : -1;
            const int nv = cam_nvar_of(s.cam_ndep[cam], 0);
            const double *src = s.cam_tab + s.cam_off[cam] + (size_t)blockIdx.x * nv * 512;
            for (int t = tid; t < nv * 512; t += blockDim.x) scam[t] = src[t];
            if (lens >= 0) {
                const int nvl = 1 + s.lens_ndep[lens];
                const double *ls = s.lens_tab + s.lens_off[lens];
                for (int t = tid; t < nvl * kLensRec; t += blockDim.x) slens[t] = ls[t];
            }
            __syncthreads();
        }
        int nb = min(kLensTileMB, m1 - mk0);
        if (GENERAL)
            for (int u = 1; u < nb; ++u)
                if (s.mkr_cam[mk0 + u] != cam) { nb = u; break; }
        // ---- evaluations of nb markers ----
        for (int u = 0; u < nb; ++u) {
            const int mk = mk0 + u, k = s.mkr_k[mk];
            double ex = 0.0, ey = 0.0, d = 0.0;
            if (live && e <= k) {
                int cv = 0, bv = 0, lv = 0;
                if (e > 0) { const PlanEntryD en = s.plan[(size_t)mk * s.kmax + e - 1]; cv = en.cam_var; bv = en.bnd_var; lv = en.lens_var; }
                double cm[14], bp[3];
#pragma unroll
                for (int jp = 0; jp < 7; ++jp) { const double2 t = scam2[(cv * 8 + jp) * 32 + lane]; cm[2 * jp] = t.x; cm[2 * jp + 1] = t.y; }
                const int b = s.mkr_bnd[mk];
                const double *bsrc = s.bnd_tab + s.bnd_off[b] + (size_t)bv * kBndRec;
                bp[0] = bsrc[0]; bp[1] = bsrc[1]; bp[2] = bsrc[2];
                const size_t i = (size_t)mk * s.F + f;
                ObsCtx o;
                o.marker = mk; o.frame = f; o.cam = cam; o.bnd = b; o.lens = lens; o.lens_frame = f;
                o.mx = s.obs_mx[i]; o.my = s.obs_my[i]; o.w = sqrt(s.obs_w[i]);
                const double *rec = GENERAL ? (lens >= 0 ? slens + lv * kLensRec : nullptr) : slens + lv * kLensRec;
                residual_from(s, o, cm, bp, rec, ex, ey, d);
            }
            sres[(u * NE + e) * 32 + lane] = make_double2(ex, ey);
            sd[(u * NE + e) * 32 + lane] = d;
        }
        __syncthreads();
        // ---- quotients and stores: warp 0 the base outputs, warp e > 0 column e - 1 ----
        if (live) {
            for (int u = 0; u < nb; ++u) {
                const int mk = mk0 + u, k = s.mkr_k[mk];
                const PlanEntryD *pl = s.plan + (size_t)mk * s.kmax;
                const size_t i = (size_t)mk * s.F + f;
                const double2 base = sres[(u * NE) * 32 + lane];
                if (e == 0) {
                    const double d = sd[(u * NE) * 32 + lane];
                    if (fvec) { fvec[2 * i] = base.x; fvec[2 * i + 1] = base.y; }
                    if (dist) dist[i] = d;
                    if (dist_last) {
                        double dl = d;
                        for (int c = 0; c < k; ++c) if (pl[c].attr == s.A - 1) dl = sd[(u * NE + c + 1) * 32 + lane];
                        dist_last[i] = dl;
                    }
                } else if (e - 1 < s.kmax) {
                    const int c = e - 1;
                    double jx = 0.0, jy = 0.0;
                    int32_t pid = -1;
                    if (c < k) {
                        const int attr = pl[c].attr;
                        pid = s.attr_param_base[attr] + (s.attr_anim[attr] ? f : 0);
                        const double inv_delta = 1.0 / s.par_delta_a[pid];
                        const double2 a = sres[(u * NE + e) * 32 + lane];
                        jx = (a.x - base.x) * inv_delta;
                        jy = (a.y - base.y) * inv_delta;
                    }
                    *reinterpret_cast<double2 *>(&jac[(size_t)c * m + 2 * i]) = make_double2(jx, jy);
                    if (cols) cols[i * s.kmax + c] = pid;
                }
            }
        }
        __syncthreads();
        mk0 += nb;
    }
}

// frame-blocked copy of the marker tracks for k_jacobian_rows: block fb holds, marker by
// marker, kRowFrames x-values, kRowFrames y-values, kRowFrames weights (one contiguous run
// per CTA tile).  Built once per handle from the observation arrays already in HBM.
__global__ void __launch_bounds__(256) k_block_markers(const double *__restrict__ mx, const double *__restrict__ my,
                                                       const double *__restrict__ w, int M, int F,
                                                       double *__restrict__ blk) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= (long long)M * F) return;
    const int mk = (int)(i / F), f = (int)(i - (long long)mk * F);
    double *dst = &blk[(((size_t)(f / kRowFrames) * M + mk) * 3) * kRowFrames + (f % kRowFrames)];
    dst[0] = mx[i]; dst[kRowFrames] = my[i]; dst[2 * kRowFrames] = w[i];
}

// ---- launchers -------------------------------------------------------------
static inline unsigned blocks_for(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

mmb_status launch_variants(DeviceProblem *p, const double *d_x, bool with_perturbed) {
    const long long flat = (p->ds.vblk_all ? 0 : p->ds.cam_jobs_total) + p->ds.bnd_jobs_total + p->ds.lens_jobs_total;
    const unsigned grid = (unsigned)p->ds.vblk_count + (flat ? blocks_for(flat, 32) : 0u);
    if (grid == 0) return MMB_OK;
    const unsigned bs = p->ds.vblk_count ? kVarThreads : 32;   // only the block path needs full CTAs
    k_variants<<<grid, bs, 0, p->stream>>>(p->ds, d_x, with_perturbed ? 0 : 1);
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    return MMB_OK;
}

mmb_status launch_residuals(DeviceProblem *p, double *d_fvec, double *d_dist) {
    k_residuals<<<blocks_for(p->hp.nobs, 256), 256, 0, p->stream>>>(p->ds, d_fvec, d_dist);
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    return MMB_OK;
}

// stiffness / smoothness rows: values (row != nullptr) and / or their Jacobian entries
__global__ void k_extra_rows(DevScene s, const double *x, double *row, double *jext, int32_t *pid) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < s.n_extra) extra_row_thread(s, x, e, row, jext, pid);
}
mmb_status launch_extra_rows(DeviceProblem *p, const double *d_x, double *d_row, double *d_jext, int32_t *d_pid) {
    if (p->hp.n_extra == 0) return MMB_OK;
    k_extra_rows<<<(p->hp.n_extra + 63) / 64, 64, 0, p->stream>>>(p->ds, d_x, d_row, d_jext, d_pid);
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    return MMB_OK;
}

mmb_status launch_jacobian(DeviceProblem *p, double *d_fvec, double *d_dist, double *d_jac, int32_t *d_cols, double *d_dist_last) {
    if (!p->cam_plans.empty() && !d_cols) {
        static int rows_mc = -1;
        if (rows_mc < 0) { const char *e = getenv("MMB_JAC_ROWS_MC"); rows_mc = e ? atoi(e) : 1; }
        if (rows_mc) {
            constexpr int W = 4, MINB = 4;
            const TiledPlan &t0 = p->cam_plans[0];
            const int nv = t0.VC + 1, nrec = (t0.VB + 1) * kBndRec;
            const int fb = (p->hp.F + kRowFrames - 1) / kRowFrames;
            const bool fixed = t0.VC == 6 && t0.VB == 3;
            const bool sparse = t0.obs_base != nullptr;
            // (sparse visibility keeps no marker block in shared memory; larger marker tiles were no faster:
            // 10^7 observations at 0.25 visibility, 0.588 ms with 64 markers per CTA, 0.594 with 128, 0.633 with 192)
            const int mt_cap = 64;
            auto smem_for = [&](int mt) { return (size_t)256 + sizeof(double) * ((size_t)nv * kCamRec * kRowFrames + (size_t)mt * (nrec + (sparse ? 0 : 3 * kRowFrames))); };
            auto kern = sparse ? (fixed ? (d_dist_last ? k_jacobian_rows<W, MINB, true, 6, 3, true> : k_jacobian_rows<W, MINB, false, 6, 3, true>)
                                        : (d_dist_last ? k_jacobian_rows<W, MINB, true, 0, 0, true> : k_jacobian_rows<W, MINB, false, 0, 0, true>))
                               : (fixed ? (d_dist_last ? k_jacobian_rows<W, MINB, true, 6, 3> : k_jacobian_rows<W, MINB, false, 6, 3>)
                                        : (d_dist_last ? k_jacobian_rows<W, MINB, true> : k_jacobian_rows<W, MINB, false>));
            int m_most = 0;
            for (const TiledPlan &t : p->cam_plans) m_most = std::max(m_most, t.m_count);
            int &mt = p->rows_mt[d_dist_last ? 1 : 0];
            if (mt == 0) {
                int occ = 1, sm_count = 1;
                MMB_CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, p->device));
                MMB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_for(mt_cap)));
                MMB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * W, smem_for(sparse ? mt_cap : 32)));
                const int slots = std::max(1, occ) * sm_count;
                int G = std::max(1, slots / fb);
                G = std::min(G, (m_most + 1) / 2);
                mt = (m_most + G - 1) / G;
                mt = std::min(mt_cap, (mt + 1) & ~1);
                static int mt_force = -1;
                if (mt_force < 0) { const char *e = getenv("MMB_ROWS_MT"); mt_force = e ? atoi(e) : 0; }
                if (mt_force > 0) mt = std::min(mt_cap, (mt_force + 1) & ~1);
            }
            // one launch for all the cameras (blockIdx.z): the tail of one camera's tiles overlaps
            // the next camera's instead of draining the device between launches
            const int G = (m_most + mt - 1) / mt;
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(fb, G, (unsigned)p->cam_plans.size()); cfg.blockDim = dim3(32 * W); cfg.dynamicSmemBytes = smem_for(mt); cfg.stream = p->stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            const CamSlice *slices = p->d_cam_slices;
            MMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, p->ds, t0, mt, d_fvec, d_dist, d_jac, d_dist_last, slices));
            ++p->launches;
            return MMB_OK;
        }
    }
    // a sparse-visibility scene can only take the rows kernel (the other tiled kernels index a dense grid)
    const bool tiled_sparse = p->tiled_ok && p->tiled.obs_base != nullptr;
    const bool tiled_sparse_rows = tiled_sparse && !d_cols && p->tiled.uniform && p->tiled.mk_blk &&
                                   p->hp.cam_fs[p->tiled.cam] == p->hp.F && p->hp.F > 1 && (p->hp.cam_off[p->tiled.cam] % 2) == 0 &&
                                   !(getenv("MMB_JAC_ROWS") && atoi(getenv("MMB_JAC_ROWS")) == 0);
    if (p->tiled_ok && (!tiled_sparse || tiled_sparse_rows)) {
        static int variant = -1;
        if (variant < 0) { const char *e = getenv("MMB_TILE"); variant = e ? atoi(e) : 4; }
        const int fblocks = (p->hp.F + kTileFrames - 1) / kTileFrames;
        int m_done = 0;
        // animated camera, uniform layout: camera rows in shared memory, one resident wave
        static int rows_mode = -1, rows_g = 0;
        if (rows_mode < 0) {
            const char *e = getenv("MMB_JAC_ROWS"); rows_mode = e ? atoi(e) : 1;
            const char *g = getenv("MMB_ROWS_G"); rows_g = g ? atoi(g) : 0;
        }
        if (rows_mode && p->tiled.uniform && !d_cols && p->tiled.mk_blk && p->hp.cam_fs[p->tiled.cam] == p->hp.F &&
            p->hp.F > 1 && (p->hp.cam_off[p->tiled.cam] % 2) == 0) {
            constexpr int W = 4, MINB = 4;
            const int nv = p->tiled.VC + 1, nrec = (p->tiled.VB + 1) * kBndRec;
            const int fb = (p->hp.F + kRowFrames - 1) / kRowFrames;
            const bool sparse = p->tiled.obs_base != nullptr;
            const int mt_cap = 64;
            auto smem_for = [&](int mt) { return (size_t)256 + sizeof(double) * ((size_t)nv * kCamRec * kRowFrames + (size_t)mt * (nrec + (sparse ? 0 : 3 * kRowFrames))); };
            // the common shape (6 camera attributes, 3 bundle attributes) has its own instance
            static int fixed_mode = -1;
            if (fixed_mode < 0) { const char *e = getenv("MMB_ROWS_FIXED"); fixed_mode = e ? atoi(e) : 1; }
            const bool fixed = fixed_mode && p->tiled.VC == 6 && p->tiled.VB == 3;
            auto kern = sparse ? (fixed ? (d_dist_last ? k_jacobian_rows<W, MINB, true, 6, 3, true> : k_jacobian_rows<W, MINB, false, 6, 3, true>)
                                        : (d_dist_last ? k_jacobian_rows<W, MINB, true, 0, 0, true> : k_jacobian_rows<W, MINB, false, 0, 0, true>))
                               : (fixed ? (d_dist_last ? k_jacobian_rows<W, MINB, true, 6, 3> : k_jacobian_rows<W, MINB, false, 6, 3>)
                                        : (d_dist_last ? k_jacobian_rows<W, MINB, true> : k_jacobian_rows<W, MINB, false>));
            int &mt = p->rows_mt[d_dist_last ? 1 : 0];
            if (mt == 0) {  // one wave: CTAs = resident slots of the device (or as close as the marker count allows)
                int occ = 1, sm_count = 1;
                MMB_CUDA_CHECK(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, p->device));
                MMB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_for(mt_cap)));
                MMB_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * W, smem_for(sparse ? mt_cap : 32)));
                const int slots = std::max(1, occ) * sm_count;
                int G = rows_g > 0 ? rows_g : std::max(1, slots / fb);
                G = std::min(G, (p->hp.M + 1) / 2);
                mt = (p->hp.M + G - 1) / G;
                mt = std::min(mt_cap, (mt + 1) & ~1);
            }
            const int G = (p->hp.M + mt - 1) / mt;
            // programmatic dependent launch: the prologue (barrier setup, marker-track fetch)
            // may start while the preceding kernel of the stream (entity variants) drains
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(fb, G); cfg.blockDim = dim3(32 * W); cfg.dynamicSmemBytes = smem_for(mt); cfg.stream = p->stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            MMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, p->ds, p->tiled, mt, d_fvec, d_dist, d_jac, d_dist_last, (const CamSlice *)nullptr));
            ++p->launches;
            return MMB_OK;
        }
        // full tiles of a uniform-layout scene: fast instance
        if (p->tiled.uniform && !d_cols) {
#define MMB_FAST(MPC, MINB)                                                                                          \
    {                                                                                                                \
        const int tiles = p->hp.M / MPC;                                                                             \
        if (tiles > 0) {                                                                                             \
            if (d_dist_last)                                                                                         \
                k_jacobian_tiled_fast<MPC, MINB, true><<<dim3(fblocks, tiles), kTileFrames, 0, p->stream>>>(         \
                    p->ds, p->tiled, d_fvec, d_dist, d_jac, d_dist_last);                                            \
            else                                                                                                     \
                k_jacobian_tiled_fast<MPC, MINB, false><<<dim3(fblocks, tiles), kTileFrames, 0, p->stream>>>(        \
                    p->ds, p->tiled, d_fvec, d_dist, d_jac, d_dist_last);                                            \
            ++p->launches;                                                                                           \
        }                                                                                                            \
        m_done = tiles * MPC;                                                                                        \
    }
            switch (variant) {
                case 1: MMB_FAST(1, 6) break;
                case 2: MMB_FAST(3, 4) break;
                case 3: MMB_FAST(4, 3) break;
                case 0: MMB_FAST(2, 4) break;
                default: MMB_FAST(4, 4) break;
            }
#undef MMB_FAST
        }
        if (m_done < p->hp.M) {  // tail markers, non-uniform layouts, column-id output
            constexpr int MPC = 2;
            const dim3 grid(fblocks, (p->hp.M - m_done + MPC - 1) / MPC);
            k_jacobian_tiled<MPC, 4><<<grid, kTileFrames, 0, p->stream>>>(p->ds, p->tiled, d_fvec, d_dist, d_jac, d_cols,
                                                                          d_dist_last, m_done);
            ++p->launches;
        }
        MMB_CUDA_CHECK(cudaGetLastError());
        return MMB_OK;
    }
    static int lens_tile_mode = -1, lens_tile_mt = 0;
    if (lens_tile_mode < 0) {
        const char *e = getenv("MMB_LENS_TILE"); lens_tile_mode = e ? atoi(e) : 1;
        const char *g = getenv("MMB_LENS_TILE_MT"); lens_tile_mt = g ? atoi(g) : 16;
    }
    if (p->lens_tile_ok && lens_tile_mode) {
        const HostProblem &hp = p->hp;
        int nv = 1, nvl = 1;
        for (int c = 0; c < hp.C; ++c) {
            nv = std::max(nv, 1 + (int)hp.cam_deps[c].size());
            const int l = hp.cam_lens[c];
            if (l >= 0) nvl = std::max(nvl, 1 + (int)hp.lens_deps[l].size());
        }
        const int NE = hp.kmax + 1;
        const size_t smem = sizeof(double) * ((size_t)nv * 512 + ((nvl * kLensRec + 1) & ~1) + (size_t)kLensTileMB * NE * 32 * 3);
        // up to 18 evaluations: 56 registers, two CTAs per SM; more: one CTA of up to 24 warps
        static int small_mode = -1;
        if (small_mode < 0) { const char *e = getenv("MMB_LENS_TILE_SMALL"); small_mode = e ? atoi(e) : 1; }
        const bool general = !(hp.C == 1 && hp.cam_lens[0] >= 0 && hp.lens_model[hp.cam_lens[0]] != 0);
        auto kern = (NE <= 18 && small_mode)
                        ? (general ? k_jacobian_lens<18, 2, true> : k_jacobian_lens<18, 2, false>)
                        : (general ? k_jacobian_lens<kLensTileMaxEval, 1, true> : k_jacobian_lens<kLensTileMaxEval, 1, false>);
        MMB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int mt = std::max(1, lens_tile_mt);
        const dim3 grid((hp.F + 31) / 32, (hp.M + mt - 1) / mt);
        kern<<<grid, 32 * NE, smem, p->stream>>>(p->ds, mt, d_fvec, d_dist, d_jac, d_cols, d_dist_last, nv, nvl);
        ++p->launches;
        MMB_CUDA_CHECK(cudaGetLastError());
        return MMB_OK;
    }
    k_jacobian<<<blocks_for(p->hp.nobs, 256), 256, 0, p->stream>>>(p->ds, d_fvec, d_dist, d_jac, d_cols, d_dist_last);
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    return MMB_OK;
}

// ---- handle ----------------------------------------------------------------
void lm_workspace_destroy(LmWorkspace *w);

DeviceProblem::~DeviceProblem() {
    const bool timing = getenv("MMB_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    cudaSetDevice(device);
    if (lm) lm_workspace_destroy(lm);
    const auto t1 = std::chrono::steady_clock::now();
    if (comm) { if (stream) cudaStreamSynchronize(stream); if (comm_owned) delete comm; }
    arena.release();  // stream-ordered frees: back to the pool once the stream's work is done
    if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); }
    if (timing)
        fprintf(stderr, "[mmb timing] destroy: LM workspace %.3f ms, scene buffers %.3f ms\n",
                1e3 * std::chrono::duration<double>(t1 - t0).count(),
                1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count());
}

mmb_status ensure_jacobian_buffers(DeviceProblem *p) {
    if (p->d_jac) return MMB_OK;
    const size_t k = (size_t)std::max(p->hp.kmax, 1);
    MMB_CUDA_CHECK(p->arena.alloc(&p->d_jac, k * (size_t)p->hp.m));
    MMB_CUDA_CHECK(p->arena.alloc(&p->d_cols, k * (size_t)p->hp.nobs));
    // the tiled kernel never touches unused columns: they stay 0 / -1
    MMB_CUDA_CHECK(cudaMemsetAsync(p->d_jac, 0, sizeof(double) * k * (size_t)p->hp.m, p->stream));
    MMB_CUDA_CHECK(cudaMemsetAsync(p->d_cols, 0xff, sizeof(int32_t) * k * (size_t)p->hp.nobs, p->stream));
    return MMB_OK;
}

// frame-blocked copy of the marker tracks (one contiguous run per tile of k_jacobian_rows) and,
// for sparse visibility, the (frame block, marker) -> first observation / frame mask index
static mmb_status build_marker_blocks(DeviceProblem *p, TiledPlan *t) {
    const HostProblem &hp = p->hp;
    DeviceArena &A = p->arena;
    const int FB = (hp.F + kRowFrames - 1) / kRowFrames;
    // (a sparse scene reads the observation arrays themselves: the blocked copy would hold, and
    // the kernel would fetch, every (marker, frame) slot whether or not it has an observation)
    const size_t nblk = hp.dense_grid ? (size_t)FB * hp.M * 3 * kRowFrames : 2;
    double *d_blk = nullptr;
    MMB_CUDA_CHECK(A.flush());   // the kernels below read tables uploaded so far
    MMB_CUDA_CHECK(A.alloc(&d_blk, nblk));
    MMB_CUDA_CHECK(cudaMemsetAsync(d_blk, 0, sizeof(double) * nblk, p->stream));
    t->obs_base = nullptr; t->obs_mask = nullptr;
    if (hp.dense_grid) {
        k_block_markers<<<(unsigned)(((long long)hp.M * hp.F + 255) / 256), 256, 0, p->stream>>>(p->ds.obs_mx, p->ds.obs_my, p->ds.obs_w, hp.M, hp.F, d_blk);
    } else {
        std::vector<int32_t> base((size_t)FB * hp.M, 0);
        std::vector<uint32_t> mask((size_t)FB * hp.M, 0u);
        for (int i = hp.nobs - 1; i >= 0; --i) {   // descending: the first observation of a block is written last
            const size_t key = (size_t)(hp.err_frame[i] / kRowFrames) * hp.M + hp.err_marker[i];
            mask[key] |= 1u << (hp.err_frame[i] % kRowFrames);
            base[key] = i;
        }
        int32_t *d_b = nullptr; uint32_t *d_m = nullptr;
        MMB_CUDA_CHECK(A.upload(&d_b, base));
        MMB_CUDA_CHECK(A.upload(&d_m, mask));
        MMB_CUDA_CHECK(cudaStreamSynchronize(p->stream));   // the host vectors go out of scope
        t->obs_base = d_b; t->obs_mask = d_m;
    }
    ++p->launches;
    MMB_CUDA_CHECK(cudaGetLastError());
    t->mk_blk = d_blk;
    return MMB_OK;
}

mmb_status device_problem_create(const mmb_scene_desc *s, const mmb_options *opt, DeviceProblem **out) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_last_error("no CUDA device: this library has no CPU path");
        return MMB_ERR_NO_DEVICE;
    }
    if (opt->device < 0 || opt->device >= ndev) {
        set_last_error("device ordinal out of range");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    const bool timing = getenv("MMB_TIMING") != nullptr;
    const auto tc0 = std::chrono::steady_clock::now();
    DeviceProblem *p = new DeviceProblem();
    mmb_status st = build_host_problem(s, opt, &p->hp);
    const auto tc1 = std::chrono::steady_clock::now();
    if (st != MMB_OK) {
        set_last_error(p->hp.error);
        delete p;
        return st;
    }
    p->device = opt->device;
#define CK(expr)                                                               \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            delete p;                                                          \
            return MMB_ERR_CUDA;                                               \
        }                                                                      \
    } while (0)
    CK(cudaSetDevice(p->device));
    CK(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    {   // keep freed device memory in the pool for the next handle (see DeviceArena)
        cudaMemPool_t pool = nullptr;
        CK(cudaDeviceGetDefaultMemPool(&pool, p->device));
        unsigned long long keep = ~0ull;
        CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    p->arena.bind(p->stream);
    CK(p->arena.begin_batch());   // small tables: one slab, one staging block, one copy (DeviceArena)
    HostProblem &hp = p->hp;
    DevScene &d = p->ds;
    DeviceArena &A = p->arena;
    d.F = hp.F; d.T = hp.T; d.C = hp.C; d.B = hp.B; d.M = hp.M; d.L = hp.L; d.A = hp.A;
    d.n = hp.n; d.nobs = hp.nobs; d.kmax = std::max(hp.kmax, 1);
    d.central = opt->auto_diff_type == 1;
    d.dense_grid = hp.dense_grid;
    d.any_lens = hp.any_lens;
    d.lens_literal = opt->lens_index_literal;
    d.delta = opt->delta;
    d.lmdif = opt->solver_type == 1;
    d.fd_eps = std::sqrt(std::max(std::fabs(opt->delta), 2.220446049250313e-16));
    d.image_width = opt->image_width;
    d.maya_dag = hp.maya_dag; d.cur_frame = hp.cur_frame; d.n_extra = hp.n_extra;
    CK(A.upload(&p->d_stat0, hp.stat0));
    CK(A.upload(&p->d_anim0, hp.anim0));
    d.stat = p->d_stat0;  // the attribute store is read-only: solved values come from x
    d.anim = p->d_anim0;
    int32_t *tmp_i = nullptr; double *tmp_d = nullptr; uint8_t *tmp_b = nullptr; int64_t *tmp_l = nullptr;
#define UP_I(field, vec) CK(A.upload(&tmp_i, vec)); d.field = tmp_i
#define UP_D(field, vec) CK(A.upload(&tmp_d, vec)); d.field = tmp_d
#define UP_L(field, vec) CK(A.upload(&tmp_l, vec)); d.field = tmp_l
    UP_I(tfm_parent, hp.tfm_parent); UP_I(tfm_roo, hp.tfm_roo); UP_I(tfm_attrs, hp.tfm_attrs);
    UP_I(cam_tfm, hp.cam_tfm); UP_I(cam_attrs, hp.cam_attrs); UP_I(cam_fit, hp.cam_fit);
    UP_I(cam_rw, hp.cam_rw); UP_I(cam_rh, hp.cam_rh); UP_I(cam_lens, hp.cam_lens);
    UP_I(bnd_tfm, hp.bnd_tfm); UP_I(mkr_cam, hp.mkr_cam); UP_I(mkr_bnd, hp.mkr_bnd);
    UP_I(lens_model, hp.lens_model); UP_I(lens_attrs, hp.lens_attrs); UP_D(lens_cam, hp.lens_cam);
    UP_I(attr_ref, hp.attr_ref); UP_I(attr_param_base, hp.attr_param_base);
    UP_I(stat_attr, hp.stat_attr); UP_I(anim_attr, hp.anim_attr);
    CK(A.upload(&tmp_b, hp.attr_anim)); d.attr_anim = tmp_b;
    {
        std::vector<uint8_t> plain(std::max(hp.A, 1), 0);
        const double fm = 3.40282346638528859811704183484516925e+38;
        for (int a = 0; a < hp.A; ++a)
            plain[a] = hp.attr_min[a] <= -fm && hp.attr_max[a] >= fm && hp.attr_scl[a] == 1.0 && hp.attr_off[a] == 0.0;
        CK(A.upload(&tmp_b, plain)); d.attr_plain = tmp_b;
    }
    UP_D(attr_min, hp.attr_min); UP_D(attr_max, hp.attr_max);
    UP_D(attr_off, hp.attr_off); UP_D(attr_scl, hp.attr_scl);
    UP_I(par_attr, hp.par_attr); UP_I(par_frame, hp.par_frame);
    if (hp.dense_grid) { d.err_marker = nullptr; d.err_frame = nullptr; }
    else { UP_I(err_marker, hp.err_marker); UP_I(err_frame, hp.err_frame); }
    UP_D(obs_mx, hp.obs_mx); UP_D(obs_my, hp.obs_my); UP_D(obs_w, hp.obs_w);
    CK(A.alloc(&d.cam_tab, (size_t)hp.cam_tab_size));
    d.cam_fwd = nullptr;
    if (hp.maya_dag) CK(A.alloc(&d.cam_fwd, (size_t)hp.cam_tab_size / 4 + 4));
    {
        std::vector<int32_t> e0(1, 0);
        CK(A.upload(&tmp_i, hp.n_extra ? hp.ex_attr : e0)); d.ex_attr = tmp_i;
        CK(A.upload(&tmp_i, hp.n_extra ? hp.ex_w : e0)); d.ex_w = tmp_i;
        CK(A.upload(&tmp_i, hp.n_extra ? hp.ex_var : e0)); d.ex_var = tmp_i;
        CK(A.upload(&tmp_i, hp.n_extra ? hp.ex_val : e0)); d.ex_val = tmp_i;
    }
    CK(A.alloc(&d.bnd_tab, (size_t)hp.bnd_tab_size));
    CK(A.alloc(&d.lens_tab, (size_t)hp.lens_tab_size));
    UP_L(cam_off, hp.cam_off); UP_L(bnd_off, hp.bnd_off); UP_L(lens_off, hp.lens_off);
    UP_I(cam_fs, hp.cam_fs); UP_I(bnd_fs, hp.bnd_fs); UP_I(lens_fs, hp.lens_fs);
    std::vector<int32_t> cnd(hp.C), bnd(hp.B), lnd(hp.L);
    for (int c = 0; c < hp.C; ++c) cnd[c] = (int32_t)hp.cam_deps[c].size();
    for (int b = 0; b < hp.B; ++b) bnd[b] = (int32_t)hp.bnd_deps[b].size();
    for (int l = 0; l < hp.L; ++l) lnd[l] = (int32_t)hp.lens_deps[l].size();
    UP_I(cam_ndep, cnd); UP_I(bnd_ndep, bnd); UP_I(lens_ndep, lnd);
    UP_I(cam_job_ent, hp.cam_job_ent); UP_I(cam_job_var, hp.cam_job_var); UP_I(cam_job_attr, hp.cam_job_attr);
    UP_L(cam_job_start, hp.cam_job_start);
    UP_I(cam_thread_job, hp.cam_thread_job); UP_I(bnd_thread_job, hp.bnd_thread_job); UP_I(lens_thread_job, hp.lens_thread_job);
    d.cam_njobs = (int)hp.cam_job_ent.size(); d.cam_jobs_total = hp.cam_jobs_total;
    {   // block path: animated cameras whose transform has no parent
        std::vector<int32_t> job0(hp.C, 0), vcam, vfb;
        std::vector<uint8_t> blk(hp.C, 0);
        for (size_t j = 0; j < hp.cam_job_ent.size(); ++j)
            if (hp.cam_job_var[j] == 0) job0[hp.cam_job_ent[j]] = (int32_t)j;
        const char *e = getenv("MMB_VARIANTS_BLOCK");
        const bool enabled = !(e && atoi(e) == 0);
        int nblk = 0;
        for (int c = 0; c < hp.C; ++c) {
            if (!enabled || hp.cam_fs[c] != hp.F || hp.F < 2 || hp.tfm_parent[hp.cam_tfm[c]] >= 0) continue;
            blk[c] = 1; ++nblk;
            for (int fb = 0; fb < (hp.F + 31) / 32; ++fb) { vcam.push_back(c); vfb.push_back(fb); }
        }
        d.vblk_count = (int)vcam.size();
        d.vblk_all = nblk == hp.C;
        if (vcam.empty()) { vcam.push_back(0); vfb.push_back(0); }
        UP_I(cam_job0, job0); UP_I(vblk_cam, vcam); UP_I(vblk_fb, vfb);
        CK(A.upload(&tmp_b, blk)); d.cam_blk = tmp_b;
    }
    UP_I(bnd_job_ent, hp.bnd_job_ent); UP_I(bnd_job_var, hp.bnd_job_var); UP_I(bnd_job_attr, hp.bnd_job_attr);
    UP_L(bnd_job_start, hp.bnd_job_start);
    d.bnd_njobs = (int)hp.bnd_job_ent.size(); d.bnd_jobs_total = hp.bnd_jobs_total;
    UP_I(lens_job_ent, hp.lens_job_ent); UP_I(lens_job_var, hp.lens_job_var); UP_I(lens_job_attr, hp.lens_job_attr);
    UP_L(lens_job_start, hp.lens_job_start);
    d.lens_njobs = (int)hp.lens_job_ent.size(); d.lens_jobs_total = hp.lens_jobs_total;
    {
        static_assert(sizeof(PlanEntry) == sizeof(PlanEntryD), "plan entry layout");
        PlanEntryD *dp = nullptr;
        CK(A.alloc(&dp, hp.plan.size()));
        if (!hp.plan.empty())
            CK(cudaMemcpyAsync(dp, hp.plan.data(), hp.plan.size() * sizeof(PlanEntry), cudaMemcpyHostToDevice, p->stream));
        d.plan = dp;
    }
    UP_I(mkr_k, hp.mkr_k);
    CK(A.alloc(&d.par_delta_a, (size_t)hp.n));
    CK(A.alloc(&d.par_delta_b, (size_t)hp.n));
    // ---- register-tiled Jacobian fast path (jacobian_tiled.cuh) ---------------
    {
        size_t VB = 0;
        bool bundles_static = true;
        for (int b = 0; b < hp.B; ++b) {
            VB = std::max(VB, hp.bnd_deps[b].size());
            if (hp.bnd_fs[b] != 1) bundles_static = false;
        }
        const size_t VC = hp.cam_deps[0].size();
        bool eligible = opt->auto_diff_type == 0 && !hp.any_lens && hp.C == 1 && !hp.maya_dag &&
                        !hp.cam_mscale_varies[0] && bundles_static && VC <= (size_t)kTileMaxVC &&
                        VB <= (size_t)kTileMaxVB && hp.kmax <= 127;
        for (int mk = 0; eligible && mk < hp.M; ++mk)
            for (int c = 0; c < hp.mkr_k[mk]; ++c) {
                const PlanEntry &e = hp.plan[(size_t)mk * hp.kmax + c];
                if ((e.cam_var != 0) == (e.bnd_var != 0) || e.lens_var != 0) eligible = false;
            }
        bool ok = eligible && hp.F >= 32;
        const char *force = getenv("MMB_JACOBIAN_PATH");
        if (force && !strcmp(force, "generic")) ok = false;
        if (force && !strcmp(force, "tiled")) ok = eligible;
        if (ok) {
            std::vector<int8_t> ccol((size_t)hp.M * std::max<size_t>(VC, 1), -1), bcol((size_t)hp.M * std::max<size_t>(VB, 1), -1);
            std::vector<int32_t> battr((size_t)hp.M * std::max<size_t>(VB, 1), -1);
            for (int mk = 0; mk < hp.M; ++mk)
                for (int c = 0; c < hp.mkr_k[mk]; ++c) {
                    const PlanEntry &e = hp.plan[(size_t)mk * hp.kmax + c];
                    if (e.cam_var) ccol[(size_t)mk * VC + (e.cam_var - 1)] = (int8_t)c;
                    else { bcol[(size_t)mk * VB + (e.bnd_var - 1)] = (int8_t)c; battr[(size_t)mk * VB + (e.bnd_var - 1)] = e.attr; }
                }
            TiledPlan &t = p->tiled;
            t.cam = 0; t.VC = (int)VC; t.VB = (int)VB;
            t.m_base = 0; t.m_count = hp.M;
            t.uniform = 1;
            for (int mk = 1; mk < hp.M && t.uniform; ++mk) {
                for (size_t v = 0; v < VC; ++v) if (ccol[(size_t)mk * VC + v] != ccol[v]) t.uniform = 0;
                for (size_t v = 0; v < VB; ++v) if (bcol[(size_t)mk * VB + v] != bcol[v]) t.uniform = 0;
            }
            for (int v = 0; v < kTileMaxVC; ++v) t.cam_col_u[v] = (size_t)v < VC ? ccol[v] : (int8_t)-1;
            for (int v = 0; v <= kTileMaxVB; ++v) t.bnd_col_u[v] = (size_t)v < VB ? bcol[v] : (int8_t)-1;
            for (int v = 0; v < kTileMaxVC; ++v) t.cam_coloff_u[v] = (long long)std::max<int>(t.cam_col_u[v], 0) * hp.nobs;
            for (int v = 0; v <= kTileMaxVB; ++v) t.bnd_coloff_u[v] = (long long)std::max<int>(t.bnd_col_u[v], 0) * hp.nobs;
            // uniform layout also needs every marker to own all VC / VB columns
            for (size_t v = 0; v < VC; ++v) if (ccol[v] < 0) t.uniform = 0;
            for (size_t v = 0; v < VB; ++v) if (bcol[v] < 0) t.uniform = 0;
            std::vector<int32_t> cattr(hp.cam_deps[0].begin(), hp.cam_deps[0].end());
            if (cattr.empty()) cattr.push_back(0);
            int8_t *d8 = nullptr;
            CK(A.upload(&tmp_i, cattr)); t.cam_dep_attr = tmp_i;
            CK(A.upload(&d8, ccol)); t.mkr_cam_col = d8;
            CK(A.upload(&d8, bcol)); t.mkr_bnd_col = d8;
            CK(A.upload(&tmp_i, battr)); t.mkr_bnd_attr = tmp_i;
            t.bnd_contig = 1;
            t.bnd_rec0 = hp.bnd_off[hp.mkr_bnd[0]];
            for (int mk = 0; mk < hp.M && t.bnd_contig; ++mk) {
                const int b = hp.mkr_bnd[mk];
                if (hp.bnd_fs[b] != 1 || hp.bnd_nvar[b] != (int)VB + 1 ||
                    hp.bnd_off[b] != t.bnd_rec0 + (int64_t)mk * (int64_t)(VB + 1) * kBndRec) t.bnd_contig = 0;
            }
            if (t.bnd_rec0 % 2) t.bnd_contig = 0;  // 16-byte alignment of the bulk copy
            t.mk_blk = nullptr; t.obs_base = nullptr; t.obs_mask = nullptr;
            // sparse visibility: only the rows kernel (uniform column layout) knows the index
            if (hp.dense_grid || t.uniform) {
                const mmb_status bs = build_marker_blocks(p, &t);
                if (bs != MMB_OK) { delete p; return bs; }
            }
            t.last_attr_cam_var = -1;
            for (size_t v = 0; v < VC; ++v) if (hp.cam_deps[0][v] == hp.A - 1) t.last_attr_cam_var = (int)v;
            p->tiled_ok = hp.dense_grid || (t.uniform && t.mk_blk);
        }
    }
    // ---- the rows kernel for several cameras (cfg5 shape): one launch per camera ---------------
    // Markers in camera-major order, every camera animated with the same number of solved
    // attributes, every marker the same column layout (its camera's attributes, then its bundle's).
    if (hp.C > 1 && opt->auto_diff_type == 0 && !hp.any_lens && !hp.maya_dag && hp.F >= 32 && hp.kmax <= 127) {
        const size_t VC = hp.cam_deps[0].size();
        size_t VB = 0;
        bool ok = VC >= 1 && VC <= (size_t)kTileMaxVC;
        for (int b = 0; b < hp.B; ++b) { VB = std::max(VB, hp.bnd_deps[b].size()); if (hp.bnd_fs[b] != 1) ok = false; }
        ok = ok && VB <= (size_t)kTileMaxVB;
        for (int c = 0; ok && c < hp.C; ++c)
            if (hp.cam_deps[c].size() != VC || hp.cam_fs[c] != hp.F || hp.cam_mscale_varies[c] || (hp.cam_off[c] % 2) != 0) ok = false;
        for (int mk = 1; ok && mk < hp.M; ++mk) if (hp.mkr_cam[mk] < hp.mkr_cam[mk - 1]) ok = false;
        std::vector<int8_t> ccol((size_t)hp.M * std::max<size_t>(VC, 1), -1), bcol((size_t)hp.M * std::max<size_t>(VB, 1), -1);
        std::vector<int32_t> battr((size_t)hp.M * std::max<size_t>(VB, 1), -1);
        for (int mk = 0; ok && mk < hp.M; ++mk)
            for (int c = 0; c < hp.mkr_k[mk]; ++c) {
                const PlanEntry &e = hp.plan[(size_t)mk * hp.kmax + c];
                if ((e.cam_var != 0) == (e.bnd_var != 0) || e.lens_var != 0) { ok = false; break; }
                if (e.cam_var) ccol[(size_t)mk * VC + (e.cam_var - 1)] = (int8_t)c;
                else { bcol[(size_t)mk * VB + (e.bnd_var - 1)] = (int8_t)c; battr[(size_t)mk * VB + (e.bnd_var - 1)] = e.attr; }
            }
        for (int mk = 0; ok && mk < hp.M; ++mk) {   // uniform layout, every column present
            for (size_t v = 0; v < VC; ++v) if (ccol[(size_t)mk * VC + v] != ccol[v] || ccol[v] < 0) ok = false;
            for (size_t v = 0; v < VB; ++v) if (bcol[(size_t)mk * VB + v] != bcol[v] || bcol[v] < 0) ok = false;
        }
        const char *force = getenv("MMB_JACOBIAN_PATH");
        if (force && !strcmp(force, "generic")) ok = false;
        if (ok) {
            int8_t *d8 = nullptr;
            TiledPlan base{};
            base.VC = (int)VC; base.VB = (int)VB; base.uniform = 1;
            for (int v = 0; v < kTileMaxVC; ++v) base.cam_col_u[v] = (size_t)v < VC ? ccol[v] : (int8_t)-1;
            for (int v = 0; v <= kTileMaxVB; ++v) base.bnd_col_u[v] = (size_t)v < VB ? bcol[v] : (int8_t)-1;
            for (int v = 0; v < kTileMaxVC; ++v) base.cam_coloff_u[v] = (long long)std::max<int>(base.cam_col_u[v], 0) * hp.nobs;
            for (int v = 0; v <= kTileMaxVB; ++v) base.bnd_coloff_u[v] = (long long)std::max<int>(base.bnd_col_u[v], 0) * hp.nobs;
            CK(A.upload(&d8, ccol)); base.mkr_cam_col = d8;
            CK(A.upload(&d8, bcol)); base.mkr_bnd_col = d8;
            CK(A.upload(&tmp_i, battr)); base.mkr_bnd_attr = tmp_i;
            base.cam_dep_attr = nullptr;
            {
                const mmb_status bs = build_marker_blocks(p, &base);
                if (bs != MMB_OK) { delete p; return bs; }
            }
            int mk = 0;
            while (mk < hp.M) {
                const int c = hp.mkr_cam[mk];
                int end = mk;
                while (end < hp.M && hp.mkr_cam[end] == c) ++end;
                TiledPlan t = base;
                t.cam = c; t.m_base = mk; t.m_count = end - mk;
                t.bnd_contig = 1;
                t.bnd_rec0 = hp.bnd_off[hp.mkr_bnd[mk]];
                for (int q = mk; q < end && t.bnd_contig; ++q) {
                    const int b = hp.mkr_bnd[q];
                    if (hp.bnd_nvar[b] != (int)VB + 1 || hp.bnd_off[b] != t.bnd_rec0 + (int64_t)(q - mk) * (int64_t)(VB + 1) * kBndRec) t.bnd_contig = 0;
                }
                if (t.bnd_rec0 % 2) t.bnd_contig = 0;
                t.last_attr_cam_var = -1;
                for (size_t v = 0; v < VC; ++v) if (hp.cam_deps[c][v] == hp.A - 1) t.last_attr_cam_var = (int)v;
                p->cam_plans.push_back(t);
                mk = end;
            }
            std::vector<CamSlice> sl;
            for (const TiledPlan &t : p->cam_plans) sl.push_back(CamSlice{t.cam, t.m_base, t.m_count, t.bnd_contig, t.last_attr_cam_var, t.bnd_rec0});
            CamSlice *d_sl = nullptr;
            CK(A.upload(&d_sl, sl));
            p->d_cam_slices = d_sl;
        }
    }
    {   // evaluation-per-thread tiled kernel (k_jacobian_lens): lens in the loop and / or several cameras
        bool ok = hp.dense_grid && opt->auto_diff_type == 0 && !opt->lens_index_literal && hp.F >= 2 && !hp.maya_dag &&
                  hp.kmax + 1 <= kLensTileMaxEval && hp.kmax >= 1;
        size_t nv = 1, nvl = 1;
        for (int c = 0; ok && c < hp.C; ++c) {
            if (hp.cam_fs[c] != hp.F) ok = false;
            nv = std::max(nv, 1 + hp.cam_deps[c].size());
            const int l = hp.cam_lens[c];
            if (l >= 0 && hp.lens_model[l] != 0) {
                if (hp.lens_fs[l] != 1) ok = false;
                nvl = std::max(nvl, 1 + hp.lens_deps[l].size());
            }
        }
        if (sizeof(double) * (nv * 512 + 26 * nvl + (size_t)kLensTileMB * (hp.kmax + 1) * 96) > 200 * 1024) ok = false;
        for (int b = 0; ok && b < hp.B; ++b) if (hp.bnd_fs[b] != 1) ok = false;
        p->lens_tile_ok = ok;
    }
    CK(A.alloc(&p->d_x, (size_t)hp.n));
    CK(A.alloc(&p->d_fvec, (size_t)hp.m_total));
    CK(A.alloc(&p->d_dist, (size_t)hp.nobs));
    CK(A.end_batch());
#undef CK
#undef UP_I
#undef UP_D
#undef UP_L
    if (timing) {
        cudaDeviceSynchronize();
        const auto tc2 = std::chrono::steady_clock::now();
        fprintf(stderr, "[mmb timing] create: host index maps / plans %.3f ms, device alloc + upload %.3f ms (%zu allocations, %.1f MB)\n",
                1e3 * std::chrono::duration<double>(tc1 - tc0).count(), 1e3 * std::chrono::duration<double>(tc2 - tc1).count(),
                A.ptrs.size(), A.bytes / 1e6);
    }
    *out = p;
    return MMB_OK;
}

}  // namespace mmb
