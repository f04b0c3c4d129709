// Linear-algebra kernels of the device LM loop (sm_100a).
//
// Normal equations in block form (SURVEY.md section 8a "structure note"):
//   eliminated side  E: animated parameters, one dense na x na block U_f per
//                       frame (a frame's observations only touch its own
//                       animated parameters),
//   reduced side     S: the ns static parameters (bundles, focal, lens ...),
//   coupling         W_f (na x ns) per frame.
//   (U + par D_E^2) p_E + W p_S = g_E ;  W^T p_E + (V + par D_S^2) p_S = g_S
// Schur complement on E:  S = V + par D_S^2 - sum_f Y_f^T Y_f,
//   Y_f = L_f^-1 W_f,  L_f L_f^T = U_f + par D_f^2,  z_f = L_f^-1 g_f.
// The right-hand side rides along as an augmented row/column (index ns), so
// the SYRK also forms g_S - Y^T z and the Cholesky also forward-solves it.
// The SYRK and the Cholesky trailing update are the dense FP64 contractions
// and run on the tensor cores (DMMA, mma.sync m8n8k4 f64).
#pragma once
#include <cuda_runtime.h>

#include "eval_bodies.cuh"
#include "lm_state.h"
#include "reduce.cuh"

namespace mmb {

constexpr int kNB = 64;        // Cholesky block

struct LmDims {
    int F, na, ns, NS1, NSP, KE, KEP, n, nobs, m, kmax;
};

// partial[blockIdx] = sum of squares of v[0..n) (grid-stride); stage 2 sums
// the partials in a control kernel -> deterministic.
__global__ void k_sumsq_partial(const double *__restrict__ v, size_t n, double *__restrict__ partial) {
    __shared__ double sh[32];
    double acc = 0.0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const double t = v[i];
        acc += t * t;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// ---- assembly of J^T J and J^T f from the block Jacobian ---------------------
// jac[c*m + 2i + {0,1}], plan gives the attr of column c of marker(i).
// Columns whose solver attribute is the same for EVERY marker (the camera's animated attributes,
// focal length, lens coefficients): their mutual products are sums over all markers of a frame
// (or over all observations), i.e. heavily contended addresses for atomics.  With a shared-column
// mask they are left to k_assemble_shared (register accumulation, one atomic per CTA and entry)
// and this kernel only does the pairs that involve a per-marker (bundle) column.
struct AsmShared {
    int ns;          // number of shared columns (0: none, k_assemble does everything)
    unsigned mask;   // bit c set: column c is shared
    int col[16];     // the shared columns
    int slot[16];    // their slot in the animated / static class
    int anim[16];
    int pbase[16];   // first parameter of the column's solver attribute (+ frame when animated)
};

// ---- order-independent accumulation of the normal equations -------------------------------
// J^T J and J^T f entries are sums of up to Nobs products whose order on the GPU depends on
// the schedule.  Floating-point atomics made two runs of the same solve differ (1e-11
// relative), and the LM iteration amplifies that.  Here every contribution is split into two
// 64-bit fixed-point limbs on a per-entry grid and accumulated with INTEGER atomics, which
// are associative: the sums, hence the whole solve, are bit-identical from run to run.
//   grid of entry (p, q): bound 2^E >= |jx_p jx_q + jy_p jy_q| of one observation, from the
//   column maxima cmax[] (k_colmax; cmax[n] = max |f| for the gradient entries);
//   hi = rint(v / 2^(E-35)),  lo = rint((v - hi 2^(E-35)) / 2^(E-71)):  71 bits below the bound.
//   Sums of partials over G observations stay below G 2^35 (hi) and 2^35 per add (lo):
//   no overflow for Nobs < 2^27 per GPU.
// Entries with at most two contributions (W of a camera attribute x a bundle attribute: one
// marker) keep plain FP atomics onto zero: a + b = b + a.
constexpr int kDetHi = 35, kDetLo = 36;
constexpr int kDetEmax = 900;
__device__ __forceinline__ double pow2i(int k) { return __longlong_as_double((long long)(k + 1023) << 52); }
// exponent e with |x| < 2^e from the bit pattern of |x| (0 / denormal -> -1022, inf / nan -> 1025)
__device__ __forceinline__ int det_exp_of(unsigned long long bits) { return (int)((bits >> 52) & 0x7ffull) - 1022; }
__device__ __forceinline__ int det_exp(int ea, int eb) { return max(-kDetEmax, min(kDetEmax, ea + eb + 1)); }
__device__ __forceinline__ void det_add(long long *acc, double v, int E, int *poison) {
    if (v == 0.0) return;
    const double h = rint(v * pow2i(kDetHi - E));
    if (!(fabs(h) < 4.0e18)) { *poison = 1; return; }   // non-finite Jacobian entry (or beyond the bound)
    const double r = fma(-h, pow2i(E - kDetHi), v);      // exact
    const long long hi = (long long)h;
    const long long lo = __double2ll_rn(r * pow2i(kDetHi + kDetLo - E));
    if (hi) atomicAdd(reinterpret_cast<unsigned long long *>(acc), (unsigned long long)hi);
    if (lo) atomicAdd(reinterpret_cast<unsigned long long *>(acc) + 1, (unsigned long long)lo);
}
__device__ __forceinline__ double det_value(const long long *acc, int E) {
    return fma((double)acc[0], pow2i(E - kDetHi), (double)acc[1] * pow2i(E - kDetHi - kDetLo));
}

struct DetBufs {
    const unsigned long long *cmax;  // [n + 1] bit patterns of the column maxima, [n] = max |f|
    long long *Ui, *gEi, *Vi, *Wci;  // two limbs per entry; Wci [KE][ncs] (contended static columns of W)
    const int32_t *cs_of_slot;       // [ns] static slot -> contended column index or -1
    int ncs;
    int *poison;
};

// column maxima: cmax[pid] = max over the observations of max(|jx|, |jy|); cmax[n] = max |f|
__global__ void __launch_bounds__(256) k_colmax(DevScene s, LmDims d, const double *__restrict__ jac,
                                                const double *__restrict__ fvec, unsigned long long *cmax) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < s.nobs;
    int marker = -1, frame = 0;
    if (live) {
        if (s.dense_grid) { marker = i / s.F; frame = i - marker * s.F; }
        else { marker = s.err_marker[i]; frame = s.err_frame[i]; }
    }
    const int m0 = __shfl_sync(0xffffffffu, marker, 0);
    const bool uniform = __all_sync(0xffffffffu, marker == m0 || !live) && m0 >= 0;
    const int lane = threadIdx.x & 31;
    const int pm = live ? marker : (uniform ? m0 : 0);
    const int k = (live || uniform) ? s.mkr_k[pm] : 0;
    const PlanEntryD *pl = s.plan + (size_t)pm * s.kmax;
    const size_t m = (size_t)d.m;
    double fa = live ? fmax(fabs(fvec[2 * (size_t)i]), fabs(fvec[2 * (size_t)i + 1])) : 0.0;
    if (live && (fvec[2 * (size_t)i] != fvec[2 * (size_t)i] || fvec[2 * (size_t)i + 1] != fvec[2 * (size_t)i + 1]))
        fa = __longlong_as_double(0x7ff8000000000000ll);
    for (int c = 0; c < k; ++c) {
        const int a = pl[c].attr;
        const bool an = s.attr_anim[a];
        double v = 0.0;
        if (live) {
            const double jx = jac[(size_t)c * m + 2 * (size_t)i], jy = jac[(size_t)c * m + 2 * (size_t)i + 1];
            v = (jx != jx || jy != jy) ? __longlong_as_double(0x7ff8000000000000ll) : fmax(fabs(jx), fabs(jy));
        }
        if (!an && uniform) {
            // one static parameter for the whole warp
            unsigned long long b = (unsigned long long)__double_as_longlong(v);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, b, o); b = t > b ? t : b; }
            // most warps find a larger maximum already there: read first (a stale read only costs an
            // atomic that changes nothing), so the few shared addresses are not hammered by every warp
            unsigned long long *dst = &cmax[s.attr_param_base[a]];
            if (lane == 0 && b > *reinterpret_cast<volatile unsigned long long *>(dst)) atomicMax(dst, b);
        } else if (live && v != 0.0) {
            unsigned long long *dst = &cmax[s.attr_param_base[a] + (an ? frame : 0)];
            const unsigned long long b = (unsigned long long)__double_as_longlong(v);
            if (b > *reinterpret_cast<volatile unsigned long long *>(dst)) atomicMax(dst, b);
        }
    }
    unsigned long long fb = (unsigned long long)__double_as_longlong(fa);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, fb, o); fb = t > fb ? t : fb; }
    if (lane == 0 && fb > *reinterpret_cast<volatile unsigned long long *>(&cmax[d.n])) atomicMax(&cmax[d.n], fb);
}

__global__ void __launch_bounds__(256) k_assemble(DevScene s, LmDims d, const double *__restrict__ jac,
                                                  const double *__restrict__ fvec,
                                                  const int32_t *__restrict__ attr_slot, DetBufs db, double *W,
                                                  unsigned shared_mask) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < s.nobs;
    int marker = -1, frame = 0;
    if (live) {
        if (s.dense_grid) { marker = i / s.F; frame = i - marker * s.F; }
        else { marker = s.err_marker[i]; frame = s.err_frame[i]; }
    }
    // do all lanes of this warp look at the same marker (-> same static cols)?
    const int m0 = __shfl_sync(0xffffffffu, marker, 0);
    const bool uniform = __all_sync(0xffffffffu, marker == m0 || !live) && m0 >= 0;
    const int lane = threadIdx.x & 31;
    const int pm = live ? marker : (uniform ? m0 : 0);
    const int k = (live || uniform) ? s.mkr_k[pm] : 0;
    const PlanEntryD *pl = s.plan + (size_t)pm * s.kmax;
    const size_t m = (size_t)d.m;
    const double fx = live ? fvec[2 * (size_t)i] : 0.0, fy = live ? fvec[2 * (size_t)i + 1] : 0.0;
    const int ef = det_exp_of(db.cmax[d.n]);
    for (int c1 = 0; c1 < k; ++c1) {
        const int a1 = pl[c1].attr;
        const bool an1 = s.attr_anim[a1];
        const int s1 = attr_slot[a1];
        const int e1 = det_exp_of(db.cmax[s.attr_param_base[a1] + (an1 ? frame : 0)]);
        const double jx1 = live ? jac[(size_t)c1 * m + 2 * (size_t)i] : 0.0;
        const double jy1 = live ? jac[(size_t)c1 * m + 2 * (size_t)i + 1] : 0.0;
        const bool sh1 = (shared_mask >> c1) & 1u;
        // gradient
        const double gv = jx1 * fx + jy1 * fy;
        const int eg = det_exp(e1, ef);
        if (sh1) {
        } else if (an1) {
            if (live) det_add(&db.gEi[2 * ((size_t)frame * d.na + s1)], gv, eg, db.poison);
        } else if (uniform) {
            const double r = warp_sum(gv);
            if (lane == 0) det_add(&db.Vi[2 * ((size_t)d.ns * d.NSP + s1)], r, eg, db.poison);
        } else if (live) {
            det_add(&db.Vi[2 * ((size_t)d.ns * d.NSP + s1)], gv, eg, db.poison);
        }
        for (int c2 = 0; c2 <= c1; ++c2) {
            if (sh1 && ((shared_mask >> c2) & 1u)) continue;
            const int a2 = pl[c2].attr;
            const bool an2 = s.attr_anim[a2];
            const int s2 = attr_slot[a2];
            const int E = det_exp(e1, det_exp_of(db.cmax[s.attr_param_base[a2] + (an2 ? frame : 0)]));
            const double jx2 = live ? jac[(size_t)c2 * m + 2 * (size_t)i] : 0.0;
            const double jy2 = live ? jac[(size_t)c2 * m + 2 * (size_t)i + 1] : 0.0;
            const double v = jx1 * jx2 + jy1 * jy2;
            if (an1 && an2) {
                if (live) {
                    const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                    det_add(&db.Ui[2 * (((size_t)frame * d.na + r) * d.na + cc)], v, E, db.poison);
                }
            } else if (an1 != an2) {
                if (live) {
                    const int ja = an1 ? s1 : s2, js = an1 ? s2 : s1;
                    const int cs = db.cs_of_slot[js];
                    if (cs >= 0) det_add(&db.Wci[2 * (((size_t)frame * d.na + ja) * db.ncs + cs)], v, E, db.poison);
                    else atomicAdd(&W[((size_t)frame * d.na + ja) * d.NSP + js], v);   // <= 2 contributions
                }
            } else {
                const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                if (uniform) {
                    const double t = warp_sum(v);
                    if (lane == 0) det_add(&db.Vi[2 * ((size_t)r * d.NSP + cc)], t, E, db.poison);
                } else if (live) {
                    det_add(&db.Vi[2 * ((size_t)r * d.NSP + cc)], v, E, db.poison);
                }
            }
        }
    }
}

// Products among the NS shared columns (and their gradient entries) for dense-grid scenes
// (observation i = marker * F + frame).  CTA = (32 frames, group of mt markers), lane = frame,
// 4 warps; the NS (NS + 1) / 2 + NS sums are dealt to the warps, every warp walks all markers of
// the group with its share of the sums in registers.  Per-frame sums (animated x animated ->
// U, animated x static -> W, animated gradient -> gE) end as one atomic per CTA, frame and entry;
// static x static sums are reduced over the frames of the warp first.
template <int NS, int WID>
__device__ __forceinline__ void assemble_shared_body(const AsmShared &sh, const LmDims &d, const double2 *__restrict__ jac2,
                                                     const double2 *__restrict__ f2, int F, int frame, bool live,
                                                     int m_begin, int m_end, const DetBufs &db, double *W) {
    constexpr int P = NS * (NS + 1) / 2 + NS;   // pairs (c1 >= c2) then gradient entries
    constexpr int PW = (P + 3) / 4;             // sums per warp
    constexpr int P0 = WID * PW, P1 = (P0 + PW < P) ? P0 + PW : P;
    double acc[PW];
#pragma unroll
    for (int u = 0; u < PW; ++u) acc[u] = 0.0;
    const size_t mh = (size_t)d.m / 2;          // double2 elements per Jacobian column
    for (int mk = m_begin; mk < m_end; ++mk) {
        const size_t i = (size_t)mk * F + frame;
        double2 jv[NS];
#pragma unroll
        for (int c = 0; c < NS; ++c) jv[c] = live ? __ldg(&jac2[(size_t)sh.col[c] * mh + i]) : make_double2(0.0, 0.0);
        const double2 fv = live ? __ldg(&f2[i]) : make_double2(0.0, 0.0);
        int pidx = 0;
#pragma unroll
        for (int c1 = 0; c1 < NS; ++c1)
#pragma unroll
            for (int c2 = 0; c2 <= c1; ++c2, ++pidx)
                if (pidx >= P0 && pidx < P1) acc[pidx - P0] = fma(jv[c1].x, jv[c2].x, fma(jv[c1].y, jv[c2].y, acc[pidx - P0]));
#pragma unroll
        for (int c1 = 0; c1 < NS; ++c1, ++pidx)
            if (pidx >= P0 && pidx < P1) acc[pidx - P0] = fma(jv[c1].x, fv.x, fma(jv[c1].y, fv.y, acc[pidx - P0]));
    }
    const int lane = threadIdx.x & 31;
    const int ef = det_exp_of(db.cmax[d.n]);
    int ex[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c) ex[c] = det_exp_of(db.cmax[sh.pbase[c] + (sh.anim[c] ? frame : 0)]);
    int pidx = 0;
#pragma unroll
    for (int c1 = 0; c1 < NS; ++c1)
#pragma unroll
        for (int c2 = 0; c2 <= c1; ++c2, ++pidx)
            if (pidx >= P0 && pidx < P1) {
                const double v = acc[pidx - P0];
                const int s1 = sh.slot[c1], s2 = sh.slot[c2];
                const bool an1 = sh.anim[c1], an2 = sh.anim[c2];
                const int E = det_exp(ex[c1], ex[c2]);
                if (an1 && an2) {
                    const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                    if (live) det_add(&db.Ui[2 * (((size_t)frame * d.na + r) * d.na + cc)], v, E, db.poison);
                } else if (an1 != an2) {
                    const int ja = an1 ? s1 : s2, js = an1 ? s2 : s1;
                    if (live) {
                        const int cs = db.cs_of_slot[js];
                        if (cs >= 0) det_add(&db.Wci[2 * (((size_t)frame * d.na + ja) * db.ncs + cs)], v, E, db.poison);
                        else atomicAdd(&W[((size_t)frame * d.na + ja) * d.NSP + js], v);
                    }
                } else {
                    const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                    const double t = warp_sum(v);
                    if (lane == 0) det_add(&db.Vi[2 * ((size_t)r * d.NSP + cc)], t, E, db.poison);
                }
            }
#pragma unroll
    for (int c1 = 0; c1 < NS; ++c1, ++pidx)
        if (pidx >= P0 && pidx < P1) {
            const double v = acc[pidx - P0];
            const int E = det_exp(ex[c1], ef);
            if (sh.anim[c1]) {
                if (live) det_add(&db.gEi[2 * ((size_t)frame * d.na + sh.slot[c1])], v, E, db.poison);
            } else {
                const double t = warp_sum(v);
                if (lane == 0) det_add(&db.Vi[2 * ((size_t)d.ns * d.NSP + sh.slot[c1])], t, E, db.poison);
            }
        }
}

template <int NS>
__global__ void __launch_bounds__(128) k_assemble_shared(AsmShared sh, LmDims d, int F, int M, int mt,
                                                         const double *__restrict__ jac, const double *__restrict__ fvec,
                                                         DetBufs db, double *W) {
    const int frame = blockIdx.x * 32 + (threadIdx.x & 31);
    const bool live = frame < F;
    const int m_begin = blockIdx.y * mt, m_end = min(M, m_begin + mt);
    const double2 *jac2 = reinterpret_cast<const double2 *>(jac);
    const double2 *f2 = reinterpret_cast<const double2 *>(fvec);
    const int fr = live ? frame : 0;
    switch (threadIdx.x >> 5) {
        case 0: assemble_shared_body<NS, 0>(sh, d, jac2, f2, F, fr, live, m_begin, m_end, db, W); break;
        case 1: assemble_shared_body<NS, 1>(sh, d, jac2, f2, F, fr, live, m_begin, m_end, db, W); break;
        case 2: assemble_shared_body<NS, 2>(sh, d, jac2, f2, F, fr, live, m_begin, m_end, db, W); break;
        default: assemble_shared_body<NS, 3>(sh, d, jac2, f2, F, fr, live, m_begin, m_end, db, W); break;
    }
}

// Limbs -> FP64: U (lower triangles), gE, V (lower triangle + right-hand side row ns, the rest
// zero) and the contended static columns of W.  One thread per entry of the four regions.
__global__ void __launch_bounds__(256) k_assemble_finalize(LmDims d, DetBufs db, const int32_t *__restrict__ slot_of_cs,
                                                           const int32_t *__restrict__ i2p, double *__restrict__ U,
                                                           double *__restrict__ gE, double *__restrict__ V,
                                                           double *__restrict__ W) {
    const size_t nU = (size_t)d.F * d.na * d.na, nG = (size_t)d.KE, nV = (size_t)d.NSP * d.NSP, nW = (size_t)d.KE * db.ncs;
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const bool bad = *db.poison != 0;
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    const int ef = det_exp_of(db.cmax[d.n]);
    auto ex = [&](int q_int) { return det_exp_of(db.cmax[i2p[q_int]]); };
    if (t < nU) {
        const int f = (int)(t / ((size_t)d.na * d.na)), rc = (int)(t - (size_t)f * d.na * d.na);
        const int r = rc / d.na, c = rc - r * d.na;
        double v = 0.0;
        if (c <= r) {
            const long long *a = &db.Ui[2 * t];
            if (a[0] | a[1]) v = det_value(a, det_exp(ex(f * d.na + r), ex(f * d.na + c)));
            if (bad) v = nan;
        }
        U[t] = v;
        return;
    }
    t -= nU;
    if (t < nG) {
        const long long *a = &db.gEi[2 * t];
        double v = (a[0] | a[1]) ? det_value(a, det_exp(ex((int)t), ef)) : 0.0;
        gE[t] = bad ? nan : v;
        return;
    }
    t -= nG;
    if (t < nV) {
        const int r = (int)(t / d.NSP), c = (int)(t - (size_t)r * d.NSP);
        double v = 0.0;
        if (c <= r && r <= d.ns && c < d.ns) {
            const long long *a = &db.Vi[2 * t];
            if (a[0] | a[1]) v = det_value(a, det_exp(ex(d.KE + c), r < d.ns ? ex(d.KE + r) : ef));
            if (bad) v = nan;
        }
        V[t] = v;
        return;
    }
    t -= nV;
    if (t < nW) {
        const size_t row = t / db.ncs;
        const int cs = (int)(t - row * db.ncs), js = slot_of_cs[cs];
        const long long *a = &db.Wci[2 * t];
        double v = (a[0] | a[1]) ? det_value(a, det_exp(ex((int)row), ex(d.KE + js))) : 0.0;
        W[row * d.NSP + js] = bad ? nan : v;
    }
}

// ---- stiffness / smoothness rows in the normal equations --------------------------------------
// Each of the few extra rows has ONE non-zero Jacobian entry jext[e] in column pid[e]: it adds
// jext^2 to that parameter's diagonal entry, jext * row to its gradient entry, (jext p)^2 to ||J p||^2.
__global__ void k_extra_colmax(int n_extra, int n, const double *jext, const int32_t *pid, const double *frow,
                               unsigned long long *cmax) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_extra) return;
    const double j = jext[e], r = frow[e];
    const unsigned long long nanb = 0x7ff8000000000000ull;
    atomicMax(&cmax[pid[e]], j != j ? nanb : (unsigned long long)__double_as_longlong(fabs(j)));
    atomicMax(&cmax[n], r != r ? nanb : (unsigned long long)__double_as_longlong(fabs(r)));
}
__global__ void k_extra_assemble(LmDims d, int n_extra, const double *jext, const int32_t *pid, const double *frow,
                                 const int32_t *p2i, DetBufs db) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_extra) return;
    const double j = jext[e], r = frow[e];
    const int ej = det_exp_of(db.cmax[pid[e]]), ef = det_exp_of(db.cmax[d.n]);
    const int q = p2i[pid[e]];
    if (q < d.KE) {
        const int f = q / d.na, a = q - f * d.na;
        det_add(&db.Ui[2 * (((size_t)f * d.na + a) * d.na + a)], j * j, det_exp(ej, ej), db.poison);
        det_add(&db.gEi[2 * (size_t)q], j * r, det_exp(ej, ef), db.poison);
    } else {
        const int sl = q - d.KE;
        det_add(&db.Vi[2 * ((size_t)sl * d.NSP + sl)], j * j, det_exp(ej, ej), db.poison);
        det_add(&db.Vi[2 * ((size_t)d.ns * d.NSP + sl)], j * r, det_exp(ej, ef), db.poison);
    }
}
__global__ void k_extra_jp(int n_extra, const double *jext, const int32_t *pid, const double *p_par, double *part_jp) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double a = 0.0;
    for (int e = 0; e < n_extra; ++e) { const double t = jext[e] * p_par[pid[e]]; a += t * t; }
    part_jp[0] += a;
}

// ||J p||^2 partials straight from the block Jacobian (MINPACK uses R P^T p).
// p is indexed by parameter id (reference order).
__global__ void __launch_bounds__(256) k_jp_partial(DevScene s, LmDims d, const double *__restrict__ jac,
                                                    const double *__restrict__ p, double *__restrict__ partial) {
    __shared__ double sh[32];
    double acc = 0.0;
    const size_t m = (size_t)d.m;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < s.nobs; i += gridDim.x * blockDim.x) {
        int marker, frame;
        if (s.dense_grid) { marker = i / s.F; frame = i - marker * s.F; }
        else { marker = s.err_marker[i]; frame = s.err_frame[i]; }
        const int k = s.mkr_k[marker];
        const PlanEntryD *pl = s.plan + (size_t)marker * s.kmax;
        double sx = 0.0, sy = 0.0;
        for (int c = 0; c < k; ++c) {
            const int a = pl[c].attr;
            const int pid = s.attr_param_base[a] + (s.attr_anim[a] ? frame : 0);
            const double pv = p[pid];
            sx += jac[(size_t)c * m + 2 * (size_t)i] * pv;
            sy += jac[(size_t)c * m + 2 * (size_t)i + 1] * pv;
        }
        acc += sx * sx + sy * sy;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// ---- per-frame block factorisation ---------------------------------------------
// One CTA per frame:  L_f = chol(U_f + par D_f^2);  Y_f = L_f^-1 W_f ;
// z_f = L_f^-1 g_f stored in the augmented column ns of Y.
// The same for a compile-time block size (a camera: NA = 6): every thread factors the block in
// registers (no barriers, reciprocal pivots), then streams its columns of W through the forward
// substitution -- the kernel is then the HBM stream W -> Y it should be.
template <int NA>
__global__ void __launch_bounds__(128) k_block_factor_fixed(LmDims d, LmState *st, const double *__restrict__ U,
                                                            const double *__restrict__ gE, const double *__restrict__ W,
                                                            const double *__restrict__ diag, const int32_t *__restrict__ i2p,
                                                            double *__restrict__ Lf, double *__restrict__ Y) {
    const int f = blockIdx.x;
    const double par = st->par_cur;
    double L[NA][NA], inv[NA], z[NA];
#pragma unroll
    for (int r = 0; r < NA; ++r) {
#pragma unroll
        for (int c = 0; c <= r; ++c) L[r][c] = U[((size_t)f * NA + r) * NA + c];
        const double dd = diag[i2p[f * NA + r]];
        L[r][r] += par * dd * dd;
        z[r] = gE[(size_t)f * NA + r];
    }
    int fail = 0;
#pragma unroll
    for (int j = 0; j < NA; ++j) {
        double pv = L[j][j];
        if (!(pv > 0.0)) { fail = 1; pv = 1.0; }
        else pv = sqrt(pv);
        L[j][j] = pv;
        inv[j] = 1.0 / pv;
#pragma unroll
        for (int r = j + 1; r < NA; ++r) L[r][j] *= inv[j];
#pragma unroll
        for (int c = j + 1; c < NA; ++c)
#pragma unroll
            for (int r = c; r < NA; ++r) L[r][c] -= L[r][j] * L[c][j];
    }
#pragma unroll
    for (int r = 0; r < NA; ++r) {
        double v = z[r];
#pragma unroll
        for (int c = 0; c < r; ++c) v -= L[r][c] * z[c];
        z[r] = v * inv[r];
    }
    if (threadIdx.x == 0) {
        if (fail) atomicExch(&st->chol_fail, 1);
#pragma unroll
        for (int r = 0; r < NA; ++r)
#pragma unroll
            for (int c = 0; c < NA; ++c) Lf[(size_t)f * NA * NA + r * NA + c] = c <= r ? L[r][c] : 0.0;
    }
    for (int j = threadIdx.x; j < d.NSP; j += blockDim.x) {
        double y[NA];
        if (j < d.ns) {
#pragma unroll
            for (int r = 0; r < NA; ++r) y[r] = W[((size_t)f * NA + r) * d.NSP + j];
#pragma unroll
            for (int r = 0; r < NA; ++r) {
                double v = y[r];
#pragma unroll
                for (int c = 0; c < r; ++c) v -= L[r][c] * y[c];
                y[r] = v * inv[r];
            }
        } else {
#pragma unroll
            for (int r = 0; r < NA; ++r) y[r] = (j == d.ns) ? z[r] : 0.0;
        }
#pragma unroll
        for (int r = 0; r < NA; ++r) Y[((size_t)f * NA + r) * d.NSP + j] = y[r];
    }
}

__global__ void __launch_bounds__(128) k_block_factor(LmDims d, LmState *st, const double *__restrict__ U,
                                                      const double *__restrict__ gE, const double *__restrict__ W,
                                                      const double *__restrict__ diag, const int32_t *__restrict__ i2p,
                                                      double *__restrict__ Lf, double *__restrict__ Y) {
    extern __shared__ double sm[];
    const int na = d.na, f = blockIdx.x;
    double *A = sm;             // na*na
    double *z = sm + na * na;   // na
    __shared__ int fail;
    const double par = st->par_cur;
    if (threadIdx.x == 0) fail = 0;
    for (int t = threadIdx.x; t < na * na; t += blockDim.x) {
        const int r = t / na, c = t - r * na;
        double v = (c <= r) ? U[((size_t)f * na + r) * na + c] : 0.0;
        if (r == c) {
            const double dd = diag[i2p[f * na + r]];
            v += par * dd * dd;
        }
        A[t] = v;
    }
    for (int t = threadIdx.x; t < na; t += blockDim.x) z[t] = gE[(size_t)f * na + t];
    __syncthreads();
    // right-looking Cholesky in shared memory
    for (int j = 0; j < na; ++j) {
        if (threadIdx.x == 0) {
            const double pv = A[j * na + j];
            if (!(pv > 0.0)) { fail = 1; A[j * na + j] = 1.0; }
            else A[j * na + j] = sqrt(pv);
        }
        __syncthreads();
        const double dj = A[j * na + j];
        for (int r = j + 1 + threadIdx.x; r < na; r += blockDim.x) A[r * na + j] /= dj;
        __syncthreads();
        const int rem = na - j - 1;
        for (int t = threadIdx.x; t < rem * rem; t += blockDim.x) {
            const int r = j + 1 + t / rem, c = j + 1 + t % rem;
            if (c <= r) A[r * na + c] -= A[r * na + j] * A[c * na + j];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (fail) atomicExch(&st->chol_fail, 1);
        // z = L^-1 g
        for (int r = 0; r < na; ++r) {
            double v = z[r];
            for (int c = 0; c < r; ++c) v -= A[r * na + c] * z[c];
            z[r] = v / A[r * na + r];
        }
    }
    for (int t = threadIdx.x; t < na * na; t += blockDim.x) Lf[(size_t)f * na * na + t] = A[t];
    __syncthreads();
    // Y columns: forward substitution, one thread per column
    for (int j = threadIdx.x; j < d.NSP; j += blockDim.x) {
        double y[kMaxNA];
        if (j < d.ns) {
            for (int r = 0; r < na; ++r) {
                double v = W[((size_t)f * na + r) * d.NSP + j];
                for (int c = 0; c < r; ++c) v -= A[r * na + c] * y[c];
                y[r] = v / A[r * na + r];
            }
        } else {
            for (int r = 0; r < na; ++r) y[r] = (j == d.ns) ? z[r] : 0.0;
        }
        for (int r = 0; r < na; ++r) Y[((size_t)f * na + r) * d.NSP + j] = y[r];
    }
}

// S = V (lower, incl. rhs row ns) + par D_S^2 on the diagonal; padding = identity.
// damp = 0 (sharded solve): the shard's part of V only; the damping and the padding
// diagonal are applied by k_schur_damp after the all-reduce of S.
__global__ void k_schur_init(LmDims d, const LmState *st, const double *__restrict__ V,
                             const double *__restrict__ diag, const int32_t *__restrict__ i2p,
                             double *__restrict__ S, int damp) {
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= (size_t)d.NSP * d.NSP) return;
    const int r = (int)(t / d.NSP), c = (int)(t - (size_t)r * d.NSP);
    double v = 0.0;
    if (c <= r && r <= d.ns) v = V[t];
    if (r == c) {
        if (r < d.ns) {
            const double dd = diag[i2p[d.KE + r]];
            if (damp) v += st->par_cur * dd * dd;
        } else if (r == d.ns) v = 0.0;  // augmented pivot: unused
        else v = damp ? 1.0 : 0.0;
    }
    S[t] = v;
}
__global__ void k_schur_damp(LmDims d, const LmState *st, const double *__restrict__ diag,
                             const int32_t *__restrict__ i2p, double *__restrict__ S) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= d.NSP) return;
    double *s = &S[(size_t)r * d.NSP + r];
    if (r < d.ns) {
        const double dd = diag[i2p[d.KE + r]];
        *s += st->par_cur * dd * dd;
    } else *s = (r == d.ns) ? 0.0 : 1.0;
}

// Sharded solve: only the lower triangle of S (by 64-wide block rows: row r keeps columns
// 0 .. 64 (r / 64 + 1) - 1) and its right-hand side row ns travel over NVLink.  Packed length:
__host__ __device__ inline size_t tri_pack_offset(int r) {   // first packed element of row r
    const size_t b = (size_t)(r >> 6), within = (size_t)(r & 63);
    return 64 * 64 * (b * (b + 1) / 2) + within * 64 * (b + 1);
}
__host__ __device__ inline size_t tri_pack_count(int rows) { return tri_pack_offset(rows); }
// pack = 1: S -> buf; pack = 0: buf -> S, and the damping / padding diagonal of k_schur_damp
__global__ void __launch_bounds__(256) k_tri_pack(LmDims d, const LmState *st, const double *__restrict__ diag,
                                                  const int32_t *__restrict__ i2p, double *__restrict__ S,
                                                  double *__restrict__ buf, int rows, int pack) {
    const int r = blockIdx.x;
    if (r < rows) {
        const int width = 64 * ((r >> 6) + 1);
        double *row = S + (size_t)r * d.NSP;
        double *dst = buf + tri_pack_offset(r);
        if (pack) { for (int c = threadIdx.x; c < width; c += blockDim.x) dst[c] = row[c]; }
        else { for (int c = threadIdx.x; c < width; c += blockDim.x) row[c] = dst[c]; }
    }
    __syncthreads();   // the diagonal entry of this row is in place before it is damped
    if (!pack && threadIdx.x == 0 && r < d.NSP) {
        double *sd = &S[(size_t)r * d.NSP + r];
        if (r < d.ns) { const double dd = diag[i2p[d.KE + r]]; *sd += st->par_cur * dd * dd; }
        else *sd = (r == d.ns) ? 0.0 : 1.0;
    }
}

// ---- DMMA helpers ------------------------------------------------------------------
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int kTS = 68;  // shared tile row stride (doubles): conflict-free fragments

// acc[4][4][2] += sA^T sB over KT k-steps; sA/sB are [KT][kTS] (k-major) tiles of
// 64 columns.  Warp (wr, wc) owns rows wr*32.., cols wc*32.. of the 64x64 tile.
template <int KT>
__device__ __forceinline__ void tile_mma(const double *sA, const double *sB, int wr, int wc, int lane,
                                         double (&acc)[4][4][2]) {
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int k0 = 0; k0 < KT; k0 += 4) {
        double a[4], b[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            a[t] = sA[(k0 + q) * kTS + wr * 32 + t * 8 + g];
            b[t] = sB[(k0 + q) * kTS + wc * 32 + t * 8 + g];
        }
#pragma unroll
        for (int ti = 0; ti < 4; ++ti)
#pragma unroll
            for (int tj = 0; tj < 4; ++tj) dmma884(acc[ti][tj][0], acc[ti][tj][1], a[ti], b[tj]);
    }
}

// C(64x64 tile at bi,bj) -= acc ; row-major C with leading dimension ld
__device__ __forceinline__ void tile_store_sub(double *C, int ld, int bi, int bj, int wr, int wc, int lane,
                                               const double (&acc)[4][4][2]) {
    const int g = lane >> 2, q = lane & 3;
#pragma unroll
    for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = 0; tj < 4; ++tj) {
            const int r = bi * 64 + wr * 32 + ti * 8 + g;
            const int c = bj * 64 + wc * 32 + tj * 8 + q * 2;
            double2 *dst = reinterpret_cast<double2 *>(&C[(size_t)r * ld + c]);
            double2 v = *dst;
            v.x -= acc[ti][tj][0];
            v.y -= acc[ti][tj][1];
            *dst = v;
        }
}

__device__ __forceinline__ void tile_index(int t, int &bi, int &bj) {
    // linear index over the lower triangle (bi >= bj), row by row
    int r = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
    while ((r + 1) * (r + 2) / 2 <= t) ++r;
    while (r * (r + 1) / 2 > t) --r;
    bi = r;
    bj = t - r * (r + 1) / 2;
}

// S(lower tiles) -= Y^T Y, Y is [KEP][NSP] row-major (k-major operands).
// grid.x = lower-triangle tiles of the NSP/64 block grid; grid.y = K splits.  The CTAs of one
// tile (its K splits) form a thread-block CLUSTER: every CTA leaves its partial tile in its own
// shared memory and the cluster folds them through distributed shared memory in rank order,
// each CTA owning a slice of the tile -- a fixed summation order (two runs give the same bits,
// where the FP atomics of the earlier version did not) and no extra trip through L2.
// Sharded solve with the Cholesky beside the SYRK: the reduced system is the sum over the ranks of
// their partial tiles.  Every rank's SYRK stores its partial tile into EVERY rank's peer-mapped
// arena (posted stores over NVLink; two buffers, alternating with the factorisation sequence
// number) and then the tile's flag; every rank's Cholesky, when it comes to a tile, waits for the
// flags of all ranks and adds the ranks' partial tiles in rank order, all from its own memory -- the
// all-reduce of the reduced system, tile by tile, inside the factorisation, with no remote load on
// the factorisation's chain.  (Pulling the partial tiles from the peers' arenas instead was
// measured: at 8 GPUs the serial remote loads cost more than the packed all-reduce they replaced.)
struct PeerTiles {
    int world, rank;           // world == 0: not in use
    int ntiles, maxtiles;      // tiles of this system; tiles a flag array has room for
    double *tiles[8];          // every rank's [2][world source ranks][ntiles][64 x 64]
    unsigned *flags[8];        // every rank's [2][8 source ranks][maxtiles]
    double *my_tiles;          // = tiles[rank], flags[rank] (no dynamic indexing of the parameter arrays)
    unsigned *my_flags;
    const unsigned *seq;       // this rank's factorisation sequence number: the epoch of its flags
    const double *diag;        // damping of the diagonal (added once, after the sum)
    const int32_t *i2p;
};

// Tiles are numbered column by column, as k_chol_tiles numbers them, and (ready != nullptr) each
// finished tile is published with the epoch the factorisation of this step waits for
// (ready[bj * nb + bi] = *epoch_ctr + 1): the tile-dataflow Cholesky runs BESIDE this kernel and
// picks every tile up as soon as it is complete, instead of starting when the last one is.
constexpr int kSyrkKT = 16;
template <bool PEER>
__global__ void __launch_bounds__(128, 4) k_schur_syrk(LmDims d, const double *__restrict__ Y, double *__restrict__ S,
                                                       int *ready, const int *epoch_ctr, PeerTiles pt) {
    __shared__ __align__(16) double smem[4 * kSyrkKT * kTS];   // sA[2], sB[2]; reused as the 64x64 partial tile
    double *sA[2] = {smem, smem + kSyrkKT * kTS};
    double *sB[2] = {smem + 2 * kSyrkKT * kTS, smem + 3 * kSyrkKT * kTS};
    const int nb = d.NSP / 64;
    int bi, bj = 0;
    {   // column-major tile numbering: column j holds nb - j tiles
        int t = blockIdx.x;
        while (t >= nb - bj) { t -= nb - bj; ++bj; }
        bi = bj + t;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wr = warp >> 1, wc = warp & 1;
    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
    const int ksteps = d.KEP / kSyrkKT;
    const int per = (ksteps + gridDim.y - 1) / gridDim.y;
    const int ks0 = blockIdx.y * per, ks1 = min(ksteps, ks0 + per);
    // each thread moves 8 doubles of each operand tile per step (16 rows x 64 cols).  The
    // global loads of step ks+1 are issued before the DMMAs of step ks and land in shared
    // memory after them, so the L2 latency hides behind the tensor pipe.
    double2 ra[4], rb[4];
    auto fetch = [&](int ks) {
        const int kbase = ks * kSyrkKT;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int idx = threadIdx.x + t * 128;  // 0..511 double2 slots
            const int row = idx >> 5, c2 = (idx & 31) * 2;
            ra[t] = *reinterpret_cast<const double2 *>(&Y[(size_t)(kbase + row) * d.NSP + bi * 64 + c2]);
            if (bi != bj) rb[t] = *reinterpret_cast<const double2 *>(&Y[(size_t)(kbase + row) * d.NSP + bj * 64 + c2]);
        }
    };
    auto commit = [&](int buf) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int idx = threadIdx.x + t * 128;
            const int row = idx >> 5, c2 = (idx & 31) * 2;
            *reinterpret_cast<double2 *>(&sA[buf][row * kTS + c2]) = ra[t];
            if (bi != bj) *reinterpret_cast<double2 *>(&sB[buf][row * kTS + c2]) = rb[t];
        }
    };
    if (ks0 < ks1) { fetch(ks0); commit(0); }
    __syncthreads();
    for (int ks = ks0; ks < ks1; ++ks) {
        const int buf = (ks - ks0) & 1;
        if (ks + 1 < ks1) fetch(ks + 1);
        tile_mma<kSyrkKT>(sA[buf], bi != bj ? sB[buf] : sA[buf], wr, wc, lane, acc);
        if (ks + 1 < ks1) commit(buf ^ 1);
        __syncthreads();
    }
    // peer mode: the result tile (S - Y^T Y of this rank) goes to this rank's arena, S stays as it is
    unsigned pepoch = 0;
    size_t pslot = 0;   // this rank's slot of the tile in every arena
    if (PEER) {
        pepoch = __ldcg(pt.seq);
        pslot = (((size_t)(pepoch & 1u) * pt.world + pt.rank) * pt.ntiles + blockIdx.x) * 4096;
    }
    auto peer_publish = [&]() {   // one thread, after every writer of the tile has fenced and arrived
        __threadfence_system();
#pragma unroll
        for (int q = 0; q < 8; ++q)
            if (q < pt.world) {
                unsigned *f = pt.flags[q] + ((size_t)(pepoch & 1u) * 8 + pt.rank) * pt.maxtiles + blockIdx.x;
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f), "r"(pepoch) : "memory");
            }
    };
    if (gridDim.y == 1) {
        if (PEER) {
            const int g = lane >> 2, q = lane & 3;
#pragma unroll
            for (int ti = 0; ti < 4; ++ti)
#pragma unroll
                for (int tj = 0; tj < 4; ++tj) {
                    const int r = wr * 32 + ti * 8 + g, c = wc * 32 + tj * 8 + q * 2;
                    const double2 v = *reinterpret_cast<const double2 *>(&S[(size_t)(bi * 64 + r) * d.NSP + bj * 64 + c]);
                    const double2 o = make_double2(v.x - acc[ti][tj][0], v.y - acc[ti][tj][1]);
#pragma unroll
                    for (int z = 0; z < 8; ++z)
                        if (z < pt.world) *reinterpret_cast<double2 *>(pt.tiles[z] + pslot + r * 64 + c) = o;
                }
            __threadfence_system();
            __syncthreads();
            if (threadIdx.x == 0) peer_publish();
            return;
        }
        tile_store_sub(S, d.NSP, bi, bj, wr, wc, lane, acc);
        if (!PEER && ready) {
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) {
                const int epoch = __ldcg(epoch_ctr) + 1;
                asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(&ready[bj * nb + bi]), "r"(epoch) : "memory");
            }
        }
        return;
    }
    // partial tile -> own shared memory, value v of thread t at [v][t] (conflict-free)
#pragma unroll
    for (int ti = 0; ti < 4; ++ti)
#pragma unroll
        for (int tj = 0; tj < 4; ++tj) {
            smem[((ti * 4 + tj) * 2 + 0) * 128 + threadIdx.x] = acc[ti][tj][0];
            smem[((ti * 4 + tj) * 2 + 1) * 128 + threadIdx.x] = acc[ti][tj][1];
        }
    unsigned rank, nrank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(nrank));
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    // this CTA folds values [v0, v1) of every thread position over the ranks, in rank order
    const int per_rank = 32 / (int)nrank;   // nrank in {2, 4, 8}
    const int v0 = (int)rank * per_rank;
    const unsigned base = (unsigned)__cvta_generic_to_shared(smem);
    const int g = lane >> 2, q = lane & 3;
    for (int v = v0; v < v0 + per_rank; v += 2) {   // (v, v + 1): the two adjacent columns of one accumulator pair
        double sum0 = 0.0, sum1 = 0.0;
        for (unsigned r = 0; r < nrank; ++r) {
            unsigned remote;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(base + (unsigned)((v * 128 + threadIdx.x) * 8)), "r"(r));
            double x0, x1;
            asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(x0) : "r"(remote));
            asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(x1) : "r"(remote + 128u * 8u));
            sum0 += x0; sum1 += x1;
        }
        const int ti = v >> 3, tj = (v >> 1) & 3;
        const int rt = wr * 32 + ti * 8 + g, ct = wc * 32 + tj * 8 + q * 2;
        double2 *sp = reinterpret_cast<double2 *>(&S[(size_t)(bi * 64 + rt) * d.NSP + bj * 64 + ct]);
        const double2 sv = *sp;
        const double2 o = make_double2(sv.x - sum0, sv.y - sum1);
        if (PEER) {
#pragma unroll
            for (int z = 0; z < 8; ++z)
                if (z < pt.world) *reinterpret_cast<double2 *>(pt.tiles[z] + pslot + rt * 64 + ct) = o;
        }
        else *sp = o;
    }
    // nobody leaves while a peer may still read its shared memory; after the barrier every rank's
    // slice of the tile is written (fenced to the device before the arrival), so rank 0 publishes
    if (!PEER && ready) __threadfence();
    if (PEER) __threadfence_system();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (PEER && rank == 0 && threadIdx.x == 0) peer_publish();
    if (!PEER && ready && rank == 0 && threadIdx.x == 0) {
        const int epoch = __ldcg(epoch_ctr) + 1;
        __threadfence();
        asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(&ready[bj * nb + bi]), "r"(epoch) : "memory");
    }
}

// ---- blocked Cholesky of S (lower, row-major, ld = NSP) ----------------------------
// sqrt(x) and 1/sqrt(x) from the hardware seed (~22 bits) + two Newton steps (<= 2 ulp):
// the IEEE sqrt and division sequences would sit on the critical path of every pivot.
__device__ __forceinline__ void sqrt_rsqrt(double x, double &s, double &inv) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double h = 0.5 * x;
    double e = fma(-h * y, y, 0.5);
    y = fma(y, e, y);
    e = fma(-h * y, y, 0.5);
    y = fma(y, e, y);
    // 1/sqrt(x) is what the next pivot waits for: two Newton steps from the 22-bit seed
    // (<= 2 ulp); sqrt(x) gets its correction off the critical path
    inv = y;
    s = x * y;
    s = fma(0.5 * y, fma(-s, s, x), s);
}

// 1/x from the hardware seed and ONE cubically convergent step (e = 1 - x r; r += r e (1 + e)):
// three dependent operations after the seed instead of the four of two Newton steps.
__device__ __forceinline__ double rcp_fast(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    return fma(r * e, 1.0 + e, r);
}

// Cholesky of an 8x8 block held in registers (lower part of a[i][j]); g0 = global index of
// its first row.  The augmented row (global index ns) gets no pivot.
// (An LDL^T-style chain -- next pivot from 1/d_j only, square roots off the chain -- was
// measured slower here: 0.82 vs 0.72 us per block; in-order issue does not overlap the two.)
__device__ __forceinline__ void chol8_regs(double (&a)[8][8], double (&dinv)[8], int g0, int ns, int &fail) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        double sq = 1.0, inv = 1.0;
        const double pv = a[j][j];
        if (g0 + j != ns) {
            if (!(pv > 1e-290)) fail = 1;
            else sqrt_rsqrt(pv, sq, inv);
        }
        a[j][j] = sq;
        dinv[j] = inv;
#pragma unroll
        for (int i = j + 1; i < 8; ++i) a[i][j] *= inv;
#pragma unroll
        for (int c = j + 1; c < 8; ++c)
#pragma unroll
            for (int i = c; i < 8; ++i) a[i][c] = fma(-a[i][j], a[c][j], a[i][c]);
    }
}

// Step kb, kernel 1 (256 threads per CTA), 8x8 sub-blocking inside the 64x64 blocks:
// every CTA factors the diagonal block in shared memory (redundantly: the factorisation
// is the sequential part of the step, repeating it costs no time) -- per sub-block column
// every thread factors the 8x8 pivot block in registers, the rows below are solved one per
// thread and the trailing part is updated by all threads: two barriers per 8 pivots.
// CTA 0 stores the factor to `Ld`; CTA r > 0 then solves its 64-row panel block
// X L_kk^T = A_ik the same way.  The pivot of the augmented row (global index ns) is
// skipped: that row is the right-hand side being forward-solved.
constexpr int kCS = 65;   // shared stride of the 64x64 tiles (conflict-free column walks)
constexpr size_t kCholPanelSmem = sizeof(double) * (2 * 64 * kCS + 64);
__global__ void __launch_bounds__(256) k_chol_panel(LmDims d, int kb, LmState *st, double *__restrict__ S,
                                                    double *__restrict__ Ld) {
    extern __shared__ __align__(16) double sm[];
    double *T = sm;                   // diagonal block -> its factor L, row-major
    double *X = sm + 64 * kCS;        // panel block
    double *dinv_s = X + 64 * kCS;    // 1 / L[c][c]
    const int tid = threadIdx.x;
    const int base = kb * 64;
    const int rb = (kb + blockIdx.x) * 64;
    for (int t = tid; t < 64 * 64; t += 256) {
        const int rr = t >> 6, c = t & 63;
        T[rr * kCS + c] = S[(size_t)(base + rr) * d.NSP + base + c];
        if (blockIdx.x) X[rr * kCS + c] = S[(size_t)(rb + rr) * d.NSP + base + c];
    }
    __syncthreads();
    int fail = 0;
#pragma unroll
    for (int sblk = 0; sblk < 8; ++sblk) {
        const int k0 = sblk * 8, nrest = 56 - k0;
        double a[8][8], dinv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j <= i; ++j) a[i][j] = T[(k0 + i) * kCS + k0 + j];
        chol8_regs(a, dinv, base + k0, d.ns, fail);
        __syncthreads();  // every thread holds the pivot block before anyone overwrites it
        if (tid < nrest) {
            // row below the pivot block: x L_ss^T = a
            const int r = k0 + 8 + tid;
            double x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = T[r * kCS + k0 + k];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                x[k] *= dinv[k];
#pragma unroll
                for (int k2 = k + 1; k2 < 8; ++k2) x[k2] = fma(-x[k], a[k2][k], x[k2]);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) T[r * kCS + k0 + k] = x[k];
        } else if (tid < nrest + 8) {
            const int i = tid - nrest;
#pragma unroll
            for (int ii = 0; ii < 8; ++ii)
                if (ii == i) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) T[(k0 + ii) * kCS + k0 + j] = j <= ii ? a[ii][j] : 0.0;
                    dinv_s[k0 + ii] = dinv[ii];
                }
        }
        __syncthreads();
        if (nrest > 0) {
            for (int e = tid; e < nrest * nrest; e += 256) {
                const int rr = e / nrest, cc = e - rr * nrest;
                if (cc > rr) continue;
                const double *pr = &T[(k0 + 8 + rr) * kCS + k0], *pc = &T[(k0 + 8 + cc) * kCS + k0];
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 8; ++k) acc = fma(pr[k], pc[k], acc);
                T[(k0 + 8 + rr) * kCS + k0 + 8 + cc] -= acc;
            }
            __syncthreads();
        }
    }
    if (blockIdx.x == 0) {
        if (tid == 0 && fail) atomicExch(&st->chol_fail, 1);
        for (int t = tid; t < 64 * 64; t += 256) {
            const int rr = t >> 6, c = t & 63;
            Ld[t] = c <= rr ? T[rr * kCS + c] : 0.0;
        }
        return;
    }
    // panel rows: X L^T = A, sub-block column by sub-block column
    const int r = tid & 63, q = tid >> 6;
#pragma unroll
    for (int sblk = 0; sblk < 8; ++sblk) {
        const int k0 = sblk * 8;
        if (tid < 64) {
            double x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = X[r * kCS + k0 + k];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                x[k] *= dinv_s[k0 + k];
#pragma unroll
                for (int k2 = k + 1; k2 < 8; ++k2) x[k2] = fma(-x[k], T[(k0 + k2) * kCS + k0 + k], x[k2]);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) X[r * kCS + k0 + k] = x[k];
        }
        __syncthreads();
        if (sblk < 7) {
            double xr[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) xr[k] = X[r * kCS + k0 + k];
            for (int c = k0 + 8 + q; c < 64; c += 4) {
                const double *pc = &T[c * kCS + k0];
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 8; ++k) acc = fma(xr[k], pc[k], acc);
                X[r * kCS + c] -= acc;
            }
            __syncthreads();
        }
    }
    for (int t = tid; t < 64 * 64; t += 256) {
        const int rr = t >> 6, c = t & 63;
        S[(size_t)(rb + rr) * d.NSP + base + c] = X[rr * kCS + c];
    }
}

// Step kb, kernel 2: trailing update A_ij -= X_i X_j^T (i >= j > kb) on the
// tensor cores; CTA 0 also copies the factored diagonal block back into S.
__global__ void __launch_bounds__(128) k_chol_trailing(LmDims d, int kb, double *__restrict__ S,
                                                       const double *__restrict__ Ld) {
    __shared__ double sA[16 * kTS];
    __shared__ double sB[16 * kTS];
    const int base = kb * 64;
    if (blockIdx.x == 0) {
        for (int t = threadIdx.x; t < 64 * 64; t += blockDim.x) {
            const int r = t >> 6, c = t & 63;
            S[(size_t)(base + r) * d.NSP + base + c] = Ld[t];
        }
    }
    const int nt = d.NSP / 64 - kb - 1;
    if (nt <= 0) return;
    int ti, tj;
    tile_index(blockIdx.x, ti, tj);
    if (ti >= nt) return;
    const int bi = kb + 1 + ti, bj = kb + 1 + tj;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wr = warp >> 1, wc = warp & 1;
    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
    for (int k0 = 0; k0 < 64; k0 += 16) {
        // operands k-major: sA[k][i] = X_i[i][k0 + k]
        for (int t = threadIdx.x; t < 64 * 16; t += blockDim.x) {
            const int r = t >> 4, c = t & 15;
            sA[c * kTS + r] = S[(size_t)(bi * 64 + r) * d.NSP + base + k0 + c];
            sB[c * kTS + r] = S[(size_t)(bj * 64 + r) * d.NSP + base + k0 + c];
        }
        __syncthreads();
        tile_mma<16>(sA, sB, wr, wc, lane, acc);
        __syncthreads();
    }
    tile_store_sub(S, d.NSP, bi, bj, wr, wc, lane, acc);
}

// ---- tile-dataflow Cholesky: the whole factorisation of S in ONE launch ------------------
// The step-by-step kernels above cost two launches per 64-column block step and every step
// waits for the slowest tile of the previous one.  Here every 64x64 tile (i, j), i >= j, of
// the lower triangle has its own CTA (left-looking, owner computes):
//   A.  acc = sum_{k<j} L(i,k) L(j,k)^T on the tensor cores, consuming each pair of tiles as
//       soon as their owners publish them (per-tile epoch flags, acquire / release);
//   B.  X = S(i,j) - acc in shared memory;
//   C.  i == j: Cholesky of X (8x8 sub-blocked, register pivot blocks), publish L(j,j);
//       i >  j: wait for L(j,j), X <- X L(j,j)^-T, publish L(i,j).
// The critical path per block column is potrf -> trsm of the tile below -> one tile update,
// everything else overlaps.  CTAs are numbered column by column, so a tile only ever waits
// for tiles of LOWER blockIdx: under in-order block dispatch no cooperative launch is
// needed and the kernel can sit in a CUDA-graph body.  Tiles published by other CTAs are
// read with ld.global.cg (L2): an SM's L1 may still hold the pre-factorisation lines.
constexpr int kPS = 12;   // staged 8-column panel stride: (12 g + q) mod 16 distinct -> conflict-free fragments
constexpr int kOS = 68;   // operand tile stride: (stride mod 16) == 4 -> conflict-free DMMA fragment loads
constexpr size_t kCholTilesSmem = sizeof(double) * (2 * 64 * kOS + 64) + 16;

// timeline of the critical path (diagonal and first sub-diagonal tile of every block column),
// %globaltimer ns: [col][diag?0:1][0 last operand flag seen, 1 operands in smem, 2 update done,
// 3 factor/solve done, 4 (sub-diagonal) diagonal flag seen, 5 published]; mmb_debug_chol_trace
__device__ unsigned long long g_chol_trace[64 * 2 * 6];
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// MINB = 1: no register cap (254 registers, one CTA per SM) -- fastest block-column chain, for
// systems whose factorisation is that chain (cfg3: 24 block columns); MINB = 2: 128 registers,
// two CTAs per SM -- for large systems (cfg5: 235 block columns), which are throughput-bound.
// syrk_ready != nullptr: launched BESIDE k_schur_syrk with a small grid of resident CTAs; a tile of
// S is read once the SYRK has published it (see k_schur_syrk).
// Wait until *f == epoch.  BOUNDED (the walking grid beside the SYRK: its CTAs must all be resident
// for a tile to find its operands; should the device ever not hold them together, the wait ends
// after 2 s, LmState.dataflow_timeout is set and the host goes back to one kernel after the other).
template <bool BOUNDED>
__device__ __forceinline__ void chol_wait_flag(const int *f, int epoch, LmState *st) {
    int v;
    if (!BOUNDED) {
        do { asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory"); } while (v != epoch);
        return;
    }
    const unsigned long long t0 = gtimer();
    unsigned spins = 0;
    for (;;) {
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (v == epoch) return;
        if ((++spins & 1023u) == 0u && (gtimer() - t0 > 2000000000ull || *(volatile int *)&st->dataflow_timeout)) { atomicExch(&st->dataflow_timeout, 1); return; }
    }
}

// PERSIST: a grid smaller than the number of tiles walks them (the launch beside the SYRK);
// PEER: sharded, the tile of S is the sum of the ranks' partial tiles (PeerTiles).
template <int MINB, bool PERSIST = false, bool PEER = false>
__global__ void __launch_bounds__(256, MINB) k_chol_tiles(LmDims d, LmState *st, double *__restrict__ S, int *tile_flags,
                                                    const int *epoch_ctr, double *__restrict__ Dinv, const int *syrk_ready,
                                                    PeerTiles pt) {
    extern __shared__ __align__(16) double sm[];
    double *sA = sm;                 // L(i,k) during phase A; afterwards T (stride kCS)
    double *sB = sm + 64 * kOS;      // L(j,k) during phase A; afterwards X (stride kCS)
    double *dinv_s = sB + 64 * kOS;  // 1 / L(j,j)[c][c]
    const int nb = d.NSP / 64;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int epoch = __ldcg(epoch_ctr) + 1;
    const size_t ld = (size_t)d.NSP;
    int *flags = tile_flags;
    const int wr = warp >> 1, wc = warp & 1;   // warp tile: rows wr*16.., cols wc*32..
    const int g = lane >> 2, q = lane & 3;
    // gridDim.x = number of tiles: one tile per CTA.  A smaller grid (every CTA resident at once:
    // the launch beside the SYRK) walks the tiles round-robin; a CTA takes its tiles in increasing
    // order and a tile only waits for tiles of lower number (and for the SYRK, which waits for
    // nothing), so the lowest unfinished tile can always proceed.
    int tile = blockIdx.x;
    do {
    if (PERSIST && tile != (int)blockIdx.x) __syncthreads();   // phase E of the previous tile is done with the shared tiles
    // column-major tile numbering: column j holds nb - j tiles
    int j = 0, t = tile;
    while (t >= nb - j) { t -= nb - j; ++j; }
    const int i = j + t;
    const bool diag = i == j;
    const bool traced = tid == 0 && i <= j + 1 && j < 64;
    unsigned long long *tr = &g_chol_trace[(j * 2 + (diag ? 0 : 1)) * 6];
    if (PERSIST && !PEER && syrk_ready) {
        // this tile of S is complete when the SYRK has published it (bounded wait: a flag that never
        // comes is reported, not spun on for ever)
        if (tid == 0) {
            const int *f = &syrk_ready[j * nb + i];
            int v;
            const unsigned long long t0 = gtimer();
            unsigned spins = 0;
            for (;;) {
                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                if (v == epoch) break;
                if ((++spins & 1023u) == 0u && (gtimer() - t0 > 2000000000ull || *(volatile int *)&st->dataflow_timeout)) { atomicExch(&st->dataflow_timeout, 1); break; }
            }
        }
        __syncthreads();
    }
    // ---- A: left-looking update on the tensor cores ----
    double acc[2][4][2];
    double2 sij[2][4];   // this tile of S, fetched before the waits: off the critical path
    if (PEER) {
        // sharded: the tile is the sum of the ranks' partial tiles, each there once its flag is
        const unsigned pep = __ldcg(pt.seq);
        const unsigned pb = pep & 1u;
        if (tid < pt.world) {
            const unsigned *f = pt.my_flags + ((size_t)pb * 8 + tid) * pt.maxtiles + tile;
            unsigned v;
            const unsigned long long t0 = gtimer();
            unsigned spins = 0;
            for (;;) {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
                if ((int)(v - pep) >= 0) break;
                if ((++spins & 1023u) == 0u && (gtimer() - t0 > 4000000000ull || *(volatile int *)&st->dataflow_timeout)) { atomicExch(&st->dataflow_timeout, 1); break; }
            }
        }
        __syncthreads();
        const double par = st->par_cur;
        const double *mine = pt.my_tiles + ((size_t)pb * pt.world * pt.ntiles + tile) * 4096;
        const size_t rstride = (size_t)pt.ntiles * 4096;   // from one source rank's slot to the next
#pragma unroll
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                // the fragment from every rank first (independent loads), then the sum in rank order
                acc[a][b][0] = acc[a][b][1] = 0.0;
                const int r = wr * 16 + a * 8 + g, c = wc * 32 + b * 8 + q * 2;
                double2 u[8];
#pragma unroll
                for (int rk = 0; rk < 8; ++rk)
                    if (rk < pt.world) u[rk] = __ldcg(reinterpret_cast<const double2 *>(mine + rk * rstride + r * 64 + c));
                double2 v = u[0];
#pragma unroll
                for (int rk = 1; rk < 8; ++rk)
                    if (rk < pt.world) { v.x += u[rk].x; v.y += u[rk].y; }
                if (diag && (r == c || r == c + 1)) {   // damping / padding of the diagonal, after the sum
                    const int R = i * 64 + r;
                    double x = (r == c) ? v.x : v.y;
                    if (R < d.ns) { const double dd = pt.diag[pt.i2p[d.KE + R]]; x += par * dd * dd; }
                    else x = (R == d.ns) ? 0.0 : 1.0;
                    if (r == c) v.x = x; else v.y = x;
                }
                sij[a][b] = v;
            }
    } else {
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            acc[a][b][0] = acc[a][b][1] = 0.0;
            const int r = wr * 16 + a * 8 + g, c = wc * 32 + b * 8 + q * 2;
            sij[a][b] = __ldcg(reinterpret_cast<const double2 *>(&S[(size_t)(i * 64 + r) * ld + j * 64 + c]));
        }
    }
    for (int k = 0; k < j; ++k) {
        if (tid == 0) {
            const int *f1 = &flags[k * nb + i], *f2 = &flags[k * nb + j];
            chol_wait_flag<PERSIST>(f1, epoch, st);
            if (!diag) chol_wait_flag<PERSIST>(f2, epoch, st);
            if (traced && k == j - 1) tr[0] = gtimer();
        }
        __syncthreads();  // also: everyone is done with the previous k's operands
        // 64 x 64 doubles per operand = 2048 double2; 8 per thread
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = tid + u * 256;
            const int r = idx >> 5, c2 = (idx & 31) * 2;
            const double2 va = __ldcg(reinterpret_cast<const double2 *>(&S[(size_t)(i * 64 + r) * ld + k * 64 + c2]));
            *reinterpret_cast<double2 *>(&sA[r * kOS + c2]) = va;
            if (!diag) {
                const double2 vb = __ldcg(reinterpret_cast<const double2 *>(&S[(size_t)(j * 64 + r) * ld + k * 64 + c2]));
                *reinterpret_cast<double2 *>(&sB[r * kOS + c2]) = vb;
            }
        }
        __syncthreads();
        if (traced && k == j - 1) tr[1] = gtimer();
        const double *pB = diag ? sA : sB;
#pragma unroll 4
        for (int k0 = 0; k0 < 64; k0 += 4) {
            double a[2], b[4];
#pragma unroll
            for (int u = 0; u < 2; ++u) a[u] = sA[(wr * 16 + u * 8 + g) * kOS + k0 + q];
#pragma unroll
            for (int u = 0; u < 4; ++u) b[u] = pB[(wc * 32 + u * 8 + g) * kOS + k0 + q];
#pragma unroll
            for (int ti = 0; ti < 2; ++ti)
#pragma unroll
                for (int tj = 0; tj < 4; ++tj) dmma884(acc[ti][tj][0], acc[ti][tj][1], a[ti], b[tj]);
        }
    }
    __syncthreads();
    // ---- B: X = S(i,j) - acc (stride kCS) ----
    double *T = sA, *X = sB;
    double *dst = diag ? T : X;
#pragma unroll
    for (int ti = 0; ti < 2; ++ti)
#pragma unroll
        for (int tj = 0; tj < 4; ++tj) {
            const int r = wr * 16 + ti * 8 + g, c = wc * 32 + tj * 8 + q * 2;
            const double2 v = sij[ti][tj];
            dst[r * kCS + c] = v.x - acc[ti][tj][0];
            dst[r * kCS + c + 1] = v.y - acc[ti][tj][1];
        }
    __syncthreads();
    if (traced) tr[2] = gtimer();
    const int base = j * 64;
    if (diag) {
        // ---- C: factor the diagonal tile (as k_chol_panel) ----
        int fail = 0;
#pragma unroll
        for (int sblk = 0; sblk < 8; ++sblk) {
            const int k0 = sblk * 8, nrest = 56 - k0;
            // the pivot block is factored in registers by the threads that use the result (the
            // panel rows and the write-back: tid < 64); the other six warps stay off the FP64 pipe
            double a[8][8], dinv[8];
            if (tid < 64) {
#pragma unroll
                for (int ii = 0; ii < 8; ++ii)
#pragma unroll
                    for (int jj = 0; jj <= ii; ++jj) a[ii][jj] = T[(k0 + ii) * kCS + k0 + jj];
                chol8_regs(a, dinv, base + k0, d.ns, fail);
            }
            __syncthreads();
            if (tid < nrest) {
                const int r = k0 + 8 + tid;
                double x[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = T[r * kCS + k0 + k];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    x[k] *= dinv[k];
#pragma unroll
                    for (int k2 = k + 1; k2 < 8; ++k2) x[k2] = fma(-x[k], a[k2][k], x[k2]);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) T[r * kCS + k0 + k] = x[k];
                // second copy for the tensor-core update below: stride 12 makes the m8n8k4
                // operand fragments bank-conflict free (stride 65 is 4-way conflicted there);
                // the X region is free in the diagonal CTA
#pragma unroll
                for (int k = 0; k < 8; ++k) X[tid * kPS + k] = x[k];
            } else if (tid < nrest + 8) {
                const int ii0 = tid - nrest;
#pragma unroll
                for (int ii = 0; ii < 8; ++ii)
                    if (ii == ii0) {
#pragma unroll
                        for (int jj = 0; jj < 8; ++jj) T[(k0 + ii) * kCS + k0 + jj] = jj <= ii ? a[ii][jj] : 0.0;
                        dinv_s[k0 + ii] = dinv[ii];
                    }
            }
            __syncthreads();
            if (nrest > 0) {
                // rank-8 update of the lower triangle below the pivot block on the tensor cores:
                // 8x8 output tiles (rt >= ct) dealt to the warps (at most 4 each), operands of
                // all of a warp's tiles fetched (from the staged panel) before the first m8n8k4.
                // The scalar version was a chain of 8 dependent FMAs (~20 cycles each) per entry.
                const int nt = nrest >> 3, ntile = nt * (nt + 1) / 2;
                const double *P = X;   // staged panel, row stride kPS
                double pa[4][2], pb[4][2], cacc[4][2];
                int roff[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int tI = warp + 8 * u;
                    int rt = 0, ct = tI;
#pragma unroll
                    for (int z = 0; z < 6; ++z)
                        if (ct > rt) { ct -= rt + 1; ++rt; }
                    roff[u] = (k0 + 8 + rt * 8 + g) * kCS + k0 + 8 + ct * 8 + 2 * q;
                    if (tI < ntile) {
#pragma unroll
                        for (int kk = 0; kk < 2; ++kk) {
                            pa[u][kk] = P[(rt * 8 + g) * kPS + kk * 4 + q];
                            pb[u][kk] = P[(ct * 8 + g) * kPS + kk * 4 + q];
                        }
                        cacc[u][0] = T[roff[u]];
                        cacc[u][1] = T[roff[u] + 1];
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (warp + 8 * u < ntile) {
                        double c0 = 0.0, c1 = 0.0;
                        dmma884(c0, c1, pa[u][0], pb[u][0]);
                        dmma884(c0, c1, pa[u][1], pb[u][1]);
                        T[roff[u]] = cacc[u][0] - c0;
                        T[roff[u] + 1] = cacc[u][1] - c1;
                    }
                __syncthreads();
            }
        }
        if (tid == 0 && fail) atomicExch(&st->chol_fail, 1);
        if (traced) tr[3] = gtimer();
        for (int e = tid; e < 64 * 64; e += 256) {
            const int rr = e >> 6, c = e & 63;
            S[(size_t)(base + rr) * ld + base + c] = c <= rr ? T[rr * kCS + c] : 0.0;
        }
    } else {
        // ---- D: X <- X L(j,j)^-T once the diagonal tile is published ----
        if (tid == 0) {
            const int *f = &flags[j * nb + j];
            chol_wait_flag<PERSIST>(f, epoch, st);
            if (traced) tr[4] = gtimer();
        }
        __syncthreads();
        // L(j,j) with the fragment-friendly row stride kOS (it is only read by broadcast and as
        // m8n8k4 operands here); 16-byte loads; 1/diag by the thread that fetched the diagonal
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = tid + u * 256;
            const int rr = idx >> 5, c2 = (idx & 31) * 2;
            const double2 v = __ldcg(reinterpret_cast<const double2 *>(&S[(size_t)(base + rr) * ld + base + c2]));
            *reinterpret_cast<double2 *>(&T[rr * kOS + c2]) = v;
            if (c2 == (rr & ~1)) dinv_s[rr] = rcp_fast((rr & 1) ? v.y : v.x);
        }
        __syncthreads();
        const int r = tid & 63;
#pragma unroll
        for (int sblk = 0; sblk < 8; ++sblk) {
            const int k0 = sblk * 8;
            if (tid < 64) {
                double x[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = X[r * kCS + k0 + k];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    x[k] *= dinv_s[k0 + k];
#pragma unroll
                    for (int k2 = k + 1; k2 < 8; ++k2) x[k2] = fma(-x[k], T[(k0 + k2) * kOS + k0 + k], x[k2]);
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) X[r * kCS + k0 + k] = x[k];
            }
            __syncthreads();
            if (sblk < 7) {
                // X[:, k0+8..] -= X[:, k0..k0+8) L(j,j)[k0+8.., k0..k0+8)^T on the tensor cores:
                // warp w owns the 8 rows w*8.. (one A fragment) and walks the 7 - sblk column tiles
                const int nct = 7 - sblk;
                const double xa0 = X[(warp * 8 + g) * kCS + k0 + q], xa1 = X[(warp * 8 + g) * kCS + k0 + 4 + q];
                double tb[7][2], xc[7][2];
#pragma unroll
                for (int ct = 0; ct < 7; ++ct)
                    if (ct < nct) {
                        tb[ct][0] = T[(k0 + 8 + ct * 8 + g) * kOS + k0 + q];
                        tb[ct][1] = T[(k0 + 8 + ct * 8 + g) * kOS + k0 + 4 + q];
                        xc[ct][0] = X[(warp * 8 + g) * kCS + k0 + 8 + ct * 8 + 2 * q];
                        xc[ct][1] = X[(warp * 8 + g) * kCS + k0 + 8 + ct * 8 + 2 * q + 1];
                    }
#pragma unroll
                for (int ct = 0; ct < 7; ++ct)
                    if (ct < nct) {
                        double c0 = 0.0, c1 = 0.0;
                        dmma884(c0, c1, xa0, tb[ct][0]);
                        dmma884(c0, c1, xa1, tb[ct][1]);
                        X[(warp * 8 + g) * kCS + k0 + 8 + ct * 8 + 2 * q] = xc[ct][0] - c0;
                        X[(warp * 8 + g) * kCS + k0 + 8 + ct * 8 + 2 * q + 1] = xc[ct][1] - c1;
                    }
                __syncthreads();
            }
        }
        const int rb = i * 64;
        if (traced) tr[3] = gtimer();
        for (int e = tid; e < 64 * 64; e += 256) {
            const int rr = e >> 6, c = e & 63;
            S[(size_t)(rb + rr) * ld + base + c] = X[rr * kCS + c];
        }
    }
    // publish
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        int *f = &flags[j * nb + i];
        asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(f), "r"(epoch) : "memory");
        if (traced) tr[5] = gtimer();
    }
    // ---- E (diagonal tiles, after publishing: nobody waits for this): inverse of L(j,j) ----
    // The triangular solves that follow (k_forward_wave / k_backward_wave) are chains of 24
    // dependent 64x64 block solves; with L(j,j)^-1 at hand each block solve is a 64x64
    // matrix-vector product instead of 64 dependent substitution steps.
    if (diag && Dinv) {
        double *T = sA, *X = sB;
        if (tid < 64) {  // inverses of the eight 8x8 diagonal blocks, one column per thread
            const int b8 = (tid >> 3) * 8, c = tid & 7;
            double x[8];
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) {
                double sacc = (ii == c) ? 1.0 : 0.0;
#pragma unroll
                for (int k = 0; k < ii; ++k) sacc = fma(-T[(b8 + ii) * kCS + b8 + k], x[k], sacc);
                x[ii] = sacc * dinv_s[b8 + ii];
            }
#pragma unroll
            for (int ii = 0; ii < 8; ++ii) X[(b8 + ii) * kCS + b8 + c] = x[ii];
        }
        __syncthreads();
        // off-diagonal 8x8 blocks by distance from the diagonal:
        //   Inv(i,j) = -Inv(i,i) * sum_{k=j}^{i-1} L(i,k) Inv(k,j); the inner sum is parked in the
        //   mirrored (upper) block (j,i), which the result never uses
        for (int dd = 1; dd < 8; ++dd) {
            const int cnt = (8 - dd) * 64;
            for (int e = tid; e < cnt; e += 256) {
                const int jb = e >> 6, r = (e >> 3) & 7, c = e & 7, ib = jb + dd;
                double s0 = 0.0, s1 = 0.0;
                for (int kb = jb; kb < ib; ++kb) {
#pragma unroll
                    for (int t = 0; t < 8; t += 2) {
                        s0 = fma(T[(8 * ib + r) * kCS + 8 * kb + t], X[(8 * kb + t) * kCS + 8 * jb + c], s0);
                        s1 = fma(T[(8 * ib + r) * kCS + 8 * kb + t + 1], X[(8 * kb + t + 1) * kCS + 8 * jb + c], s1);
                    }
                }
                X[(8 * jb + r) * kCS + 8 * ib + c] = s0 + s1;
            }
            __syncthreads();
            for (int e = tid; e < cnt; e += 256) {
                const int jb = e >> 6, r = (e >> 3) & 7, c = e & 7, ib = jb + dd;
                double s0 = 0.0;
#pragma unroll
                for (int t = 0; t < 8; ++t)
                    if (t <= r) s0 = fma(X[(8 * ib + r) * kCS + 8 * ib + t], X[(8 * jb + t) * kCS + 8 * ib + c], s0);
                X[(8 * ib + r) * kCS + 8 * jb + c] = -s0;
            }
            __syncthreads();
        }
        double *out = Dinv + (size_t)j * 4096;
        for (int e = tid; e < 64 * 64; e += 256) {
            const int rr = e >> 6, c = e & 63;
            out[e] = c <= rr ? X[rr * kCS + c] : 0.0;
        }
    }
    } while (PERSIST && (tile += gridDim.x) < nb * (nb + 1) / 2);   // tiles of this CTA
}

// ---- triangular solves with L_S (single CTA, 1024 threads) ---------------------------
// Both sweeps stage the 64x64 diagonal block in shared memory (the dependent
// substitution then never waits on global memory) and read the off-diagonal
// panel with fully coalesced warp-per-row accesses.
constexpr int kDS = 65;  // shared stride of the staged diagonal block

// forward: solve L w = b for rows [0, ns).  b and w may alias.
__global__ void __launch_bounds__(1024) k_forward_s(LmDims d, const double *__restrict__ S, const double *b, double *w) {
    extern __shared__ double sm[];
    double *D = sm;               // 64 x kDS
    double *xs = sm + 64 * kDS;   // 64: rhs then solution of the block
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ns = d.ns;
    for (int c0 = 0; c0 < ns; c0 += kNB) {
        for (int t = threadIdx.x; t < 64 * 64; t += blockDim.x) {
            const int r = t >> 6, c = t & 63;
            const bool in = (c <= r) && (c0 + r < ns);  // rows >= ns (rhs row, padding): identity
            D[r * kDS + c] = in ? S[(size_t)(c0 + r) * d.NSP + c0 + c] : (r == c ? 1.0 : 0.0);
        }
        // rows c0 .. c0+63: rhs = b[row] - L[row, 0:c0] . w[0:c0]   (two rows per warp)
        for (int rr = warp; rr < kNB; rr += 32) {
            const int row = c0 + rr;
            double a = 0.0;
            if (row < ns) {
                const double *Lr = S + (size_t)row * d.NSP;
                for (int c = lane; c < c0; c += 32) a += Lr[c] * w[c];
            }
            a = warp_sum(a);
            if (lane == 0) xs[rr] = (row < ns) ? b[row] - a : 0.0;
        }
        __syncthreads();
        if (warp == 0) {
            // 64x64 lower-triangular solve out of shared memory; lanes own rows lane, lane+32
            for (int j = 0; j < kNB; ++j) {
                const double xj = xs[j] / D[j * kDS + j];
                __syncwarp();
                if (lane == 0) xs[j] = xj;
                for (int r = lane; r < kNB; r += 32)
                    if (r > j) xs[r] -= D[r * kDS + j] * xj;
                __syncwarp();
            }
        }
        __syncthreads();
        for (int r = threadIdx.x; r < kNB; r += blockDim.x)
            if (c0 + r < ns) w[c0 + r] = xs[r];
        __syncthreads();
    }
}

// backward: solve L^T p = y where y is the augmented row ns of S.
__global__ void __launch_bounds__(1024) k_backward_s(LmDims d, const double *__restrict__ S, double *__restrict__ p) {
    extern __shared__ double sm[];
    double *D = sm;                     // 64 x kDS
    double *xs = sm + 64 * kDS;         // 64
    double *part = xs + 64;             // 32 x 64 per-warp partial column sums
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ns = d.ns;
    const int nblk = (ns + kNB - 1) / kNB;
    for (int bb = nblk - 1; bb >= 0; --bb) {
        const int c0 = bb * kNB;
        for (int t = threadIdx.x; t < 64 * 64; t += blockDim.x) {
            const int r = t >> 6, c = t & 63;
            const bool in = (c <= r) && (c0 + r < ns);
            D[r * kDS + c] = in ? S[(size_t)(c0 + r) * d.NSP + c0 + c] : (r == c ? 1.0 : 0.0);
        }
        // sum over rows c >= c0+64 of L[c][c0 + j] p[c], j = lane, lane+32
        double a0 = 0.0, a1 = 0.0;
        for (int c = c0 + kNB + warp; c < ns; c += 32) {
            const double pc = p[c];
            a0 += S[(size_t)c * d.NSP + c0 + lane] * pc;
            a1 += S[(size_t)c * d.NSP + c0 + 32 + lane] * pc;
        }
        part[warp * 64 + lane] = a0;
        part[warp * 64 + lane + 32] = a1;
        __syncthreads();
        if (threadIdx.x < kNB) {
            double t = 0.0;
            for (int w2 = 0; w2 < 32; ++w2) t += part[w2 * 64 + threadIdx.x];
            const int col = c0 + threadIdx.x;
            xs[threadIdx.x] = (col < ns) ? S[(size_t)ns * d.NSP + col] - t : 0.0;
        }
        __syncthreads();
        if (warp == 0) {
            // solve L_kk^T x = xs: column sweep from the bottom; lanes own rows lane, lane+32
            for (int j = kNB - 1; j >= 0; --j) {
                const double xj = xs[j] / D[j * kDS + j];
                __syncwarp();
                if (lane == 0) xs[j] = xj;
                for (int r = lane; r < kNB; r += 32)
                    if (r < j) xs[r] -= D[j * kDS + r] * xj;
                __syncwarp();
            }
        }
        __syncthreads();
        for (int r = threadIdx.x; r < kNB; r += blockDim.x)
            if (c0 + r < ns) p[c0 + r] = xs[r];
        __syncthreads();
    }
}

// ---- wavefront triangular solves (one CTA per 64-row block, one launch) --------------------
// The single-CTA kernels above walk the 24 blocks of cfg3 one after another and pull the
// whole factor (9 MB) through one SM.  Here CTA b owns block b of the solution: it folds
// the blocks already solved by the other CTAs into its right-hand side as soon as each is
// published (a 64x64 GEMV, operands prefetched while it waits), then solves its own
// diagonal block and publishes.  The critical path is nblk x (flag + GEMV + 64-step
// substitution in one warp); every CTA must be resident (cooperative launch, the host falls
// back to the single-CTA kernels when the grid does not fit).  Flags carry an epoch so they
// need no reset between launches; the epoch counter lives in device memory (advanced by
// k_ctrl_lmpar_decide / k_ctrl_lmpar_update) so that the launch arguments are the same every time -- the launches can
// be replayed from a CUDA graph.  A CTA only ever waits for CTAs of lower blockIdx, so
// the plain (non-cooperative) launch used inside graph bodies cannot deadlock under
// in-order block dispatch either.
__device__ __forceinline__ void wave_publish(int *flag, int epoch) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
    }
}
__device__ __forceinline__ void wave_wait(const int *flag, int epoch) {
    if (threadIdx.x == 0) {
        int v;
        do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        } while (v != epoch);
    }
    __syncthreads();
}

// 64x64 substitution in one warp out of the staged diagonal block D (stride kDS); lanes own
// entries lane and lane + 32 of the vector xs (shared).  lower = true: L x = b, else L^T x = b.
template <bool LOWER>
__device__ __forceinline__ void wave_block_solve(const double *D, const double *dinv, double *xs, int lane) {
    double x0 = xs[lane], x1 = xs[lane + 32];
    if (LOWER) {
#pragma unroll 8
        for (int j = 0; j < 64; ++j) {
            const double src = j < 32 ? x0 : x1;
            const double xj = __shfl_sync(0xffffffffu, src, j & 31) * dinv[j];
            if (lane == (j & 31)) { if (j < 32) x0 = xj; else x1 = xj; }
            if (lane > j) x0 = fma(-D[lane * kDS + j], xj, x0);
            if (lane + 32 > j) x1 = fma(-D[(lane + 32) * kDS + j], xj, x1);
        }
    } else {
#pragma unroll 8
        for (int j = 63; j >= 0; --j) {
            const double src = j < 32 ? x0 : x1;
            const double xj = __shfl_sync(0xffffffffu, src, j & 31) * dinv[j];
            if (lane == (j & 31)) { if (j < 32) x0 = xj; else x1 = xj; }
            if (lane < j) x0 = fma(-D[j * kDS + lane], xj, x0);
            if (lane + 32 < j) x1 = fma(-D[j * kDS + lane + 32], xj, x1);
        }
    }
    xs[lane] = x0; xs[lane + 32] = x1;
}

// Dinv != nullptr: stage the INVERSE of the diagonal block (k_chol_tiles phase E) instead; the
// identity rows of the mask are their own inverse
__device__ __forceinline__ void wave_stage_diag(LmDims d, const double *__restrict__ S, int c0, double *D, double *dinv,
                                                const double *__restrict__ Dinv = nullptr) {
    const double *inv = Dinv ? Dinv + (size_t)(c0 / kNB) * 4096 : nullptr;
    for (int t = threadIdx.x; t < 64 * 64; t += blockDim.x) {
        const int r = t >> 6, c = t & 63;
        const bool in = (c <= r) && (c0 + r < d.ns);  // rows >= ns (rhs row, padding): identity
        const double v = inv ? inv[t] : S[(size_t)(c0 + r) * d.NSP + c0 + c];
        D[r * kDS + c] = in ? v : (r == c ? 1.0 : 0.0);
    }
    __syncthreads();
    if (!inv && threadIdx.x < 64) dinv[threadIdx.x] = 1.0 / D[threadIdx.x * kDS + threadIdx.x];
}

// backward: L^T p = y, y = augmented row ns of S.  blockIdx 0 owns the LAST block.
__global__ void __launch_bounds__(256) k_backward_wave(LmDims d, const double *__restrict__ S, double *p, int *flags,
                                                       const int *epoch_ctr, int slot, const double *__restrict__ Dinv) {
    extern __shared__ double sm[];
    const int epoch = __ldcg(epoch_ctr) + slot;
    double *D = sm;                // 64 x kDS
    double *dinv = D + 64 * kDS;   // 64
    double *acc = dinv + 64;       // 64: right-hand side, then solution
    double *pi = acc + 64;         // 64
    double *part = pi + 64;        // 4 x 64
    const int nblk = (d.ns + kNB - 1) / kNB;
    const int b = nblk - 1 - blockIdx.x, c0 = b * kNB;
    const int tid = threadIdx.x, g = tid >> 6, c = tid & 63;
    wave_stage_diag(d, S, c0, D, dinv, Dinv);
    if (tid < 64) acc[tid] = (c0 + tid < d.ns) ? S[(size_t)d.ns * d.NSP + c0 + tid] : 0.0;
    // block (i, b) of L, rows g, g+4, ... of block row i, column c of block column b
    double cur[16], nxt[16];
    auto fetch = [&](int i, double (&v)[16]) {
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = S[(size_t)(i * kNB + g + 4 * k) * d.NSP + c0 + c];
    };
    if (nblk - 1 > b) fetch(nblk - 1, cur);
    for (int i = nblk - 1; i > b; --i) {
        if (i - 1 > b) fetch(i - 1, nxt);
        wave_wait(&flags[i], epoch);
        if (tid < 64) pi[tid] = __ldcg(&p[i * kNB + tid]);
        __syncthreads();
        double sacc = 0.0;
#pragma unroll
        for (int k = 0; k < 16; ++k) sacc = fma(cur[k], pi[g + 4 * k], sacc);
        part[g * 64 + c] = sacc;
        __syncthreads();
        if (tid < 64) acc[tid] -= (part[tid] + part[64 + tid]) + (part[128 + tid] + part[192 + tid]);
#pragma unroll
        for (int k = 0; k < 16; ++k) cur[k] = nxt[k];
    }
    __syncthreads();
    if (Dinv) {  // x = L(b,b)^-T acc: x_c = sum_k inv[k][c] acc[k], four partial sums per entry
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k += 2) {
            s0 = fma(D[(g + 4 * k) * kDS + c], acc[g + 4 * k], s0);
            s1 = fma(D[(g + 4 * k + 4) * kDS + c], acc[g + 4 * k + 4], s1);
        }
        part[g * 64 + c] = s0 + s1;
        __syncthreads();
        if (tid < 64) acc[tid] = (part[tid] + part[64 + tid]) + (part[128 + tid] + part[192 + tid]);
    } else if (tid < 32) {
        wave_block_solve<false>(D, dinv, acc, tid);
    }
    __syncthreads();
    if (tid < 64) p[c0 + tid] = (c0 + tid < d.ns) ? acc[tid] : 0.0;
    wave_publish(&flags[b], epoch);
}

// forward: L w = rhs for rows [0, ns).  rhs and w may alias.  blockIdx b owns block b.
__global__ void __launch_bounds__(256) k_forward_wave(LmDims d, const double *__restrict__ S, const double *rhs, double *w,
                                                      int *flags, const int *epoch_ctr, int slot,
                                                      const double *__restrict__ Dinv) {
    extern __shared__ double sm[];
    const int epoch = __ldcg(epoch_ctr) + slot;
    double *D = sm;
    double *dinv = D + 64 * kDS;
    double *acc = dinv + 64;
    double *wi = acc + 64;
    const int b = blockIdx.x, c0 = b * kNB;
    const int tid = threadIdx.x, r = tid >> 2, q = tid & 3;
    wave_stage_diag(d, S, c0, D, dinv, Dinv);
    if (tid < 64) acc[tid] = (c0 + tid < d.ns) ? rhs[c0 + tid] : 0.0;
    __syncthreads();
    // block (b, i) of L: row r of block row b, columns q*16 .. q*16+15 of block column i
    const bool live = c0 + r < d.ns;
    double cur[16], nxt[16];
    auto fetch = [&](int i, double (&v)[16]) {
        const double2 *src = reinterpret_cast<const double2 *>(&S[(size_t)(c0 + r) * d.NSP + i * kNB + q * 16]);
#pragma unroll
        for (int k = 0; k < 8; ++k) { const double2 t = src[k]; v[2 * k] = t.x; v[2 * k + 1] = t.y; }
    };
    if (b > 0) fetch(0, cur);
    for (int i = 0; i < b; ++i) {
        if (i + 1 < b) fetch(i + 1, nxt);
        wave_wait(&flags[i], epoch);
        if (tid < 64) wi[tid] = __ldcg(&w[i * kNB + tid]);
        __syncthreads();
        double sacc = 0.0;
#pragma unroll
        for (int k = 0; k < 16; ++k) sacc = fma(cur[k], wi[q * 16 + k], sacc);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
        if (q == 0 && live) acc[r] -= sacc;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) cur[k] = nxt[k];
    }
    __syncthreads();
    if (Dinv) {  // x = L(b,b)^-1 acc: row r by four threads, 16 columns each
        double s0 = 0.0, s1 = 0.0;
#pragma unroll
        for (int k = 0; k < 16; k += 2) {
            s0 = fma(D[r * kDS + q * 16 + k], acc[q * 16 + k], s0);
            s1 = fma(D[r * kDS + q * 16 + k + 1], acc[q * 16 + k + 1], s1);
        }
        double sacc = s0 + s1;
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
        __syncthreads();
        if (q == 0) acc[r] = sacc;
    } else if (tid < 32) {
        wave_block_solve<true>(D, dinv, acc, tid);
    }
    __syncthreads();
    if (tid < 64 && c0 + tid < d.ns) w[c0 + tid] = acc[tid];
    wave_publish(&flags[b], epoch);
}
constexpr size_t kWaveSmemBwd = sizeof(double) * (64 * kDS + 64 * 3 + 4 * 64);
constexpr size_t kWaveSmemFwd = sizeof(double) * (64 * kDS + 64 * 3);

// ---- eliminated-side back substitution -------------------------------------------------
// p_f = L_f^-T (z_f - Y_f p_S); CTA per frame.  p_int = [p_E (KE) | p_S (ns)]
// compile-time block size: the NA dot products of a frame in ONE pass over its rows of Y
// (all loads in flight together, one reduction stage) instead of NA passes with a block
// reduction each
template <int NA>
__global__ void __launch_bounds__(128) k_backsolve_e_fixed(LmDims d, const double *__restrict__ Lf,
                                                           const double *__restrict__ Y, double *__restrict__ p_int) {
    __shared__ double part[4][NA];
    const int f = blockIdx.x, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const double *pS = p_int + d.KE;
    const double *yr = Y + (size_t)f * NA * d.NSP;
    double acc[NA];
#pragma unroll
    for (int a = 0; a < NA; ++a) acc[a] = 0.0;
    for (int j = threadIdx.x; j < d.ns; j += blockDim.x) {
        const double pj = pS[j];
#pragma unroll
        for (int a = 0; a < NA; ++a) acc[a] = fma(yr[(size_t)a * d.NSP + j], pj, acc[a]);
    }
#pragma unroll
    for (int a = 0; a < NA; ++a) {
        const double v = warp_sum(acc[a]);
        if (lane == 0) part[w][a] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double r[NA];
#pragma unroll
        for (int a = 0; a < NA; ++a) r[a] = yr[(size_t)a * d.NSP + d.ns] - ((part[0][a] + part[1][a]) + (part[2][a] + part[3][a]));
        const double *L = Lf + (size_t)f * NA * NA;
#pragma unroll
        for (int a = NA - 1; a >= 0; --a) {
            double v = r[a];
#pragma unroll
            for (int c = a + 1; c < NA; ++c) v -= L[c * NA + a] * r[c];
            r[a] = v / L[a * NA + a];
        }
#pragma unroll
        for (int a = 0; a < NA; ++a) p_int[(size_t)f * NA + a] = r[a];
    }
}

__global__ void __launch_bounds__(128) k_backsolve_e(LmDims d, const double *__restrict__ Lf,
                                                     const double *__restrict__ Y, double *__restrict__ p_int) {
    __shared__ double sh[32];
    __shared__ double r[kMaxNA];
    const int na = d.na, f = blockIdx.x;
    const double *pS = p_int + d.KE;
    for (int a = 0; a < na; ++a) {
        const double *yr = Y + ((size_t)f * na + a) * d.NSP;
        double acc = 0.0;
        for (int j = threadIdx.x; j < d.ns; j += blockDim.x) acc += yr[j] * pS[j];
        acc = block_sum(acc, sh);
        if (threadIdx.x == 0) r[a] = yr[d.ns] - acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double *L = Lf + (size_t)f * na * na;
        for (int a = na - 1; a >= 0; --a) {
            double v = r[a];
            for (int c = a + 1; c < na; ++c) v -= L[c * na + a] * r[c];
            r[a] = v / L[a * na + a];
        }
        for (int a = 0; a < na; ++a) p_int[(size_t)f * na + a] = r[a];
    }
}

// w_f = L_f^-1 q_f with q = D^2 p (internal order; the 1/||D p|| factor of lmpar's Newton
// correction is applied to the squared norm at the end); one thread per frame.
__global__ void k_qsolve_e(LmDims d, const double *__restrict__ Lf, const double *__restrict__ p_int,
                           const double *__restrict__ diag, const int32_t *__restrict__ i2p, double *__restrict__ w_int) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= d.F) return;
    const int na = d.na;
    const double *L = Lf + (size_t)f * na * na;
    double w[kMaxNA];
    for (int a = 0; a < na; ++a) {
        const double dd = diag[i2p[f * na + a]];
        double v = dd * dd * p_int[(size_t)f * na + a];
        for (int c = 0; c < a; ++c) v -= L[a * na + c] * w[c];
        w[a] = v / L[a * na + a];
    }
    for (int a = 0; a < na; ++a) w_int[(size_t)f * na + a] = w[a];
}

// t[j] partials of Y^T w: grid (ceil(ns/128), row chunks); partial[chunk][j]
__global__ void __launch_bounds__(128) k_gemv_t_partial(LmDims d, const double *__restrict__ Y, const double *__restrict__ w_int,
                                                        double *__restrict__ partial) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int per = (d.KE + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * per, r1 = min(d.KE, r0 + per);
    if (j >= d.ns) return;
    // eight independent accumulators: the loads of eight rows are in flight together and no
    // FMA waits for the previous one (a single chain ran at HBM latency per row)
    double acc[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    int r = r0;
    for (; r + 8 <= r1; r += 8) {
        double y[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) y[u] = Y[(size_t)(r + u) * d.NSP + j];
#pragma unroll
        for (int u = 0; u < 8; ++u) acc[u] = fma(y[u], w_int[r + u], acc[u]);
    }
    for (; r < r1; ++r) acc[0] = fma(Y[(size_t)r * d.NSP + j], w_int[r], acc[0]);
    partial[(size_t)blockIdx.y * d.NSP + j] = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
}

// tvec[j] = sum_chunks partial[chunk][j]  (= (Y^T w_E)[j] of this shard)
__global__ void k_gemv_fold(LmDims d, int nchunks, const double *__restrict__ partial, double *__restrict__ tvec) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d.ns) return;
    double v = 0.0;
    for (int c = 0; c < nchunks; ++c) v += partial[(size_t)c * d.NSP + j];
    tvec[j] = v;
}
// b_S[j] = D_j^2 p_j - (Y^T w_E)[j]
__global__ void k_qrhs_s(LmDims d, const double *__restrict__ tvec, const double *__restrict__ p_int,
                         const double *__restrict__ diag, const int32_t *__restrict__ i2p, double *__restrict__ b) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d.ns) return;
    const double dd = diag[i2p[d.KE + j]];
    b[j] = dd * dd * p_int[d.KE + j] - tvec[j];
}

}  // namespace mmb
