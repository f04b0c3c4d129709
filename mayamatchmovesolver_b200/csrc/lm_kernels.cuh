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
__global__ void __launch_bounds__(256) k_assemble(DevScene s, LmDims d, const double *__restrict__ jac,
                                                  const double *__restrict__ fvec,
                                                  const int32_t *__restrict__ attr_slot, double *U, double *gE,
                                                  double *W, double *V) {
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
    for (int c1 = 0; c1 < k; ++c1) {
        const int a1 = pl[c1].attr;
        const bool an1 = s.attr_anim[a1];
        const int s1 = attr_slot[a1];
        const double jx1 = live ? jac[(size_t)c1 * m + 2 * (size_t)i] : 0.0;
        const double jy1 = live ? jac[(size_t)c1 * m + 2 * (size_t)i + 1] : 0.0;
        // gradient
        const double gv = jx1 * fx + jy1 * fy;
        if (an1) {
            if (live) atomicAdd(&gE[(size_t)frame * d.na + s1], gv);
        } else if (uniform) {
            const double r = warp_sum(gv);
            if (lane == 0) atomicAdd(&V[(size_t)d.ns * d.NSP + s1], r);
        } else if (live) {
            atomicAdd(&V[(size_t)d.ns * d.NSP + s1], gv);
        }
        for (int c2 = 0; c2 <= c1; ++c2) {
            const int a2 = pl[c2].attr;
            const bool an2 = s.attr_anim[a2];
            const int s2 = attr_slot[a2];
            const double jx2 = live ? jac[(size_t)c2 * m + 2 * (size_t)i] : 0.0;
            const double jy2 = live ? jac[(size_t)c2 * m + 2 * (size_t)i + 1] : 0.0;
            const double v = jx1 * jx2 + jy1 * jy2;
            if (an1 && an2) {
                if (live) {
                    const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                    atomicAdd(&U[((size_t)frame * d.na + r) * d.na + cc], v);
                }
            } else if (an1 != an2) {
                if (live) {
                    const int ja = an1 ? s1 : s2, js = an1 ? s2 : s1;
                    atomicAdd(&W[((size_t)frame * d.na + ja) * d.NSP + js], v);
                }
            } else {
                const int r = s1 > s2 ? s1 : s2, cc = s1 > s2 ? s2 : s1;
                if (uniform) {
                    const double t = warp_sum(v);
                    if (lane == 0) atomicAdd(&V[(size_t)r * d.NSP + cc], t);
                } else if (live) {
                    atomicAdd(&V[(size_t)r * d.NSP + cc], v);
                }
            }
        }
    }
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
// grid.x = lower-triangle tiles of the NSP/64 block grid; grid.y = K splits
// (partial sums are subtracted with atomics only when grid.y > 1).
constexpr int kSyrkKT = 16;
__global__ void __launch_bounds__(128) k_schur_syrk(LmDims d, const double *__restrict__ Y, double *__restrict__ S) {
    __shared__ double sA[2][kSyrkKT * kTS];
    __shared__ double sB[2][kSyrkKT * kTS];
    int bi, bj;
    tile_index(blockIdx.x, bi, bj);
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
    // each thread loads 8 doubles of each operand tile per step: 16 rows x 64 cols
    auto load = [&](int buf, int ks) {
        const int kbase = ks * kSyrkKT;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int idx = threadIdx.x + t * 128;  // 0..511 double2 slots
            const int row = idx >> 5, c2 = (idx & 31) * 2;
            const double2 va = *reinterpret_cast<const double2 *>(&Y[(size_t)(kbase + row) * d.NSP + bi * 64 + c2]);
            const double2 vb = *reinterpret_cast<const double2 *>(&Y[(size_t)(kbase + row) * d.NSP + bj * 64 + c2]);
            sA[buf][row * kTS + c2] = va.x; sA[buf][row * kTS + c2 + 1] = va.y;
            sB[buf][row * kTS + c2] = vb.x; sB[buf][row * kTS + c2 + 1] = vb.y;
        }
    };
    if (ks0 < ks1) load(0, ks0);
    __syncthreads();
    for (int ks = ks0; ks < ks1; ++ks) {
        const int buf = (ks - ks0) & 1;
        if (ks + 1 < ks1) load(buf ^ 1, ks + 1);
        tile_mma<kSyrkKT>(sA[buf], sB[buf], wr, wc, lane, acc);
        __syncthreads();
    }
    if (gridDim.y == 1) {
        tile_store_sub(S, d.NSP, bi, bj, wr, wc, lane, acc);
    } else {
        const int g = lane >> 2, q = lane & 3;
#pragma unroll
        for (int ti = 0; ti < 4; ++ti)
#pragma unroll
            for (int tj = 0; tj < 4; ++tj) {
                const int r = bi * 64 + wr * 32 + ti * 8 + g, c = bj * 64 + wc * 32 + tj * 8 + q * 2;
                atomicAdd(&S[(size_t)r * d.NSP + c], -acc[ti][tj][0]);
                atomicAdd(&S[(size_t)r * d.NSP + c + 1], -acc[ti][tj][1]);
            }
    }
}

// ---- blocked Cholesky of S (lower, row-major, ld = NSP) ----------------------------
// Step kb, kernel 1: every CTA factors the diagonal block redundantly in shared
// memory; CTA 0 stores it to `Ld`; CTA r>0 solves its 64-row panel block
// X = A_ik L_kk^-T in place.  The pivot of the augmented row (global index ns)
// is skipped: that row is the right-hand side being forward-solved.
constexpr int kCS = 65;  // shared stride for 64x64 tiles in the panel kernel
__global__ void __launch_bounds__(256) k_chol_panel(LmDims d, int kb, LmState *st, double *__restrict__ S,
                                                    double *__restrict__ Ld) {
    extern __shared__ double sm[];
    double *L = sm;                 // 64 x kCS
    double *X = sm + 64 * kCS;      // 64 x kCS
    __shared__ int fail;
    const int tid = threadIdx.x;
    if (tid == 0) fail = 0;
    const int base = kb * 64;
    for (int t = tid; t < 64 * 64; t += blockDim.x) {
        const int r = t >> 6, c = t & 63;
        L[r * kCS + c] = (c <= r) ? S[(size_t)(base + r) * d.NSP + base + c] : 0.0;
    }
    __syncthreads();
    for (int j = 0; j < 64; ++j) {
        if (tid == 0) {
            const int gj = base + j;
            double pv = L[j * kCS + j];
            if (gj == d.ns) pv = 1.0;  // augmented row: no pivot
            else if (!(pv > 0.0)) { fail = 1; pv = 1.0; }
            else pv = sqrt(pv);
            L[j * kCS + j] = pv;
        }
        __syncthreads();
        const double dj = L[j * kCS + j];
        if (tid > j && tid < 64) L[tid * kCS + j] /= dj;
        __syncthreads();
        // trailing update of the lower triangle: rows r > j, cols j < c <= r
        for (int t = tid; t < 64 * 64; t += blockDim.x) {
            const int r = t >> 6, c = t & 63;
            if (c > j && c <= r) L[r * kCS + c] -= L[r * kCS + j] * L[c * kCS + j];
        }
        __syncthreads();
    }
    if (blockIdx.x == 0) {
        if (tid == 0 && fail) atomicExch(&st->chol_fail, 1);
        for (int t = tid; t < 64 * 64; t += blockDim.x) Ld[t] = L[(t >> 6) * kCS + (t & 63)];
        return;
    }
    const int rb = (kb + blockIdx.x) * 64;
    for (int t = tid; t < 64 * 64; t += blockDim.x) {
        const int r = t >> 6, c = t & 63;
        X[r * kCS + c] = S[(size_t)(rb + r) * d.NSP + base + c];
    }
    __syncthreads();
    // x L^T = a  ->  x[c] = (a[c] - sum_{j<c} x[j] L[c][j]) / L[c][c]; 4 threads per row
    {
        const int r = tid >> 2, part = tid & 3;
        for (int c = 0; c < 64; ++c) {
            double sacc = 0.0;
            for (int j = part; j < c; j += 4) sacc += X[r * kCS + j] * L[c * kCS + j];
            sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
            sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
            if (part == 0) X[r * kCS + c] = (X[r * kCS + c] - sacc) / L[c * kCS + c];
            __syncwarp();
        }
    }
    __syncthreads();
    for (int t = tid; t < 64 * 64; t += blockDim.x) {
        const int r = t >> 6, c = t & 63;
        S[(size_t)(rb + r) * d.NSP + base + c] = X[r * kCS + c];
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

// ---- eliminated-side back substitution -------------------------------------------------
// p_f = L_f^-T (z_f - Y_f p_S); CTA per frame.  p_int = [p_E (KE) | p_S (ns)]
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
    double acc = 0.0;
    for (int r = r0; r < r1; ++r) acc += Y[(size_t)r * d.NSP + j] * w_int[r];
    partial[(size_t)blockIdx.y * d.NSP + j] = acc;
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
