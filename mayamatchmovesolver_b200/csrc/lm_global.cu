// Device Levenberg-Marquardt driver for the global (all frames at once) solve.
//
// Replaces solve_3d_cminpack_lmder (adjust_cminpack_lmder.cpp:64-198) and the
// tail of solveFrames (adjust_base.cpp:1076-1287).  The control law is
// MINPACK's lmder/lmpar evaluated by single-block
// control kernels on the device; the linear algebra is the Schur/Cholesky
// path of lm_kernels.cuh.
//
// Unsharded solve: the whole LM loop is ONE CUDA graph.  Its conditional nodes
//   WHILE(!done) { IF(need_jacobian){J, assembly}  WHILE(!lmpar_done){factor, solve}
//                  IF(!done){trial, ratio test, accept} }
// are steered by tiny kernels that copy LmState flags into the conditional
// handles (cudaGraphSetConditional), so the host launches the graph once and
// reads the result once: no host round trip per iteration.
// Sharded solve (NCCL all-reduces between the kernels): the host enqueues the
// same kernels and reads the flag words of LmState to know what comes next.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "comm.h"
#include "lm.h"
#include "lm_kernels.cuh"

namespace mmb {

constexpr int kRedBlocks = 592;  // 148 SMs x 4 partial-sum blocks
// One contiguous reduction scratch, so a sharded solve all-reduces each stage in one call:
//   [part_f kRedBlocks][part_jp kRedBlocks][scal kRedScal (sums)][mx kRedMax (maxima)]
constexpr int kRedScal = 8, kRedMax = 4;
constexpr int kStatBlocks = 148;

// One pinned (mapped) host block per workspace: [LmState mirror | cancel flag | staging].
// Pinning is slow (~1 ms), so the block of a destroyed workspace is kept for the next one.
struct PinnedBlock { void *ptr = nullptr; size_t bytes = 0; };
static std::mutex g_pinned_mu;
static PinnedBlock g_pinned_spare;
static cudaError_t pinned_acquire(size_t bytes, PinnedBlock *out) {
    {
        std::lock_guard<std::mutex> lk(g_pinned_mu);
        if (g_pinned_spare.ptr && g_pinned_spare.bytes >= bytes) {
            *out = g_pinned_spare;
            g_pinned_spare = PinnedBlock{};
            return cudaSuccess;
        }
    }
    out->bytes = bytes;
    return cudaHostAlloc(&out->ptr, bytes, cudaHostAllocMapped | cudaHostAllocPortable);
}
static void pinned_release(PinnedBlock b) {
    if (!b.ptr) return;
    {
        std::lock_guard<std::mutex> lk(g_pinned_mu);
        if (!g_pinned_spare.ptr || g_pinned_spare.bytes < b.bytes) std::swap(g_pinned_spare, b);
    }
    if (b.ptr) cudaFreeHost(b.ptr);
}

struct LmWorkspace {
    LmDims d{};
    DeviceArena arena;
    LmState *st = nullptr;     // device
    PinnedBlock pinned;        // backs h_st, h_cancel, h_stage
    LmState *h_st = nullptr;   // pinned host mirror
    int32_t *attr_slot = nullptr, *i2p = nullptr;
    AsmShared asm_shared{};    // columns shared by every marker (k_assemble_shared)
    // order-independent assembly (lm_kernels.cuh "order-independent accumulation")
    unsigned long long *cmax = nullptr;   // [n + 1]
    long long *det_limbs = nullptr;       // Ui | gEi | Vi | Wci, zeroed with one memset per Jacobian
    size_t det_limb_count = 0;
    int32_t *cs_of_slot = nullptr, *slot_of_cs = nullptr;
    int ncs = 0;
    int *poison = nullptr;
    DetBufs det{};
    // stiffness / smoothness rows: their one Jacobian entry each and its parameter
    double *jext = nullptr;
    int32_t *ext_pid = nullptr, *p2i = nullptr;
    int *d_zero = nullptr;       // an int that stays 0: the cancel word k_cond polls in a sharded graph (see lm_graph_build)
    bool tri_is_peer = false;
    double *tri_buf = nullptr;   // sharded solve: packed lower triangle of S for the all-reduce (allocated on first use)
    double *U = nullptr, *gE = nullptr, *W = nullptr, *Y = nullptr, *V = nullptr, *S = nullptr;
    double *Lf = nullptr, *Ld = nullptr;
    double *Dinv = nullptr;    // [NSP/64][64*64] inverses of the factor's diagonal blocks (k_chol_tiles phase E)
    double *diag = nullptr, *acnorm = nullptr;        // [n] by parameter id
    double *p_int = nullptr, *w_int = nullptr;        // [KE + NSP] internal order
    double *p_par = nullptr, *x2 = nullptr;           // [n] parameter order
    double *fvec2 = nullptr, *dist2 = nullptr;
    double *part_f = nullptr, *part_jp = nullptr;     // [kRedBlocks]
    double *scal = nullptr, *mx = nullptr;            // [kRedScal], [kRedMax]
    double *sdiag = nullptr;                          // [2 ns]: diagonal of V, then J^T f of the static side
    double *tvec = nullptr;                           // [ns]: Y^T w
    double *stat_part = nullptr;                      // [3 kStatBlocks]
    double *part_gemv = nullptr;                      // [chunks * NSP]
    int gemv_chunks = 1;
    long long n_allreduce = 0;
    int *wave_flags = nullptr;   // [NSP / 64] epoch flags of the wavefront triangular solves
    int *wave_ctr = nullptr;     // device epoch counter (k_ctrl_lmpar_decide / _update add 2 per lmpar step)
    int *tile_flags = nullptr;   // [(NSP/64)^2] epoch flags of the tile-dataflow Cholesky
    int *syrk_ready = nullptr;   // [(NSP/64)^2] epoch flags of the SYRK's tiles (Cholesky running beside the SYRK)
    cudaStream_t side = nullptr; // the stream of that Cholesky
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int chol_beside = 0;         // CTAs of the Cholesky launched beside the SYRK (0: one after the other)
    PeerTiles ptiles{};          // sharded solve: partial tiles through the peer arenas (world == 0: packed all-reduce instead)
    unsigned *peer_seq = nullptr; // this rank's factorisation sequence number (in its arena)
    bool ptiles_decided = false;
    bool chol_tiles = true;      // one-launch tile-dataflow Cholesky (MMB_CHOL=steps: two launches per block step)
    double *h_stage = nullptr;   // pinned staging: [max(2 n, kRedBlocks + n)]
    int *h_cancel = nullptr;     // pinned, mapped: set by mmb_request_cancel, polled by the loop-head kernel
    int *d_cancel = nullptr;     // device alias of h_cancel
    // the device-resident loop
    bool capturing = false;      // enqueue helpers are recording into a graph
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    bool graph_failed = false;
    int body_launches[5] = {0, 0, 0, 0, 0};  // kernels per execution of: loop head/tail, jacobian body, lmpar body, trial body, Newton-correction body
    int fixed_launches = 0;               // prologue + epilogue kernels
    bool wave_ok = false;        // every block of the wavefront kernels fits on the device at once
    double init_avg = 0, init_min = 0, init_max = 0;
    // optional phase timing (MMB_LM_PROFILE=1): schur, cholesky, back-substitution, q-solve
    bool profile = false;
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    double phase_ms[4] = {0, 0, 0, 0};
    int phase_n = 0;
    ~LmWorkspace() {
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
        if (side) cudaStreamDestroy(side);
        if (graph_exec) cudaGraphExecDestroy(graph_exec);
        if (graph) cudaGraphDestroy(graph);
        pinned_release(pinned);
    }
};

void lm_workspace_destroy(LmWorkspace *w) { delete w; }

// ---- control kernels (one block of 1024 threads) ---------------------------------------
// Sums over all parameters are split in two: the eliminated side (animated parameters:
// local to a shard, so its part is all-reduced when the solve is sharded) goes to the
// reduction scratch `scal` / `mx`; the reduced side (static parameters: replicated and
// identical on every rank) goes to st->part_s and is added after the all-reduce.
__device__ double sum_partials(const double *part, int n, double *sh) {
    double a = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) a += part[i];
    return block_sum(a, sh);
}

__global__ void k_ctrl_init(LmState *st, const double *part_f, int nparts) {
    __shared__ double sh[32];
    const double ss = sum_partials(part_f, nparts, sh);
    if (threadIdx.x == 0) {
        st->fnorm = sqrt(ss);
        st->par = 0.0; st->par_cur = 0.0;
        st->iter = 1; st->nfev = 1; st->njev = 0; st->info = 0;
        st->need_jacobian = 1; st->done = 0; st->accepted = 0;
        st->delta = 0.0; st->xnorm = 0.0; st->gnorm = 0.0;
        st->chol_fail_total = 0; st->n_factor = 0; st->dataflow_timeout = 0;
        // lmder input checks (info = 0: improper input parameters)
        if (st->n <= 0 || st->m < st->n || st->ftol < 0.0 || st->xtol < 0.0 || st->gtol < 0.0 ||
            st->maxfev <= 0 || st->factor <= 0.0) {
            st->info = 0; st->done = 1;
        }
    }
}

// diagonal and right-hand side of the static block: what the column norms / gradient test
// need from V.  A sharded solve all-reduces these 2 ns values instead of the whole block.
__global__ void k_prep_gather(LmDims d, const double *__restrict__ V, double *__restrict__ sdiag) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d.ns) return;
    sdiag[j] = V[(size_t)j * d.NSP + j];
    sdiag[d.ns + j] = V[(size_t)d.ns * d.NSP + j];
}

// After assembly: column norms and diag; partial sums of xnorm^2, ||g/diag||^2 and the
// gradient test's maximum.
__global__ void __launch_bounds__(1024) k_ctrl_prep(LmDims d, LmState *st, const double *U, const double *gE,
                                                    const double *sdiag, const int32_t *i2p, const double *x,
                                                    double *diag, double *acnorm, double *scal, double *mx) {
    __shared__ double sh[32];
    __shared__ double wmax[2][32];
    const int n = d.n;
    const bool first = st->iter == 1;
    const double fnorm = st->fnorm;
    double xsE = 0.0, gsE = 0.0, xsS = 0.0, gsS = 0.0, gmaxE = 0.0, gmaxS = 0.0;
    for (int q = threadIdx.x; q < n; q += blockDim.x) {
        const int pid = i2p[q];
        const bool elim = q < d.KE;
        double h, g;
        if (elim) {
            const int f = q / d.na, a = q - f * d.na;
            h = U[((size_t)f * d.na + a) * d.na + a];
            g = gE[q];
        } else {
            const int s = q - d.KE;
            h = sdiag[s];
            g = sdiag[d.ns + s];
        }
        const double ac = sqrt(fmax(h, 0.0));
        acnorm[pid] = ac;
        double dg = diag[pid];
        if (st->mode != 2) {
            if (first) dg = (ac == 0.0) ? 1.0 : ac;
        }
        if (first) {
            const double t = dg * x[pid];
            if (elim) xsE += t * t; else xsS += t * t;
        }
        if (fnorm != 0.0 && ac != 0.0) {
            const double t = fabs(g / fnorm) / ac;
            if (elim) gmaxE = fmax(gmaxE, t); else gmaxS = fmax(gmaxS, t);
        }
        if (st->mode != 2) dg = fmax(dg, ac);
        diag[pid] = dg;
        const double t2 = g / dg;
        if (elim) gsE += t2 * t2; else gsS += t2 * t2;
    }
    xsE = block_sum(xsE, sh); gsE = block_sum(gsE, sh);
    xsS = block_sum(xsS, sh); gsS = block_sum(gsS, sh);
    for (int o = 16; o > 0; o >>= 1) {
        gmaxE = fmax(gmaxE, __shfl_xor_sync(0xffffffffu, gmaxE, o));
        gmaxS = fmax(gmaxS, __shfl_xor_sync(0xffffffffu, gmaxS, o));
    }
    if ((threadIdx.x & 31) == 0) { wmax[0][threadIdx.x >> 5] = gmaxE; wmax[1][threadIdx.x >> 5] = gmaxS; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ge = 0.0, gs2 = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ge = fmax(ge, wmax[0][w]); gs2 = fmax(gs2, wmax[1][w]); }
        scal[0] = xsE; scal[1] = gsE; mx[0] = ge;
        st->part_s[0] = xsS; st->part_s[1] = gsS; st->part_s[2] = gs2;
    }
}

// xnorm / delta on the first iteration, the gtol test (info 4), start of lmpar
__global__ void k_ctrl_prep_decide(LmState *st, const double *scal, const double *mx) {
    const bool first = st->iter == 1;
    const double g = fmax(mx[0], st->part_s[2]);
    st->njev += 1;
    st->nfev += st->nfev_per_jac;
    if (first) {
        st->xnorm = sqrt(scal[0] + st->part_s[0]);
        st->delta = st->factor * st->xnorm;
        if (st->delta == 0.0) st->delta = st->factor;
    }
    st->gnorm = g;
    st->gnorm_scaled = sqrt(scal[1] + st->part_s[1]);
    st->need_jacobian = 0;
    if (g <= st->gtol) { st->info = 4; st->done = 1; }
    st->lmpar_iter = 0; st->lmpar_done = 0; st->chol_fail = 0; st->par_cur = 0.0;
}

// After a factorisation + solve: eliminated-side part of ||D p||^2 and the failure flag.
__global__ void __launch_bounds__(1024) k_lmpar_partial(LmDims d, LmState *st, const double *p_int,
                                                        const double *diag, const int32_t *i2p, double *scal) {
    __shared__ double sh[32];
    double dxE = 0.0, dxS = 0.0;
    for (int q = threadIdx.x; q < d.n; q += blockDim.x) {
        const double t = diag[i2p[q]] * p_int[q];
        if (q < d.KE) dxE += t * t;
        else dxS += t * t;
    }
    dxE = block_sum(dxE, sh); dxS = block_sum(dxS, sh);
    if (threadIdx.x == 0) {
        scal[0] = dxE; scal[1] = st->chol_fail ? 1.0 : 0.0;
        st->part_s[0] = dxS;
    }
}

// eliminated-side part of ||L^-1 D^2 p||^2 (the Newton-correction vector, still unscaled by 1/||D p||)
__global__ void __launch_bounds__(1024) k_lmpar_partial_w(LmDims d, const double *w_int, double *scal) {
    __shared__ double sh[32];
    double wE = 0.0;
    for (int q = threadIdx.x; q < d.KE; q += blockDim.x) { const double w = w_int[q]; wE += w * w; }
    wE = block_sum(wE, sh);
    if (threadIdx.x == 0) scal[0] = wE;
}

// lmpar step (MINPACK lmpar restated on the normal equations), first half: everything ||D p||
// decides.  lmpar ends here at most steps (Gauss-Newton step inside the trust region, or
// | ||D p|| - delta | <= 0.1 delta); only a step that goes on needs the Newton correction
// ||L^-1 D^2 p||, i.e. another pass over the eliminated side and a triangular solve with the
// reduced factor: lmpar_corr asks for it and k_ctrl_lmpar_update finishes the step.
__global__ void k_ctrl_lmpar_decide(LmState *st, const double *scal, int *wave_ctr) {
    const double p1 = 0.1, p001 = 0.001, dwarf = 2.2250738585072014e-308;
    const double delta = st->delta;
    const bool fail = scal[1] > 0.0;  // a pivot <= 0 on any rank
    if (fail) st->chol_fail_total += 1;
    st->chol_fail = 0;
    st->n_factor += 1;
    const double dx2 = scal[0] + st->part_s[0];
    st->dx2 = dx2;
    st->dxnorm = fail ? INFINITY : sqrt(dx2);
    st->lmpar_corr = 0;
    st->qnorm2 = 0.0;
    if (st->lmpar_iter == 0) {
        // Gauss-Newton probe at par = 0
        const double dxnorm = st->dxnorm;
        const double fp = dxnorm - delta;
        st->fp = fp;
        if (!fail && fp <= p1 * delta) { st->par = 0.0; st->lmpar_done = 1; *wave_ctr += 2; return; }
        if (!fail && dx2 > 0.0) { st->lmpar_corr = 1; return; }   // parl needs the correction
        double paru = st->gnorm_scaled / delta;
        if (paru == 0.0) paru = dwarf / fmin(delta, p1);
        double par = st->par;
        par = fmax(par, 0.0);
        par = fmin(par, paru);
        if (par == 0.0) par = fail ? 0.0 : st->gnorm_scaled / dxnorm;
        if (par == 0.0) par = fmax(dwarf, p001 * paru);
        st->parl = 0.0; st->paru = paru; st->par = par; st->par_cur = par;
        st->lmpar_iter = 1;
        *wave_ctr += 2;
        return;
    }
    if (fail) {
        // damped system not positive definite (numerical): raise the damping
        st->parl = fmax(st->parl, st->par);
        st->par = fmax(st->par * 10.0, p001 * st->paru);
        st->par_cur = st->par;
        st->lmpar_iter += 1;
        if (st->lmpar_iter > 30) { st->lmpar_done = 1; st->info = -2; st->done = 1; }
        *wave_ctr += 2;
        return;
    }
    const double temp = st->fp;
    const double fp = st->dxnorm - delta;
    st->fp = fp;
    if (fabs(fp) <= p1 * delta || (st->parl == 0.0 && fp <= temp && temp < 0.0) || st->lmpar_iter >= 10) {
        st->lmpar_done = 1;
        *wave_ctr += 2;
        return;
    }
    st->lmpar_corr = 1;
}

// second half, for a step that goes on: the next par from the Newton correction
__global__ void __launch_bounds__(1024) k_ctrl_lmpar_update(LmDims d, LmState *st, const double *w_int, const double *scal,
                                                            int *wave_ctr) {
    __shared__ double sh[32];
    if (!st->lmpar_corr) return;   // (the host-driven loop enqueues both halves unconditionally)
    double a = 0.0;
    for (int q = d.KE + threadIdx.x; q < d.n; q += blockDim.x) {
        const double t = w_int[q];
        a += t * t;
    }
    a = block_sum(a, sh);
    if (threadIdx.x != 0) return;
    *wave_ctr += 2;  // the two wavefront solves of this step are done: retire their epochs
    st->lmpar_corr = 0;
    const double p1 = 0.1, p001 = 0.001, dwarf = 2.2250738585072014e-308;
    const double delta = st->delta;
    const double dx2 = st->dx2;
    // ||L^-1 (D^2 p / dxnorm)||^2: the solves ran without the 1/dxnorm factor
    const double qn2 = !(dx2 > 0.0) ? 0.0 : (scal[0] + a) / dx2;
    st->qnorm2 = qn2;
    const double fp = st->fp;
    if (st->lmpar_iter == 0) {
        const double dxnorm = st->dxnorm;
        double parl = 0.0;
        if (qn2 > 0.0) parl = ((fp / delta) / qn2);
        double paru = st->gnorm_scaled / delta;
        if (paru == 0.0) paru = dwarf / fmin(delta, p1);
        double par = st->par;
        par = fmax(par, parl);
        par = fmin(par, paru);
        if (par == 0.0) par = st->gnorm_scaled / dxnorm;
        if (par == 0.0) par = fmax(dwarf, p001 * paru);
        st->parl = parl; st->paru = paru; st->par = par; st->par_cur = par;
        st->lmpar_iter = 1;
        return;
    }
    const double parc = ((fp / delta) / qn2);
    if (fp > 0.0) st->parl = fmax(st->parl, st->par);
    if (fp < 0.0) st->paru = fmin(st->paru, st->par);
    double par = fmax(st->parl, st->par + parc);
    if (par == 0.0) par = fmax(dwarf, p001 * st->paru);
    st->par = par; st->par_cur = par;
    st->lmpar_iter += 1;
}

// trial point: x2 = x - p; partial sums of pnorm^2 = ||D p||^2 and of ||D x2||^2
// scal[2] carries this rank's cancel request through the trial's all-reduce, so that every rank
// of a sharded solve leaves the loop at the same point (k_ctrl_accept)
__global__ void __launch_bounds__(1024) k_ctrl_trial(LmDims d, LmState *st, const double *p_int, const int32_t *i2p,
                                                     const double *diag, const double *x, double *x2, double *p_par,
                                                     double *scal, const int *cancel) {
    __shared__ double sh[32];
    double pE = 0.0, pS = 0.0, xE = 0.0, xS = 0.0;
    for (int q = threadIdx.x; q < d.n; q += blockDim.x) {
        const int pid = i2p[q];
        const double step = -p_int[q];
        p_par[pid] = step;
        const double xn = x[pid] + step;
        x2[pid] = xn;
        const double t = diag[pid] * step, u = diag[pid] * xn;
        if (q < d.KE) { pE += t * t; xE += u * u; } else { pS += t * t; xS += u * u; }
    }
    pE = block_sum(pE, sh); pS = block_sum(pS, sh); xE = block_sum(xE, sh); xS = block_sum(xS, sh);
    if (threadIdx.x == 0) {
        scal[0] = pE; scal[1] = xE;
        scal[2] = *(const volatile int *)cancel ? 1.0 : 0.0;
        st->part_s[0] = pS; st->part_s[1] = xS;
    }
}

// ratio test, delta / par update, accept / reject, convergence tests
__global__ void __launch_bounds__(1024) k_ctrl_accept(LmDims d, LmState *st, const double *part_f, const double *part_jp,
                                                      int nparts, const double *scal) {
    __shared__ double sh[32];
    const double ss1 = sum_partials(part_f, nparts, sh);
    const double jp2 = sum_partials(part_jp, nparts, sh);
    if (threadIdx.x != 0) return;
    const double p1 = 0.1, p5 = 0.5, p25 = 0.25, p75 = 0.75, p0001 = 1e-4, epsmch = 2.220446049250313e-16;
    const double pnorm = sqrt(scal[0] + st->part_s[0]);
    const double xs = scal[1] + st->part_s[1];
    st->pnorm = pnorm;
    if (st->iter == 1) st->delta = fmin(st->delta, pnorm);
    const double fnorm = st->fnorm, fnorm1 = sqrt(ss1);
    st->fnorm1 = fnorm1;
    st->jp_sumsq = jp2;
    st->nfev += 1;
    double actred = -1.0;
    if (p1 * fnorm1 < fnorm) { const double t = fnorm1 / fnorm; actred = 1.0 - t * t; }
    const double temp1 = sqrt(jp2) / fnorm;
    const double temp2 = (sqrt(st->par) * pnorm) / fnorm;
    const double prered = temp1 * temp1 + temp2 * temp2 / p5;
    const double dirder = -(temp1 * temp1 + temp2 * temp2);
    double ratio = 0.0;
    if (prered != 0.0) ratio = actred / prered;
    if (ratio <= p25) {
        double temp;
        if (actred >= 0.0) temp = p5; else temp = p5 * dirder / (dirder + p5 * actred);
        if (p1 * fnorm1 >= fnorm || temp < p1) temp = p1;
        st->delta = temp * fmin(st->delta, pnorm / p1);
        st->par /= temp;
    } else if (st->par == 0.0 || ratio >= p75) {
        st->delta = pnorm / p5;
        st->par = p5 * st->par;
    }
    st->accepted = 0;
    if (ratio >= p0001) {
        st->accepted = 1;
        st->xnorm = sqrt(xs);
        st->fnorm = fnorm1;
        st->iter += 1;
    }
    st->ratio = ratio; st->actred = actred; st->prered = prered;
    int info = 0;
    const bool small_red = fabs(actred) <= st->ftol && prered <= st->ftol && p5 * ratio <= 1.0;
    if (small_red) info = 1;
    if (st->delta <= st->xtol * st->xnorm) info = 2;
    if (small_red && info == 2) info = 3;
    if (info == 0) {
        if (st->nfev >= st->maxfev) info = 5;
        if (fabs(actred) <= epsmch && prered <= epsmch && p5 * ratio <= 1.0) info = 6;
        if (st->delta <= epsmch * st->xnorm) info = 7;
        if (st->gnorm <= epsmch) info = 8;
    }
    if (info != 0) { st->info = info; st->done = 1; }
    if (scal[2] > 0.0) { st->cancelled = 1; st->done = 1; }   // some rank was asked to stop
    // next pass: a new Jacobian after an accepted step, else lmpar again
    st->need_jacobian = st->accepted;
    st->lmpar_iter = 0; st->lmpar_done = 0; st->par_cur = 0.0; st->chol_fail = 0;
    (void)d;
}

__global__ void k_accept_copy(const LmState *st, size_t n, size_t m, size_t nobs, const double *x2, double *x,
                              const double *f2, double *f, const double *d2, double *dd) {
    if (!st->accepted) return;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < m; i += (size_t)gridDim.x * blockDim.x) {
        f[i] = f2[i];
        if (i < n) x[i] = x2[i];
        if (i < nobs) dd[i] = d2[i];
    }
}

// compute_error_stats (adjust_base.cpp:346-372): avg / min / max of finite distances.
// Stage 1: per-block sum / min / max; stage 2 folds them into scal / mx (sum and count;
// max and -min, so a sharded solve needs one sum and one max all-reduce); stage 3
// writes LmState.
__global__ void __launch_bounds__(256) k_error_stats_partial(const double *__restrict__ dist, int nobs, double *part) {
    __shared__ double sh[32];
    __shared__ double smin[8], smax[8];
    double a = 0.0, mn = 1.7976931348623157e308, mx = -0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nobs; i += gridDim.x * blockDim.x) {
        const double e = dist[i];
        if (!isfinite(e)) continue;
        a += e; mn = fmin(mn, e); mx = fmax(mx, e);
    }
    a = block_sum(a, sh);
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = mn; smax[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { mn = fmin(mn, smin[w]); mx = fmax(mx, smax[w]); }
        part[3 * blockIdx.x] = a; part[3 * blockIdx.x + 1] = mn; part[3 * blockIdx.x + 2] = mx;
    }
}
__global__ void k_error_stats_fold(const double *part, int nparts, int nobs, double *scal, double *mxv) {
    if (threadIdx.x != 0) return;
    double a = 0.0, mn = 1.7976931348623157e308, mx = -0.0;
    for (int b = 0; b < nparts; ++b) { a += part[3 * b]; mn = fmin(mn, part[3 * b + 1]); mx = fmax(mx, part[3 * b + 2]); }
    scal[0] = a; scal[1] = (double)nobs;
    mxv[0] = mx; mxv[1] = -mn;
}
__global__ void k_error_stats_final(LmState *st, const double *scal, const double *mxv, int to_init) {
    st->err_avg = scal[0] / scal[1]; st->err_max = mxv[0]; st->err_min = -mxv[1];
    if (to_init) st->init_avg = st->err_avg;
}

// ---- phase boundaries of the device-resident loop ---------------------------------------
// One thread: stamps the phase timers (%globaltimer) and copies LmState flags into the
// conditional handles of the graph (the device-side half of cudaGraphConditionalHandle).
enum { kCondStart = 0, kCondLoopHead, kCondAfterJac, kCondInLmpar, kCondAfterLmpar, kCondCorr, kCondAfterTrial };
__global__ void k_cond(LmState *st, const int *cancel, int where, cudaGraphConditionalHandle h0,
                       cudaGraphConditionalHandle h1) {
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    switch (where) {
        case kCondStart:  // after k_ctrl_init; h0 = outer WHILE
            cudaGraphSetConditional(h0, st->done ? 0u : 1u);
            break;
        case kCondLoopHead:  // h0 = IF(jacobian)
            if (*(const volatile int *)cancel) { st->cancelled = 1; st->done = 1; }
            st->n_body_loop += 1;
            st->t_mark = now;
            cudaGraphSetConditional(h0, (!st->done && st->need_jacobian) ? 1u : 0u);
            break;
        case kCondAfterJac:  // h0 = WHILE(lmpar)
            st->t_jac_ns += now - st->t_mark;
            st->t_mark = now;
            cudaGraphSetConditional(h0, (!st->done && !st->lmpar_done) ? 1u : 0u);
            break;
        case kCondInLmpar:  // tail of the lmpar body; h0 = WHILE(lmpar)
            st->n_body_lmpar += 1;
            cudaGraphSetConditional(h0, (!st->done && !st->lmpar_done) ? 1u : 0u);
            break;
        case kCondCorr:  // inside the lmpar body, after the decision; h0 = IF(Newton correction)
            if (st->lmpar_corr) st->n_body_corr += 1;
            cudaGraphSetConditional(h0, st->lmpar_corr ? 1u : 0u);
            break;
        case kCondAfterLmpar:  // h0 = IF(trial); h1 = outer WHILE (the trial body may be skipped)
            st->t_lin_ns += now - st->t_mark;
            st->t_mark = now;
            cudaGraphSetConditional(h0, st->done ? 0u : 1u);
            cudaGraphSetConditional(h1, st->done ? 0u : 1u);
            break;
        default:  // kCondAfterTrial: tail of the trial body; h0 = outer WHILE
            st->n_body_trial += 1;
            st->t_func_ns += now - st->t_mark;
            cudaGraphSetConditional(h0, st->done ? 0u : 1u);
            break;
    }
}

// debug: the Cholesky critical-path timeline of the last factorisation (see g_chol_trace)
mmb_status lm_chol_trace(unsigned long long *out, int count) {
    MMB_CUDA_CHECK(cudaDeviceSynchronize());
    MMB_CUDA_CHECK(cudaMemcpyFromSymbol(out, g_chol_trace, sizeof(unsigned long long) * std::min(count, 64 * 2 * 6)));
    return MMB_OK;
}

// ---- workspace ------------------------------------------------------------------------------
static mmb_status lm_workspace_create(DeviceProblem *p) {
    if (p->lm) return MMB_OK;
    const HostProblem &hp = p->hp;
    if (hp.na > kMaxNA) {
        set_last_error("more than 64 animated solver attributes per frame are not supported");
        return MMB_ERR_UNSUPPORTED;
    }
    if ((long long)hp.nobs >= (1ll << 27)) {
        // the fixed-point limbs of the normal equations hold sums of up to 2^27 contributions
        set_last_error("more than 2^27 observations per GPU: shard the frames over more ranks");
        return MMB_ERR_UNSUPPORTED;
    }
    LmWorkspace *w = new LmWorkspace();
    LmDims &d = w->d;
    d.F = hp.F; d.na = hp.na; d.ns = hp.ns; d.NS1 = hp.ns + 1;
    d.NSP = ((d.NS1 + 63) / 64) * 64;
    d.KE = hp.F * hp.na; d.KEP = ((d.KE + kSyrkKT - 1) / kSyrkKT) * kSyrkKT;
    d.n = hp.n; d.nobs = hp.nobs; d.m = hp.m; d.kmax = p->ds.kmax;
    DeviceArena &A = w->arena;
    A.bind(p->stream);
#define CKW(expr)                                                                \
    do {                                                                         \
        cudaError_t _e = (expr);                                                 \
        if (_e != cudaSuccess) {                                                 \
            set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e));  \
            delete w;                                                            \
            return MMB_ERR_CUDA;                                                 \
        }                                                                        \
    } while (0)
    CKW(A.alloc(&w->st, 1));
    {
        const size_t st_bytes = (sizeof(LmState) + 255) / 256 * 256;
        CKW(pinned_acquire(st_bytes + 256 + sizeof(double) * (size_t)(2 * d.n + kRedBlocks), &w->pinned));
        char *base = (char *)w->pinned.ptr;
        w->h_st = (LmState *)base;
        w->h_cancel = (int *)(base + st_bytes);
        w->h_stage = (double *)(base + st_bytes + 256);
        *w->h_cancel = 0;
        CKW(cudaHostGetDevicePointer((void **)&w->d_cancel, w->h_cancel, 0));
    }
    {   // columns whose attribute is the same for every marker (dense-grid scenes only)
        AsmShared &as = w->asm_shared;
        as.ns = 0; as.mask = 0;
        const char *e = getenv("MMB_ASM_SHARED");
        const bool enabled = !(e && atoi(e) == 0);
        if (enabled && hp.dense_grid && hp.M > 1 && hp.kmax <= 32) {
            const int k0 = hp.mkr_k[0];
            bool same_k = true;
            for (int mk = 1; mk < hp.M && same_k; ++mk) same_k = hp.mkr_k[mk] == k0;
            if (same_k) {
                for (int c = 0; c < k0 && as.ns < 16; ++c) {
                    const int a0 = hp.plan[c].attr;
                    bool same = a0 >= 0;
                    for (int mk = 1; mk < hp.M && same; ++mk) same = hp.plan[(size_t)mk * hp.kmax + c].attr == a0;
                    if (same) {
                        as.col[as.ns] = c; as.slot[as.ns] = hp.attr_slot[a0]; as.anim[as.ns] = hp.attr_anim[a0] ? 1 : 0;
                        as.pbase[as.ns] = hp.attr_param_base[a0];
                        as.mask |= 1u << c;
                        ++as.ns;
                    }
                }
            }
            const int ok[] = {6, 7, 9, 11, 13, 15};
            bool supported = false;
            for (int v : ok) supported = supported || v == as.ns;
            if (!supported) { as.ns = 0; as.mask = 0; }
        }
    }
    CKW(A.upload(&w->attr_slot, hp.attr_slot));
    CKW(A.upload(&w->i2p, hp.int_to_par));
    CKW(A.upload(&w->p2i, hp.par_to_int));
    CKW(A.alloc(&w->d_zero, 1));
    CKW(cudaMemsetAsync(w->d_zero, 0, sizeof(int), p->stream));
    CKW(A.alloc(&w->jext, (size_t)std::max(hp.n_extra, 1)));
    CKW(A.alloc(&w->ext_pid, (size_t)std::max(hp.n_extra, 1)));
    {
        // Which static columns of W collect three or more contributions per entry?  Entry
        // (frame, animated slot a, static slot s) gets one per marker whose plan holds both.
        std::vector<int32_t> cs_of_slot(std::max(d.ns, 1), -1), slot_of_cs;
        if (d.ns && d.na) {
            std::vector<int32_t> cnt((size_t)d.na * d.ns, 0);
            std::vector<uint8_t> hot(d.ns, 0);
            std::vector<int> as_, ss_;
            for (int mk = 0; mk < hp.M; ++mk) {
                as_.clear(); ss_.clear();
                for (int c = 0; c < hp.mkr_k[mk]; ++c) {
                    const int a = hp.plan[(size_t)mk * hp.kmax + c].attr;
                    (hp.attr_anim[a] ? as_ : ss_).push_back(hp.attr_slot[a]);
                }
                for (int sl : ss_)
                    for (int al : as_)
                        if (++cnt[(size_t)al * d.ns + sl] >= 3) hot[sl] = 1;
            }
            for (int sl = 0; sl < d.ns; ++sl)
                if (hot[sl]) { cs_of_slot[sl] = (int32_t)slot_of_cs.size(); slot_of_cs.push_back(sl); }
        }
        w->ncs = (int)slot_of_cs.size();
        if (slot_of_cs.empty()) slot_of_cs.push_back(0);
        CKW(A.upload(&w->cs_of_slot, cs_of_slot));
        CKW(A.upload(&w->slot_of_cs, slot_of_cs));
        const size_t nU = (size_t)d.F * d.na * d.na, nG = (size_t)d.KE, nV = (size_t)d.NSP * d.NSP, nW = (size_t)d.KE * w->ncs;
        w->det_limb_count = 2 * (nU + nG + nV + nW);
        CKW(A.alloc(&w->det_limbs, w->det_limb_count));
        CKW(A.alloc(&w->cmax, (size_t)d.n + 1));
        CKW(A.alloc(&w->poison, 1));
        w->det.cmax = w->cmax;
        w->det.Ui = w->det_limbs;
        w->det.gEi = w->det.Ui + 2 * nU;
        w->det.Vi = w->det.gEi + 2 * nG;
        w->det.Wci = w->det.Vi + 2 * nV;
        w->det.cs_of_slot = w->cs_of_slot;
        w->det.ncs = w->ncs;
        w->det.poison = w->poison;
    }
    CKW(A.alloc(&w->U, (size_t)d.F * d.na * d.na));
    CKW(A.alloc(&w->gE, (size_t)d.KE));
    CKW(A.alloc(&w->W, (size_t)d.KE * d.NSP));
    CKW(A.alloc(&w->Y, (size_t)d.KEP * d.NSP));
    CKW(cudaMemsetAsync(w->Y, 0, sizeof(double) * std::max<size_t>((size_t)d.KEP * d.NSP, 1), p->stream));
    CKW(A.alloc(&w->V, (size_t)d.NSP * d.NSP));
    CKW(A.alloc(&w->S, (size_t)d.NSP * d.NSP));
    CKW(A.alloc(&w->Lf, (size_t)d.F * d.na * d.na));
    CKW(A.alloc(&w->Ld, (size_t)64 * 64));
    CKW(A.alloc(&w->diag, (size_t)d.n));
    CKW(A.alloc(&w->acnorm, (size_t)d.n));
    CKW(A.alloc(&w->p_int, (size_t)d.KE + d.NSP));
    CKW(A.alloc(&w->w_int, (size_t)d.KE + d.NSP));
    CKW(A.alloc(&w->p_par, (size_t)d.n));
    CKW(A.alloc(&w->x2, (size_t)d.n));
    CKW(A.alloc(&w->fvec2, (size_t)hp.m_total));
    CKW(A.alloc(&w->dist2, (size_t)d.nobs));
    CKW(A.alloc(&w->part_f, (size_t)2 * kRedBlocks + kRedScal + kRedMax));
    w->part_jp = w->part_f + kRedBlocks;
    w->scal = w->part_jp + kRedBlocks;
    w->mx = w->scal + kRedScal;
    CKW(A.alloc(&w->sdiag, (size_t)2 * std::max(d.ns, 1)));
    CKW(A.alloc(&w->tvec, (size_t)std::max(d.ns, 1)));
    CKW(A.alloc(&w->stat_part, (size_t)3 * kStatBlocks));
    w->gemv_chunks = std::max(1, std::min(256, d.KE / 48));   // ~1100 CTAs of 128 threads at cfg3
    CKW(A.alloc(&w->part_gemv, (size_t)w->gemv_chunks * d.NSP));
    CKW(cudaFuncSetAttribute(k_chol_panel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholPanelSmem));
    CKW(cudaFuncSetAttribute(k_backward_s, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(double) * (64 * kDS + 64 + 32 * 64)));
    {
        const int nblk = std::max(1, (d.ns + kNB - 1) / kNB);
        CKW(A.alloc(&w->wave_flags, (size_t)d.NSP / 64 + 1));
        CKW(cudaMemsetAsync(w->wave_flags, 0, sizeof(int) * ((size_t)d.NSP / 64 + 1), p->stream));
        {
            const size_t nt = ((size_t)d.NSP / 64) * ((size_t)d.NSP / 64);
            CKW(A.alloc(&w->tile_flags, nt));
            CKW(cudaMemsetAsync(w->tile_flags, 0, sizeof(int) * nt, p->stream));
            CKW(A.alloc(&w->syrk_ready, nt));
            CKW(cudaMemsetAsync(w->syrk_ready, 0, sizeof(int) * nt, p->stream));
            CKW(cudaFuncSetAttribute(k_chol_tiles<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholTilesSmem));
            CKW(cudaFuncSetAttribute(k_chol_tiles<1, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholTilesSmem));
            CKW(cudaFuncSetAttribute(k_chol_tiles<1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholTilesSmem));
            CKW(cudaFuncSetAttribute(k_chol_tiles<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCholTilesSmem));
            const char *e = getenv("MMB_CHOL");
            w->chol_tiles = !(e && !strcmp(e, "steps"));
            // Cholesky beside the SYRK (chain-bound sizes only): that many resident CTAs, one per SM.
            // cfg3, linear algebra of a 15-iteration solve: 25.5 ms one after the other; beside, with
            // 16 / 20 / 24 / 28 / 32 / 40 / 48 CTAs: 23.5 / 21.5 / 21.3 / 21.3 / 21.7 / 22.3 / 23.2 ms.
            const char *eb = getenv("MMB_CHOL_BESIDE");
            w->chol_beside = eb ? atoi(eb) : 24;
            // chain-bound sizes only: at 47 block columns (8 cameras x 200 frames x 1000 bundles) the
            // factorisation is throughput-bound, 24 CTAs are too few for it and the sequence is faster
            // (13.75 vs 14.55 ms of linear algebra for three trials)
            if (!w->chol_tiles || d.KE == 0 || d.NSP / 64 > 32) w->chol_beside = 0;
            if (w->chol_beside > 0) {
                int sms = 0;
                CKW(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device));
                w->chol_beside = std::min(w->chol_beside, sms);
                CKW(cudaStreamCreateWithFlags(&w->side, cudaStreamNonBlocking));
                CKW(cudaEventCreateWithFlags(&w->ev_fork, cudaEventDisableTiming));
                CKW(cudaEventCreateWithFlags(&w->ev_join, cudaEventDisableTiming));
            }
            const char *ei = getenv("MMB_WAVE_INV");
            if (w->chol_tiles && !(ei && atoi(ei) == 0)) CKW(A.alloc(&w->Dinv, ((size_t)d.NSP / 64) * 4096));
        }
        CKW(A.alloc(&w->wave_ctr, 1));
        CKW(cudaMemsetAsync(w->wave_ctr, 0, sizeof(int), p->stream));
        CKW(cudaFuncSetAttribute(k_backward_wave, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWaveSmemBwd));
        CKW(cudaFuncSetAttribute(k_forward_wave, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWaveSmemFwd));
        int coop = 0, sms = 0, occ_b = 0, occ_f = 0;
        CKW(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, p->device));
        CKW(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device));
        CKW(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_b, k_backward_wave, 256, kWaveSmemBwd));
        CKW(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_f, k_forward_wave, 256, kWaveSmemFwd));
        const char *e = getenv("MMB_WAVE");
        w->wave_ok = coop && nblk <= sms * std::min(occ_b, occ_f) && !(e && atoi(e) == 0);
    }
#undef CKW
    { const char *e = getenv("MMB_LM_PROFILE"); w->profile = e && atoi(e) != 0; }
    if (w->profile) for (auto &e : w->ev) cudaEventCreate(&e);
    p->lm = w;
    return MMB_OK;
}

__global__ void k_peer_seq_next(unsigned *seq) { *seq += 1u; }

// ---- enqueue helpers ------------------------------------------------------------------------
#define LAUNCH_CHECK()                                  \
    do {                                                \
        ++p->launches;                                  \
        MMB_CUDA_CHECK(cudaGetLastError());             \
    } while (0)

static mmb_status read_state(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    MMB_CUDA_CHECK(cudaMemcpyAsync(w->h_st, w->st, sizeof(LmState), cudaMemcpyDeviceToHost, p->stream));
    MMB_CUDA_CHECK(cudaStreamSynchronize(p->stream));
    ++p->solve_host_syncs;
    return MMB_OK;
}

// the exchange step of a sharded solve: in-place sum (or max) over the ranks; nothing to do
// for an unsharded handle
static mmb_status all_reduce(DeviceProblem *p, double *buf, size_t count, bool is_max = false) {
    if (!p->comm || p->comm->world == 1 || count == 0) return MMB_OK;
    ++p->lm->n_allreduce;
    return p->comm->allreduce(buf, count, is_max, p->stream);
}

static mmb_status enqueue_residual_eval(DeviceProblem *p, const double *d_x, double *d_f, double *d_dist) {
    LmWorkspace *w = p->lm;
    mmb_status st;
    if ((st = launch_variants(p, d_x, false)) != MMB_OK) return st;
    if ((st = launch_residuals(p, d_f, d_dist)) != MMB_OK) return st;
    if ((st = launch_extra_rows(p, d_x, d_f + p->hp.m, nullptr, nullptr)) != MMB_OK) return st;
    k_sumsq_partial<<<kRedBlocks, 256, 0, p->stream>>>(d_f, (size_t)p->hp.m_total, w->part_f);
    LAUNCH_CHECK();
    return MMB_OK;
}

// avg / min / max of `dist` over every shard -> LmState err_*
static mmb_status enqueue_error_stats(DeviceProblem *p, const double *dist, bool to_init = false) {
    LmWorkspace *w = p->lm;
    mmb_status st;
    k_error_stats_partial<<<kStatBlocks, 256, 0, p->stream>>>(dist, p->hp.nobs, w->stat_part);
    LAUNCH_CHECK();
    k_error_stats_fold<<<1, 32, 0, p->stream>>>(w->stat_part, kStatBlocks, p->hp.nobs, w->scal, w->mx);
    LAUNCH_CHECK();
    if ((st = all_reduce(p, w->scal, 2)) != MMB_OK) return st;
    if ((st = all_reduce(p, w->mx, 2, true)) != MMB_OK) return st;
    k_error_stats_final<<<1, 1, 0, p->stream>>>(w->st, w->scal, w->mx, to_init ? 1 : 0);
    LAUNCH_CHECK();
    return MMB_OK;
}

static mmb_status enqueue_jacobian_and_assembly(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    const LmDims &d = w->d;
    mmb_status st;
    if ((st = launch_variants(p, p->d_x, true)) != MMB_OK) return st;
    if ((st = launch_jacobian(p, nullptr, nullptr, p->d_jac, nullptr, w->dist2)) != MMB_OK) return st;
    const int nx = p->hp.n_extra;
    if ((st = launch_extra_rows(p, p->d_x, nullptr, w->jext, w->ext_pid)) != MMB_OK) return st;
    // column maxima -> the fixed-point grids; limbs and the plain-FP part of W start at zero
    MMB_CUDA_CHECK(cudaMemsetAsync(w->cmax, 0, sizeof(unsigned long long) * ((size_t)d.n + 1), p->stream));
    MMB_CUDA_CHECK(cudaMemsetAsync(w->poison, 0, sizeof(int), p->stream));
    MMB_CUDA_CHECK(cudaMemsetAsync(w->det_limbs, 0, sizeof(long long) * w->det_limb_count, p->stream));
    if (d.KE && d.ns) MMB_CUDA_CHECK(cudaMemsetAsync(w->W, 0, sizeof(double) * (size_t)d.KE * d.NSP, p->stream));
    k_colmax<<<(d.nobs + 255) / 256, 256, 0, p->stream>>>(p->ds, d, p->d_jac, p->d_fvec, w->cmax);
    LAUNCH_CHECK();
    if (nx) {
        k_extra_colmax<<<(nx + 63) / 64, 64, 0, p->stream>>>(nx, d.n, w->jext, w->ext_pid, p->d_fvec + d.m, w->cmax);
        LAUNCH_CHECK();
    }
    k_assemble<<<(d.nobs + 255) / 256, 256, 0, p->stream>>>(p->ds, d, p->d_jac, p->d_fvec, w->attr_slot, w->det, w->W,
                                                            w->asm_shared.mask);
    if (w->asm_shared.ns) {
        LAUNCH_CHECK();
        const HostProblem &hp = p->hp;
        // marker groups sized for about two CTAs per SM slot
        const int fb = (hp.F + 31) / 32;
        int G = std::max(1, std::min(hp.M, (148 * 8 + fb - 1) / fb));
        const int mt = (hp.M + G - 1) / G;
        G = (hp.M + mt - 1) / mt;
        const dim3 grid(fb, G);
#define MMB_ASM_SHARED(NS_)                                                                                          \
    case NS_:                                                                                                        \
        k_assemble_shared<NS_><<<grid, 128, 0, p->stream>>>(w->asm_shared, d, hp.F, hp.M, mt, p->d_jac, p->d_fvec,   \
                                                            w->det, w->W);                                           \
        break;
        switch (w->asm_shared.ns) {
            MMB_ASM_SHARED(6) MMB_ASM_SHARED(7) MMB_ASM_SHARED(9) MMB_ASM_SHARED(11) MMB_ASM_SHARED(13) MMB_ASM_SHARED(15)
            default: break;
        }
#undef MMB_ASM_SHARED
    }
    LAUNCH_CHECK();
    if (nx) {
        k_extra_assemble<<<(nx + 63) / 64, 64, 0, p->stream>>>(d, nx, w->jext, w->ext_pid, p->d_fvec + d.m, w->p2i, w->det);
        LAUNCH_CHECK();
    }
    {
        const size_t total = w->det_limb_count / 2;
        k_assemble_finalize<<<(unsigned)((total + 255) / 256), 256, 0, p->stream>>>(d, w->det, w->slot_of_cs, w->i2p, w->U, w->gE,
                                                                                    w->V, w->W);
        LAUNCH_CHECK();
    }
    if (d.ns) {
        k_prep_gather<<<(d.ns + 255) / 256, 256, 0, p->stream>>>(d, w->V, w->sdiag);
        LAUNCH_CHECK();
        if ((st = all_reduce(p, w->sdiag, (size_t)2 * d.ns)) != MMB_OK) return st;
    }
    k_ctrl_prep<<<1, 1024, 0, p->stream>>>(d, w->st, w->U, w->gE, w->sdiag, w->i2p, p->d_x, w->diag, w->acnorm, w->scal, w->mx);
    LAUNCH_CHECK();
    if ((st = all_reduce(p, w->scal, 2)) != MMB_OK) return st;
    if ((st = all_reduce(p, w->mx, 1, true)) != MMB_OK) return st;
    k_ctrl_prep_decide<<<1, 1, 0, p->stream>>>(w->st, w->scal, w->mx);
    LAUNCH_CHECK();
    return MMB_OK;
}

// triangular solves with the factor of the reduced system
static mmb_status enqueue_backward_s(DeviceProblem *p, double *pS) {
    LmWorkspace *w = p->lm;
    LmDims d = w->d;
    if (w->wave_ok && d.ns > 0) {
        const int nblk = (d.ns + kNB - 1) / kNB;
        const double *S = w->S;
        int *flags = w->wave_flags;
        const int *ctr = w->wave_ctr;
        const double *dinv = w->Dinv;
        int slot = 1;
        if (w->capturing) {  // graph body: plain launch (see the note above k_backward_wave)
            k_backward_wave<<<nblk, 256, kWaveSmemBwd, p->stream>>>(d, S, pS, flags, ctr, slot, dinv);
            LAUNCH_CHECK();
            return MMB_OK;
        }
        void *args[] = {&d, &S, &pS, &flags, &ctr, &slot, &dinv};
        MMB_CUDA_CHECK(cudaLaunchCooperativeKernel((void *)k_backward_wave, dim3(nblk), dim3(256), args, kWaveSmemBwd, p->stream));
        ++p->launches;
        return MMB_OK;
    }
    k_backward_s<<<1, 1024, sizeof(double) * (64 * kDS + 64 + 32 * 64), p->stream>>>(d, w->S, pS);
    LAUNCH_CHECK();
    return MMB_OK;
}
static mmb_status enqueue_forward_s(DeviceProblem *p, double *wS) {
    LmWorkspace *w = p->lm;
    LmDims d = w->d;
    if (w->wave_ok && d.ns > 0) {
        const int nblk = (d.ns + kNB - 1) / kNB;
        const double *S = w->S;
        const double *rhs = wS;
        int *flags = w->wave_flags;
        const int *ctr = w->wave_ctr;
        const double *dinv = w->Dinv;
        int slot = 2;
        if (w->capturing) {
            k_forward_wave<<<nblk, 256, kWaveSmemFwd, p->stream>>>(d, S, rhs, wS, flags, ctr, slot, dinv);
            LAUNCH_CHECK();
            return MMB_OK;
        }
        void *args[] = {&d, &S, &rhs, &wS, &flags, &ctr, &slot, &dinv};
        MMB_CUDA_CHECK(cudaLaunchCooperativeKernel((void *)k_forward_wave, dim3(nblk), dim3(256), args, kWaveSmemFwd, p->stream));
        ++p->launches;
        return MMB_OK;
    }
    k_forward_s<<<1, 1024, sizeof(double) * (64 * kDS + 64), p->stream>>>(d, w->S, wS, wS);
    LAUNCH_CHECK();
    return MMB_OK;
}

// one factorisation + solve + Newton-correction solve at st->par_cur, then the lmpar control step
static mmb_status enqueue_lmpar_step(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    const LmDims &d = w->d;
    cudaStream_t s = p->stream;
    const bool sharded = p->comm && p->comm->world > 1;
    mmb_status rs;
    const bool prof = w->profile && !w->capturing;
    if (prof) cudaEventRecord(w->ev[0], s);
    if (d.KE) {
        const size_t sm = sizeof(double) * ((size_t)d.na * d.na + d.na);
        if (d.na == 6)
            k_block_factor_fixed<6><<<d.F, 128, 0, s>>>(d, w->st, w->U, w->gE, w->W, w->diag, w->i2p, w->Lf, w->Y);
        else
            k_block_factor<<<d.F, 128, sm, s>>>(d, w->st, w->U, w->gE, w->W, w->diag, w->i2p, w->Lf, w->Y);
        LAUNCH_CHECK();
    }
    // (graph bodies only: the two launches are independent branches there.  Launched from the host,
    // two streams may share one hardware queue, and the queue entry that signals the end of the
    // Cholesky would hold back the SYRK launch behind it -- the SYRK the Cholesky is waiting for.)
    const bool peer_tiles = sharded && w->ptiles.world > 0;
    const bool beside = w->chol_beside > 0 && (!sharded || peer_tiles) && d.KE && w->capturing;
    {
        const size_t tot = (size_t)d.NSP * d.NSP;
        k_schur_init<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(d, w->st, w->V, w->diag, w->i2p, w->S, sharded ? 0 : 1);
        LAUNCH_CHECK();
        if (beside && peer_tiles) {   // a new epoch for the tile flags (every rank counts alike)
            k_peer_seq_next<<<1, 1, 0, s>>>(w->peer_seq);
            LAUNCH_CHECK();
        }
    }
    const int nb = d.NSP / 64;
    if (d.KE) {
        const int tiles = nb * (nb + 1) / 2;
        // 300 tiles on 148 SMs leave some SMs with 3 CTAs and most with 2 (measured 0.72 ms); 16 x
        // 148 CTAs of an eighth of the K range balance the tensor pipe (0.50 ms).  K splits add
        // their partial sums with atomics.
        int ksplit = 1;
        static int ks_target = -1;
        if (ks_target < 0) { const char *e = getenv("MMB_SYRK_CTAS"); ks_target = e ? atoi(e) : 2368; }
        // the K splits of a tile are one thread-block cluster (portable size <= 8)
        while (tiles * ksplit < ks_target && ksplit < 8 && (d.KEP / kSyrkKT) / (ksplit * 2) >= 8) ksplit *= 2;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(tiles, ksplit);
        cfg.blockDim = dim3(128);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = ksplit; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = ksplit > 1 ? 1 : 0;
        if (beside) {
            // The factorisation is a latency chain over the block columns that needs few SMs; the
            // SYRK needs them all.  Launch the Cholesky (a few resident CTAs walking the tiles column
            // by column) on a second stream and let it take each tile when the SYRK publishes it.
            // The uncapped instance (254 registers x 256 threads) fills the register file of its SM:
            // those SMs run the chain undisturbed, the SYRK has the others, and it never waits, so
            // the pair cannot dead-lock.  (Sharing SMs -- the 128-register instance plus two SYRK
            // CTAs -- was measured slower: the chain's FP64 operations queue behind the SYRK's DMMAs.)
            MMB_CUDA_CHECK(cudaEventRecord(w->ev_fork, s));
            MMB_CUDA_CHECK(cudaStreamWaitEvent(w->side, w->ev_fork, 0));
            if (peer_tiles)
                k_chol_tiles<1, true, true><<<std::min(w->chol_beside, tiles), 256, kCholTilesSmem, w->side>>>(
                    d, w->st, w->S, w->tile_flags, w->wave_ctr, w->Dinv, (const int *)nullptr, w->ptiles);
            else
                k_chol_tiles<1, true, false><<<std::min(w->chol_beside, tiles), 256, kCholTilesSmem, w->side>>>(
                    d, w->st, w->S, w->tile_flags, w->wave_ctr, w->Dinv, (const int *)w->syrk_ready, PeerTiles{});
            LAUNCH_CHECK();
            MMB_CUDA_CHECK(cudaEventRecord(w->ev_join, w->side));
        }
        // (test hook: flags that never come, to exercise the bounded wait and the fall-back)
        int *ready = (beside && !peer_tiles && !getenv("MMB_TEST_DROP_SYRK_FLAGS")) ? w->syrk_ready : nullptr;
        const int *ectr = w->wave_ctr;
        if (beside && peer_tiles)
            MMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_schur_syrk<true>, d, (const double *)w->Y, w->S, ready, ectr, w->ptiles));
        else
            MMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k_schur_syrk<false>, d, (const double *)w->Y, w->S, ready, ectr, PeerTiles{}));
        LAUNCH_CHECK();
        if (beside) MMB_CUDA_CHECK(cudaStreamWaitEvent(s, w->ev_join, 0));
    }
    if (sharded && !beside) {
        // the one data-path collective of the solve: rows 0..ns of the reduced system
        // (lower triangle + right-hand side row), summed over the frame shards
        // (lower triangle by 64-wide block rows + right-hand side row), packed into one buffer:
        // half the bytes of the rows 0..ns rectangle (cfg3 9.6 instead of 18.9 MB, cfg5 0.9 instead of 1.8 GB)
        const int rows = d.ns + 1;
        k_tri_pack<<<d.NSP, 256, 0, s>>>(d, w->st, w->diag, w->i2p, w->S, w->tri_buf, rows, 1);
        LAUNCH_CHECK();
        if ((rs = all_reduce(p, w->tri_buf, tri_pack_count(rows))) != MMB_OK) return rs;
        k_tri_pack<<<d.NSP, 256, 0, s>>>(d, w->st, w->diag, w->i2p, w->S, w->tri_buf, rows, 0);
        LAUNCH_CHECK();
    }
    if (prof) cudaEventRecord(w->ev[1], s);
    if (beside) {
        // factorised beside the SYRK above
    } else if (w->chol_tiles) {
        static int chol_minb = -1;
        if (chol_minb < 0) { const char *e = getenv("MMB_CHOL_MINB"); chol_minb = e ? atoi(e) : 0; }
        if (nb > 64 || chol_minb == 2)
            k_chol_tiles<2><<<nb * (nb + 1) / 2, 256, kCholTilesSmem, s>>>(d, w->st, w->S, w->tile_flags, w->wave_ctr, w->Dinv, (const int *)nullptr, PeerTiles{});
        else
            k_chol_tiles<1><<<nb * (nb + 1) / 2, 256, kCholTilesSmem, s>>>(d, w->st, w->S, w->tile_flags, w->wave_ctr, w->Dinv, (const int *)nullptr, PeerTiles{});
        LAUNCH_CHECK();
    } else {
        for (int kb = 0; kb < nb; ++kb) {
            k_chol_panel<<<nb - kb, 256, kCholPanelSmem, s>>>(d, kb, w->st, w->S, w->Ld);
            LAUNCH_CHECK();
            const int nt = nb - kb - 1;
            k_chol_trailing<<<std::max(1, nt * (nt + 1) / 2), 128, 0, s>>>(d, kb, w->S, w->Ld);
            LAUNCH_CHECK();
        }
    }
    if (prof) cudaEventRecord(w->ev[2], s);
    if ((rs = enqueue_backward_s(p, w->p_int + d.KE)) != MMB_OK) return rs;
    if (d.KE) {
        if (d.na == 6) k_backsolve_e_fixed<6><<<d.F, 128, 0, s>>>(d, w->Lf, w->Y, w->p_int);
        else k_backsolve_e<<<d.F, 128, 0, s>>>(d, w->Lf, w->Y, w->p_int);
        LAUNCH_CHECK();
    }
    if (prof) cudaEventRecord(w->ev[3], s);
    k_lmpar_partial<<<1, 1024, 0, s>>>(d, w->st, w->p_int, w->diag, w->i2p, w->scal);
    LAUNCH_CHECK();
    if ((rs = all_reduce(p, w->scal, 2)) != MMB_OK) return rs;
    k_ctrl_lmpar_decide<<<1, 1, 0, s>>>(w->st, w->scal, w->wave_ctr);
    LAUNCH_CHECK();
    if (prof) {
        cudaEventRecord(w->ev[4], s);
        cudaEventSynchronize(w->ev[4]);
        for (int i = 0; i < 4; ++i) { float ms = 0; cudaEventElapsedTime(&ms, w->ev[i], w->ev[i + 1]); w->phase_ms[i] += ms; }
        ++w->phase_n;
    }
    return MMB_OK;
}

// second half of an lmpar step that goes on (st->lmpar_corr): Newton correction
// w = L^-1 D^2 p (scaled by 1/||D p|| in k_ctrl_lmpar_update), then the next par
static mmb_status enqueue_lmpar_correction(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    const LmDims &d = w->d;
    cudaStream_t s = p->stream;
    mmb_status rs;
    const bool prof = w->profile && !w->capturing;
    if (prof) cudaEventRecord(w->ev[3], s);
    if (d.KE) {
        k_qsolve_e<<<(d.F + 127) / 128, 128, 0, s>>>(d, w->Lf, w->p_int, w->diag, w->i2p, w->w_int);
        LAUNCH_CHECK();
        if (d.ns) {
            k_gemv_t_partial<<<dim3((d.ns + 127) / 128, w->gemv_chunks), 128, 0, s>>>(d, w->Y, w->w_int, w->part_gemv);
            LAUNCH_CHECK();
        }
    }
    if (d.ns) {
        k_gemv_fold<<<(d.ns + 127) / 128, 128, 0, s>>>(d, d.KE ? w->gemv_chunks : 0, w->part_gemv, w->tvec);
        LAUNCH_CHECK();
    }
    k_lmpar_partial_w<<<1, 1024, 0, s>>>(d, w->w_int, w->scal);
    LAUNCH_CHECK();
    if ((rs = all_reduce(p, w->scal, 1)) != MMB_OK) return rs;
    if (d.ns) {
        if ((rs = all_reduce(p, w->tvec, (size_t)d.ns)) != MMB_OK) return rs;
        k_qrhs_s<<<(d.ns + 127) / 128, 128, 0, s>>>(d, w->tvec, w->p_int, w->diag, w->i2p, w->w_int + d.KE);
        LAUNCH_CHECK();
        if ((rs = enqueue_forward_s(p, w->w_int + d.KE)) != MMB_OK) return rs;
    }
    k_ctrl_lmpar_update<<<1, 1024, 0, s>>>(d, w->st, w->w_int, w->scal, w->wave_ctr);
    LAUNCH_CHECK();
    if (prof) {
        cudaEventRecord(w->ev[4], s);
        cudaEventSynchronize(w->ev[4]);
        float ms = 0; cudaEventElapsedTime(&ms, w->ev[3], w->ev[4]); w->phase_ms[3] += ms;
    }
    return MMB_OK;
}

static mmb_status enqueue_trial(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    const LmDims &d = w->d;
    cudaStream_t s = p->stream;
    k_ctrl_trial<<<1, 1024, 0, s>>>(d, w->st, w->p_int, w->i2p, w->diag, p->d_x, w->x2, w->p_par, w->scal, w->d_cancel);
    LAUNCH_CHECK();
    mmb_status st = enqueue_residual_eval(p, w->x2, w->fvec2, w->dist2);
    if (st != MMB_OK) return st;
    k_jp_partial<<<kRedBlocks, 256, 0, s>>>(p->ds, d, p->d_jac, w->p_par, w->part_jp);
    LAUNCH_CHECK();
    if (p->hp.n_extra) {
        k_extra_jp<<<1, 32, 0, s>>>(p->hp.n_extra, w->jext, w->ext_pid, w->p_par, w->part_jp);
        LAUNCH_CHECK();
    }
    // squared norms of the trial residual and of J p, ||D p||^2, ||D x2||^2: one all-reduce
    if ((st = all_reduce(p, w->part_f, (size_t)2 * kRedBlocks + 3)) != MMB_OK) return st;
    k_ctrl_accept<<<1, 1024, 0, s>>>(d, w->st, w->part_f, w->part_jp, kRedBlocks, w->scal);
    LAUNCH_CHECK();
    k_accept_copy<<<296, 256, 0, s>>>(w->st, (size_t)d.n, (size_t)p->hp.m_total, (size_t)d.nobs, w->x2, p->d_x, w->fvec2, p->d_fvec, w->dist2, p->d_dist);
    LAUNCH_CHECK();
    return MMB_OK;
}

using clk = std::chrono::steady_clock;
static double secs(clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); }

void lm_request_cancel(DeviceProblem *p) {
    p->cancel = 1;
    if (p->lm && p->lm->h_cancel) *(volatile int *)p->lm->h_cancel = 1;
}

// prologue of both drivers: the acceptOnlyBetter reference measurement, the first function
// evaluation and the lmder input checks
static mmb_status enqueue_prologue(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    mmb_status st;
    // initial measureErrors when acceptOnlyBetter (adjust_base.cpp:1080-1103): at the
    // scene's own values, before any parameter is set.
    if (p->hp.opt.accept_only_better) {
        if ((st = launch_variants(p, nullptr, false)) != MMB_OK) return st;
        if ((st = launch_residuals(p, p->d_fvec, p->d_dist)) != MMB_OK) return st;
        if ((st = enqueue_error_stats(p, p->d_dist, true)) != MMB_OK) return st;
    }
    if ((st = enqueue_residual_eval(p, p->d_x, p->d_fvec, p->d_dist)) != MMB_OK) return st;
    // errorDistanceList of the last measureErrors call so far (the epilogue reads dist2: a solve
    // that stops before its first Jacobian or trial must not report stale pool memory)
    MMB_CUDA_CHECK(cudaMemcpyAsync(w->dist2, p->d_dist, sizeof(double) * (size_t)p->hp.nobs, cudaMemcpyDeviceToDevice, p->stream));
    if ((st = all_reduce(p, w->part_f, kRedBlocks)) != MMB_OK) return st;
    k_ctrl_init<<<1, 1024, 0, p->stream>>>(w->st, w->part_f, kRedBlocks);
    LAUNCH_CHECK();
    return MMB_OK;
}

// epilogue: error statistics of the last measureErrors call and ||fvec||^2 partials.
// errorDistanceList holds whatever the LAST measureErrors call left there (a rejected
// trial or the last FD perturbation included): dist2 tracks that.
static mmb_status enqueue_epilogue(DeviceProblem *p) {
    LmWorkspace *w = p->lm;
    mmb_status st;
    if ((st = enqueue_error_stats(p, w->dist2)) != MMB_OK) return st;
    k_sumsq_partial<<<kRedBlocks, 256, 0, p->stream>>>(p->d_fvec, (size_t)p->hp.m_total, w->part_f);
    LAUNCH_CHECK();
    return all_reduce(p, w->part_f, kRedBlocks);
}

// ---- the device-resident loop: one CUDA graph with conditional nodes -------------------
// Adds a conditional node after whatever the capturing stream has recorded so far and
// makes it the stream's new dependency; returns the (empty) body graph.
static mmb_status capture_conditional(cudaStream_t s, cudaGraphConditionalHandle h, cudaGraphConditionalNodeType type,
                                      cudaGraph_t *body) {
    cudaStreamCaptureStatus status;
    unsigned long long id = 0;
    cudaGraph_t g = nullptr;
    const cudaGraphNode_t *deps = nullptr;
    size_t ndeps = 0;
    MMB_CUDA_CHECK(cudaStreamGetCaptureInfo(s, &status, &id, &g, &deps, &ndeps));
    cudaGraphNodeParams params = {cudaGraphNodeTypeConditional};
    params.type = cudaGraphNodeTypeConditional;
    params.conditional.handle = h;
    params.conditional.type = type;
    params.conditional.size = 1;
    cudaGraphNode_t node;
    MMB_CUDA_CHECK(cudaGraphAddNode(&node, g, deps, ndeps, &params));
    MMB_CUDA_CHECK(cudaStreamUpdateCaptureDependencies(s, &node, 1, cudaStreamSetCaptureDependencies));
    *body = params.conditional.phGraph_out[0];
    return MMB_OK;
}

// handle_owner: which graph the handles of the nested conditionals are created on
// (0 = the top-level graph, 1 = the loop body that holds their nodes)
static mmb_status lm_graph_build(DeviceProblem *p, int handle_owner) {
    LmWorkspace *w = p->lm;
    cudaStream_t s = p->stream;
    // A sharded solve must not act on a rank-local cancel word at the loop head (the ranks would
    // leave the collective sequence at different points): there the request only travels through
    // the trial's all-reduce (k_ctrl_trial / k_ctrl_accept) and k_cond polls a word that stays 0.
    const int *cond_cancel = (p->comm && p->comm->world > 1) ? w->d_zero : w->d_cancel;
    mmb_status st = MMB_OK;
    cudaGraph_t g = nullptr, body_loop = nullptr, body_jac = nullptr, body_lmpar = nullptr, body_trial = nullptr, body_corr = nullptr, tmp = nullptr;
    cudaGraphConditionalHandle h_loop, h_jac, h_lmpar, h_trial, h_corr;
    MMB_CUDA_CHECK(cudaGraphCreate(&g, 0));
    w->graph = g;
    MMB_CUDA_CHECK(cudaGraphConditionalHandleCreate(&h_loop, g, 0, 0));
    const long long launches_before = p->launches;
    w->capturing = true;
    auto fail = [&](mmb_status e) {
        cudaGraph_t dead = nullptr;
        cudaStreamEndCapture(s, &dead);  // leaves capture mode whatever state it is in
        cudaGetLastError();
        w->capturing = false;
        p->launches = launches_before;
        return e;
    };
#define GB(expr)                                                                      \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            set_last_error(std::string("LM graph: " #expr ": ") + cudaGetErrorString(_e)); \
            return fail(MMB_ERR_CUDA);                                                \
        }                                                                             \
    } while (0)
#define GS(expr)                                     \
    do {                                             \
        if ((st = (expr)) != MMB_OK) return fail(st); \
    } while (0)
    // top level: prologue, WHILE(!done), epilogue
    long long l0 = p->launches;
    GB(cudaStreamBeginCaptureToGraph(s, g, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    GS(enqueue_prologue(p));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondStart, h_loop, h_loop);
    ++p->launches;
    GS(capture_conditional(s, h_loop, cudaGraphCondTypeWhile, &body_loop));
    GS(enqueue_epilogue(p));
    GB(cudaStreamEndCapture(s, &tmp));
    w->fixed_launches = (int)(p->launches - l0);
    // loop body: IF(need_jacobian) ; WHILE(!lmpar_done) ; IF(!done) trial
    l0 = p->launches;
    {
        cudaGraph_t owner = handle_owner ? body_loop : g;
        cudaError_t e1 = cudaGraphConditionalHandleCreate(&h_jac, owner, 0, 0);
        cudaError_t e2 = cudaGraphConditionalHandleCreate(&h_lmpar, owner, 0, 0);
        cudaError_t e3 = cudaGraphConditionalHandleCreate(&h_trial, owner, 0, 0);
        if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
            set_last_error(std::string("LM graph: cudaGraphConditionalHandleCreate: ") +
                           cudaGetErrorString(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3)));
            cudaGetLastError();
            w->capturing = false;
            p->launches = launches_before;
            return MMB_ERR_CUDA;
        }
    }
    GB(cudaStreamBeginCaptureToGraph(s, body_loop, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondLoopHead, h_jac, h_jac);
    GS(capture_conditional(s, h_jac, cudaGraphCondTypeIf, &body_jac));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondAfterJac, h_lmpar, h_lmpar);
    GS(capture_conditional(s, h_lmpar, cudaGraphCondTypeWhile, &body_lmpar));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondAfterLmpar, h_trial, h_loop);
    p->launches += 3;
    GS(capture_conditional(s, h_trial, cudaGraphCondTypeIf, &body_trial));
    GB(cudaStreamEndCapture(s, &tmp));
    w->body_launches[0] = (int)(p->launches - l0);
    // Jacobian + assembly
    l0 = p->launches;
    GB(cudaStreamBeginCaptureToGraph(s, body_jac, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    GS(enqueue_jacobian_and_assembly(p));
    GB(cudaStreamEndCapture(s, &tmp));
    w->body_launches[1] = (int)(p->launches - l0);
    // one lmpar step
    l0 = p->launches;
    {
        cudaError_t e1 = cudaGraphConditionalHandleCreate(&h_corr, handle_owner ? body_lmpar : g, 0, 0);
        if (e1 != cudaSuccess) {
            set_last_error(std::string("LM graph: cudaGraphConditionalHandleCreate: ") + cudaGetErrorString(e1));
            cudaGetLastError();
            w->capturing = false;
            p->launches = launches_before;
            return MMB_ERR_CUDA;
        }
    }
    GB(cudaStreamBeginCaptureToGraph(s, body_lmpar, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    GS(enqueue_lmpar_step(p));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondCorr, h_corr, h_corr);
    ++p->launches;
    GS(capture_conditional(s, h_corr, cudaGraphCondTypeIf, &body_corr));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondInLmpar, h_lmpar, h_lmpar);
    ++p->launches;
    GB(cudaStreamEndCapture(s, &tmp));
    w->body_launches[2] = (int)(p->launches - l0);
    // the Newton correction of an lmpar step that goes on
    l0 = p->launches;
    GB(cudaStreamBeginCaptureToGraph(s, body_corr, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    GS(enqueue_lmpar_correction(p));
    GB(cudaStreamEndCapture(s, &tmp));
    w->body_launches[4] = (int)(p->launches - l0);
    // trial point, ratio test, accept / reject
    l0 = p->launches;
    GB(cudaStreamBeginCaptureToGraph(s, body_trial, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    GS(enqueue_trial(p));
    k_cond<<<1, 1, 0, s>>>(w->st, cond_cancel, kCondAfterTrial, h_loop, h_loop);
    ++p->launches;
    GB(cudaStreamEndCapture(s, &tmp));
    w->body_launches[3] = (int)(p->launches - l0);
#undef GB
#undef GS
    w->capturing = false;
    p->launches = launches_before;  // nothing ran yet
    const auto ti0 = clk::now();
    cudaError_t e = cudaGraphInstantiate(&w->graph_exec, g, 0);
    if (getenv("MMB_LM_GRAPH_VERBOSE")) fprintf(stderr, "[mmb] LM graph: cudaGraphInstantiate %.3f ms\n", 1e3 * secs(ti0, clk::now()));
    if (e != cudaSuccess) {
        set_last_error(std::string("LM graph: cudaGraphInstantiate: ") + cudaGetErrorString(e));
        cudaGetLastError();
        w->graph_exec = nullptr;
        return MMB_ERR_CUDA;
    }
    return MMB_OK;
}

// host-driven loop (sharded solves: NCCL collectives sit between the kernels)
static mmb_status lm_run_host_loop(DeviceProblem *p, mmb_result *res, double *t_jac, double *t_func, double *t_lin) {
    LmWorkspace *w = p->lm;
    mmb_status st;
    auto t0 = clk::now();
    if ((st = enqueue_prologue(p)) != MMB_OK) return st;
    if ((st = read_state(p)) != MMB_OK) return st;
    *t_func += secs(t0, clk::now());
    int guard = 0;
    while (!w->h_st->done) {
        // a cancel request is seen by k_ctrl_trial and travels through the trial's all-reduce:
        // every rank gets `done` in the same iteration (no rank-local exit between collectives)
        if (++guard > 100000) break;
        if (w->h_st->need_jacobian) {
            t0 = clk::now();
            if ((st = enqueue_jacobian_and_assembly(p)) != MMB_OK) return st;
            if ((st = read_state(p)) != MMB_OK) return st;
            *t_jac += secs(t0, clk::now());
            if (w->h_st->done) break;
        }
        t0 = clk::now();
        while (!w->h_st->lmpar_done && !w->h_st->done) {
            if ((st = enqueue_lmpar_step(p)) != MMB_OK) return st;
            if ((st = read_state(p)) != MMB_OK) return st;
            if (w->h_st->lmpar_corr) {
                if ((st = enqueue_lmpar_correction(p)) != MMB_OK) return st;
                if ((st = read_state(p)) != MMB_OK) return st;
            }
        }
        *t_lin += secs(t0, clk::now());
        if (w->h_st->done) break;
        t0 = clk::now();
        if ((st = enqueue_trial(p)) != MMB_OK) return st;
        if ((st = read_state(p)) != MMB_OK) return st;
        *t_func += secs(t0, clk::now());
    }
    return enqueue_epilogue(p);
}

mmb_status lm_solve(DeviceProblem *p, mmb_result *res, double *x_out, double *attr_out, double *dist_out) {
    std::memset(res, 0, sizeof(*res));
    const HostProblem &hp = p->hp;
    // pre-checks of solveFrames (adjust_base.cpp:844-882)
    if (hp.nobs == 0 || hp.n == 0) { set_last_error("no used markers or attributes"); return MMB_ERR_NOT_ENOUGH_OBJECTS; }
    if (hp.n > hp.m_total) { set_last_error("more parameters than errors"); return MMB_ERR_MORE_PARAMS_THAN_ERRORS; }
    if ((hp.maya_dag || hp.n_extra) && p->comm && p->comm->world > 1) {
        set_last_error("sceneGraphMode=1 / stiffness rows: global solves on one GPU only");
        return MMB_ERR_UNSUPPORTED;
    }
    MMB_CUDA_CHECK(cudaSetDevice(p->device));
    mmb_status st;
    const auto t_begin = clk::now();
    p->cancel = 0;
    p->solve_host_syncs = 0;
    p->solve_graph_launches = 0;
    if ((st = ensure_jacobian_buffers(p)) != MMB_OK) return st;
    if ((st = lm_workspace_create(p)) != MMB_OK) return st;
    if (getenv("MMB_TIMING")) {
        cudaDeviceSynchronize();
        fprintf(stderr, "[mmb timing] solve: Jacobian buffers + LM workspace %.3f ms\n", 1e3 * secs(t_begin, clk::now()));
    }
    LmWorkspace *w = p->lm;
    *(volatile int *)w->h_cancel = p->cancel;  // a request that raced the workspace creation

    // initial parameters (get_initial_parameters, adjust_base.cpp:260-295); diag =
    // paramWeightList = 1.0 (adjust_relationships.cpp:320).  One pinned staging buffer,
    // asynchronous uploads.
    double *stage = w->h_stage;  // [2 n]: x0, ones
    for (int i = 0; i < hp.n; ++i) {
        const int a = hp.par_attr[i], f = hp.par_frame[i];
        const int32_t ref = hp.attr_ref[a];
        const double v = ref >= 0 ? hp.stat0[ref] : hp.anim0[(size_t)(-ref - 1) * hp.F + f];
        stage[i] = bound_external_to_internal(v, hp.attr_min[a], hp.attr_max[a], hp.attr_off[a], hp.attr_scl[a]);
        stage[hp.n + i] = 1.0;
    }
    std::vector<double> x0(stage, stage + hp.n);
    MMB_CUDA_CHECK(cudaMemcpyAsync(p->d_x, stage, sizeof(double) * hp.n, cudaMemcpyHostToDevice, p->stream));
    MMB_CUDA_CHECK(cudaMemcpyAsync(w->diag, stage + hp.n, sizeof(double) * hp.n, cudaMemcpyHostToDevice, p->stream));
    LmState h{};
    h.n = hp.n; h.m = hp.m_total; h.maxfev = hp.opt.iter_max; h.mode = hp.opt.auto_param_scale == 1 ? 1 : 2;
    h.ftol = hp.opt.eps1; h.xtol = hp.opt.eps2; h.gtol = hp.opt.eps3; h.factor = hp.opt.tau * 100.0;
    // lmdif: fdjac2 calls fcn once per parameter of the WHOLE problem (all shards)
    const int n_all = p->frames_global > hp.F ? hp.ns + hp.na * p->frames_global : hp.n;
    h.nfev_per_jac = hp.opt.solver_type == 1 ? n_all : 0;
    *w->h_st = h;
    MMB_CUDA_CHECK(cudaMemcpyAsync(w->st, w->h_st, sizeof(LmState), cudaMemcpyHostToDevice, p->stream));

    const bool sharded = p->comm && p->comm->world > 1;
    bool peer_graph = false;
    if (sharded) {
        // the packed triangle lives in the communicator's peer-mapped arena when there is one (the
        // all-reduce then runs in place over NVLink peer loads / stores, and is graph-capturable)
        const size_t tri_n = tri_pack_count(w->d.ns + 1);
        double *pb = p->comm->peer_buffer(tri_n);
        if (pb) w->tri_buf = pb;
        else if (!w->tri_buf || w->tri_is_peer) MMB_CUDA_CHECK(w->arena.alloc(&w->tri_buf, tri_n));
        w->tri_is_peer = pb != nullptr;
        peer_graph = pb && p->comm->capturable(std::max<size_t>((size_t)2 * w->d.ns, (size_t)2 * kRedBlocks + 3));
    }
    int graph_mode = 1;
    { const char *e = getenv("MMB_LM_GRAPH"); if (e) graph_mode = atoi(e); }
    // sharded: with the peer-memory all-reduce (plain kernels + flag barriers over NVLink) the
    // collectives are captured into the graph bodies like the kernels around them and the sharded
    // solve is ONE graph launch too -- every rank builds and launches the same graph.  NCCL's own
    // kernels cannot sit in a conditional body (cudaGraphInstantiate: invalid argument, measured),
    // and the in-process "local" backend orders its reduction with host barriers.
    static int sharded_graph = -1;
    if (sharded_graph < 0) { const char *e = getenv("MMB_LM_GRAPH_SHARDED"); sharded_graph = e ? atoi(e) : 1; }
    const bool graph_capable = !sharded || (sharded_graph && peer_graph);
    bool use_graph = graph_mode && graph_capable && !w->profile && !w->graph_failed;
    if (sharded && !w->ptiles_decided) {
        // Reduced system exchanged tile by tile inside the factorisation (PeerTiles) instead of packed
        // and all-reduced: only if every rank can (first solve of a handle; one small collective).
        w->ptiles_decided = true;
        static int tiles_mode = -1;
        if (tiles_mode < 0) { const char *e = getenv("MMB_PEER_TILES"); tiles_mode = e ? atoi(e) : 1; }
        const int nbt = w->d.NSP / 64, T = nbt * (nbt + 1) / 2;
        PeerBig pbg{};
        // every rank stores its partial tiles into every arena: the traffic grows with the number of
        // ranks.  Linear algebra of a 15-iteration cfg3-per-GPU solve, tiles vs packed all-reduce:
        // 23.9 vs 27.5 ms at 2 GPUs, 25.3 vs 26.2 ms at 8 -- where the rest of the step lost more than
        // that to the 77 MB per trial and rank on the wire; so up to 4 ranks (MMB_PEER_TILES=2: any).
        const bool can = tiles_mode && use_graph && w->chol_beside > 0 && w->d.KE > 0 && T <= kPeerFlagTiles &&
                         p->comm->peer_big(&pbg) && (pbg.world <= 4 || tiles_mode >= 2) &&
                         (size_t)2 * pbg.world * T * 4096 <= pbg.doubles;
        double cannot = can ? 0.0 : 1.0;
        MMB_CUDA_CHECK(cudaMemcpyAsync(w->scal, &cannot, sizeof(double), cudaMemcpyHostToDevice, p->stream));
        if ((st = all_reduce(p, w->scal, 1, true)) != MMB_OK) return st;
        MMB_CUDA_CHECK(cudaMemcpyAsync(&cannot, w->scal, sizeof(double), cudaMemcpyDeviceToHost, p->stream));
        MMB_CUDA_CHECK(cudaStreamSynchronize(p->stream));
        if (cannot == 0.0) {
            PeerTiles &t = w->ptiles;
            t.world = pbg.world; t.rank = pbg.rank; t.ntiles = T; t.maxtiles = kPeerFlagTiles;
            for (int q = 0; q < 8; ++q) { t.tiles[q] = pbg.base[q]; t.flags[q] = pbg.flags[q]; }
            t.my_tiles = pbg.base[pbg.rank]; t.my_flags = pbg.flags[pbg.rank];
            t.seq = pbg.seq; t.diag = w->diag; t.i2p = w->i2p;
            w->peer_seq = pbg.seq;
        }
    }
    if (use_graph && !w->graph_exec) {
        const auto tg0 = clk::now();
        mmb_status gs = lm_graph_build(p, 0);
        if (gs != MMB_OK) {
            if (w->graph) { cudaGraphDestroy(w->graph); w->graph = nullptr; }
            const std::string first = get_last_error();
            gs = lm_graph_build(p, 1);
            if (gs != MMB_OK) set_last_error(first + " | " + get_last_error());
        }
        if (gs != MMB_OK) {
            if (w->graph) { cudaGraphDestroy(w->graph); w->graph = nullptr; }
            // not fatal: the host-driven loop runs the same kernels
            if (getenv("MMB_LM_GRAPH_STRICT")) return MMB_ERR_CUDA;
            if (getenv("MMB_LM_GRAPH_VERBOSE")) fprintf(stderr, "[mmb] %s -> host-driven loop\n", get_last_error().c_str());
            w->graph_failed = true;
            use_graph = false;
        } else if (getenv("MMB_LM_GRAPH_VERBOSE")) {
            const auto tg1 = clk::now();
            cudaGraphUpload(w->graph_exec, p->stream);
            cudaStreamSynchronize(p->stream);
            fprintf(stderr, "[mmb] LM graph: build+instantiate %.3f ms, upload %.3f ms\n", 1e3 * secs(tg0, tg1), 1e3 * secs(tg1, clk::now()));
        }
    }
    double t_jac = 0, t_func = 0, t_lin = 0;
    if (use_graph) {
        MMB_CUDA_CHECK(cudaGraphLaunch(w->graph_exec, p->stream));
        ++p->solve_graph_launches;
    } else {
        MMB_CUDA_CHECK(cudaStreamSynchronize(p->stream));  // h_st is reused as the read-back mirror
        ++p->solve_host_syncs;
        if ((st = lm_run_host_loop(p, res, &t_jac, &t_func, &t_lin)) != MMB_OK) return st;
    }
    // results (adjust_cminpack_lmder.cpp:185-197, adjust_base.cpp:1201-1282)
    double *part = stage;            // [kRedBlocks]
    double *x = stage + kRedBlocks;  // [n]
    MMB_CUDA_CHECK(cudaMemcpyAsync(part, w->part_f, sizeof(double) * kRedBlocks, cudaMemcpyDeviceToHost, p->stream));
    MMB_CUDA_CHECK(cudaMemcpyAsync(x, p->d_x, sizeof(double) * hp.n, cudaMemcpyDeviceToHost, p->stream));
    if (dist_out)
        MMB_CUDA_CHECK(cudaMemcpyAsync(dist_out, w->dist2, sizeof(double) * hp.nobs, cudaMemcpyDeviceToHost, p->stream));
    if ((st = read_state(p)) != MMB_OK) return st;
    const LmState &f = *w->h_st;
    if (sharded && p->comm->peer_error()) { set_last_error("peer all-reduce: a barrier timed out (ranks out of step)"); return MMB_ERR_COMM; }
    if (f.dataflow_timeout) {
        // never seen; if the two branches of the graph were ever not run side by side, go back to
        // one after the other for the rest of this handle's life and solve again
        if (w->chol_beside > 0 && use_graph && !sharded) {
            w->chol_beside = 0;
            cudaGraphExecDestroy(w->graph_exec); w->graph_exec = nullptr;
            cudaGraphDestroy(w->graph); w->graph = nullptr;
            return lm_solve(p, res, x_out, attr_out, dist_out);
        }
        set_last_error("Cholesky beside the SYRK: a tile of the reduced system never arrived");
        return sharded ? MMB_ERR_COMM : MMB_ERR_CUDA;
    }
    if (f.cancelled) res->user_interrupted = 1;
    if (use_graph) {
        t_jac = 1e-9 * (double)f.t_jac_ns; t_lin = 1e-9 * (double)f.t_lin_ns; t_func = 1e-9 * (double)f.t_func_ns;
        p->launches += w->fixed_launches + (long long)f.n_body_loop * w->body_launches[0] +
                       (long long)f.njev * w->body_launches[1] + (long long)f.n_body_lmpar * w->body_launches[2] +
                       (long long)f.n_body_trial * w->body_launches[3] + (long long)f.n_body_corr * w->body_launches[4];
    }
    double ss = 0.0;
    for (int i = 0; i < kRedBlocks; ++i) ss += part[i];
    res->error_final = std::sqrt(ss);
    res->success = 1;  // success = iterNum > 0: the first function evaluation always runs
    res->reason_num = res->user_interrupted ? -1 : f.info;
    res->iteration_num = f.nfev;
    // iflag = 1 calls: the first evaluation + one per trial (lmdif's nfev also counts fdjac2's calls)
    res->iteration_function_num = f.nfev - f.njev * f.nfev_per_jac;
    res->iteration_jacobian_num = f.njev * hp.n * (hp.opt.auto_diff_type == 1 ? 2 : 1);
    res->iteration_attempt_num = f.njev;
    res->error_avg = f.err_avg; res->error_min = f.err_min; res->error_max = f.err_max;
    bool better = true;
    if (hp.opt.accept_only_better) better = f.err_avg <= f.init_avg;
    res->error_is_better = better ? 1 : 0;
    for (int i = 0; i < hp.n; ++i) {
        const int a = hp.par_attr[i];
        if (x_out) x_out[i] = x[i];
        if (attr_out)
            attr_out[i] = bound_internal_to_external(better ? x[i] : x0[i], hp.attr_min[a], hp.attr_max[a], hp.attr_off[a], hp.attr_scl[a]);
    }
    res->timer_solve_s = secs(t_begin, clk::now());
    res->timer_jacobian_s = t_jac;
    res->timer_function_s = t_func;
    res->timer_linear_s = t_lin;
    if (w->profile && w->phase_n) {
        fprintf(stderr, "[mmb lm profile] %d lmpar steps: schur %.3f ms, cholesky %.3f ms, backsolve %.3f ms, qsolve %.3f ms (per step)\n",
                w->phase_n, w->phase_ms[0] / w->phase_n, w->phase_ms[1] / w->phase_n, w->phase_ms[2] / w->phase_n, w->phase_ms[3] / w->phase_n);
        w->phase_n = 0; for (double &v : w->phase_ms) v = 0;
    }
    return MMB_OK;
}

}  // namespace mmb
