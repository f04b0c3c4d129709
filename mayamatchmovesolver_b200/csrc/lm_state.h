// Device-resident Levenberg-Marquardt state.
//
// The control law is MINPACK lmder/lmpar (cminpack 1.3.8 as called from
// adjust_cminpack_lmder.cpp:114-184), restated on the
// damped normal equations (J^T J + par D^2) p = J^T f so that the block
// structure of J^T J can be used (Schur complement + Cholesky).  Every
// decision (lmpar iteration, ratio test, delta / par update, accept / reject,
// convergence tests) is taken by single-thread control kernels on the device;
// the host only reads the flags below to know what to enqueue next.
#pragma once
#include <stdint.h>

namespace mmb {

struct LmState {
    // problem constants
    int n, m, maxfev, mode;
    double ftol, xtol, gtol, factor;
    // lmder scalars
    double fnorm, fnorm1, xnorm, delta, par, gnorm, pnorm, ratio, actred, prered;
    double par_cur;         // damping the factorisation kernels use right now
    double jp_sumsq;        // ||J p||^2
    int iter, nfev, njev, info;
    int accepted;           // last trial accepted (ratio >= 1e-4)
    int need_jacobian;      // outer loop must run before the next inner pass
    int done;               // solve finished (info set or aborted)
    // lmpar scalars
    double parl, paru, fp, dxnorm, gnorm_scaled, qnorm2;
    int lmpar_iter;         // 0 = Gauss-Newton probe pending
    int lmpar_done;
    int chol_fail;          // a pivot was <= 0 in the current factorisation
    int chol_fail_total;
    // reduced-side (replicated) parts of sums whose eliminated-side parts are all-reduced
    double part_s[4];
    // statistics
    int n_factor;
    double err_avg, err_min, err_max;  // compute_error_stats (adjust_base.cpp:346)
};

}  // namespace mmb
