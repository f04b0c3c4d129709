// Device-resident Levenberg-Marquardt state.
//
// The control law is MINPACK lmder/lmpar (cminpack 1.3.8 as called from
// adjust_cminpack_lmder.cpp:114-184), restated on the
// damped normal equations (J^T J + par D^2) p = J^T f so that the block
// structure of J^T J can be used (Schur complement + Cholesky).  Every
// decision (lmpar iteration, ratio test, delta / par update, accept / reject,
// convergence tests) is taken by single-thread control kernels on the device;
// the host either replays one CUDA graph whose conditional (WHILE / IF) nodes
// are steered by these flags -- no host round trip inside the solve -- or, for
// a sharded solve (NCCL collectives between the kernels), reads the flags to
// know what to enqueue next.
#pragma once
#include <stdint.h>

namespace mmb {

struct LmState {
    // problem constants
    int n, m, maxfev, mode;
    int nfev_per_jac;  // lmdif: fdjac2 costs n function calls per Jacobian (0 for lmder)
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
    int lmpar_corr;         // the lmpar step goes on: the Newton-correction solve is wanted
    double dx2;             // ||D p||^2 of the current lmpar step
    int chol_fail;          // a pivot was <= 0 in the current factorisation
    int chol_fail_total;
    // reduced-side (replicated) parts of sums whose eliminated-side parts are all-reduced
    double part_s[4];
    // statistics
    int n_factor;
    double err_avg, err_min, err_max;  // compute_error_stats (adjust_base.cpp:346)
    double init_avg;                   // err_avg of the scene's own values (acceptOnlyBetter)
    // device-resident loop (CUDA graph with conditional nodes): how often each body ran,
    // and the device time of each phase (%globaltimer, ns) stamped at the phase boundaries
    int n_body_jac, n_body_lmpar, n_body_trial, n_body_loop, n_body_corr;
    int cancelled;                     // the host's cancel flag was seen set
    int dataflow_timeout;              // a tile of the SYRK never arrived at the Cholesky running beside it
    unsigned long long t_mark, t_jac_ns, t_lin_ns, t_func_ns;
};

}  // namespace mmb
