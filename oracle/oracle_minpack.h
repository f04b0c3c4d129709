// TEST INFRASTRUCTURE ONLY -- see oracle_math.h for the rules.
//
// Restatement of MINPACK-1 lmder / qrfac / qrsolv / lmpar / enorm (More,
// Garbow, Hillstrom, ANL-80-74), the algorithm the reference reaches through
// the third-party dependency cminpack v1.3.8
// (share/cmake/modules/Findcminpack.cmake:173-212; NOT vendored under
// /root/reference).  Call site being matched:
// src/mmSolver/adjust/adjust_cminpack_lmder.cpp:114-184.
// Written from the published algorithm description;
// pinned in tests/ against scipy.optimize.leastsq, which wraps the original
// Fortran MINPACK lmder (iterates, nfev, info, final x).
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <vector>

namespace orc {

// fcn(m, n, x, fvec, fjac, ldfjac, iflag) -> <0 aborts.  iflag 1: fvec,
// iflag 2: fjac (column-major, ldfjac rows).
typedef int (*lmder_fcn)(void *ud, int m, int n, const double *x, double *fvec,
                         double *fjac, int ldfjac, int iflag);

inline double enorm(int n, const double *x) {
    // MINPACK enorm: three accumulators to avoid over/underflow.
    const double rdwarf = 3.834e-20, rgiant = 1.304e19;
    double s1 = 0, s2 = 0, s3 = 0, x1max = 0, x3max = 0;
    const double agiant = rgiant / (double)n;
    for (int i = 0; i < n; ++i) {
        const double xabs = std::fabs(x[i]);
        if (xabs > rdwarf && xabs < agiant) { s2 += xabs * xabs; }
        else if (xabs <= rdwarf) {
            if (xabs > x3max) { double r = x3max / xabs; s3 = 1.0 + s3 * r * r; x3max = xabs; }
            else if (xabs != 0.0) { double r = xabs / x3max; s3 += r * r; }
        } else {
            if (xabs > x1max) { double r = x1max / xabs; s1 = 1.0 + s1 * r * r; x1max = xabs; }
            else { double r = xabs / x1max; s1 += r * r; }
        }
    }
    if (s1 != 0.0) return x1max * std::sqrt(s1 + (s2 / x1max) / x1max);
    if (s2 != 0.0) {
        if (s2 >= x3max) return std::sqrt(s2 * (1.0 + (x3max / s2) * (x3max * s3)));
        return std::sqrt(x3max * ((s2 / x3max) + (x3max * s3)));
    }
    return x3max * std::sqrt(s3);
}

// Householder QR with column pivoting. a: m x n column-major (lda).
inline void qrfac(int m, int n, double *a, int lda, bool pivot, int *ipvt,
                  double *rdiag, double *acnorm, double *wa) {
    const double epsmch = DBL_EPSILON;
    for (int j = 0; j < n; ++j) {
        acnorm[j] = enorm(m, &a[j * lda]);
        rdiag[j] = acnorm[j]; wa[j] = rdiag[j];
        if (pivot) ipvt[j] = j;
    }
    const int minmn = std::min(m, n);
    for (int j = 0; j < minmn; ++j) {
        if (pivot) {
            int kmax = j;
            for (int k = j; k < n; ++k) if (rdiag[k] > rdiag[kmax]) kmax = k;
            if (kmax != j) {
                for (int i = 0; i < m; ++i) std::swap(a[i + j * lda], a[i + kmax * lda]);
                rdiag[kmax] = rdiag[j]; wa[kmax] = wa[j];
                std::swap(ipvt[j], ipvt[kmax]);
            }
        }
        double ajnorm = enorm(m - j, &a[j + j * lda]);
        if (ajnorm != 0.0) {
            if (a[j + j * lda] < 0.0) ajnorm = -ajnorm;
            for (int i = j; i < m; ++i) a[i + j * lda] /= ajnorm;
            a[j + j * lda] += 1.0;
            for (int k = j + 1; k < n; ++k) {
                double sum = 0.0;
                for (int i = j; i < m; ++i) sum += a[i + j * lda] * a[i + k * lda];
                const double temp = sum / a[j + j * lda];
                for (int i = j; i < m; ++i) a[i + k * lda] -= temp * a[i + j * lda];
                if (pivot && rdiag[k] != 0.0) {
                    double t = a[j + k * lda] / rdiag[k];
                    rdiag[k] *= std::sqrt(std::max(0.0, 1.0 - t * t));
                    double t2 = rdiag[k] / wa[k];
                    if (0.05 * (t2 * t2) <= epsmch) {
                        rdiag[k] = enorm(m - j - 1, &a[j + 1 + k * lda]);
                        wa[k] = rdiag[k];
                    }
                }
            }
        }
        rdiag[j] = -ajnorm;
    }
}

// Solve R z = Q^T b with diagonal D appended (least squares), MINPACK qrsolv.
inline void qrsolv(int n, double *r, int ldr, const int *ipvt, const double *diag,
                   const double *qtb, double *x, double *sdiag, double *wa) {
    for (int j = 0; j < n; ++j) {
        for (int i = j; i < n; ++i) r[i + j * ldr] = r[j + i * ldr];
        x[j] = r[j + j * ldr]; wa[j] = qtb[j];
    }
    for (int j = 0; j < n; ++j) {
        const int l = ipvt[j];
        if (diag[l] != 0.0) {
            for (int k = j; k < n; ++k) sdiag[k] = 0.0;
            sdiag[j] = diag[l];
            double qtbpj = 0.0;
            for (int k = j; k < n; ++k) {
                if (sdiag[k] == 0.0) continue;
                double cs, sn;
                if (std::fabs(r[k + k * ldr]) < std::fabs(sdiag[k])) {
                    const double cotan = r[k + k * ldr] / sdiag[k];
                    sn = 0.5 / std::sqrt(0.25 + 0.25 * (cotan * cotan)); cs = sn * cotan;
                } else {
                    const double tn = sdiag[k] / r[k + k * ldr];
                    cs = 0.5 / std::sqrt(0.25 + 0.25 * (tn * tn)); sn = cs * tn;
                }
                r[k + k * ldr] = cs * r[k + k * ldr] + sn * sdiag[k];
                double temp = cs * wa[k] + sn * qtbpj;
                qtbpj = -sn * wa[k] + cs * qtbpj; wa[k] = temp;
                for (int i = k + 1; i < n; ++i) {
                    temp = cs * r[i + k * ldr] + sn * sdiag[i];
                    sdiag[i] = -sn * r[i + k * ldr] + cs * sdiag[i];
                    r[i + k * ldr] = temp;
                }
            }
        }
        sdiag[j] = r[j + j * ldr];
        r[j + j * ldr] = x[j];
    }
    int nsing = n;
    for (int j = 0; j < n; ++j) {
        if (sdiag[j] == 0.0 && nsing == n) nsing = j;
        if (nsing < n) wa[j] = 0.0;
    }
    for (int k = 0; k < nsing; ++k) {
        const int j = nsing - k - 1;
        double sum = 0.0;
        for (int i = j + 1; i < nsing; ++i) sum += r[i + j * ldr] * wa[i];
        wa[j] = (wa[j] - sum) / sdiag[j];
    }
    for (int j = 0; j < n; ++j) x[ipvt[j]] = wa[j];
}

inline void lmpar(int n, double *r, int ldr, const int *ipvt, const double *diag,
                  const double *qtb, double delta, double *par, double *x,
                  double *sdiag, double *wa1, double *wa2) {
    const double p1 = 0.1, p001 = 0.001, dwarf = DBL_MIN;
    int nsing = n;
    for (int j = 0; j < n; ++j) {
        wa1[j] = qtb[j];
        if (r[j + j * ldr] == 0.0 && nsing == n) nsing = j;
        if (nsing < n) wa1[j] = 0.0;
    }
    for (int k = 0; k < nsing; ++k) {
        const int j = nsing - k - 1;
        wa1[j] /= r[j + j * ldr];
        const double temp = wa1[j];
        for (int i = 0; i < j; ++i) wa1[i] -= r[i + j * ldr] * temp;
    }
    for (int j = 0; j < n; ++j) x[ipvt[j]] = wa1[j];
    int iter = 0;
    for (int j = 0; j < n; ++j) wa2[j] = diag[j] * x[j];
    double dxnorm = enorm(n, wa2);
    double fp = dxnorm - delta;
    if (fp <= p1 * delta) { if (iter == 0) *par = 0.0; return; }
    double parl = 0.0;
    if (nsing >= n) {
        for (int j = 0; j < n; ++j) { const int l = ipvt[j]; wa1[j] = diag[l] * (wa2[l] / dxnorm); }
        for (int j = 0; j < n; ++j) {
            double sum = 0.0;
            for (int i = 0; i < j; ++i) sum += r[i + j * ldr] * wa1[i];
            wa1[j] = (wa1[j] - sum) / r[j + j * ldr];
        }
        const double temp = enorm(n, wa1);
        parl = ((fp / delta) / temp) / temp;
    }
    for (int j = 0; j < n; ++j) {
        double sum = 0.0;
        for (int i = 0; i <= j; ++i) sum += r[i + j * ldr] * qtb[i];
        wa1[j] = sum / diag[ipvt[j]];
    }
    const double gnorm = enorm(n, wa1);
    double paru = gnorm / delta;
    if (paru == 0.0) paru = dwarf / std::min(delta, p1);
    *par = std::max(*par, parl);
    *par = std::min(*par, paru);
    if (*par == 0.0) *par = gnorm / dxnorm;
    for (;;) {
        ++iter;
        if (*par == 0.0) *par = std::max(dwarf, p001 * paru);
        const double temp = std::sqrt(*par);
        for (int j = 0; j < n; ++j) wa1[j] = temp * diag[j];
        qrsolv(n, r, ldr, ipvt, wa1, qtb, x, sdiag, wa2);
        for (int j = 0; j < n; ++j) wa2[j] = diag[j] * x[j];
        dxnorm = enorm(n, wa2);
        const double temp2 = fp;
        fp = dxnorm - delta;
        if (std::fabs(fp) <= p1 * delta || (parl == 0.0 && fp <= temp2 && temp2 < 0.0) || iter == 10)
            return;
        for (int j = 0; j < n; ++j) { const int l = ipvt[j]; wa1[j] = diag[l] * (wa2[l] / dxnorm); }
        for (int j = 0; j < n; ++j) {
            wa1[j] /= sdiag[j];
            const double t = wa1[j];
            for (int i = j + 1; i < n; ++i) wa1[i] -= r[i + j * ldr] * t;
        }
        const double t3 = enorm(n, wa1);
        const double parc = ((fp / delta) / t3) / t3;
        if (fp > 0.0) parl = std::max(parl, *par);
        if (fp < 0.0) paru = std::min(paru, *par);
        *par = std::max(parl, *par + parc);
    }
}

struct LmderTrace {  // optional per-outer-iteration record (for tests)
    std::vector<double> fnorm, par, delta;
};

inline int lmder(lmder_fcn fcn, void *ud, int m, int n, double *x, double *fvec,
                 double *fjac, int ldfjac, double ftol, double xtol, double gtol,
                 int maxfev, double *diag, int mode, double factor, int *nfev,
                 int *njev, int *ipvt, double *qtf, double *wa1, double *wa2,
                 double *wa3, double *wa4, LmderTrace *trace = nullptr, bool dif = false,
                 double epsfcn = 0.0) {
    // dif = true is MINPACK lmdif (adjust_cminpack_lmdif.cpp:114-184): the same control law,
    // the Jacobian by fdjac2 -- n calls of fcn with iflag = 2 that only fill fvec, forward
    // differences with the relative step h = sqrt(max(epsfcn, epsmch)) * |x_j| (h = that
    // factor when x_j = 0) -- and nfev += n per Jacobian.
    const double epsmch = DBL_EPSILON;
    const double p1 = 0.1, p5 = 0.5, p25 = 0.25, p75 = 0.75, p0001 = 1e-4;
    int info = 0, iflag = 0;
    *nfev = 0; *njev = 0;
    double delta = 0, xnorm = 0, par = 0, fnorm = 0, gnorm = 0, ratio = 0;
    if (n <= 0 || m < n || ldfjac < m || ftol < 0.0 || xtol < 0.0 || gtol < 0.0 ||
        maxfev <= 0 || factor <= 0.0)
        return 0;
    if (mode == 2)
        for (int j = 0; j < n; ++j) if (diag[j] <= 0.0) return 0;
    iflag = fcn(ud, m, n, x, fvec, fjac, ldfjac, 1);
    *nfev = 1;
    if (iflag < 0) return iflag;
    fnorm = enorm(m, fvec);
    par = 0.0;
    int iter = 1;
    for (;;) {
        if (!dif) {
            iflag = fcn(ud, m, n, x, fvec, fjac, ldfjac, 2);
        } else {
            const double eps = std::sqrt(std::max(epsfcn, epsmch));
            for (int j = 0; j < n; ++j) {
                const double temp = x[j];
                double h = eps * std::fabs(temp);
                if (h == 0.0) h = eps;
                x[j] = temp + h;
                iflag = fcn(ud, m, n, x, wa4, nullptr, ldfjac, 2);
                x[j] = temp;
                if (iflag < 0) break;
                for (int i = 0; i < m; ++i) fjac[i + (size_t)j * ldfjac] = (wa4[i] - fvec[i]) / h;
            }
            *nfev += n;
        }
        ++(*njev);
        if (iflag < 0) return iflag;
        qrfac(m, n, fjac, ldfjac, true, ipvt, wa1, wa2, wa3);
        if (iter == 1) {
            if (mode != 2)
                for (int j = 0; j < n; ++j) { diag[j] = wa2[j]; if (wa2[j] == 0.0) diag[j] = 1.0; }
            for (int j = 0; j < n; ++j) wa3[j] = diag[j] * x[j];
            xnorm = enorm(n, wa3);
            delta = factor * xnorm;
            if (delta == 0.0) delta = factor;
        }
        for (int i = 0; i < m; ++i) wa4[i] = fvec[i];
        for (int j = 0; j < n; ++j) {
            if (fjac[j + j * ldfjac] != 0.0) {
                double sum = 0.0;
                for (int i = j; i < m; ++i) sum += fjac[i + j * ldfjac] * wa4[i];
                const double temp = -sum / fjac[j + j * ldfjac];
                for (int i = j; i < m; ++i) wa4[i] += fjac[i + j * ldfjac] * temp;
            }
            fjac[j + j * ldfjac] = wa1[j];
            qtf[j] = wa4[j];
        }
        gnorm = 0.0;
        if (fnorm != 0.0) {
            for (int j = 0; j < n; ++j) {
                const int l = ipvt[j];
                if (wa2[l] != 0.0) {
                    double sum = 0.0;
                    for (int i = 0; i <= j; ++i) sum += fjac[i + j * ldfjac] * (qtf[i] / fnorm);
                    gnorm = std::max(gnorm, std::fabs(sum / wa2[l]));
                }
            }
        }
        if (gnorm <= gtol) info = 4;
        if (info != 0) break;
        if (mode != 2) for (int j = 0; j < n; ++j) diag[j] = std::max(diag[j], wa2[j]);
        do {
            lmpar(n, fjac, ldfjac, ipvt, diag, qtf, delta, &par, wa1, wa2, wa3, wa4);
            for (int j = 0; j < n; ++j) {
                wa1[j] = -wa1[j];
                wa2[j] = x[j] + wa1[j];
                wa3[j] = diag[j] * wa1[j];
            }
            const double pnorm = enorm(n, wa3);
            if (iter == 1) delta = std::min(delta, pnorm);
            iflag = fcn(ud, m, n, wa2, wa4, fjac, ldfjac, 1);
            ++(*nfev);
            if (iflag < 0) return iflag;
            const double fnorm1 = enorm(m, wa4);
            double actred = -1.0;
            if (p1 * fnorm1 < fnorm) { const double d = fnorm1 / fnorm; actred = 1.0 - d * d; }
            for (int j = 0; j < n; ++j) {
                wa3[j] = 0.0;
                const int l = ipvt[j];
                const double temp = wa1[l];
                for (int i = 0; i <= j; ++i) wa3[i] += fjac[i + j * ldfjac] * temp;
            }
            const double temp1 = enorm(n, wa3) / fnorm;
            const double temp2 = (std::sqrt(par) * pnorm) / fnorm;
            const double prered = temp1 * temp1 + temp2 * temp2 / p5;
            const double dirder = -(temp1 * temp1 + temp2 * temp2);
            ratio = 0.0;
            if (prered != 0.0) ratio = actred / prered;
            if (ratio <= p25) {
                double temp;
                if (actred >= 0.0) temp = p5; else temp = p5 * dirder / (dirder + p5 * actred);
                if (p1 * fnorm1 >= fnorm || temp < p1) temp = p1;
                delta = temp * std::min(delta, pnorm / p1);
                par /= temp;
            } else if (par == 0.0 || ratio >= p75) {
                delta = pnorm / p5;
                par = p5 * par;
            }
            if (ratio >= p0001) {
                for (int j = 0; j < n; ++j) { x[j] = wa2[j]; wa2[j] = diag[j] * x[j]; }
                for (int i = 0; i < m; ++i) fvec[i] = wa4[i];
                xnorm = enorm(n, wa2);
                fnorm = fnorm1;
                ++iter;
                if (trace) { trace->fnorm.push_back(fnorm); trace->par.push_back(par); trace->delta.push_back(delta); }
            }
            if (std::fabs(actred) <= ftol && prered <= ftol && p5 * ratio <= 1.0) info = 1;
            if (delta <= xtol * xnorm) info = 2;
            if (std::fabs(actred) <= ftol && prered <= ftol && p5 * ratio <= 1.0 && info == 2) info = 3;
            if (info != 0) break;
            if (*nfev >= maxfev) info = 5;
            if (std::fabs(actred) <= epsmch && prered <= epsmch && p5 * ratio <= 1.0) info = 6;
            if (delta <= epsmch * xnorm) info = 7;
            if (gnorm <= epsmch) info = 8;
            if (info != 0) break;
        } while (ratio < p0001);
        if (info != 0) break;
    }
    return info;
}

}  // namespace orc
