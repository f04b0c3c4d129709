// TEST INFRASTRUCTURE ONLY -- see oracle_core.h.
//
// Block-sparse CPU restatement of the solve (BASELINE.md "B2"): the SAME arithmetic per
// evaluation as the faithful path of oracle.cpp (one marker-frame re-evaluated through the
// same scene / lens functions, forward differences exactly as adjust_solveFunc.cpp:305-479
// forms them) and the SAME MINPACK lmder / lmpar control law, but
//   * a Jacobian column is only evaluated on the observations it can touch
//     (findErrorToParameterRelationship, adjust_relationships.cpp:565-617), and
//   * the damped normal equations are solved through the Schur complement of the per-frame
//     blocks instead of a dense pivoted QR of the m x n Jacobian (60 GB at cfg3).
// This is what makes full-size (cfg3 / cfg4) CPU solves possible: the full-size golden
// vectors under tests/golden/ come from here, and bench.py times it as the multi-core CPU
// baseline.  It is independent code from the CUDA kernels (scalar loops, OpenMP over
// observations / rows), validated against the dense oracle at small sizes
// (tests/test_oracle_cpu.py).  Forward differences only.
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

#include "oracle_core.h"

namespace orcs {

struct Override { int32_t ref; int frame; double value; bool on; };

struct Getter {
    const orc_problem *p;
    Override ov;
    inline double operator()(int32_t ref, int f) const {
        if (ov.on && ref == ov.ref && (ref >= 0 || f == ov.frame)) return ov.value;
        return attr_get(p, ref, f);
    }
};

static M4 world_of(const Getter &G, int t, int f) {
    const orc_problem *p = G.p;
    int chain[64], nc = 0;
    while (t >= 0 && nc < 64) { chain[nc++] = t; t = p->tfm_parent[t]; }
    M4 w = m4_identity();
    for (int i = nc - 1; i >= 0; --i) {
        const int32_t *a = &p->tfm_attrs[chain[i] * 9];
        const M4 local = local_matrix(G(a[0], f), G(a[1], f), G(a[2], f), G(a[3], f), G(a[4], f), G(a[5], f),
                                      G(a[6], f), G(a[7], f), G(a[8], f), p->tfm_roo[chain[i]]);
        w = (i == nc - 1) ? local : m4_mul(w, local);
    }
    return w;
}

// one marker-frame through evaluate_scene + measureErrors of oracle_core.h
static void obs_eval(const Getter &G, int i, double &ex, double &ey, double &dist) {
    const orc_problem *p = G.p;
    const int mk = p->errorToMarker[i].first, f = p->errorToMarker[i].second;
    const int c = p->mkr_cam[mk], b = p->mkr_bnd[mk];
    const int32_t *ca = &p->cam_attrs[c * 8];
    const double focal = G(ca[2], f), sx_mm = G(ca[0], f), sy_mm = G(ca[1], f);
    const M4 proj = projection_matrix(focal, sx_mm * MM_TO_INCH, sy_mm * MM_TO_INCH, G(ca[3], f) * MM_TO_INCH,
                                      G(ca[4], f) * MM_TO_INCH, (double)p->cam_rw[c], (double)p->cam_rh[c],
                                      p->cam_fit[c], 0.1, G(ca[6], f), G(ca[7], f));
    double px, py;
    reproject_normalised(world_of(G, p->cam_tfm[c], f), proj, world_of(G, p->bnd_tfm[b], f), px, py);
    double mx = p->mkr_x[(size_t)mk * p->F + f], my = p->mkr_y[(size_t)mk * p->F + f];
    scale_marker_film_fit(p->cam_fit[c], sx_mm / sy_mm, (double)p->cam_rw[c] / (double)p->cam_rh[c], mx, my);
    const int l = p->cam_lens[c];
    if (l >= 0 && p->lens_model[l] != 0) {
        LensEval L; L.model = p->lens_model[l];
        for (int k = 0; k < 6; ++k) L.cam[k] = p->lens_cam[l * 6 + k];
        for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k) {
            const int32_t r = p->lens_attrs[l * MMB_LENS_MAX_COEFFS + k];
            L.c[k] = (r == INT32_MIN) ? 0.0 : G(r, f);
        }
        lens_prepare(L);
        double ox = px, oy = py;
        lens_distort(L, px, py, ox, oy);
        if (std::isfinite(ox)) px = ox;
        if (std::isfinite(oy)) py = oy;
    }
    const double w = std::sqrt(p->markerWeight[i]);
    const double dx = std::fabs(mx - px), dy = std::fabs(my - py);
    ex = dx * p->opt.image_width * w;
    ey = dy * p->opt.image_width * w;
    dist = std::sqrt(dx * dx + dy * dy) * p->opt.image_width;
}

struct Sparse {
    orc_problem *p;
    int n, nobs, m, F, na, ns, kmax;
    std::vector<int> attr_base, attr_slot;
    std::vector<uint8_t> attr_anim;
    std::vector<std::vector<int>> mkr_attrs;   // marker -> solver attrs it depends on
    std::vector<double> jac;                   // [nobs][kmax][2]
    std::vector<int> jcol;                     // [nobs][kmax] parameter id or -1
    // normal equations
    std::vector<double> U, gE, W, V, gS;       // U [F][na][na], W [F*na][ns], V [ns][ns]
    std::vector<double> Lf, Y, z, S, rhs;      // factors
    int internal(int pid) const {
        const int a = p->paramToAttr[pid].first, f = p->paramToAttr[pid].second;
        return attr_anim[a] ? f * na + attr_slot[a] : F * na + attr_slot[a];
    }
};

static void sparse_init(Sparse &s, orc_problem *p) {
    s.p = p; s.n = p->n; s.nobs = p->nobs; s.m = p->m; s.F = p->F;
    s.attr_base.assign(p->A, 0); s.attr_slot.assign(p->A, 0); s.attr_anim.assign(p->A, 0);
    s.na = s.ns = 0;
    int base = 0;
    for (int a = 0; a < p->A; ++a) {
        s.attr_base[a] = base;
        if (p->attr_ref[a] < 0) { s.attr_anim[a] = 1; s.attr_slot[a] = s.na++; base += p->F; }
        else { s.attr_slot[a] = s.ns++; base += 1; }
    }
    s.mkr_attrs.assign(p->M, {});
    s.kmax = 0;
    for (int mk = 0; mk < p->M; ++mk) {
        for (int a = 0; a < p->A; ++a) if (p->markerToAttr[mk][a]) s.mkr_attrs[mk].push_back(a);
        s.kmax = std::max(s.kmax, (int)s.mkr_attrs[mk].size());
    }
    s.jac.assign((size_t)s.nobs * std::max(s.kmax, 1) * 2, 0.0);
    s.jcol.assign((size_t)s.nobs * std::max(s.kmax, 1), -1);
}

// solveFunc iflag = 1 at x (setParameters + measureErrors over every observation)
static void sparse_residuals(Sparse &s, const double *x, double *fvec, double *dist) {
    orc_problem *p = s.p;
    setParameters(p, x);
    ++p->funcEvalNum; ++p->iterNum;
    Getter G{p, {0, 0, 0.0, false}};
#pragma omp parallel for schedule(static)
    for (int i = 0; i < s.nobs; ++i) obs_eval(G, i, fvec[2 * i], fvec[2 * i + 1], dist[i]);
}

// solveFunc iflag = 2: forward differences, column by column, only where the column lives.
// `dist` ends up as the reference leaves errorDistanceList: the last measureErrors call is
// the perturbation of the LAST parameter, which rewrites the entries it evaluates.
static void sparse_jacobian(Sparse &s, const double *x, const double *fvec, double *dist) {
    orc_problem *p = s.p;
    setParameters(p, x);
    const double delta = p->opt.delta;
    p->funcEvalNum += s.n; p->jacIterNum += s.n;
    const int last = s.n - 1;
    const int last_attr = p->paramToAttr[last].first, last_frame = p->paramToAttr[last].second;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < s.nobs; ++i) {
        const int mk = p->errorToMarker[i].first, f = p->errorToMarker[i].second;
        const std::vector<int> &as = s.mkr_attrs[mk];
        for (int c = 0; c < (int)as.size(); ++c) {
            const int a = as[c];
            const int pid = s.attr_base[a] + (s.attr_anim[a] ? f : 0);
            const double dA = calculateParameterDelta(x[pid], delta, 1, p->attr_min[a], p->attr_max[a]);
            const double real = boundInternalToExternal(x[pid] + dA, p->attr_min[a], p->attr_max[a], p->attr_off[a], p->attr_scl[a]);
            Getter G{p, {p->attr_ref[a], f, real, true}};
            double ex, ey, d;
            obs_eval(G, i, ex, ey, d);
            const double inv = 1.0 / dA;
            const size_t o = ((size_t)i * s.kmax + c);
            s.jac[2 * o] = (ex - fvec[2 * i]) * inv;
            s.jac[2 * o + 1] = (ey - fvec[2 * i + 1]) * inv;
            s.jcol[o] = pid;
            if (a == last_attr && (last_frame < 0 || last_frame == f)) dist[i] = d;
        }
        for (int c = (int)as.size(); c < s.kmax; ++c) s.jcol[(size_t)i * s.kmax + c] = -1;
    }
}

// J^T J and J^T f in block form
static void sparse_assemble(Sparse &s, const double *fvec) {
    const int F = s.F, na = s.na, ns = s.ns;
    s.U.assign((size_t)F * na * na, 0.0); s.gE.assign((size_t)F * na, 0.0);
    s.W.assign((size_t)F * na * ns, 0.0); s.V.assign((size_t)ns * ns, 0.0); s.gS.assign(ns, 0.0);
    for (int i = 0; i < s.nobs; ++i) {
        const double fx = fvec[2 * i], fy = fvec[2 * i + 1];
        for (int c1 = 0; c1 < s.kmax; ++c1) {
            const int p1 = s.jcol[(size_t)i * s.kmax + c1];
            if (p1 < 0) continue;
            const int q1 = s.internal(p1);
            const double jx1 = s.jac[2 * ((size_t)i * s.kmax + c1)], jy1 = s.jac[2 * ((size_t)i * s.kmax + c1) + 1];
            const double g = jx1 * fx + jy1 * fy;
            if (q1 < F * na) s.gE[q1] += g; else s.gS[q1 - F * na] += g;
            for (int c2 = 0; c2 <= c1; ++c2) {
                const int p2 = s.jcol[(size_t)i * s.kmax + c2];
                if (p2 < 0) continue;
                const int q2 = s.internal(p2);
                const double v = jx1 * s.jac[2 * ((size_t)i * s.kmax + c2)] + jy1 * s.jac[2 * ((size_t)i * s.kmax + c2) + 1];
                const bool e1 = q1 < F * na, e2 = q2 < F * na;
                if (e1 && e2) {
                    const int f = q1 / na, r = q1 % na, cc = q2 % na;  // same frame by construction
                    s.U[((size_t)f * na + r) * na + cc] += v;
                    if (r != cc) s.U[((size_t)f * na + cc) * na + r] += v;
                } else if (e1 != e2) {
                    const int qe = e1 ? q1 : q2, qs = (e1 ? q2 : q1) - F * na;
                    s.W[(size_t)qe * ns + qs] += v;
                } else {
                    const int r = q1 - F * na, cc = q2 - F * na;
                    s.V[(size_t)r * ns + cc] += v;
                    if (r != cc) s.V[(size_t)cc * ns + r] += v;
                }
            }
        }
    }
}

// Factor and solve (J^T J + par D^2) p = J^T f via the Schur complement of the per-frame
// blocks; also the Newton-correction quantity of lmpar.  Returns false when a pivot fails.
// p_int / w_int: internal order [F*na | ns]; dint = diag in internal order.
static bool sparse_solve(Sparse &s, double par, const std::vector<double> &dint, std::vector<double> &p_int,
                         double &dx2, double &qnorm2) {
    const int F = s.F, na = s.na, ns = s.ns, KE = F * na;
    bool ok = true;
    s.Lf.assign((size_t)F * na * na, 0.0);
    std::vector<double> Yt((size_t)ns * KE);   // Y transposed: [ns][KE] for contiguous dot products
    s.z.assign(KE, 0.0);
#pragma omp parallel for schedule(static) reduction(&& : ok)
    for (int f = 0; f < F; ++f) {
        double *L = &s.Lf[(size_t)f * na * na];
        for (int r = 0; r < na; ++r)
            for (int c = 0; c <= r; ++c) {
                double v = s.U[((size_t)f * na + r) * na + c];
                if (r == c) v += par * dint[f * na + r] * dint[f * na + r];
                L[r * na + c] = v;
            }
        for (int j = 0; j < na; ++j) {
            double pv = L[j * na + j];
            for (int k = 0; k < j; ++k) pv -= L[j * na + k] * L[j * na + k];
            if (!(pv > 0.0)) { ok = false; pv = 1.0; }
            const double d = std::sqrt(pv);
            L[j * na + j] = d;
            for (int r = j + 1; r < na; ++r) {
                double v = L[r * na + j];
                for (int k = 0; k < j; ++k) v -= L[r * na + k] * L[j * na + k];
                L[r * na + j] = v / d;
            }
        }
        for (int r = 0; r < na; ++r) {
            double v = s.gE[(size_t)f * na + r];
            for (int c = 0; c < r; ++c) v -= L[r * na + c] * s.z[(size_t)f * na + c];
            s.z[(size_t)f * na + r] = v / L[r * na + r];
        }
        for (int j = 0; j < ns; ++j)
            for (int r = 0; r < na; ++r) {
                double v = s.W[((size_t)f * na + r) * ns + j];
                for (int c = 0; c < r; ++c) v -= L[r * na + c] * Yt[(size_t)j * KE + f * na + c];
                Yt[(size_t)j * KE + f * na + r] = v / L[r * na + r];
            }
    }
    // S = V + par D_S^2 - Y^T Y (lower), rhs = g_S - Y^T z
    s.S.assign((size_t)ns * ns, 0.0); s.rhs.assign(ns, 0.0);
#pragma omp parallel for schedule(dynamic, 8)
    for (int i = 0; i < ns; ++i) {
        const double *yi = &Yt[(size_t)i * KE];
        for (int j = 0; j <= i; ++j) {
            const double *yj = &Yt[(size_t)j * KE];
            double acc = 0.0;
#pragma omp simd reduction(+ : acc)
            for (int k = 0; k < KE; ++k) acc += yi[k] * yj[k];
            double v = s.V[(size_t)i * ns + j] - acc;
            if (i == j) v += par * dint[KE + i] * dint[KE + i];
            s.S[(size_t)i * ns + j] = v;
        }
        double acc = 0.0;
#pragma omp simd reduction(+ : acc)
        for (int k = 0; k < KE; ++k) acc += yi[k] * s.z[k];
        s.rhs[i] = s.gS[i] - acc;
    }
    // Cholesky of S (row-oriented, lower)
    for (int j = 0; j < ns; ++j) {
        double *Sj = &s.S[(size_t)j * ns];
        double pv = Sj[j];
        for (int k = 0; k < j; ++k) pv -= Sj[k] * Sj[k];
        if (!(pv > 1e-290)) { ok = false; pv = 1.0; }
        const double d = std::sqrt(pv);
        Sj[j] = d;
#pragma omp parallel for schedule(static) if (ns - j > 256)
        for (int r = j + 1; r < ns; ++r) {
            double *Sr = &s.S[(size_t)r * ns];
            double v = Sr[j];
#pragma omp simd reduction(- : v)
            for (int k = 0; k < j; ++k) v -= Sr[k] * Sj[k];
            Sr[j] = v / d;
        }
    }
    std::vector<double> y(ns), pS(ns);
    for (int r = 0; r < ns; ++r) {
        double v = s.rhs[r];
        for (int c = 0; c < r; ++c) v -= s.S[(size_t)r * ns + c] * y[c];
        y[r] = v / s.S[(size_t)r * ns + r];
    }
    for (int r = ns - 1; r >= 0; --r) {
        double v = y[r];
        for (int c = r + 1; c < ns; ++c) v -= s.S[(size_t)c * ns + r] * pS[c];
        pS[r] = v / s.S[(size_t)r * ns + r];
    }
    p_int.assign(KE + ns, 0.0);
    for (int j = 0; j < ns; ++j) p_int[KE + j] = pS[j];
    std::vector<double> wE(KE, 0.0);
    for (int f = 0; f < F; ++f) {
        const double *L = &s.Lf[(size_t)f * na * na];
        double r[64];
        for (int a = 0; a < na; ++a) {
            double acc = 0.0;
            for (int j = 0; j < ns; ++j) acc += Yt[(size_t)j * KE + f * na + a] * pS[j];
            r[a] = s.z[(size_t)f * na + a] - acc;
        }
        for (int a = na - 1; a >= 0; --a) {
            double v = r[a];
            for (int c = a + 1; c < na; ++c) v -= L[c * na + a] * r[c];
            r[a] = v / L[a * na + a];
        }
        for (int a = 0; a < na; ++a) p_int[(size_t)f * na + a] = r[a];
        // w_f = L_f^-1 D_f^2 p_f
        for (int a = 0; a < na; ++a) {
            const double dd = dint[f * na + a];
            double v = dd * dd * r[a];
            for (int c = 0; c < a; ++c) v -= L[a * na + c] * wE[(size_t)f * na + c];
            wE[(size_t)f * na + a] = v / L[a * na + a];
        }
    }
    dx2 = 0.0;
    for (int q = 0; q < KE + ns; ++q) { const double t = dint[q] * p_int[q]; dx2 += t * t; }
    // w_S = L_S^-1 (D_S^2 p_S - Y^T w_E)
    std::vector<double> wS(ns);
    double wn = 0.0;
    for (int k = 0; k < KE; ++k) wn += wE[k] * wE[k];
    for (int r = 0; r < ns; ++r) {
        double acc = 0.0;
        const double *yr = &Yt[(size_t)r * KE];
        for (int k = 0; k < KE; ++k) acc += yr[k] * wE[k];
        double v = dint[KE + r] * dint[KE + r] * pS[r] - acc;
        for (int c = 0; c < r; ++c) v -= s.S[(size_t)r * ns + c] * wS[c];
        wS[r] = v / s.S[(size_t)r * ns + r];
        wn += wS[r] * wS[r];
    }
    qnorm2 = (ok && dx2 > 0.0) ? wn / dx2 : 0.0;
    return ok;
}

// lmder + lmpar on the normal equations
static int sparse_lmder(Sparse &s, std::vector<double> &x, std::vector<double> &fvec, std::vector<double> &dist,
                        int *nfev_out, int *njev_out, int max_outer) {
    orc_problem *p = s.p;
    const int n = s.n, m = s.m, KE = s.F * s.na;
    const double ftol = p->opt.eps1, xtol = p->opt.eps2, gtol = p->opt.eps3, factor = p->opt.tau * 100.0;
    const int maxfev = p->opt.iter_max, mode = p->opt.auto_param_scale == 1 ? 1 : 2;
    const double epsmch = 2.220446049250313e-16, dwarf = 2.2250738585072014e-308;
    int info = 0, nfev = 0, njev = 0, iter = 1;
    std::vector<double> diag(n, 1.0), acnorm(n), dint(n), p_int, x2(n), f2(m), d2(s.nobs), step(n);
    double delta = 0.0, xnorm = 0.0, par = 0.0;
    *nfev_out = *njev_out = 0;
    if (n <= 0 || m < n || ftol < 0 || xtol < 0 || gtol < 0 || maxfev <= 0 || factor <= 0) return 0;
    sparse_residuals(s, x.data(), fvec.data(), dist.data());
    nfev = 1;
    double fnorm = enorm(m, fvec.data());
    for (;;) {
        sparse_jacobian(s, x.data(), fvec.data(), dist.data());
        ++njev;
        sparse_assemble(s, fvec.data());
        std::vector<double> g(n);
        for (int pid = 0; pid < n; ++pid) {
            const int q = s.internal(pid);
            double h;
            if (q < KE) { const int f = q / s.na, a = q % s.na; h = s.U[((size_t)f * s.na + a) * s.na + a]; g[pid] = s.gE[q]; }
            else { const int r = q - KE; h = s.V[(size_t)r * s.ns + r]; g[pid] = s.gS[r]; }
            acnorm[pid] = std::sqrt(std::max(h, 0.0));
        }
        if (iter == 1) {
            if (mode != 2) for (int j = 0; j < n; ++j) diag[j] = acnorm[j] == 0.0 ? 1.0 : acnorm[j];
            double t = 0.0;
            for (int j = 0; j < n; ++j) t += (diag[j] * x[j]) * (diag[j] * x[j]);
            xnorm = std::sqrt(t);
            delta = factor * xnorm;
            if (delta == 0.0) delta = factor;
        }
        double gnorm = 0.0;
        if (fnorm != 0.0)
            for (int j = 0; j < n; ++j)
                if (acnorm[j] != 0.0) gnorm = std::max(gnorm, std::fabs(g[j] / fnorm) / acnorm[j]);
        if (gnorm <= gtol) { info = 4; break; }
        if (mode != 2) for (int j = 0; j < n; ++j) diag[j] = std::max(diag[j], acnorm[j]);
        double gs = 0.0;
        for (int j = 0; j < n; ++j) { const double t = g[j] / diag[j]; gs += t * t; }
        const double gnorm_scaled = std::sqrt(gs);
        for (int pid = 0; pid < n; ++pid) dint[s.internal(pid)] = diag[pid];
        double ratio = 0.0;
        do {
            // ---- lmpar ----
            double dx2 = 0.0, qn2 = 0.0, parl = 0.0, paru = 0.0, fp = 0.0;
            bool ok = sparse_solve(s, 0.0, dint, p_int, dx2, qn2);
            double dxnorm = ok ? std::sqrt(dx2) : INFINITY;
            fp = dxnorm - delta;
            if (ok && fp <= 0.1 * delta) {
                par = 0.0;
            } else {
                if (ok && qn2 > 0.0) parl = (fp / delta) / qn2;
                paru = gnorm_scaled / delta;
                if (paru == 0.0) paru = dwarf / std::min(delta, 0.1);
                par = std::max(par, parl);
                par = std::min(par, paru);
                if (par == 0.0) par = ok ? gnorm_scaled / dxnorm : 0.0;
                if (par == 0.0) par = std::max(dwarf, 0.001 * paru);
                for (int it = 1;; ++it) {
                    ok = sparse_solve(s, par, dint, p_int, dx2, qn2);
                    if (!ok) {
                        parl = std::max(parl, par);
                        par = std::max(par * 10.0, 0.001 * paru);
                        if (it > 30) return -2;
                        continue;
                    }
                    dxnorm = std::sqrt(dx2);
                    const double temp = fp;
                    fp = dxnorm - delta;
                    if (std::fabs(fp) <= 0.1 * delta || (parl == 0.0 && fp <= temp && temp < 0.0) || it >= 10) break;
                    const double parc = (fp / delta) / qn2;
                    if (fp > 0.0) parl = std::max(parl, par);
                    if (fp < 0.0) paru = std::min(paru, par);
                    par = std::max(parl, par + parc);
                    if (par == 0.0) par = std::max(dwarf, 0.001 * paru);
                }
            }
            // ---- trial ----
            double pn2 = 0.0;
            for (int pid = 0; pid < n; ++pid) {
                step[pid] = -p_int[s.internal(pid)];
                x2[pid] = x[pid] + step[pid];
                const double t = diag[pid] * step[pid];
                pn2 += t * t;
            }
            const double pnorm = std::sqrt(pn2);
            if (iter == 1) delta = std::min(delta, pnorm);
            sparse_residuals(s, x2.data(), f2.data(), d2.data());
            ++nfev;
            const double fnorm1 = enorm(m, f2.data());
            double jp2 = 0.0;
            for (int i = 0; i < s.nobs; ++i) {
                double sx = 0.0, sy = 0.0;
                for (int c = 0; c < s.kmax; ++c) {
                    const int pid = s.jcol[(size_t)i * s.kmax + c];
                    if (pid < 0) continue;
                    sx += s.jac[2 * ((size_t)i * s.kmax + c)] * step[pid];
                    sy += s.jac[2 * ((size_t)i * s.kmax + c) + 1] * step[pid];
                }
                jp2 += sx * sx + sy * sy;
            }
            double actred = -1.0;
            if (0.1 * fnorm1 < fnorm) { const double t = fnorm1 / fnorm; actred = 1.0 - t * t; }
            const double temp1 = std::sqrt(jp2) / fnorm, temp2 = (std::sqrt(par) * pnorm) / fnorm;
            const double prered = temp1 * temp1 + temp2 * temp2 / 0.5;
            const double dirder = -(temp1 * temp1 + temp2 * temp2);
            ratio = prered != 0.0 ? actred / prered : 0.0;
            if (ratio <= 0.25) {
                double temp = actred >= 0.0 ? 0.5 : 0.5 * dirder / (dirder + 0.5 * actred);
                if (0.1 * fnorm1 >= fnorm || temp < 0.1) temp = 0.1;
                delta = temp * std::min(delta, pnorm / 0.1);
                par /= temp;
            } else if (par == 0.0 || ratio >= 0.75) {
                delta = pnorm / 0.5;
                par = 0.5 * par;
            }
            dist = d2;  // errorDistanceList now holds the trial's distances, accepted or not
            if (ratio >= 1e-4) {
                x = x2; fvec = f2;
                double t = 0.0;
                for (int j = 0; j < n; ++j) t += (diag[j] * x[j]) * (diag[j] * x[j]);
                xnorm = std::sqrt(t);
                fnorm = fnorm1;
                ++iter;
            }
            const bool small = std::fabs(actred) <= ftol && prered <= ftol && 0.5 * ratio <= 1.0;
            if (small) info = 1;
            if (delta <= xtol * xnorm) info = 2;
            if (small && info == 2) info = 3;
            if (info != 0) break;
            if (nfev >= maxfev) info = 5;
            if (std::fabs(actred) <= epsmch && prered <= epsmch && 0.5 * ratio <= 1.0) info = 6;
            if (delta <= epsmch * xnorm) info = 7;
            if (gnorm <= epsmch) info = 8;
            if (info != 0) break;
        } while (ratio < 1e-4);
        if (info != 0) break;
        if (max_outer > 0 && njev >= max_outer) { info = 5; break; }
    }
    *nfev_out = nfev; *njev_out = njev;
    return info;
}

}  // namespace orcs
