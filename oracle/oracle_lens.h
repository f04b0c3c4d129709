// TEST INFRASTRUCTURE ONLY -- see oracle_math.h for the rules.
//
// CPU restatement of the lens-distortion models the solver evaluates inside
// measureErrors (adjust_measureErrors.cpp:461-479):
//   lib/cppbind/mmlens/src/lens_model_3de_*.cpp           (applyModelDistort)
//   lib/cppbind/mmlens/src/distortion_operations.h:33-96  (apply_..._once)
//   lib/cppbind/mmlens/src/distortion_structs.h:75-301    (model wrappers)
//   lib/cppbind/mmlens/include/mmlens/lib.h:36-75         (unit <-> dn)
//   ldpk: ldpk_generic_distortion_base.h:99-118           (map_inverse)
//   ldpk: ldpk_classic_3de_mixed_distortion.h:29-73
//   ldpk: ldpk_radial_decentered_distortion.h:54-72
//   ldpk: ldpk_cylindric_extender.h:69-110
//   ldpk: ldpk_generic_anamorphic_distortion.h:216-292 (degree-4)
//   ldpk: ldpk_rotation_extender.h / ldpk_squeeze_extender.h /
//         ldpk_linear_extender.h
// Pinned against the reference's own sources via oracle/_ref (tests).
#pragma once
#include <cmath>

namespace orc {

struct V2 { double x, y; };
struct Mat2 { double a, b, c, d; };  // [[a b][c d]]
inline V2 m2v(const Mat2 &m, V2 p) { return {m.a * p.x + m.b * p.y, m.c * p.x + m.d * p.y}; }
inline Mat2 m2m(const Mat2 &p, const Mat2 &q) {
    return {p.a * q.a + p.b * q.c, p.a * q.b + p.b * q.d,
            p.c * q.a + p.d * q.c, p.c * q.b + p.d * q.d};
}
inline Mat2 m2inv(const Mat2 &m) {  // ldpk_vec2d.h:172
    double det = m.a * m.d - m.b * m.c;
    return {m.d / det, -m.b / det, -m.c / det, m.a / det};
}

// A lens "prepared" for evaluation: what set_parameter()/
// initialize_parameters() leave in the LDPK objects.
struct LensEval {
    int model = 0;     // MMB_LENS_*
    double cam[6] = {3.0, 3.6, 2.4, 1.0, 0.0, 0.0};  // lens_model.h:42
    double radius = 0; // lib.h:36-43
    double c[14] = {0};
    // classic: derived coefficients (update())
    double cxx, cxy, cyx, cyy, cxxx, cxxy, cxyy, cyxx, cyyx, cyyy;
    // radial: cylindric matrices
    Mat2 cyl_m, cyl_inv;
    // anamorphic: prepare() coefficients + extender matrices
    double ax2, ay2, ax4, ax2y2, ay4, bx2, by2, bx4, bx2y2, by4;
    Mat2 pre_m, pre_inv;   // m_rotation_squeeze_xy(_rescale)_pixel_aspect
    Mat2 post_m, post_inv; // m_pixel_aspect(_rescale)_and_rotation
};

inline void lens_prepare(LensEval &L) {
    const double w = L.cam[1], h = L.cam[2];
    L.radius = std::sqrt(w * w + h * h) / 2.0;
    const double *c = L.c;
    if (L.model == 1) {  // ldpk_classic_3de_mixed_distortion.h:29-42
        const double ld = c[0], sq = c[1], cx = c[2], cy = c[3], qu = c[4];
        L.cxx = ld / sq; L.cxy = (ld + cx) / sq; L.cyx = ld + cy; L.cyy = ld;
        L.cxxx = qu / sq; L.cxxy = 2.0 * qu / sq; L.cxyy = qu / sq;
        L.cyxx = qu; L.cyyx = 2.0 * qu; L.cyyy = qu;
    } else if (L.model == 2) {  // ldpk_cylindric_extender.h:82-87
        const double q = std::sqrt(1.0 + c[7]);
        const double cs = std::cos(c[6] * M_PI / 180.0), sn = std::sin(c[6] * M_PI / 180.0);
        L.cyl_m = {cs * cs * q + sn * sn / q, (q - 1.0 / q) * cs * sn,
                   (q - 1.0 / q) * cs * sn, cs * cs / q + sn * sn * q};
        L.cyl_inv = m2inv(L.cyl_m);
    } else if (L.model == 3 || L.model == 4) {
        // generic_anamorphic_distortion<.,.,4>::prepare()
        const double cx02 = c[0], cy02 = c[1], cx22 = c[2], cy22 = c[3], cx04 = c[4],
                     cy04 = c[5], cx24 = c[6], cy24 = c[7], cx44 = c[8], cy44 = c[9];
        L.ax2 = cx02 + cx22; L.ay2 = cx02 - cx22;
        L.ax4 = cx04 + cx24 + cx44; L.ax2y2 = 2.0 * cx04 - 6.0 * cx44; L.ay4 = cx04 - cx24 + cx44;
        L.bx2 = cy02 + cy22; L.by2 = cy02 - cy22;
        L.bx4 = cy04 + cy24 + cy44; L.bx2y2 = 2.0 * cy04 - 6.0 * cy44; L.by4 = cy04 - cy24 + cy44;
        // distortion_structs.h:158-176 / 230-252
        const double phi = c[10] / 180.0 * M_PI;
        Mat2 rot = {std::cos(phi), -std::sin(phi), std::sin(phi), std::cos(phi)};
        Mat2 sqx = {c[11], 0, 0, 1.0}, sqy = {1.0, 0, 0, c[12]};
        Mat2 pa = {L.cam[3], 0, 0, 1.0};
        if (L.model == 3) {
            L.pre_m = m2m(m2m(m2m(rot, sqx), sqy), pa);
            L.post_m = m2m(pa, rot);
        } else {
            Mat2 rs = {c[13], 0, 0, 1.0};
            L.pre_m = m2m(m2m(m2m(m2m(rot, sqx), sqy), rs), pa);
            L.post_m = m2m(m2m(pa, rs), rot);
        }
        L.pre_inv = m2inv(L.pre_m);
        L.post_inv = m2inv(L.post_m);
    }
}

// The polynomial core f(p) of each model ("remove distortion").
inline V2 lens_core(const LensEval &L, V2 p) {
    const double x = p.x, y = p.y;
    if (L.model == 1) {  // ldpk_classic_3de_mixed_distortion.h:58-73
        const double x2 = x * x, y2 = y * y, x4 = x2 * x2, y4 = y2 * y2, xy2 = x2 * y2;
        return {x * (1 + L.cxx * x2 + L.cxy * y2 + L.cxxx * x4 + L.cxxy * xy2 + L.cxyy * y4),
                y * (1 + L.cyx * x2 + L.cyy * y2 + L.cyxx * x4 + L.cyyx * xy2 + L.cyyy * y4)};
    } else if (L.model == 2) {  // ldpk_radial_decentered_distortion.h:54-72
        const double c2 = L.c[0], u2 = L.c[1], v2 = L.c[2], c4 = L.c[3], u4 = L.c[4], v4 = L.c[5];
        const double x2 = x * x, y2 = y * y, xy = x * y, r2 = x2 + y2, r4 = r2 * r2;
        return {x * (1.0 + c2 * r2 + c4 * r4) + (r2 + 2.0 * x2) * (u2 + u4 * r2) + 2.0 * xy * (v2 + v4 * r2),
                y * (1.0 + c2 * r2 + c4 * r4) + (r2 + 2.0 * y2) * (v2 + v4 * r2) + 2.0 * xy * (u2 + u4 * r2)};
    } else {  // ldpk_generic_anamorphic_distortion.h:279-292
        const double x2 = x * x, x4 = x2 * x2, y2 = y * y, y4 = y2 * y2;
        return {x * (1.0 + x2 * L.ax2 + y2 * L.ay2 + x4 * L.ax4 + x2 * y2 * L.ax2y2 + y4 * L.ay4),
                y * (1.0 + x2 * L.bx2 + y2 * L.by2 + x4 * L.bx4 + x2 * y2 * L.bx2y2 + y4 * L.by4)};
    }
}

// ldpk_generic_distortion_base.h:99-118: fixed-point inverse, eps 1e-6,
// at most 20 iterations then 2 post-iterations.
inline V2 lens_core_inverse(const LensEval &L, V2 q) {
    V2 fq = lens_core(L, q);
    V2 p = {q.x - (fq.x - q.x), q.y - (fq.y - q.y)};
    for (int i = 0; i < 20; ++i) {
        V2 qi = lens_core(L, p);
        p = {p.x + q.x - qi.x, p.y + q.y - qi.y};
        const double dx = qi.x - q.x, dy = qi.y - q.y;
        if (std::sqrt(dx * dx + dy * dy) < 1e-6) break;
    }
    for (int i = 0; i < 2; ++i) {
        V2 qi = lens_core(L, p);
        p = {p.x + q.x - qi.x, p.y + q.y - qi.y};
    }
    return p;
}

inline V2 unit_to_dn(const LensEval &L, V2 u) {  // lib.h:45-58
    return {((u.x - 1.0 / 2.0) * L.cam[1] - L.cam[4]) / L.radius,
            ((u.y - 1.0 / 2.0) * L.cam[2] - L.cam[5]) / L.radius};
}
inline V2 dn_to_unit(const LensEval &L, V2 d) {  // lib.h:60-75
    double xc = d.x * L.radius, yc = d.y * L.radius;
    xc += (L.cam[1] / 2) + L.cam[4];
    yc += (L.cam[2] / 2) + L.cam[5];
    return {xc / L.cam[1], yc / L.cam[2]};
}

// applyModelUndistort: removes distortion (direct polynomial).
inline void lens_undistort(const LensEval &L, double xd, double yd, double &xu, double &yu) {
    if (L.model == 0) { xu = xd; yu = yd; return; }
    V2 p = unit_to_dn(L, {xd + 0.5, yd + 0.5});
    V2 o;
    if (L.model == 1) o = lens_core(L, p);
    else if (L.model == 2) o = m2v(L.cyl_m, lens_core(L, p));
    else o = m2v(L.pre_m, lens_core(L, m2v(L.post_inv, p)));
    V2 u = dn_to_unit(L, o);
    xu = u.x - 0.5; yu = u.y - 0.5;
}

// applyModelDistort: what measureErrors calls on the projected point
// (lens_model_3de_radial_decentered_deg_4_cylindric.cpp:77-118 and siblings).
inline void lens_distort(const LensEval &L, double xd, double yd, double &xo, double &yo) {
    if (L.model == 0) { xo = xd; yo = yd; return; }
    V2 p = unit_to_dn(L, {xd + 0.5, yd + 0.5});
    V2 o;
    if (L.model == 1) o = lens_core_inverse(L, p);
    else if (L.model == 2) o = lens_core_inverse(L, m2v(L.cyl_inv, p));
    else o = m2v(L.post_m, lens_core_inverse(L, m2v(L.pre_inv, p)));
    V2 u = dn_to_unit(L, o);
    xo = u.x - 0.5; yo = u.y - 0.5;
}

}  // namespace orc
