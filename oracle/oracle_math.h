// TEST INFRASTRUCTURE ONLY -- CPU oracle for the mmSolver solve hot path.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may build, import or call anything under oracle/.
// The product (mayamatchmovesolver_b200/) never links or loads this code.
//
// This header restates, in plain scalar FP64 C++, the scene mathematics of the
// reference's Rust crate `mmscenegraph` (not compilable here: no rustc) and
// the lens models of LDPK 2.8 as wrapped by lib/cppbind/mmlens.  Every
// function cites the reference file:line it follows (paths relative to
// /root/reference; `ldpk:` = inside external/archives/ldpk-2.8.tar).
//
// Parity pinning: scene math is pinned by the Rust KATs
// (lib/rust/mmscenegraph/tests/reprojection.rs:36-99, camera.rs:79-120); lens
// math is pinned against the reference's own mmlens C++ sources compiled into
// oracle/_ref/libmmlens_ref.so (see oracle/Makefile).
#pragma once
#include <cmath>
#include <cstring>

namespace orc {

constexpr double DEG2RAD = 0.017453292519943295;  // constant.rs DEGREES_TO_RADIANS
constexpr double MM_TO_INCH = 1.0 / 25.4;          // constant.rs MM_TO_INCH
constexpr double INCH_TO_MM = 25.4;
constexpr double MM_TO_CM = 0.1;

// 4x4 row-major: m[r*4+c]; column-vector convention like nalgebra.
struct M4 { double m[16]; };

inline M4 m4_identity() {
    M4 r; std::memset(r.m, 0, sizeof(r.m));
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0; return r;
}

inline M4 m4_mul(const M4 &a, const M4 &b) {
    M4 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0.0;
            for (int k = 0; k < 4; ++k) s += a.m[i * 4 + k] * b.m[k * 4 + j];
            r.m[i * 4 + j] = s;
        }
    return r;
}

// General 4x4 inverse by cofactors (nalgebra try_inverse for 4x4 uses the same
// closed form); returns false when singular (reprojection.rs:34-38 then
// substitutes identity).
inline bool m4_inverse(const M4 &a, M4 &out) {
    const double *m = a.m; double inv[16];
    inv[0] = m[5]*m[10]*m[15] - m[5]*m[11]*m[14] - m[9]*m[6]*m[15] + m[9]*m[7]*m[14] + m[13]*m[6]*m[11] - m[13]*m[7]*m[10];
    inv[4] = -m[4]*m[10]*m[15] + m[4]*m[11]*m[14] + m[8]*m[6]*m[15] - m[8]*m[7]*m[14] - m[12]*m[6]*m[11] + m[12]*m[7]*m[10];
    inv[8] = m[4]*m[9]*m[15] - m[4]*m[11]*m[13] - m[8]*m[5]*m[15] + m[8]*m[7]*m[13] + m[12]*m[5]*m[11] - m[12]*m[7]*m[9];
    inv[12] = -m[4]*m[9]*m[14] + m[4]*m[10]*m[13] + m[8]*m[5]*m[14] - m[8]*m[6]*m[13] - m[12]*m[5]*m[10] + m[12]*m[6]*m[9];
    inv[1] = -m[1]*m[10]*m[15] + m[1]*m[11]*m[14] + m[9]*m[2]*m[15] - m[9]*m[3]*m[14] - m[13]*m[2]*m[11] + m[13]*m[3]*m[10];
    inv[5] = m[0]*m[10]*m[15] - m[0]*m[11]*m[14] - m[8]*m[2]*m[15] + m[8]*m[3]*m[14] + m[12]*m[2]*m[11] - m[12]*m[3]*m[10];
    inv[9] = -m[0]*m[9]*m[15] + m[0]*m[11]*m[13] + m[8]*m[1]*m[15] - m[8]*m[3]*m[13] - m[12]*m[1]*m[11] + m[12]*m[3]*m[9];
    inv[13] = m[0]*m[9]*m[14] - m[0]*m[10]*m[13] - m[8]*m[1]*m[14] + m[8]*m[2]*m[13] + m[12]*m[1]*m[10] - m[12]*m[2]*m[9];
    inv[2] = m[1]*m[6]*m[15] - m[1]*m[7]*m[14] - m[5]*m[2]*m[15] + m[5]*m[3]*m[14] + m[13]*m[2]*m[7] - m[13]*m[3]*m[6];
    inv[6] = -m[0]*m[6]*m[15] + m[0]*m[7]*m[14] + m[4]*m[2]*m[15] - m[4]*m[3]*m[14] - m[12]*m[2]*m[7] + m[12]*m[3]*m[6];
    inv[10] = m[0]*m[5]*m[15] - m[0]*m[7]*m[13] - m[4]*m[1]*m[15] + m[4]*m[3]*m[13] + m[12]*m[1]*m[7] - m[12]*m[3]*m[5];
    inv[14] = -m[0]*m[5]*m[14] + m[0]*m[6]*m[13] + m[4]*m[1]*m[14] - m[4]*m[2]*m[13] - m[12]*m[1]*m[6] + m[12]*m[2]*m[5];
    inv[3] = -m[1]*m[6]*m[11] + m[1]*m[7]*m[10] + m[5]*m[2]*m[11] - m[5]*m[3]*m[10] - m[9]*m[2]*m[7] + m[9]*m[3]*m[6];
    inv[7] = m[0]*m[6]*m[11] - m[0]*m[7]*m[10] - m[4]*m[2]*m[11] + m[4]*m[3]*m[10] + m[8]*m[2]*m[7] - m[8]*m[3]*m[6];
    inv[11] = -m[0]*m[5]*m[11] + m[0]*m[7]*m[9] + m[4]*m[1]*m[11] - m[4]*m[3]*m[9] - m[8]*m[1]*m[7] + m[8]*m[3]*m[5];
    inv[15] = m[0]*m[5]*m[10] - m[0]*m[6]*m[9] - m[4]*m[1]*m[10] + m[4]*m[2]*m[9] + m[8]*m[1]*m[6] - m[8]*m[2]*m[5];
    double det = m[0]*inv[0] + m[1]*inv[4] + m[2]*inv[8] + m[3]*inv[12];
    if (det == 0.0) return false;
    double id = 1.0 / det;
    for (int i = 0; i < 16; ++i) out.m[i] = inv[i] * id;
    return true;
}

// transform.rs:338-453  calculate_matrix_with_values : T * R * S
inline M4 local_matrix(double tx, double ty, double tz, double rx, double ry,
                       double rz, double sx, double sy, double sz, int roo) {
    M4 s = m4_identity(); s.m[0] = sx; s.m[5] = sy; s.m[10] = sz;
    const double srx = std::sin(rx * DEG2RAD), crx = std::cos(rx * DEG2RAD);
    const double sry = std::sin(ry * DEG2RAD), cry = std::cos(ry * DEG2RAD);
    const double srz = std::sin(rz * DEG2RAD), crz = std::cos(rz * DEG2RAD);
    M4 X = m4_identity(), Y = m4_identity(), Z = m4_identity();
    X.m[5] = crx; X.m[6] = -srx; X.m[9] = srx; X.m[10] = crx;
    Y.m[0] = cry; Y.m[2] = sry; Y.m[8] = -sry; Y.m[10] = cry;
    Z.m[0] = crz; Z.m[1] = -srz; Z.m[4] = srz; Z.m[5] = crz;
    M4 r;
    switch (roo) {  // transform.rs:433-440
        case 0: r = m4_mul(m4_mul(Z, Y), X); break;  // XYZ
        case 1: r = m4_mul(m4_mul(X, Z), Y); break;  // YZX
        case 2: r = m4_mul(m4_mul(Y, X), Z); break;  // ZXY
        case 3: r = m4_mul(m4_mul(Y, Z), X); break;  // XZY
        case 4: r = m4_mul(m4_mul(Z, X), Y); break;  // YXZ
        default: r = m4_mul(m4_mul(X, Y), Z); break; // ZYX
    }
    M4 t = m4_identity(); t.m[3] = tx; t.m[7] = ty; t.m[11] = tz;
    return m4_mul(m4_mul(t, r), s);
}

// camera.rs:153-327 get_projection_matrix (inputs as dag.rs:120-172 passes
// them: sensor/offset in inches, focal in mm, near hard-coded by the caller).
inline M4 projection_matrix(double focal_mm, double fbw_inch, double fbh_inch,
                            double offx_inch, double offy_inch, double img_w,
                            double img_h, int film_fit, double near_cm,
                            double far_cm, double camera_scale) {
    const double film_aspect = fbw_inch / fbh_inch;
    const double image_aspect = img_w / img_h;
    // compute_frustum_coordinates camera.rs:153-175
    const double fw = fbw_inch * INCH_TO_MM, fh = fbh_inch * INCH_TO_MM;
    const double ox = offx_inch * INCH_TO_MM, oy = offy_inch * INCH_TO_MM;
    const double f2n = (near_cm / focal_mm) * camera_scale;
    double right = f2n * (0.5 * fw + ox), left = f2n * (-0.5 * fw + ox);
    double top = f2n * (0.5 * fh + oy), bottom = f2n * (-0.5 * fh + oy);
    // apply_film_fit_logic camera.rs:183-242 (note: Horizontal writes scale.x)
    double fsx = 1.0, fsy = 1.0, size_x = 0.0, size_y = 0.0;
    if (film_fit == 1) {
        fsx = image_aspect / film_aspect;
        size_x = right - left; size_y = size_x / image_aspect;
    } else if (film_fit == 2) {
        fsx = 1.0 / (image_aspect / film_aspect);
        size_y = top - bottom; size_x = size_y * image_aspect;
    } else if (film_fit == 0) {
        if (film_aspect > image_aspect) {
            fsx = film_aspect / image_aspect;
            size_y = top - bottom; size_x = size_y * image_aspect;
        } else {
            fsy = image_aspect / film_aspect;
            size_x = right - left;
            size_y = (size_x * (film_aspect / image_aspect)) / film_aspect;
        }
    } else {
        if (film_aspect > image_aspect) {
            fsy = image_aspect / film_aspect;
            size_x = right - left; size_y = (right - left) / image_aspect;
        } else {
            fsx = film_aspect / image_aspect;
            size_x = (right - left) * (image_aspect / film_aspect);
            size_y = top - bottom;
        }
    }
    right *= fsx; left *= fsx; top *= fsy; bottom *= fsy;
    // compute_projection_matrix camera.rs:245-279 (row-major constructor)
    M4 p; std::memset(p.m, 0, sizeof(p.m));
    p.m[0] = 1.0 / (size_x * 0.5) * MM_TO_CM;
    p.m[5] = 1.0 / (size_y * 0.5) * MM_TO_CM;
    p.m[8] = (right + left) / (right - left) * fsx;
    p.m[9] = (top + bottom) / (top - bottom) * fsy;
    p.m[10] = (far_cm + near_cm) / (far_cm - near_cm);
    p.m[11] = 2.0 * far_cm * near_cm / (far_cm - near_cm);
    p.m[14] = -1.0;
    return p;
}

// reprojection.rs:27-63 as called from flat.rs:319-323 with
// (cam_tfm_matrix, cam_proj_matrix, bnd_matrix): S = P * inv(camWorld) * bnd.
inline void reproject_normalised(const M4 &cam_world, const M4 &proj,
                                 const M4 &bnd_world, double &x, double &y) {
    M4 inv;
    if (!m4_inverse(cam_world, inv)) inv = m4_identity();
    M4 wp = m4_mul(proj, inv);
    M4 s = m4_mul(wp, bnd_world);
    x = s.m[3] / s.m[15] * 0.5;
    y = s.m[7] / s.m[15] * 0.5;
}

// flat.rs:73-97 scale_xy_with_film_fit
inline void scale_marker_film_fit(int film_fit, double sensor_aspect,
                                  double render_aspect, double &x, double &y) {
    if (film_fit == 1) { y *= render_aspect / sensor_aspect; }
    else if (film_fit == 2) { x *= 1.0 / (render_aspect / sensor_aspect); }
    else if (film_fit == 0) {
        if (sensor_aspect > render_aspect) x *= sensor_aspect / render_aspect;
        else y *= render_aspect / sensor_aspect;
    } else if (film_fit == 3) {
        if (sensor_aspect > render_aspect) y *= render_aspect / sensor_aspect;
        else x *= sensor_aspect / render_aspect;
    }
}

}  // namespace orc
