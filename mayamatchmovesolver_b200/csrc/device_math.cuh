// FP64 scene mathematics of the solve path, written for the CUDA kernels.
//
// What is computed follows the reference (mmSceneGraph convention, column
// vectors):
//   transform.rs:338-453   local = T * R * S, rotate orders
//   dag.rs:120-172,234-327 world = parent * local ; projection inputs
//   camera.rs:153-327      frustum, film fit, projection matrix
//   reprojection.rs:27-62  S = P * inv(camWorld) * bundleWorld, x,y * 0.5
//   flat.rs:73-97          marker film-fit scale
//   mmlens + LDPK 2.8      lens models (see lens section below)
// How it is computed is ours: a camera at a frame is reduced once to the three
// rows of P*inv(W) that the reprojection reads, so an observation costs one
// 3x4 mat-vec and one reciprocal.
//
// MMB_HD lets tests/emu compile the same per-thread bodies on the host to
// debug index logic without a GPU.  The shipped library only ever launches
// them as CUDA kernels.
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define MMB_HD __host__ __device__ __forceinline__
#else
#define MMB_HD inline
#endif

namespace mmb {

constexpr double kDeg2Rad = 0.017453292519943295;  // constant.rs
constexpr double kMmToInch = 1.0 / 25.4;
constexpr double kInchToMm = 25.4;
constexpr double kMmToCm = 0.1;
constexpr double kPi = 3.14159265358979323846;

// 3x4 affine (bottom row 0 0 0 1 implied): every T*R*S product stays affine.
struct Aff {
    double m[12];  // row-major 3x4
};

MMB_HD Aff aff_mul(const Aff &a, const Aff &b) {
    Aff r;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            double s = a.m[i * 4 + 0] * b.m[0 * 4 + j] + a.m[i * 4 + 1] * b.m[1 * 4 + j] +
                       a.m[i * 4 + 2] * b.m[2 * 4 + j];
            if (j == 3) s += a.m[i * 4 + 3];
            r.m[i * 4 + j] = s;
        }
    }
    return r;
}

MMB_HD void mul3x3x3(const double *A, const double *B, const double *C, double *R) {
    double AB[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            AB[i * 3 + j] = A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j] + A[i * 3 + 2] * B[2 * 3 + j];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            R[i * 3 + j] = AB[i * 3 + 0] * C[0 * 3 + j] + AB[i * 3 + 1] * C[1 * 3 + j] + AB[i * 3 + 2] * C[2 * 3 + j];
}

// transform.rs:338-453: T * R * S with R from the rotate order table :433-440
MMB_HD Aff local_matrix(double tx, double ty, double tz, double rx, double ry, double rz,
                        double sx, double sy, double sz, int roo) {
    double srx, crx, sry, cry, srz, crz;
    sincos(rx * kDeg2Rad, &srx, &crx);
    sincos(ry * kDeg2Rad, &sry, &cry);
    sincos(rz * kDeg2Rad, &srz, &crz);
    // 3x3 factors
    const double X[9] = {1, 0, 0, 0, crx, -srx, 0, srx, crx};
    const double Y[9] = {cry, 0, sry, 0, 1, 0, -sry, 0, cry};
    const double Z[9] = {crz, -srz, 0, srz, crz, 0, 0, 0, 1};
    double R[9];
    switch (roo) {  // R = A * B * C (transform.rs:433-440)
        case 0: mul3x3x3(Z, Y, X, R); break;  // XYZ
        case 1: mul3x3x3(X, Z, Y, R); break;  // YZX
        case 2: mul3x3x3(Y, X, Z, R); break;  // ZXY
        case 3: mul3x3x3(Y, Z, X, R); break;  // XZY
        case 4: mul3x3x3(Z, X, Y, R); break;  // YXZ
        default: mul3x3x3(X, Y, Z, R); break; // ZYX
    }
    Aff o;
    o.m[0] = R[0] * sx; o.m[1] = R[1] * sy; o.m[2] = R[2] * sz; o.m[3] = tx;
    o.m[4] = R[3] * sx; o.m[5] = R[4] * sy; o.m[6] = R[5] * sz; o.m[7] = ty;
    o.m[8] = R[6] * sx; o.m[9] = R[7] * sy; o.m[10] = R[8] * sz; o.m[11] = tz;
    return o;
}

// inverse of an affine 3x4 (reprojection.rs:34-38: singular -> identity)
MMB_HD Aff aff_inverse(const Aff &a) {
    const double *m = a.m;
    const double c00 = m[5] * m[10] - m[6] * m[9];
    const double c01 = m[6] * m[8] - m[4] * m[10];
    const double c02 = m[4] * m[9] - m[5] * m[8];
    const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
    Aff r;
    if (det == 0.0) {
        for (int i = 0; i < 12; ++i) r.m[i] = 0.0;
        r.m[0] = r.m[5] = r.m[10] = 1.0;
        return r;
    }
    const double id = 1.0 / det;
    r.m[0] = c00 * id;
    r.m[1] = (m[2] * m[9] - m[1] * m[10]) * id;
    r.m[2] = (m[1] * m[6] - m[2] * m[5]) * id;
    r.m[4] = c01 * id;
    r.m[5] = (m[0] * m[10] - m[2] * m[8]) * id;
    r.m[6] = (m[2] * m[4] - m[0] * m[6]) * id;
    r.m[8] = c02 * id;
    r.m[9] = (m[1] * m[8] - m[0] * m[9]) * id;
    r.m[10] = (m[0] * m[5] - m[1] * m[4]) * id;
    r.m[3] = -(r.m[0] * m[3] + r.m[1] * m[7] + r.m[2] * m[11]);
    r.m[7] = -(r.m[4] * m[3] + r.m[5] * m[7] + r.m[6] * m[11]);
    r.m[11] = -(r.m[8] * m[3] + r.m[9] * m[7] + r.m[10] * m[11]);
    return r;
}

// The non-zero entries of the projection matrix (camera.rs:245-279):
//   row0 = [p00 0 0 0] row1 = [0 p11 0 0] row2 = [p20 p21 p22 p23]
//   row3 = [0 0 -1 0]
struct Proj {
    double p00, p11, p20, p21;
};

// camera.rs:153-242 with the inputs dag.rs:120-172 passes (mm -> inch -> mm).
MMB_HD Proj projection_entries(double focal_mm, double sensor_w_mm, double sensor_h_mm,
                               double offx_mm, double offy_mm, double img_w, double img_h,
                               int film_fit, double near_cm, double camera_scale, int maya = 0) {
    const double fbw_inch = sensor_w_mm * kMmToInch, fbh_inch = sensor_h_mm * kMmToInch;
    const double offx_inch = offx_mm * kMmToInch, offy_inch = offy_mm * kMmToInch;
    const double film_aspect = fbw_inch / fbh_inch;
    const double image_aspect = img_w / img_h;
    const double fw = fbw_inch * kInchToMm, fh = fbh_inch * kInchToMm;
    const double ox = offx_inch * kInchToMm, oy = offy_inch * kInchToMm;
    const double f2n = (near_cm / focal_mm) * camera_scale;
    double right = f2n * (0.5 * fw + ox), left = f2n * (-0.5 * fw + ox);
    double top = f2n * (0.5 * fh + oy), bottom = f2n * (-0.5 * fh + oy);
    double fsx = 1.0, fsy = 1.0, size_x = 0.0, size_y = 0.0;
    if (film_fit == 1) {  // horizontal: camera.rs:200-201 writes scale.x, maya_camera.cpp:160-165 scale.y
        if (maya) fsy = image_aspect / film_aspect; else fsx = image_aspect / film_aspect;
        size_x = right - left; size_y = size_x / image_aspect;
    } else if (film_fit == 2) {  // vertical
        fsx = 1.0 / (image_aspect / film_aspect);
        size_y = top - bottom; size_x = size_y * image_aspect;
    } else if (film_fit == 0) {  // fill
        if (film_aspect > image_aspect) {
            fsx = film_aspect / image_aspect;
            size_y = top - bottom; size_x = size_y * image_aspect;
        } else {
            fsy = image_aspect / film_aspect;
            size_x = right - left;
            size_y = (size_x * (film_aspect / image_aspect)) / film_aspect;
        }
    } else {  // overscan
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
    Proj p;
    p.p00 = 1.0 / (size_x * 0.5) * kMmToCm;
    p.p11 = 1.0 / (size_y * 0.5) * kMmToCm;
    p.p20 = (right + left) / (right - left) * fsx;
    p.p21 = (top + bottom) / (top - bottom) * fsy;
    return p;
}

// flat.rs:73-97: factors the marker x,y are multiplied with
MMB_HD void marker_film_fit_scale(int film_fit, double sensor_aspect, double render_aspect,
                                  double &sx, double &sy) {
    sx = 1.0; sy = 1.0;
    if (film_fit == 1) sy = render_aspect / sensor_aspect;
    else if (film_fit == 2) sx = 1.0 / (render_aspect / sensor_aspect);
    else if (film_fit == 0) {
        if (sensor_aspect > render_aspect) sx = sensor_aspect / render_aspect;
        else sy = render_aspect / sensor_aspect;
    } else {
        if (sensor_aspect > render_aspect) sy = render_aspect / sensor_aspect;
        else sx = sensor_aspect / render_aspect;
    }
}

// Rows 0, 1 and 3 of P * inv(camWorld) (3x4 each; row 3 of the product is
// -row2 of inv(W) because P's last row is [0 0 -1 0]).  Note P's row 0/1
// only have their diagonal entry, and S[0,3]/S[3,3] never reads row 2.
// out[0..3] = row0, out[4..7] = row1, out[8..11] = row3.
// maya: Maya's row-vector layout (maya_camera.cpp:339-361, point * inverse(world) * proj) puts the
// film-offset entries p20, p21 on the camera-space z: x_clip = p00 x + p20 z, y_clip = p11 y + p21 z
// (in the mmscenegraph layout they only reach z_clip, which nothing reads).
MMB_HD void camera_rows(const Aff &cam_world, const Proj &p, double *out, int maya = 0) {
    const Aff inv = aff_inverse(cam_world);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        out[j] = p.p00 * inv.m[j];
        out[4 + j] = p.p11 * inv.m[4 + j];
        if (maya) { out[j] += p.p20 * inv.m[8 + j]; out[4 + j] += p.p21 * inv.m[8 + j]; }
        out[8 + j] = -inv.m[8 + j];
    }
}

// ---- lens models -----------------------------------------------------------
// Prepared lens record (24 doubles), what LDPK's set_parameter()/
// initialize_parameters() leave behind (distortion_structs.h:75-301):
//  [0] model  [1] film back radius (lib.h:36-43)  [2] fbw cm  [3] fbh cm
//  [4] lco x  [5] lco y  [6..15] polynomial coefficients
//  [16..19] 2x2 applied BEFORE the inverse polynomial (a b c d)
//  [20..23] 2x2 applied AFTER it
constexpr int kLensRec = 24;

MMB_HD void m2_inv(const double *m, double *o) {  // ldpk_vec2d.h:172
    const double det = m[0] * m[3] - m[1] * m[2];
    o[0] = m[3] / det; o[1] = -m[1] / det; o[2] = -m[2] / det; o[3] = m[0] / det;
}
MMB_HD void m2_mul(const double *p, const double *q, double *o) {
    const double a = p[0] * q[0] + p[1] * q[2], b = p[0] * q[1] + p[1] * q[3];
    const double c = p[2] * q[0] + p[3] * q[2], d = p[2] * q[1] + p[3] * q[3];
    o[0] = a; o[1] = b; o[2] = c; o[3] = d;
}

// c[14]: raw coefficients in the mmlens parameter order; cam[6]: focal_cm,
// fbw_cm, fbh_cm, pixel_aspect, lco_x_cm, lco_y_cm.
MMB_HD void lens_prepare(int model, const double *c, const double *cam, double *rec) {
    for (int i = 0; i < kLensRec; ++i) rec[i] = 0.0;
    rec[0] = (double)model;
    const double w = cam[1], h = cam[2];
    rec[1] = sqrt(w * w + h * h) / 2.0;
    rec[2] = w; rec[3] = h; rec[4] = cam[4]; rec[5] = cam[5];
    rec[16] = 1.0; rec[19] = 1.0; rec[20] = 1.0; rec[23] = 1.0;
    if (model == 1) {  // ldpk_classic_3de_mixed_distortion.h:29-42
        const double ld = c[0], sq = c[1], cx = c[2], cy = c[3], qu = c[4];
        // lens_core order: x row (x2 y2 x4 x2y2 y4) then y row
        rec[6] = ld / sq; rec[7] = (ld + cx) / sq;
        rec[8] = qu / sq; rec[9] = 2.0 * qu / sq; rec[10] = qu / sq;
        rec[11] = ld + cy; rec[12] = ld;
        rec[13] = qu; rec[14] = 2.0 * qu; rec[15] = qu;
    } else if (model == 2) {  // radial decentered deg 4 + cylindric extender
        for (int i = 0; i < 6; ++i) rec[6 + i] = c[i];
        const double q = sqrt(1.0 + c[7]);
        const double cs = cos(c[6] * kPi / 180.0), sn = sin(c[6] * kPi / 180.0);
        double cyl[4] = {cs * cs * q + sn * sn / q, (q - 1.0 / q) * cs * sn,
                         (q - 1.0 / q) * cs * sn, cs * cs / q + sn * sn * q};
        m2_inv(cyl, &rec[16]);  // distort applies inv(cyl) first
    } else if (model == 3 || model == 4) {  // anamorphic deg 4
        const double cx02 = c[0], cy02 = c[1], cx22 = c[2], cy22 = c[3], cx04 = c[4],
                     cy04 = c[5], cx24 = c[6], cy24 = c[7], cx44 = c[8], cy44 = c[9];
        rec[6] = cx02 + cx22; rec[7] = cx02 - cx22;
        rec[8] = cx04 + cx24 + cx44; rec[9] = 2.0 * cx04 - 6.0 * cx44; rec[10] = cx04 - cx24 + cx44;
        rec[11] = cy02 + cy22; rec[12] = cy02 - cy22;
        rec[13] = cy04 + cy24 + cy44; rec[14] = 2.0 * cy04 - 6.0 * cy44; rec[15] = cy04 - cy24 + cy44;
        const double phi = c[10] / 180.0 * kPi;
        double rot[4] = {cos(phi), -sin(phi), sin(phi), cos(phi)};
        double sqx[4] = {c[11], 0, 0, 1.0}, sqy[4] = {1.0, 0, 0, c[12]};
        double pa[4] = {cam[3], 0, 0, 1.0};
        double pre[4], post[4];
        m2_mul(rot, sqx, pre); m2_mul(pre, sqy, pre);
        if (model == 4) {
            double rs[4] = {c[13], 0, 0, 1.0};
            m2_mul(pre, rs, pre); m2_mul(pre, pa, pre);
            m2_mul(pa, rs, post); m2_mul(post, rot, post);
        } else {
            m2_mul(pre, pa, pre);
            m2_mul(pa, rot, post);
        }
        m2_inv(pre, &rec[16]);
        rec[20] = post[0]; rec[21] = post[1]; rec[22] = post[2]; rec[23] = post[3];
    }
}

// The polynomial core f(p) ("remove distortion") of each model.
MMB_HD void lens_core(int model, const double *k, double x, double y, double &ox, double &oy) {
    if (model == 2) {  // ldpk_radial_decentered_distortion.h:54-72
        const double c2 = k[0], u2 = k[1], v2 = k[2], c4 = k[3], u4 = k[4], v4 = k[5];
        const double x2 = x * x, y2 = y * y, xy = x * y, r2 = x2 + y2, r4 = r2 * r2;
        const double rad = 1.0 + c2 * r2 + c4 * r4;
        const double uu = u2 + u4 * r2, vv = v2 + v4 * r2;
        ox = x * rad + (r2 + 2.0 * x2) * uu + 2.0 * xy * vv;
        oy = y * rad + (r2 + 2.0 * y2) * vv + 2.0 * xy * uu;
    } else {
        // classic (ldpk_classic_3de_mixed_distortion.h:58-73) and anamorphic
        // (ldpk_generic_anamorphic_distortion.h:279-292) share one shape
        const double x2 = x * x, y2 = y * y, x4 = x2 * x2, y4 = y2 * y2, x2y2 = x2 * y2;
        ox = x * (1.0 + k[0] * x2 + k[1] * y2 + k[2] * x4 + k[3] * x2y2 + k[4] * y4);
        oy = y * (1.0 + k[5] * x2 + k[6] * y2 + k[7] * x4 + k[8] * x2y2 + k[9] * y4);
    }
}

// applyModelDistort (lens_model_3de_*.cpp) on a point in -0.5..0.5:
// unit -> diagonal-normalised (lib.h:45-58) -> pre 2x2 -> fixed-point inverse
// (ldpk_generic_distortion_base.h:99-118: eps 1e-6, <=20 iterations, 2 post
// iterations) -> post 2x2 -> unit (lib.h:60-75).  Non-finite outputs keep the
// input (adjust_measureErrors.cpp:469-478).
MMB_HD void lens_distort(const double *rec, double px, double py, double &outx, double &outy) {
    const int model = (int)rec[0];
    outx = px; outy = py;
    if (model == 0) return;
    const double radius = rec[1], w = rec[2], h = rec[3], lx = rec[4], ly = rec[5];
    const double *k = &rec[6];
    double qx = (((px + 0.5) - 0.5) * w - lx) / radius;
    double qy = (((py + 0.5) - 0.5) * h - ly) / radius;
    {
        const double tx = rec[16] * qx + rec[17] * qy, ty = rec[18] * qx + rec[19] * qy;
        qx = tx; qy = ty;
    }
    double fx, fy;
    lens_core(model, k, qx, qy, fx, fy);
    double x = qx - (fx - qx), y = qy - (fy - qy);
    for (int i = 0; i < 20; ++i) {
        lens_core(model, k, x, y, fx, fy);
        x = x + qx - fx; y = y + qy - fy;
        const double dx = fx - qx, dy = fy - qy;
        if (sqrt(dx * dx + dy * dy) < 1e-6) break;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        lens_core(model, k, x, y, fx, fy);
        x = x + qx - fx; y = y + qy - fy;
    }
    {
        const double tx = rec[20] * x + rec[21] * y, ty = rec[22] * x + rec[23] * y;
        x = tx; y = ty;
    }
    double xc = x * radius, yc = y * radius;
    xc += (w / 2) + lx; yc += (h / 2) + ly;
    const double ux = xc / w - 0.5, uy = yc / h - 0.5;
    if (isfinite(ux)) outx = ux;
    if (isfinite(uy)) outy = uy;
}

// the same without the non-finite guard: a layer inside a lens chain (the guard of
// adjust_measureErrors.cpp belongs to the solver's call site, not to the lens model)
MMB_HD void lens_distort_raw(const double *rec, double px, double py, double &outx, double &outy) {
    const int model = (int)rec[0];
    outx = px; outy = py;
    if (model == 0) return;
    const double radius = rec[1], w = rec[2], h = rec[3], lx = rec[4], ly = rec[5];
    const double *k = &rec[6];
    double qx = (((px + 0.5) - 0.5) * w - lx) / radius;
    double qy = (((py + 0.5) - 0.5) * h - ly) / radius;
    {
        const double tx = rec[16] * qx + rec[17] * qy, ty = rec[18] * qx + rec[19] * qy;
        qx = tx; qy = ty;
    }
    double fx, fy;
    lens_core(model, k, qx, qy, fx, fy);
    double x = qx - (fx - qx), y = qy - (fy - qy);
    for (int i = 0; i < 20; ++i) {
        lens_core(model, k, x, y, fx, fy);
        x = x + qx - fx; y = y + qy - fy;
        const double dx = fx - qx, dy = fy - qy;
        if (sqrt(dx * dx + dy * dy) < 1e-6) break;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        lens_core(model, k, x, y, fx, fy);
        x = x + qx - fx; y = y + qy - fy;
    }
    {
        const double tx = rec[20] * x + rec[21] * y, ty = rec[22] * x + rec[23] * y;
        x = tx; y = ty;
    }
    double xc = x * radius, yc = y * radius;
    xc += (w / 2) + lx; yc += (h / 2) + ly;
    const double ux = xc / w - 0.5, uy = yc / h - 0.5;
    outx = ux; outy = uy;
}


// adjust_base.cpp:194-220 parameterBoundFromInternalToExternal
MMB_HD double bound_int_to_ext(double value, double xmin, double xmax, double offset, double scale) {
    const double float_max = 3.40282346638528859811704183484516925e+38;
    if ((xmin <= -float_max) && (xmax >= float_max)) {
        value = (value / scale) - offset;
    } else {
        if (xmax >= float_max) value = xmin - (1.0 + sqrt(value * value + 1.0));
        else if (xmin <= -float_max) value = xmax + (1.0 - sqrt(value * value + 1.0));
        else value = xmin + ((xmax - xmin) / 2.0) * (sin(value) + 1.0);
        value = (value / scale) - offset;
    }
    value = fmax(value, xmin);
    value = fmin(value, xmax);
    return value;
}

// adjust_solveFunc.cpp:148-180 (internal value vs external bounds: kept)
MMB_HD double parameter_delta(double value, double delta, double sign, double xmin, double xmax) {
    double new_sign = sign;
    if ((value + delta) > xmax) new_sign = -1.0;
    if ((value - delta) < xmin) new_sign = 1.0;
    return delta * new_sign;
}

}  // namespace mmb
