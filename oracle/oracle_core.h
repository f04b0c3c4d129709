// TEST INFRASTRUCTURE ONLY.
//
// CPU oracle for the mmSolver LM solve hot path: a single-threaded, scalar
// FP64 restatement of the reference's algorithm with the reference's own cost
// structure (one scene evaluation per Jacobian column, dense column-major J,
// dense pivoted-QR lmder).  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may use it; the product
// path never does.
//
// Followed reference code (paths relative to /root/reference):
//   src/mmSolver/adjust/adjust_relationships.cpp:75-337,565-617  index maps
//   src/mmSolver/adjust/adjust_base.cpp:194-295,346-372,713-1287,1297-1488
//   src/mmSolver/adjust/adjust_setParameters.cpp:174-276
//   src/mmSolver/adjust/adjust_measureErrors.cpp:392-562
//   src/mmSolver/adjust/adjust_solveFunc.cpp:148-226,262-525
//   src/mmSolver/adjust/adjust_cminpack_lmder.cpp:64-198
//   lib/rust/mmscenegraph/src/scene/flat.rs:172-358, math/dag.rs:120-327
//
// Parity pinning: see tests/test_oracle_*.py (Rust reprojection KATs,
// test1.py / test_deviation_calculation.py / test_marker_weight.py scenes
// restated as arrays, scipy MINPACK cross-check, oracle/_ref lens library).
#pragma once
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <vector>

#include "../include/mmsolver_b200.h"
#include "oracle_lens.h"
#include "oracle_math.h"
#include "oracle_minpack.h"

using namespace orc;

struct orc_problem {
    // deep copy of the scene
    int F = 0, T = 0, C = 0, B = 0, M = 0, L = 0, A = 0;
    std::vector<int32_t> frames;
    std::vector<double> stat, anim;  // live attribute store (AttrDataBlock)
    std::vector<double> stat0, anim0;
    std::vector<int32_t> tfm_parent, tfm_roo, tfm_attrs;
    std::vector<int32_t> cam_tfm, cam_attrs, cam_fit, cam_rw, cam_rh, cam_lens;
    std::vector<int32_t> bnd_tfm, mkr_cam, mkr_bnd;
    std::vector<double> mkr_x, mkr_y, mkr_w;
    std::vector<uint8_t> mkr_en;
    std::vector<int32_t> lens_model, lens_attrs;
    std::vector<double> lens_cam;
    std::vector<int32_t> attr_ref;
    std::vector<double> attr_min, attr_max, attr_off, attr_scl;
    // StiffAttrs / SmoothAttrs (adjust_data.h): attrIndex + weight / variance / value refs
    struct ExtraAttr { int attrIndex; int32_t weight, variance, value; };
    std::vector<ExtraAttr> stiffAttrs, smoothAttrs;
    int numStiff = 0, numSmooth = 0;  // counted rows (weight > 0, adjust_relationships.cpp:183-219)
    int currentFrame = 0;
    mmb_options opt;
    // index maps
    std::vector<std::pair<int, int>> errorToMarker, paramToAttr;
    std::vector<double> markerWeight;         // normalised
    std::vector<std::vector<bool>> paramFrame; // n x F
    std::vector<std::vector<bool>> markerToAttr; // M x A (structural)
    int n = 0, nobs = 0, m = 0;
    // work
    std::vector<double> previousParam, errorList, errorDist;
    std::vector<double> out_points, out_markers;  // flat.rs out_*_list [M*F*2]
    std::vector<M4> world;                       // [T*F]
    int iterNum = 0, jacIterNum = 0, funcEvalNum = 0;
    std::vector<uint8_t> marker_used_in, attr_used_in;  // per marker / attribute of the caller's lists
    bool faithful_eval = true;  // evaluate whole scene on every call
    bool central_literal = false;  // see func_jacobian
};

static inline double attr_get(const orc_problem *p, int32_t ref, int f) {
    // datablock.rs:76-91 get_attr_value
    if (ref >= 0) return p->stat[ref];
    return p->anim[(size_t)(-ref - 1) * p->F + f];
}
static inline void attr_set(orc_problem *p, int32_t ref, int f, double v) {
    // datablock.rs:93-111: static ignores the frame
    if (ref >= 0) p->stat[ref] = v;
    else p->anim[(size_t)(-ref - 1) * p->F + f] = v;
}

// adjust_base.cpp:194-220
static double boundInternalToExternal(double value, double xmin, double xmax,
                                      double offset, double scale) {
    const double float_max = std::numeric_limits<float>::max();
    if ((xmin <= -float_max) && (xmax >= float_max)) {
        value = (value / scale) - offset;
        value = std::max<double>(value, xmin);
        value = std::min<double>(value, xmax);
        return value;
    } else if (xmax >= float_max) {
        value = xmin - (1.0 + std::sqrt(value * value + 1.0));
    } else if (xmin <= -float_max) {
        value = xmax + (1.0 - std::sqrt(value * value + 1.0));
    } else {
        value = xmin + ((xmax - xmin) / 2.0) * (std::sin(value) + 1.0);
    }
    value = (value / scale) - offset;
    value = std::max<double>(value, xmin);
    value = std::min<double>(value, xmax);
    return value;
}

// adjust_base.cpp:225-258 (quirks kept: `xmin <= float_max` test at :239)
static double boundExternalToInternal(double value, double xmin, double xmax,
                                      double offset, double scale) {
    value = std::max<double>(value, xmin);
    value = std::min<double>(value, xmax);
    value = (value * scale) + offset;
    xmin = (xmin * scale) + offset;
    xmax = (xmax * scale) + offset;
    const double float_max = std::numeric_limits<float>::max();
    if ((xmin <= float_max) && (xmax >= float_max)) {
        return value;
    } else if (xmax >= float_max) {
        value = std::sqrt(std::pow(((value - xmin) + 1.0), 2.0) - 1.0);
    } else if (xmin <= -float_max) {
        value = std::sqrt(std::pow((xmax - value) + 1.0, 2.0) - 1.0);
    } else {
        value = std::asin((2.0 * (value - xmin) / (xmax - xmin)) - 1.0);
    }
    return value;
}

// Which attr-store refs feed marker m?  Structural restatement of the
// marker->attr "affects" relationship (adjust_relationships.cpp:369-466 asks
// Maya's DG; here the hierarchy is explicit).
static bool marker_uses_ref(const orc_problem *p, int mk, int32_t ref) {
    auto chain_has = [&](int t) {
        while (t >= 0) {
            for (int k = 0; k < 9; ++k) if (p->tfm_attrs[t * 9 + k] == ref) return true;
            t = p->tfm_parent[t];
        }
        return false;
    };
    const int c = p->mkr_cam[mk], b = p->mkr_bnd[mk];
    if (chain_has(p->cam_tfm[c]) || chain_has(p->bnd_tfm[b])) return true;
    for (int k = 0; k < 8; ++k) if (p->cam_attrs[c * 8 + k] == ref) return true;
    const int l = p->cam_lens[c];
    if (l >= 0)
        for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k)
            if (p->lens_attrs[l * MMB_LENS_MAX_COEFFS + k] == ref) return true;
    return false;
}

// solve_v1 (adjust_base.cpp:1322-1395): markerToAttrList, then
// splitUsedMarkersAndAttributes (adjust_base.cpp:574-610, count maps as there) and the
// removal of the unused objects when removeUnusedMarkers / removeUnusedAttributes are set.
static void split_used_objects(orc_problem *p) {
    const int M = p->M, A = p->A, F = p->F;
    p->marker_used_in.assign(M, 1); p->attr_used_in.assign(A, 1);
    if (!p->opt.remove_unused_markers && !p->opt.remove_unused_attributes) return;
    std::map<size_t, int> markerIndexUsedCount, attrIndexUsedCount;
    for (int markerIndex = 0; markerIndex < M; ++markerIndex)
        for (int attrIndex = 0; attrIndex < A; ++attrIndex)
            if (marker_uses_ref(p, markerIndex, p->attr_ref[attrIndex])) {
                markerIndexUsedCount[markerIndex] += 1;
                attrIndexUsedCount[attrIndex] += 1;
            }
    if (p->opt.remove_unused_markers) {
        int k = 0;
        for (int i = 0; i < M; ++i) {
            const bool used = markerIndexUsedCount.count(i) && markerIndexUsedCount[i] > 0;
            p->marker_used_in[i] = used;
            if (!used) continue;
            p->mkr_cam[k] = p->mkr_cam[i]; p->mkr_bnd[k] = p->mkr_bnd[i];
            for (int f = 0; f < F; ++f) {
                p->mkr_x[(size_t)k * F + f] = p->mkr_x[(size_t)i * F + f];
                p->mkr_y[(size_t)k * F + f] = p->mkr_y[(size_t)i * F + f];
                p->mkr_w[(size_t)k * F + f] = p->mkr_w[(size_t)i * F + f];
                p->mkr_en[(size_t)k * F + f] = p->mkr_en[(size_t)i * F + f];
            }
            ++k;
        }
        p->M = k;
    }
    if (p->opt.remove_unused_attributes) {
        int k = 0;
        for (int a = 0; a < A; ++a) {
            const bool used = attrIndexUsedCount.count(a) && attrIndexUsedCount[a] > 0;
            p->attr_used_in[a] = used;
            if (!used) continue;
            p->attr_ref[k] = p->attr_ref[a]; p->attr_min[k] = p->attr_min[a]; p->attr_max[k] = p->attr_max[a];
            p->attr_off[k] = p->attr_off[a]; p->attr_scl[k] = p->attr_scl[a];
            ++k;
        }
        p->A = k;
    }
}

static void build_index_maps(orc_problem *p) {
    const int F = p->F;
    // countUpNumberOfErrors, adjust_relationships.cpp:75-221
    p->errorToMarker.clear(); p->markerWeight.clear();
    std::map<int, double> weightMaxPerFrame;
    for (int i = 0; i < p->M; ++i) {
        for (int j = 0; j < F; ++j) {
            const bool enable = p->mkr_en[(size_t)i * F + j] != 0;
            const double weight = p->mkr_w[(size_t)i * F + j];
            if (enable && (weight > 0.0)) {
                p->errorToMarker.push_back({i, j});
                p->markerWeight.push_back(weight);
                auto it = weightMaxPerFrame.find(j);
                double wmax = weight;
                if (it != weightMaxPerFrame.end()) { wmax = it->second; if (weight > wmax) wmax = weight; }
                weightMaxPerFrame[j] = wmax;
            }
        }
    }
    for (size_t i = 0; i < p->errorToMarker.size(); ++i)
        p->markerWeight[i] = p->markerWeight[i] / weightMaxPerFrame[p->errorToMarker[i].second];
    p->nobs = (int)p->errorToMarker.size();
    // stiffness / smoothness rows: one per entry whose weight is > 0 (adjust_relationships.cpp:183-219;
    // getValue without a time = Maya's current time)
    p->numStiff = p->numSmooth = 0;
    for (const auto &e : p->stiffAttrs) if (attr_get(p, e.weight, p->currentFrame) > 0.0) p->numStiff++;
    for (const auto &e : p->smoothAttrs) if (attr_get(p, e.weight, p->currentFrame) > 0.0) p->numSmooth++;
    p->m = 2 * p->nobs + p->numStiff + p->numSmooth;
    // countUpNumberOfUnknownParameters, adjust_relationships.cpp:223-337
    p->paramToAttr.clear(); p->paramFrame.clear();
    for (int i = 0; i < p->A; ++i) {
        if (p->attr_ref[i] < 0) {
            for (int j = 0; j < F; ++j) {
                p->paramToAttr.push_back({i, j});
                std::vector<bool> fr(F, false); fr[j] = true;
                p->paramFrame.push_back(fr);
            }
        } else {
            p->paramToAttr.push_back({i, -1});
            p->paramFrame.push_back(std::vector<bool>(F, true));
        }
    }
    p->n = (int)p->paramToAttr.size();
    p->markerToAttr.assign(p->M, std::vector<bool>(p->A, false));
    for (int i = 0; i < p->M; ++i)
        for (int j = 0; j < p->A; ++j) p->markerToAttr[i][j] = marker_uses_ref(p, i, p->attr_ref[j]);
    p->previousParam.assign(p->n, 0.0);
    p->errorList.assign(p->m, 0.0);
    p->errorDist.assign(p->nobs, 0.0);
}

// flat.rs:172-358 FlatScene::evaluate (all transforms x all frames)
static void evaluate_scene(orc_problem *p, const std::vector<bool> *frameEnable) {
    const int F = p->F;
    p->world.resize((size_t)p->T * F);
    const bool all = p->faithful_eval || frameEnable == nullptr;
    for (int i = 0; i < p->T; ++i) {
        const int32_t *a = &p->tfm_attrs[i * 9];
        for (int f = 0; f < F; ++f) {
            if (!all && !(*frameEnable)[f]) continue;
            M4 local = local_matrix(attr_get(p, a[0], f), attr_get(p, a[1], f), attr_get(p, a[2], f),
                                    attr_get(p, a[3], f), attr_get(p, a[4], f), attr_get(p, a[5], f),
                                    attr_get(p, a[6], f), attr_get(p, a[7], f), attr_get(p, a[8], f),
                                    p->tfm_roo[i]);
            const int par = p->tfm_parent[i];
            p->world[(size_t)i * F + f] = par >= 0 ? m4_mul(p->world[(size_t)par * F + f], local) : local;
        }
    }
    p->out_points.resize((size_t)p->M * F * 2);
    p->out_markers.resize((size_t)p->M * F * 2);
    for (int mk = 0; mk < p->M; ++mk) {
        const int c = p->mkr_cam[mk], b = p->mkr_bnd[mk];
        const int32_t *ca = &p->cam_attrs[c * 8];
        for (int f = 0; f < F; ++f) {
            if (!all && !(*frameEnable)[f]) continue;
            // dag.rs:120-172 compute_projection_matrix_with_attrs
            const double focal = attr_get(p, ca[2], f);
            const double sx_mm = attr_get(p, ca[0], f), sy_mm = attr_get(p, ca[1], f);
            const double sx_in = sx_mm * MM_TO_INCH, sy_in = sy_mm * MM_TO_INCH;
            const double ox_in = attr_get(p, ca[3], f) * MM_TO_INCH;
            const double oy_in = attr_get(p, ca[4], f) * MM_TO_INCH;
            const double near_clip = 0.1;  // dag.rs:152
            const double far_clip = attr_get(p, ca[6], f);
            const double cam_scale = attr_get(p, ca[7], f);
            M4 proj = projection_matrix(focal, sx_in, sy_in, ox_in, oy_in, (double)p->cam_rw[c],
                                        (double)p->cam_rh[c], p->cam_fit[c], near_clip, far_clip, cam_scale);
            double px, py;
            reproject_normalised(p->world[(size_t)p->cam_tfm[c] * F + f], proj,
                                 p->world[(size_t)p->bnd_tfm[b] * F + f], px, py);
            const size_t o = ((size_t)mk * F + f) * 2;
            p->out_points[o] = px; p->out_points[o + 1] = py;
            // flat.rs:327-348
            double mx = p->mkr_x[(size_t)mk * F + f], my = p->mkr_y[(size_t)mk * F + f];
            scale_marker_film_fit(p->cam_fit[c], sx_mm / sy_mm,
                                  (double)p->cam_rw[c] / (double)p->cam_rh[c], mx, my);
            p->out_markers[o] = mx; p->out_markers[o + 1] = my;
        }
    }
}

// ---- Maya-DAG-convention arithmetic (sceneGraphMode = 1) -----------------------------------
// maya_camera.cpp:75-414 in Maya's ROW-vector convention: MMatrix m[row][col], point * matrix.
// getProjectionMatrix :363-414 = computeFrustumCoordinates :75-98 + applyFilmFitLogic :100-205
// (Horizontal fit scales Y here, X in camera.rs:200) + computeProjectionMatrix :327-361.
static M4 maya_projection_matrix(double focal_mm, double fbw_inch, double fbh_inch, double offx_inch,
                                 double offy_inch, double img_w, double img_h, int film_fit, double near_cm,
                                 double far_cm, double camera_scale) {
    const double filmAspectRatio = fbw_inch / fbh_inch;
    const double imageAspectRatio = img_w / img_h;
    const double filmWidth_mm = fbw_inch * INCH_TO_MM, filmHeight_mm = fbh_inch * INCH_TO_MM;
    const double offsetX_mm = offx_inch * INCH_TO_MM, offsetY_mm = offy_inch * INCH_TO_MM;
    const double focal_to_near = (near_cm / focal_mm) * camera_scale;
    const double right = focal_to_near * (0.5 * filmWidth_mm + offsetX_mm);
    const double left = focal_to_near * (-0.5 * filmWidth_mm + offsetX_mm);
    const double top = focal_to_near * (0.5 * filmHeight_mm + offsetY_mm);
    const double bottom = focal_to_near * (-0.5 * filmHeight_mm + offsetY_mm);
    double filmFitScaleX = 1.0, filmFitScaleY = 1.0, screenSizeX = 0.0, screenSizeY = 0.0;
    double screenLeft = left, screenRight = right, screenTop = top, screenBottom = bottom;
    switch (film_fit) {
        default:
        case 1:
            filmFitScaleY = imageAspectRatio / filmAspectRatio;
            screenSizeX = right - left;
            screenSizeY = screenSizeX / imageAspectRatio;
            break;
        case 2:
            filmFitScaleX = 1.0 / (imageAspectRatio / filmAspectRatio);
            screenSizeY = top - bottom;
            screenSizeX = screenSizeY * imageAspectRatio;
            break;
        case 0:
            if (filmAspectRatio > imageAspectRatio) {
                filmFitScaleX = filmAspectRatio / imageAspectRatio;
                screenSizeY = top - bottom;
                screenSizeX = screenSizeY * imageAspectRatio;
            } else {
                filmFitScaleY = imageAspectRatio / filmAspectRatio;
                screenSizeX = right - left;
                screenSizeY = (screenSizeX * (filmAspectRatio / imageAspectRatio)) / filmAspectRatio;
            }
            break;
        case 3:
            if (filmAspectRatio > imageAspectRatio) {
                filmFitScaleY = imageAspectRatio / filmAspectRatio;
                screenSizeX = right - left;
                screenSizeY = (right - left) / imageAspectRatio;
            } else {
                filmFitScaleX = filmAspectRatio / imageAspectRatio;
                screenSizeX = (right - left) * (imageAspectRatio / filmAspectRatio);
                screenSizeY = top - bottom;
            }
            break;
    }
    screenRight *= filmFitScaleX; screenLeft *= filmFitScaleX;
    screenTop *= filmFitScaleY; screenBottom *= filmFitScaleY;
    M4 pm; std::memset(pm.m, 0, sizeof(pm.m));   // pm.m[r * 4 + c] = projectionMatrix[r][c]
    pm.m[0 * 4 + 0] = 1.0 / (screenSizeX * 0.5) * MM_TO_CM;
    pm.m[1 * 4 + 1] = 1.0 / (screenSizeY * 0.5) * MM_TO_CM;
    pm.m[2 * 4 + 0] = (screenRight + screenLeft) / (screenRight - screenLeft) * filmFitScaleX;
    pm.m[2 * 4 + 1] = (screenTop + screenBottom) / (screenTop - screenBottom) * filmFitScaleY;
    pm.m[2 * 4 + 2] = (far_cm + near_cm) / (far_cm - near_cm);
    pm.m[2 * 4 + 3] = -1;
    pm.m[3 * 4 + 2] = 2.0 * far_cm * near_cm / (far_cm - near_cm);
    return pm;
}
static M4 m4_transpose(const M4 &a) {
    M4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.m[i * 4 + j] = a.m[j * 4 + i];
    return r;
}
inline double gaussian(double x, double mean, double sigma) {  // adjust_measureErrors.cpp:106-109
    return std::exp(-(std::pow((x - mean), 2.0) / (2.0 * (std::pow(sigma, 2.0)))));
}

static LensEval lens_at(const orc_problem *p, int l, int f);

// adjust_measureErrors.cpp:116-387 measureErrors_mayaDag.  Maya's DG is replaced by the flat
// scene (world = T * R * S chains); everything after the world matrices is as written there.
static void measureErrors_mayaDag(orc_problem *p, const std::vector<bool> &frameIndexEnable,
                                  const std::vector<bool> &errorMeasurements, double *errors,
                                  double &error_avg, double &error_max, double &error_min) {
    const int F = p->F;
    const double imageWidth = p->opt.image_width;
    // world matrices of every transform (column convention), all frames
    p->world.resize((size_t)p->T * F);
    for (int i = 0; i < p->T; ++i) {
        const int32_t *a = &p->tfm_attrs[i * 9];
        for (int f = 0; f < F; ++f) {
            M4 local = local_matrix(attr_get(p, a[0], f), attr_get(p, a[1], f), attr_get(p, a[2], f),
                                    attr_get(p, a[3], f), attr_get(p, a[4], f), attr_get(p, a[5], f),
                                    attr_get(p, a[6], f), attr_get(p, a[7], f), attr_get(p, a[8], f), p->tfm_roo[i]);
            const int par = p->tfm_parent[i];
            p->world[(size_t)i * F + f] = par >= 0 ? m4_mul(p->world[(size_t)par * F + f], local) : local;
        }
    }
    int numberOfErrorsMeasured = 0;
    for (int i = 0; i < p->nobs; ++i) {
        const int markerIndex = p->errorToMarker[i].first, frameIndex = p->errorToMarker[i].second;
        if (!frameIndexEnable[frameIndex]) continue;
        if (!errorMeasurements[i]) continue;
        const int c = p->mkr_cam[markerIndex], b = p->mkr_bnd[markerIndex];
        const int32_t *ca = &p->cam_attrs[c * 8];
        // Camera::getWorldProjMatrix maya_camera.cpp:863-894: inverse(world) * proj (row vectors)
        const M4 worldMat = m4_transpose(p->world[(size_t)p->cam_tfm[c] * F + frameIndex]);  // Maya layout
        M4 worldInv;
        if (!m4_inverse(worldMat, worldInv)) worldInv = m4_identity();
        // Camera::getProjMatrix :741-812: film back / offsets / focal at the frame, camera scale and far
        // clip plane without a time (the current time), near clip hard-coded 0.1
        const double filmWidth = attr_get(p, ca[0], frameIndex) * MM_TO_INCH, filmHeight = attr_get(p, ca[1], frameIndex) * MM_TO_INCH;
        const double filmOffsetX = attr_get(p, ca[3], frameIndex) * MM_TO_INCH, filmOffsetY = attr_get(p, ca[4], frameIndex) * MM_TO_INCH;
        const double focal = attr_get(p, ca[2], frameIndex);
        const double cameraScale = attr_get(p, ca[7], p->currentFrame), farClip = attr_get(p, ca[6], p->currentFrame);
        const M4 projMat = maya_projection_matrix(focal, filmWidth, filmHeight, filmOffsetX, filmOffsetY, (double)p->cam_rw[c],
                                                  (double)p->cam_rh[c], p->cam_fit[c], 0.1, farClip, cameraScale);
        const M4 cameraWorldProjectionMatrix = m4_mul(worldInv, projMat);
        // getWorldPosition :814-833, getForwardDirection :840-856
        const double cam_pos[3] = {worldMat.m[3 * 4 + 0] / worldMat.m[3 * 4 + 3], worldMat.m[3 * 4 + 1] / worldMat.m[3 * 4 + 3],
                                   worldMat.m[3 * 4 + 2] / worldMat.m[3 * 4 + 3]};
        double cam_dir[3] = {-worldMat.m[2 * 4 + 0], -worldMat.m[2 * 4 + 1], -worldMat.m[2 * 4 + 2]};  // (0, 0, -1) * worldMat
        { const double l = std::sqrt(cam_dir[0] * cam_dir[0] + cam_dir[1] * cam_dir[1] + cam_dir[2] * cam_dir[2]);
          for (double &v : cam_dir) v /= l; }
        double mkr_x = p->mkr_x[(size_t)markerIndex * F + frameIndex], mkr_y = p->mkr_y[(size_t)markerIndex * F + frameIndex];
        const double filmBackAspect = filmWidth / filmHeight;
        const double renderAspect = (double)p->cam_rw[c] / (double)p->cam_rh[c];
        // applyFilmFitCorrectionScaleBackward maya_camera.cpp:207-325 (same cases as flat.rs:73-97)
        scale_marker_film_fit(p->cam_fit[c], filmBackAspect, renderAspect, mkr_x, mkr_y);
        const double mkr_weight = std::sqrt(p->markerWeight[i]);
        const M4 &bw = p->world[(size_t)p->bnd_tfm[b] * F + frameIndex];
        const double bnd[3] = {bw.m[3], bw.m[7], bw.m[11]};
        double bnd_dir[3] = {bnd[0] - cam_pos[0], bnd[1] - cam_pos[1], bnd[2] - cam_pos[2]};
        { const double l = std::sqrt(bnd_dir[0] * bnd_dir[0] + bnd_dir[1] * bnd_dir[1] + bnd_dir[2] * bnd_dir[2]);
          for (double &v : bnd_dir) v /= l; }
        // bnd_mpos * cameraWorldProjectionMatrix; cartesianize
        double q[4];
        for (int col = 0; col < 4; ++col)
            q[col] = bnd[0] * cameraWorldProjectionMatrix.m[0 * 4 + col] + bnd[1] * cameraWorldProjectionMatrix.m[1 * 4 + col] +
                     bnd[2] * cameraWorldProjectionMatrix.m[2 * 4 + col] + 1.0 * cameraWorldProjectionMatrix.m[3 * 4 + col];
        double point_x = (q[0] / q[3]) * 0.5, point_y = (q[1] / q[3]) * 0.5;
        int lm = markerIndex, lf = frameIndex;
        if (p->opt.lens_index_literal) {
            const int idx = markerIndex + frameIndex;
            lm = idx / F; lf = idx % F;
            if (lm >= p->M) lm = -1;
        }
        const int l = lm >= 0 ? p->cam_lens[p->mkr_cam[lm]] : -1;
        if (l >= 0 && p->lens_model[l] != 0) {
            LensEval L = lens_at(p, l, lf);
            double out_x = point_x, out_y = point_y;
            lens_distort(L, point_x, point_y, out_x, out_y);
            if (std::isfinite(out_x)) point_x = out_x;
            if (std::isfinite(out_y)) point_y = out_y;
        }
        // behind the camera?  adjust_measureErrors.cpp:262-270
        double behind_camera_error_factor = 1.0;
        const double cam_dot_bnd = cam_dir[0] * bnd_dir[0] + cam_dir[1] * bnd_dir[1] + cam_dir[2] * bnd_dir[2];
        if (cam_dot_bnd < 0.0) behind_camera_error_factor = 1e+6;
        const double dx = std::fabs(mkr_x - point_x) * imageWidth;
        const double dy = std::fabs(mkr_y - point_y) * imageWidth;
        errors[2 * i] = dx * mkr_weight * behind_camera_error_factor;
        errors[2 * i + 1] = dy * mkr_weight * behind_camera_error_factor;
        p->errorList[2 * i] = dx * behind_camera_error_factor;
        p->errorList[2 * i + 1] = dy * behind_camera_error_factor;
        const double ddx = mkr_x - point_x, ddy = mkr_y - point_y;
        const double d = std::sqrt((ddx * ddx) + (ddy * ddy)) * imageWidth;
        p->errorDist[i] = d;
        error_avg += d;
        if (d > error_max) error_max = d;
        if (d < error_min) error_min = d;
        ++numberOfErrorsMeasured;
    }
    if (numberOfErrorsMeasured == 0) { error_max = 0.0; error_min = 0.0; error_avg = 0.0; }
    else error_avg *= 1.0 / numberOfErrorsMeasured;
    // stiffness, then smoothness rows (:311-387): the first numStiff / numSmooth entries of the lists
    const int numberOfMarkerErrors = 2 * p->nobs;
    for (int pass = 0; pass < 2; ++pass) {
        const auto &list = pass == 0 ? p->stiffAttrs : p->smoothAttrs;
        const int count = pass == 0 ? p->numStiff : p->numSmooth;
        const int offset = pass == 0 ? 0 : p->numStiff;
        for (int i = 0; i < count; ++i) {
            const int indexIntoErrorArray = numberOfMarkerErrors + offset + i;
            const auto &e = list[i];
            const double weight = attr_get(p, e.weight, p->currentFrame);
            const double variance = attr_get(p, e.variance, p->currentFrame);
            const double value = attr_get(p, e.value, p->currentFrame);
            const double attrValue = attr_get(p, p->attr_ref[e.attrIndex], p->currentFrame);
            const double error = ((1.0 / gaussian(attrValue, value, variance)) - 1.0);
            p->errorList[indexIntoErrorArray] = error * weight;
            errors[indexIntoErrorArray] = error * weight;
        }
    }
}

static LensEval lens_at(const orc_problem *p, int l, int f) {
    LensEval L; L.model = p->lens_model[l];
    for (int k = 0; k < 6; ++k) L.cam[k] = p->lens_cam[l * 6 + k];
    for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k) {
        const int32_t r = p->lens_attrs[l * MMB_LENS_MAX_COEFFS + k];
        L.c[k] = (r == INT32_MIN) ? 0.0 : attr_get(p, r, f);
    }
    lens_prepare(L);
    return L;
}

// adjust_setParameters.cpp:174-276.  Lens coefficients live in the same attr
// store here, so the lens branch (:201-222) and the AttrDataBlock branch
// (:225-233) become one write (intended [attr*F + f] indexing).
static void setParameters(orc_problem *p, const double *x) {
    for (int i = 0; i < p->n; ++i) {
        const int ai = p->paramToAttr[i].first, fi = p->paramToAttr[i].second;
        const double real = boundInternalToExternal(x[i], p->attr_min[ai], p->attr_max[ai],
                                                    p->attr_off[ai], p->attr_scl[ai]);
        attr_set(p, p->attr_ref[ai], fi < 0 ? 0 : fi, real);
    }
    for (int j = 0; j < p->n; ++j) p->previousParam[j] = x[j];
}

// adjust_measureErrors.cpp:392-562 (mmSceneGraph variant)
static void measureErrors(orc_problem *p, const std::vector<bool> &frameEnable,
                          const std::vector<bool> &errorMeasurements, double *errors,
                          double &eavg, double &emax, double &emin) {
    eavg = 0.0; emax = -0.0; emin = std::numeric_limits<double>::max();
    if (p->opt.scene_graph_mode == 1) {   // measureErrors :539-544
        measureErrors_mayaDag(p, frameEnable, errorMeasurements, errors, eavg, emax, emin);
        return;
    }
    evaluate_scene(p, &frameEnable);
    const double imageWidth = p->opt.image_width;
    const int F = p->F;
    int measured = 0;
    for (int i = 0; i < p->nobs; ++i) {
        const int markerIndex = p->errorToMarker[i].first, frameIndex = p->errorToMarker[i].second;
        if (!frameEnable[frameIndex]) continue;
        if (!errorMeasurements[i]) continue;
        const double w = std::sqrt(p->markerWeight[i]);
        const size_t o = ((size_t)markerIndex * F + frameIndex) * 2;
        const double mkr_x = p->out_markers[o], mkr_y = p->out_markers[o + 1];
        double point_x = p->out_points[o], point_y = p->out_points[o + 1];
        // lens: adjust_measureErrors.cpp:461-479
        int lm = markerIndex, lf = frameIndex;
        if (p->opt.lens_index_literal) {  // literal [markerIndex + frameIndex] read
            const int idx = markerIndex + frameIndex;
            lm = idx / F; lf = idx % F;
            if (lm >= p->M) lm = -1;
        }
        const int l = lm >= 0 ? p->cam_lens[p->mkr_cam[lm]] : -1;
        if (l >= 0 && p->lens_model[l] != 0) {
            LensEval L = lens_at(p, l, lf);
            double ox = point_x, oy = point_y;
            lens_distort(L, point_x, point_y, ox, oy);
            if (std::isfinite(ox)) point_x = ox;
            if (std::isfinite(oy)) point_y = oy;
        }
        const double dx = std::fabs(mkr_x - point_x), dy = std::fabs(mkr_y - point_y);
        const double dxp = dx * imageWidth, dyp = dy * imageWidth;
        errors[2 * i] = dxp * w * 1.0;
        errors[2 * i + 1] = dyp * w * 1.0;
        p->errorList[2 * i] = dxp; p->errorList[2 * i + 1] = dyp;
        const double d = std::sqrt((dx * dx) + (dy * dy)) * imageWidth;
        p->errorDist[i] = d;
        eavg += d; if (d > emax) emax = d; if (d < emin) emin = d;
        ++measured;
    }
    if (measured == 0) { emax = emin = eavg = 0.0; } else { eavg *= 1.0 / measured; }
}
