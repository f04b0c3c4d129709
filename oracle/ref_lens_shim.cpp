// TEST INFRASTRUCTURE ONLY.
//
// extern "C" shim over the REFERENCE's own mmlens C++ lens models, compiled
// from the sources where they lie under /root/reference (see oracle/Makefile;
// output: oracle/_ref/libmmlens_ref.so, git-ignored).  Used by tests/ to pin
// oracle_lens.h (and through it the CUDA lens code) against the reference.
// This file is ours; no reference source is copied into the repo.
//
//   lib/cppbind/mmlens/include/mmlens/lens_model*.h
//   lib/cppbind/mmlens/src/lens_model_3de_*.cpp
#include <mmlens/lens_model.h>
#include <mmlens/lens_model_3de_anamorphic_deg_4_rotate_squeeze_xy.h>
#include <mmlens/lens_model_3de_anamorphic_deg_4_rotate_squeeze_xy_rescaled.h>
#include <mmlens/lens_model_3de_classic.h>
#include <mmlens/lens_model_3de_radial_decentered_deg_4_cylindric.h>

#include <mmlens/lens_model_passthrough.h>

#include <memory>
#include <vector>

using namespace mmlens;

static std::unique_ptr<LensModel> make_lens(int model, const double *c) {
    std::unique_ptr<LensModel> L;
    switch (model) {
        case 1:
            L.reset(new LensModel3deClassic(c[0], c[1], c[2], c[3], c[4]));
            break;
        case 2:
            L.reset(new LensModel3deRadialDecenteredDeg4Cylindric(
                c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7]));
            break;
        case 3:
            L.reset(new LensModel3deAnamorphicDeg4RotateSqueezeXY(
                c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7], c[8], c[9],
                c[10], c[11], c[12]));
            break;
        case 4:
            L.reset(new LensModel3deAnamorphicDeg4RotateSqueezeXYRescaled(
                c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7], c[8], c[9],
                c[10], c[11], c[12], c[13]));
            break;
        default:
            break;
    }
    return L;
}

extern "C" int ref_lens_apply(int model, const double *coeffs,
                              const double *cam6, int direction, int npts,
                              const double *xy_in, double *xy_out) {
    std::unique_ptr<LensModel> L = make_lens(model, coeffs);
    if (!L) {
        for (int i = 0; i < 2 * npts; ++i) xy_out[i] = xy_in[i];
        return model == 0 ? 0 : -1;
    }
    L->setFocalLength(cam6[0]);
    L->setFilmBackWidth(cam6[1]);
    L->setFilmBackHeight(cam6[2]);
    L->setPixelAspect(cam6[3]);
    L->setLensCenterOffsetX(cam6[4]);
    L->setLensCenterOffsetY(cam6[5]);
    for (int i = 0; i < npts; ++i) {
        double x = xy_in[2 * i], y = xy_in[2 * i + 1], ox = x, oy = y;
        if (direction == 0) L->applyModelUndistort(x, y, ox, oy);
        else L->applyModelDistort(x, y, ox, oy);
        xy_out[2 * i] = ox; xy_out[2 * i + 1] = oy;
    }
    return 0;
}

// Layered lenses: layer 0 is the innermost (it has no input), layer k's inputLensModel is
// layer k-1; the call goes to the last layer, as a Maya lens-node chain evaluates
// (src/mmSolver/node/MMLensModel3deNode.cpp:164: newLensModel->setInputLensModel(input)).
extern "C" int ref_lens_apply_layered(int n_layers, const int *models, const double *coeffs14,
                                      const double *cam6, int direction, int npts,
                                      const double *xy_in, double *xy_out) {
    std::vector<std::shared_ptr<LensModel>> chain;
    for (int k = 0; k < n_layers; ++k) {
        std::shared_ptr<LensModel> L;
        if (models[k] == 0) L.reset(new LensModelPassthrough());
        else L = std::shared_ptr<LensModel>(make_lens(models[k], coeffs14 + 14 * k).release());
        if (!L) return -1;
        const double *c = cam6 + 6 * k;
        L->setFocalLength(c[0]);
        L->setFilmBackWidth(c[1]);
        L->setFilmBackHeight(c[2]);
        L->setPixelAspect(c[3]);
        L->setLensCenterOffsetX(c[4]);
        L->setLensCenterOffsetY(c[5]);
        if (k > 0) L->setInputLensModel(chain[k - 1]);
        chain.push_back(L);
    }
    if (chain.empty()) return -1;
    LensModel *top = chain.back().get();
    for (int i = 0; i < npts; ++i) {
        double x = xy_in[2 * i], y = xy_in[2 * i + 1], ox = x, oy = y;
        if (direction == 0) top->applyModelUndistort(x, y, ox, oy);
        else top->applyModelDistort(x, y, ox, oy);
        xy_out[2 * i] = ox; xy_out[2 * i + 1] = oy;
    }
    return 0;
}
