// Frame-range sharding of a solve across GPUs.
//
// A shard is the same scene restricted to a contiguous range of frames: the
// animated attribute curves and the marker tracks are cut to [f0, f1), every
// static value (bundle positions, focal length, lens coefficients ...) is
// replicated.  A frame's observations touch only that frame's animated
// parameters, so
//   * per-frame solves (frameSolveMode = 1, adjust_base.cpp:1430-1484) of
//     different shards are independent: no collective at all;
//   * a global solve keeps every camera-frame block U_f local to its shard and
//     shares only the static side: one all-reduce of the reduced system per
//     lmpar trial plus a few scalars (lm_global.cu).
#pragma once
#include <cstdint>
#include <vector>

#include "../../include/mmsolver_b200.h"

namespace mmb {

// contiguous, balanced frame ranges: the first (F % world) shards get one frame more
void shard_frame_range(int F, int world, int rank, int *f0, int *f1);

// A scene description cut to frames [f0, f1); owns the cut arrays.
struct SceneSlice {
    mmb_scene_desc desc;
    std::vector<int32_t> frames;
    std::vector<double> anim, mx, my, mw;
    std::vector<uint8_t> en;
};
void slice_scene_frames(const mmb_scene_desc *full, int f0, int f1, SceneSlice *out);

// number of solver parameters of a scene with F frames (adjust_relationships.cpp:223-337)
int count_parameters(const mmb_scene_desc *s, int F);

// global parameter id of every parameter of the shard [f0, f1) of `full`
void shard_param_map(const mmb_scene_desc *full, int f0, int f1, std::vector<int32_t> *map);

// CommandResult::add / divide (adjust_results.h:89-117) over per-frame results
void aggregate_results(const mmb_result *per_frame, int count, mmb_result *out);

}  // namespace mmb
