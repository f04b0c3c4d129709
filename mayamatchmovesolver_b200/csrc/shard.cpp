// Host C++: frame-range sharding of a scene and merging of shard results.
// See shard.h.  No CUDA in this file.
#include "shard.h"

#include <algorithm>
#include <cstring>

namespace mmb {

void shard_frame_range(int F, int world, int rank, int *f0, int *f1) {
    if (world < 1) world = 1;
    rank = std::max(0, std::min(rank, world - 1));
    const int base = F / world, extra = F % world;
    const int b = rank * base + std::min(rank, extra);
    *f0 = b;
    *f1 = b + base + (rank < extra ? 1 : 0);
}

void slice_scene_frames(const mmb_scene_desc *full, int f0, int f1, SceneSlice *out) {
    const int F = full->num_frames, Fl = f1 - f0;
    out->desc = *full;
    out->frames.resize(Fl);
    for (int f = 0; f < Fl; ++f) out->frames[f] = full->frames ? full->frames[f0 + f] : f0 + f;
    out->anim.resize((size_t)full->num_anim * Fl);
    for (int a = 0; a < full->num_anim; ++a)
        std::copy(full->anim_values + (size_t)a * F + f0, full->anim_values + (size_t)a * F + f1,
                  out->anim.begin() + (size_t)a * Fl);
    const size_t M = (size_t)full->num_markers;
    out->mx.resize(M * Fl); out->my.resize(M * Fl); out->mw.resize(M * Fl); out->en.resize(M * Fl);
    for (size_t m = 0; m < M; ++m) {
        const size_t src = m * F + f0, dst = m * Fl;
        std::copy(full->mkr_x + src, full->mkr_x + src + Fl, out->mx.begin() + dst);
        std::copy(full->mkr_y + src, full->mkr_y + src + Fl, out->my.begin() + dst);
        std::copy(full->mkr_weight + src, full->mkr_weight + src + Fl, out->mw.begin() + dst);
        std::copy(full->mkr_enable + src, full->mkr_enable + src + Fl, out->en.begin() + dst);
    }
    out->desc.num_frames = Fl;
    out->desc.frames = out->frames.data();
    out->desc.anim_values = out->anim.data();
    out->desc.mkr_x = out->mx.data();
    out->desc.mkr_y = out->my.data();
    out->desc.mkr_weight = out->mw.data();
    out->desc.mkr_enable = out->en.data();
}

int count_parameters(const mmb_scene_desc *s, int F) {
    int n = 0;
    for (int a = 0; a < s->num_attrs; ++a) n += s->attr_ref[a] < 0 ? F : 1;
    return n;
}

void shard_param_map(const mmb_scene_desc *full, int f0, int f1, std::vector<int32_t> *map) {
    const int F = full->num_frames, Fl = f1 - f0;
    map->clear();
    int gbase = 0;
    for (int a = 0; a < full->num_attrs; ++a) {
        if (full->attr_ref[a] < 0) {  // animated: one parameter per frame, frame-minor
            for (int f = 0; f < Fl; ++f) map->push_back(gbase + f0 + f);
            gbase += F;
        } else {
            map->push_back(gbase);
            gbase += 1;
        }
    }
}

void aggregate_results(const mmb_result *per, int count, mmb_result *out) {
    std::memset(out, 0, sizeof(*out));
    out->success = 1;
    out->error_min = 1.7976931348623157e308;
    out->error_max = -0.0;
    out->error_is_better = 1;
    for (int i = 0; i < count; ++i) {
        const mmb_result &r = per[i];
        out->success = std::min(out->success, r.success);
        out->error_final += r.error_final;
        out->error_avg += r.error_avg;
        out->error_min = std::min(out->error_min, r.error_min);
        out->error_max = std::max(out->error_max, r.error_max);
        out->iteration_num += r.iteration_num;
        out->iteration_function_num += r.iteration_function_num;
        out->iteration_jacobian_num += r.iteration_jacobian_num;
        out->iteration_attempt_num += r.iteration_attempt_num;
        out->reason_num = r.reason_num;  // the last frame's reason survives (SolverResult::add)
        out->user_interrupted = std::max(out->user_interrupted, r.user_interrupted);
        out->error_is_better = std::min(out->error_is_better, r.error_is_better);
        out->timer_solve_s += r.timer_solve_s;
        out->timer_jacobian_s += r.timer_jacobian_s;
        out->timer_function_s += r.timer_function_s;
        out->timer_linear_s += r.timer_linear_s;
    }
    if (count > 1) {  // SolverResult::divide / TimerResult::divide
        const double inv = 1.0 / (double)count;
        out->error_final *= inv;
        out->error_avg *= inv;
        out->timer_solve_s *= inv; out->timer_jacobian_s *= inv;
        out->timer_function_s *= inv; out->timer_linear_s *= inv;
    }
}

}  // namespace mmb
