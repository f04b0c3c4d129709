// Host C++: the `key=value` strings the mmSolver command returns
// (CommandResult::appendToMStringArray, adjust_results.h:871-938), so that
// python/mmSolver/_api/solveresult.py consumes this library's results unchanged.
//
// Order (kept "for backwards compatibility", adjust_results.h:909):
//   [numberOf* counts when printStats.input]          SolverObjectCountResult  :838-860
//   success= ... user_interrupted=                      SolverResult             :123-160
//   timer_* / ticks_*                                   TimerResult              :244-286
//   error_per_marker_per_frame=name#frame#dev, error_per_frame=frame#dev
//                                                       ErrorMetricsResult       :526-565
//   solve_parameter_list=v#v#..., solve_error_list=...  SolveValuesResult        :357-375
// Numbers print as mmstring::numberToString does: operator<< of a default
// std::stringstream (6 significant digits), lists end with the split char '#'
// (adjust_defines.h:64).
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/mmsolver_b200.h"

namespace {

template <typename T>
std::string num(T v) {
    std::stringstream ss;
    ss << v;
    return ss.str();
}

const char kSplit[] = "#";

}  // namespace

extern "C" mmb_status mmb_format_results(const mmb_result *r, int32_t n_params, int32_t n_obs,
                                         const int32_t *error_to_marker, const int32_t *frames, int32_t num_frames,
                                         const char *const *marker_names, int32_t num_markers,
                                         const double *solve_parameter_list, const double *solve_error_list,
                                         const double *error_dist, int32_t flags, char *buf, int64_t cap,
                                         int64_t *needed) {
    if (!r || n_params < 0 || n_obs < 0 || num_frames < 0 || num_markers < 0) return MMB_ERR_INVALID_ARGUMENT;
    const bool want_metrics = (flags & (MMB_FMT_ERROR_PER_MARKER_PER_FRAME | MMB_FMT_ERROR_PER_FRAME)) != 0;
    const int n_stiff = (flags >> 8) & 0x7ff, n_smooth = (flags >> 19) & 0x7ff;
    if (want_metrics && (!error_to_marker || !error_dist)) return MMB_ERR_INVALID_ARGUMENT;
    try {
        std::string out;
        auto line = [&](const std::string &s) { out += s; out += '\n'; };
        if (flags & MMB_FMT_INPUT_STATS) {
            line("numberOfParameters=" + num<int>(n_params));
            line("numberOfErrors=" + num<int>(2 * n_obs + n_stiff + n_smooth));
            line("numberOfMarkerErrors=" + num<int>(2 * n_obs));
            line("numberOfAttrStiffnessErrors=" + num<int>(n_stiff));
            line("numberOfAttrSmoothnessErrors=" + num<int>(n_smooth));
        }
        line("success=" + num<int>(r->success != 0));
        line("reason_num=" + num<int>(r->reason_num));
        line("error_final=" + num<double>(r->error_final));
        line("error_final_average=" + num<double>(r->error_avg));
        line("error_final_maximum=" + num<double>(r->error_max));
        line("error_final_minimum=" + num<double>(r->error_min));
        line("iteration_num=" + num<int>(r->iteration_num));
        line("iteration_function_num=" + num<int>(r->iteration_function_num));
        line("iteration_jacobian_num=" + num<int>(r->iteration_jacobian_num));
        line("user_interrupted=" + num<int>(r->user_interrupted != 0));
        // TimerResult: parameter / error timers have no counterpart here (setParameters and
        // measureErrors are fused into the evaluation kernels); ticks are reported in ns
        line("timer_solve=" + num<double>(r->timer_solve_s));
        line("timer_function=" + num<double>(r->timer_function_s));
        line("timer_jacobian=" + num<double>(r->timer_jacobian_s));
        line("timer_parameter=" + num<double>(0.0));
        line("timer_error=" + num<double>(0.0));
        line("ticks_solve=" + num<unsigned long long>((unsigned long long)(r->timer_solve_s * 1e9)));
        line("ticks_function=" + num<unsigned long long>((unsigned long long)(r->timer_function_s * 1e9)));
        line("ticks_jacobian=" + num<unsigned long long>((unsigned long long)(r->timer_jacobian_s * 1e9)));
        line("ticks_parameter=" + num<unsigned long long>(0ull));
        line("ticks_error=" + num<unsigned long long>(0ull));
        if (want_metrics) {
            std::vector<double> fsum(num_frames, 0.0);
            std::vector<int> fcnt(num_frames, 0);
            for (int i = 0; i < n_obs; ++i) {
                const int mk = error_to_marker[2 * i], fi = error_to_marker[2 * i + 1];
                if (mk < 0 || mk >= num_markers || fi < 0 || fi >= num_frames) return MMB_ERR_INVALID_ARGUMENT;
                const double fv = frames ? (double)frames[fi] : (double)fi;
                fsum[fi] += error_dist[i];
                fcnt[fi] += 1;
                if (flags & MMB_FMT_ERROR_PER_MARKER_PER_FRAME) {
                    std::string s = "error_per_marker_per_frame=";
                    if (marker_names && marker_names[mk]) s += marker_names[mk];
                    else s += "marker" + num<int>(mk);
                    s += kSplit; s += num<double>(fv);
                    s += kSplit; s += num<double>(error_dist[i]);
                    line(s);
                }
            }
            if (flags & MMB_FMT_ERROR_PER_FRAME)
                for (int fi = 0; fi < num_frames; ++fi) {
                    if (!fcnt[fi]) continue;
                    const double fv = frames ? (double)frames[fi] : (double)fi;
                    line("error_per_frame=" + num<double>(fv) + kSplit + num<double>(fsum[fi] / fcnt[fi]));
                }
        }
        if (flags & MMB_FMT_SOLVE_VALUES) {
            std::string s = "solve_parameter_list=";
            if (solve_parameter_list)
                for (int i = 0; i < n_params; ++i) { s += num<double>(solve_parameter_list[i]); s += kSplit; }
            line(s);
            s = "solve_error_list=";
            if (solve_error_list)
                for (int i = 0; i < 2 * n_obs + n_stiff + n_smooth; ++i) { s += num<double>(solve_error_list[i]); s += kSplit; }
            line(s);
        }
        if (needed) *needed = (int64_t)out.size() + 1;
        if (buf && cap > 0) {
            const size_t nn = std::min<size_t>((size_t)cap - 1, out.size());
            std::memcpy(buf, out.data(), nn);
            buf[nn] = 0;
            if ((int64_t)out.size() + 1 > cap) return MMB_ERR_INVALID_ARGUMENT;  // truncated: call again with *needed
        }
        return MMB_OK;
    } catch (...) {
        return MMB_ERR_INTERNAL;
    }
}
