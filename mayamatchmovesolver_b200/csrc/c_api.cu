// extern "C" surface declared in include/mmsolver_b200.h.
#include <cstring>
#include <new>

#include "comm.h"
#include "device_problem.h"
#include "lm.h"
#include "shard.h"

using namespace mmb;

struct mmb_problem {
    DeviceProblem *dp;
};
struct mmb_comm {
    Comm *c;
    int device;
};

static const char *k_reasons[9] = {
    // adjust_cminpack_base.h:51-83
    "Improper input parameters",
    "Both actual and predicted relative reductions in the sum of squares are at most ftol.",
    "Relative error between two consecutive iterates is at most xtol.",
    "Conditions for info = 1 and info = 2 both hold.",
    "The cosine of the angle between fvec and any column of the Jacobian is at most gtol in absolute value.",
    "Number of calls to fcn has reached or exceeded maxfev.",
    "'ftol' is too small. No further reduction in the sum of squares is possible.",
    "'xtol' is too small. No further improvement in the approximate solution x is possible.",
    "'gtol' is too small. fvec is orthogonal to the columns of the Jacobian to machine precision."};

const char *mmb_reason_string(int info) { return (info >= 0 && info < 9) ? k_reasons[info] : "User interrupted or solver error."; }

extern "C" {

void mmb_options_default(mmb_options *o) {
    if (!o) return;
    std::memset(o, 0, sizeof(*o));
    o->iter_max = 100;        // arg_flags_solve_info.h ITERATIONS_DEFAULT_VALUE
    o->tau = 1.0;
    o->eps1 = 1e-6; o->eps2 = 1e-6; o->eps3 = 1e-6;
    o->delta = 1e-4;
    o->auto_diff_type = 0;
    o->auto_param_scale = 1;
    o->robust_loss_type = 0;
    o->robust_loss_scale = 1.0;
    o->scene_graph_mode = 2;
    o->solver_type = 2;
    o->time_eval_mode = 0;
    o->accept_only_better = 1;
    o->image_width = 2048.0;
    o->frame_solve_mode = 0;
    o->remove_unused_markers = 1;
    o->remove_unused_attributes = 1;
    o->lens_index_literal = 0;
    o->device = 0;
}

mmb_status mmb_problem_create(const mmb_scene_desc *scene, const mmb_options *opt, mmb_problem **out) {
    if (!scene || !opt || !out) return MMB_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    try {
        DeviceProblem *dp = nullptr;
        mmb_options o = *opt;
        // lmdif only has forward differences (adjust_cminpack_lmdif.cpp:103-105)
        if (o.solver_type == 1) o.auto_diff_type = 0;
        mmb_status st = device_problem_create(scene, &o, &dp);
        if (st != MMB_OK) return st;
        mmb_problem *p = new mmb_problem{dp};
        *out = p;
        return MMB_OK;
    } catch (const std::bad_alloc &) {
        set_last_error("host allocation failed");
        return MMB_ERR_INTERNAL;
    } catch (...) {
        set_last_error("unexpected exception");
        return MMB_ERR_INTERNAL;
    }
}

void mmb_problem_destroy(mmb_problem *p) {
    if (!p) return;
    delete p->dp;
    delete p;
}

mmb_status mmb_problem_sizes(const mmb_problem *p, int32_t *n, int32_t *nobs, int32_t *m) {
    if (!p) return MMB_ERR_INVALID_ARGUMENT;
    if (n) *n = p->dp->hp.n;
    if (nobs) *nobs = p->dp->hp.nobs;
    if (m) *m = p->dp->hp.m_total;
    return MMB_OK;
}

mmb_status mmb_get_index_maps(const mmb_problem *p, int32_t *e2m, int32_t *p2a) {
    if (!p) return MMB_ERR_INVALID_ARGUMENT;
    const HostProblem &hp = p->dp->hp;
    if (e2m) for (int i = 0; i < hp.nobs; ++i) { e2m[2 * i] = hp.err_marker[i]; e2m[2 * i + 1] = hp.err_frame[i]; }
    if (p2a) for (int i = 0; i < hp.n; ++i) { p2a[2 * i] = hp.par_attr[i]; p2a[2 * i + 1] = hp.par_frame[i]; }
    return MMB_OK;
}

mmb_status mmb_problem_used_objects(const mmb_problem *p, int32_t *num_markers_used, int32_t *used_marker_index,
                                    int32_t *num_attrs_used, int32_t *used_attr_index) {
    if (!p) return MMB_ERR_INVALID_ARGUMENT;
    const HostProblem &hp = p->dp->hp;
    if (num_markers_used) *num_markers_used = hp.M;
    if (num_attrs_used) *num_attrs_used = hp.A;
    if (used_marker_index) for (int i = 0; i < hp.M; ++i) used_marker_index[i] = hp.mkr_src[i];
    if (used_attr_index) for (int a = 0; a < hp.A; ++a) used_attr_index[a] = hp.attr_src[a];
    return MMB_OK;
}

mmb_status mmb_affects(const mmb_scene_desc *scene, uint8_t *marker_affects_attr, uint8_t *marker_used,
                       uint8_t *attr_used) {
    if (!scene) return MMB_ERR_INVALID_ARGUMENT;
    try {
        HostProblem hp;
        mmb_options opt;
        mmb_options_default(&opt);
        const mmb_status st = copy_scene_tables(scene, &opt, &hp);
        if (st != MMB_OK) { set_last_error(hp.error); return st; }
        std::vector<uint8_t> table, mu, au;
        marker_attr_usage(hp, marker_affects_attr ? &table : nullptr, &mu, &au);
        if (marker_affects_attr) std::copy(table.begin(), table.end(), marker_affects_attr);
        if (marker_used) std::copy(mu.begin(), mu.end(), marker_used);
        if (attr_used) std::copy(au.begin(), au.end(), attr_used);
    } catch (const std::exception &e) {
        set_last_error(e.what());
        return MMB_ERR_INTERNAL;
    }
    return MMB_OK;
}

mmb_status mmb_get_marker_weights(const mmb_problem *p, double *w) {
    if (!p || !w) return MMB_ERR_INVALID_ARGUMENT;
    const HostProblem &hp = p->dp->hp;
    for (int i = 0; i < hp.nobs; ++i) w[i] = hp.obs_w[i];
    return MMB_OK;
}

mmb_status mmb_get_initial_parameters(const mmb_problem *p, double *x) {
    if (!p || !x) return MMB_ERR_INVALID_ARGUMENT;
    const HostProblem &hp = p->dp->hp;
    for (int i = 0; i < hp.n; ++i) {
        const int a = hp.par_attr[i], f = hp.par_frame[i];
        const int32_t ref = hp.attr_ref[a];
        const double v = ref >= 0 ? hp.stat0[ref] : hp.anim0[(size_t)(-ref - 1) * hp.F + f];
        x[i] = bound_external_to_internal(v, hp.attr_min[a], hp.attr_max[a], hp.attr_off[a], hp.attr_scl[a]);
    }
    return MMB_OK;
}

static mmb_status upload_x(DeviceProblem *dp, const double *x) {
    MMB_CUDA_CHECK(cudaSetDevice(dp->device));
    if (dp->hp.n)
        MMB_CUDA_CHECK(cudaMemcpyAsync(dp->d_x, x, sizeof(double) * dp->hp.n, cudaMemcpyHostToDevice, dp->stream));
    return MMB_OK;
}

mmb_status mmb_eval_residuals(mmb_problem *p, const double *x, double *fvec, double *error_dist) {
    if (!p || !fvec) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    mmb_status st;
    if (x && (st = upload_x(dp, x)) != MMB_OK) return st;
    // x == NULL: the scene at its own attribute values, no parameter applied (the initial
    // measureErrors of solveFrames, adjust_base.cpp:1080-1103)
    if ((st = launch_variants(dp, x ? dp->d_x : nullptr, false)) != MMB_OK) return st;
    if ((st = launch_residuals(dp, dp->d_fvec, dp->d_dist)) != MMB_OK) return st;
    if ((st = launch_extra_rows(dp, x ? dp->d_x : nullptr, dp->d_fvec + dp->hp.m, nullptr, nullptr)) != MMB_OK) return st;
    MMB_CUDA_CHECK(cudaMemcpyAsync(fvec, dp->d_fvec, sizeof(double) * dp->hp.m_total, cudaMemcpyDeviceToHost, dp->stream));
    if (error_dist)
        MMB_CUDA_CHECK(cudaMemcpyAsync(error_dist, dp->d_dist, sizeof(double) * dp->hp.nobs, cudaMemcpyDeviceToHost, dp->stream));
    MMB_CUDA_CHECK(cudaStreamSynchronize(dp->stream));
    return MMB_OK;
}

static mmb_status run_jacobian(DeviceProblem *dp, const double *x, bool want_cols = true) {
    mmb_status st;
    if ((st = ensure_jacobian_buffers(dp)) != MMB_OK) return st;
    if ((st = upload_x(dp, x)) != MMB_OK) return st;
    if ((st = launch_variants(dp, dp->d_x, true)) != MMB_OK) return st;
    if ((st = launch_jacobian(dp, dp->d_fvec, dp->d_dist, dp->d_jac, want_cols ? dp->d_cols : nullptr)) != MMB_OK) return st;
    return MMB_OK;
}

mmb_status mmb_eval_jacobian_blocks(mmb_problem *p, const double *x, double *values, int32_t *col_ids,
                                    int32_t *k_max) {
    if (!p) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    if (k_max) *k_max = dp->ds.kmax;
    if (!values && !col_ids) return MMB_OK;
    if (!x) return MMB_ERR_INVALID_ARGUMENT;
    mmb_status st = run_jacobian(dp, x, col_ids != nullptr);
    if (st != MMB_OK) return st;
    const size_t k = dp->ds.kmax;
    if (values)
        MMB_CUDA_CHECK(cudaMemcpyAsync(values, dp->d_jac, sizeof(double) * k * dp->hp.m, cudaMemcpyDeviceToHost, dp->stream));
    if (col_ids)
        MMB_CUDA_CHECK(cudaMemcpyAsync(col_ids, dp->d_cols, sizeof(int32_t) * k * dp->hp.nobs, cudaMemcpyDeviceToHost, dp->stream));
    MMB_CUDA_CHECK(cudaStreamSynchronize(dp->stream));
    return MMB_OK;
}

mmb_status mmb_eval_jacobian_dense(mmb_problem *p, const double *x, double *fjac, int32_t ldfjac) {
    if (!p || !x || !fjac) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    const HostProblem &hp = dp->hp;
    if (ldfjac < hp.m_total) return MMB_ERR_INVALID_ARGUMENT;
    const size_t k = dp->ds.kmax;
    std::vector<double> vals(k * (size_t)hp.m);
    std::vector<int32_t> cols(k * (size_t)hp.nobs);
    int32_t kk = 0;
    mmb_status st = mmb_eval_jacobian_blocks(p, x, vals.data(), cols.data(), &kk);
    if (st != MMB_OK) return st;
    // layout conversion only: block columns -> cminpack's dense column-major
    for (size_t j = 0; j < (size_t)hp.n; ++j) std::memset(&fjac[j * ldfjac], 0, sizeof(double) * hp.m_total);
    for (size_t i = 0; i < (size_t)hp.nobs; ++i)
        for (size_t c = 0; c < k; ++c) {
            const int32_t pid = cols[i * k + c];
            if (pid < 0) continue;
            fjac[(size_t)pid * ldfjac + 2 * i] = vals[c * hp.m + 2 * i];
            fjac[(size_t)pid * ldfjac + 2 * i + 1] = vals[c * hp.m + 2 * i + 1];
        }
    if (hp.n_extra) {
        // stiffness / smoothness rows: one entry each (the variants of this x are still in place)
        double *d_j = nullptr; int32_t *d_p = nullptr;
        MMB_CUDA_CHECK(cudaMallocAsync((void **)&d_j, sizeof(double) * hp.n_extra, dp->stream));
        MMB_CUDA_CHECK(cudaMallocAsync((void **)&d_p, sizeof(int32_t) * hp.n_extra, dp->stream));
        st = launch_extra_rows(dp, dp->d_x, nullptr, d_j, d_p);
        std::vector<double> jv(hp.n_extra);
        std::vector<int32_t> pv(hp.n_extra);
        if (st == MMB_OK) {
            cudaMemcpyAsync(jv.data(), d_j, sizeof(double) * hp.n_extra, cudaMemcpyDeviceToHost, dp->stream);
            cudaMemcpyAsync(pv.data(), d_p, sizeof(int32_t) * hp.n_extra, cudaMemcpyDeviceToHost, dp->stream);
        }
        cudaFreeAsync(d_j, dp->stream);
        cudaFreeAsync(d_p, dp->stream);
        MMB_CUDA_CHECK(cudaStreamSynchronize(dp->stream));
        if (st != MMB_OK) return st;
        for (int e = 0; e < hp.n_extra; ++e) fjac[(size_t)pv[e] * ldfjac + hp.m + e] = jv[e];
    }
    return MMB_OK;
}

mmb_status mmb_solve(mmb_problem *p, mmb_result *result, double *x_out, double *attr_values_out,
                     double *error_dist_out) {
    if (!p || !result) return MMB_ERR_INVALID_ARGUMENT;
    try {
        if (p->dp->hp.opt.frame_solve_mode == 1)
            return mmb_solve_per_frame(p, result, nullptr, x_out, attr_values_out, error_dist_out);
        mmb_status st = lm_solve(p->dp, result, x_out, attr_values_out, error_dist_out);
        result->reason_str = mmb_reason_string(result->reason_num);
        return st;
    } catch (...) {
        set_last_error("unexpected exception in mmb_solve");
        return MMB_ERR_INTERNAL;
    }
}

mmb_status mmb_solve_per_frame(mmb_problem *p, mmb_result *result, mmb_result *per_frame_results, double *x_out,
                               double *attr_values_out, double *error_dist_out) {
    if (!p || !result) return MMB_ERR_INVALID_ARGUMENT;
    try {
        mmb_status st = lm_solve_per_frame(p->dp, result, x_out, attr_values_out, error_dist_out, per_frame_results);
        result->reason_str = mmb_reason_string(result->reason_num);
        if (per_frame_results && st == MMB_OK)
            for (int f = 0; f < p->dp->hp.F; ++f)
                per_frame_results[f].reason_str = mmb_reason_string(per_frame_results[f].reason_num);
        return st;
    } catch (...) {
        set_last_error("unexpected exception in mmb_solve_per_frame");
        return MMB_ERR_INTERNAL;
    }
}

// ---- multi-GPU: frame-range shards ----------------------------------------------------------
mmb_status mmb_shard_frames(int32_t num_frames, int32_t world, int32_t rank, int32_t *frame_begin, int32_t *frame_end) {
    if (num_frames < 0 || world < 1 || rank < 0 || rank >= world || !frame_begin || !frame_end) return MMB_ERR_INVALID_ARGUMENT;
    int f0, f1;
    shard_frame_range(num_frames, world, rank, &f0, &f1);
    *frame_begin = f0; *frame_end = f1;
    return MMB_OK;
}

mmb_status mmb_shard_param_map(const mmb_scene_desc *scene, int32_t world, int32_t rank, int32_t *n_local,
                               int32_t *global_param) {
    if (!scene || world < 1 || rank < 0 || rank >= world) return MMB_ERR_INVALID_ARGUMENT;
    if (scene->num_attrs > 0 && !scene->attr_ref) return MMB_ERR_INVALID_ARGUMENT;
    try {
        int f0, f1;
        shard_frame_range(scene->num_frames, world, rank, &f0, &f1);
        std::vector<int32_t> map;
        shard_param_map(scene, f0, f1, &map);
        if (n_local) *n_local = (int32_t)map.size();
        if (global_param) std::memcpy(global_param, map.data(), sizeof(int32_t) * map.size());
        return MMB_OK;
    } catch (...) {
        set_last_error("unexpected exception in mmb_shard_param_map");
        return MMB_ERR_INTERNAL;
    }
}

mmb_status mmb_problem_create_shard(const mmb_scene_desc *scene, const mmb_options *opt, int32_t rank, int32_t world,
                                    mmb_problem **out) {
    if (!scene || !opt || !out || world < 1 || rank < 0 || rank >= world) return MMB_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    try {
        int f0, f1;
        shard_frame_range(scene->num_frames, world, rank, &f0, &f1);
        if (f1 <= f0) {
            set_last_error("more ranks than frames: this rank's shard is empty");
            return MMB_ERR_NOT_ENOUGH_OBJECTS;
        }
        SceneSlice slice;
        slice_scene_frames(scene, f0, f1, &slice);
        mmb_status st = mmb_problem_create(&slice.desc, opt, out);
        if (st != MMB_OK) return st;
        (*out)->dp->frame_begin = f0;
        (*out)->dp->frames_global = scene->num_frames;
        return MMB_OK;
    } catch (const std::bad_alloc &) {
        set_last_error("host allocation failed");
        return MMB_ERR_INTERNAL;
    } catch (...) {
        set_last_error("unexpected exception in mmb_problem_create_shard");
        return MMB_ERR_INTERNAL;
    }
}

mmb_status mmb_problem_create_shard_local(const mmb_scene_desc *slice, const mmb_options *opt, int32_t frame_begin,
                                          int32_t frames_global, mmb_problem **out) {
    if (!slice || !opt || !out || frame_begin < 0 || frame_begin + slice->num_frames > frames_global)
        return MMB_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    mmb_status st = mmb_problem_create(slice, opt, out);
    if (st != MMB_OK) return st;
    (*out)->dp->frame_begin = frame_begin;
    (*out)->dp->frames_global = frames_global;
    return MMB_OK;
}

mmb_status mmb_comm_unique_id(uint8_t id[MMB_COMM_ID_BYTES], int32_t local_backend) {
    if (!id) return MMB_ERR_INVALID_ARGUMENT;
    try {
        return comm_unique_id(id, local_backend);
    } catch (...) {
        set_last_error("unexpected exception in mmb_comm_unique_id");
        return MMB_ERR_INTERNAL;
    }
}

mmb_status mmb_problem_attach_comm(mmb_problem *p, const uint8_t id[MMB_COMM_ID_BYTES], int32_t rank, int32_t world) {
    if (!p || !id) return MMB_ERR_INVALID_ARGUMENT;
    try {
        DeviceProblem *dp = p->dp;
        if (dp->comm) { set_last_error("handle already has a communicator"); return MMB_ERR_INVALID_ARGUMENT; }
        if (dp->hp.opt.frame_solve_mode == 1) {
            set_last_error("per-frame solves of different shards are independent: no communicator is needed");
            return MMB_ERR_INVALID_ARGUMENT;
        }
        Comm *c = nullptr;
        mmb_status st = comm_create(id, rank, world, dp->device, &c);
        if (st != MMB_OK) return st;
        dp->comm = c;
        return MMB_OK;
    } catch (...) {
        set_last_error("unexpected exception in mmb_problem_attach_comm");
        return MMB_ERR_INTERNAL;
    }
}

mmb_status mmb_comm_create(const uint8_t id[MMB_COMM_ID_BYTES], int32_t rank, int32_t world, int32_t device,
                           mmb_comm **out) {
    if (!id || !out) return MMB_ERR_INVALID_ARGUMENT;
    *out = nullptr;
    try {
        Comm *c = nullptr;
        mmb_status st = comm_create(id, rank, world, device, &c);
        if (st != MMB_OK) return st;
        // warm-up: the first collectives of an NCCL communicator set up its connections
        // (hundreds of ms); pay that here, once, for a latency-sized and a bandwidth-sized call
        if (world > 1) {
            cudaStream_t s = nullptr;
            double *buf = nullptr;
            const size_t big = (size_t)2 << 20;  // 16 MB of doubles: the size class of the reduced system
            bool ok = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) == cudaSuccess &&
                      cudaMalloc((void **)&buf, big * sizeof(double)) == cudaSuccess &&
                      cudaMemsetAsync(buf, 0, big * sizeof(double), s) == cudaSuccess;
            if (ok) ok = c->allreduce(buf, 4, false, s) == MMB_OK && c->allreduce(buf, 4, true, s) == MMB_OK &&
                         c->allreduce(buf, 1184 + 2, false, s) == MMB_OK && c->allreduce(buf, big, false, s) == MMB_OK;
            if (ok) ok = cudaStreamSynchronize(s) == cudaSuccess;
            if (buf) cudaFree(buf);
            if (s) cudaStreamDestroy(s);
            if (!ok) { delete c; if (get_last_error().empty()) set_last_error("communicator warm-up failed"); return MMB_ERR_COMM; }
        }
        *out = new mmb_comm{c, device};
        return MMB_OK;
    } catch (...) {
        set_last_error("unexpected exception in mmb_comm_create");
        return MMB_ERR_INTERNAL;
    }
}

void mmb_comm_destroy(mmb_comm *c) {
    if (!c) return;
    delete c->c;
    delete c;
}

mmb_status mmb_problem_use_comm(mmb_problem *p, mmb_comm *c) {
    if (!p || !c) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    if (dp->comm) { set_last_error("handle already has a communicator"); return MMB_ERR_INVALID_ARGUMENT; }
    if (dp->hp.opt.frame_solve_mode == 1) {
        set_last_error("per-frame solves of different shards are independent: no communicator is needed");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    if (c->device != dp->device) { set_last_error("communicator and handle are bound to different devices"); return MMB_ERR_INVALID_ARGUMENT; }
    dp->comm = c->c;
    dp->comm_owned = false;
    return MMB_OK;
}

mmb_status mmb_aggregate_results(const mmb_result *per_frame, int32_t count, mmb_result *out) {
    if (!per_frame || !out || count < 0) return MMB_ERR_INVALID_ARGUMENT;
    aggregate_results(per_frame, count, out);
    out->reason_str = mmb_reason_string(out->reason_num);
    return MMB_OK;
}

mmb_status mmb_get_last_residuals(mmb_problem *p, double *fvec) {
    if (!p || !fvec) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    MMB_CUDA_CHECK(cudaSetDevice(dp->device));
    MMB_CUDA_CHECK(cudaMemcpyAsync(fvec, dp->d_fvec, sizeof(double) * dp->hp.m_total, cudaMemcpyDeviceToHost, dp->stream));
    MMB_CUDA_CHECK(cudaStreamSynchronize(dp->stream));
    return MMB_OK;
}

void mmb_request_cancel(mmb_problem *p) {
    if (p) lm_request_cancel(p->dp);
}

mmb_status mmb_bench_eval(mmb_problem *p, int32_t reps, int32_t flush_l2, double *ms_per_eval,
                          double *ms_jacobian_kernel, int32_t *launches) {
    if (!p || reps <= 0) return MMB_ERR_INVALID_ARGUMENT;
    DeviceProblem *dp = p->dp;
    const HostProblem &hp = dp->hp;
    std::vector<double> x(hp.n);
    mmb_get_initial_parameters(p, x.data());
    mmb_status st = MMB_OK;
    if (!dp->d_jac) {  // first call: buffers + attribute store at x0
        st = run_jacobian(dp, x.data());
        if (st != MMB_OK) return st;
        MMB_CUDA_CHECK(cudaStreamSynchronize(dp->stream));
    }
    double *flush = dp->d_flush;
    const size_t flush_n = (size_t)256 << 20 >> 3;  // 256 MiB > 126 MB L2
    if (flush_l2 && !flush) {
        MMB_CUDA_CHECK(dp->arena.alloc(&dp->d_flush, flush_n));
        flush = dp->d_flush;
    }
    if (!flush_l2) flush = nullptr;
    cudaEvent_t e0, e1, ej, e2;
    MMB_CUDA_CHECK(cudaEventCreate(&e0));
    MMB_CUDA_CHECK(cudaEventCreate(&e1));
    MMB_CUDA_CHECK(cudaEventCreate(&ej));
    MMB_CUDA_CHECK(cudaEventCreate(&e2));
    double total_ms = 0.0, jac_ms = 0.0;
    long long step_launches = 0;
    for (int r = 0; r < reps; ++r) {
        // (1) the whole evaluation step: entity variants + residual/Jacobian, nothing between them
        if (flush) MMB_CUDA_CHECK(cudaMemsetAsync(flush, r & 0xff, flush_n * sizeof(double), dp->stream));
        const long long la = dp->launches;
        MMB_CUDA_CHECK(cudaEventRecord(e0, dp->stream));
        if ((st = launch_variants(dp, dp->d_x, true)) != MMB_OK) return st;
        if ((st = launch_jacobian(dp, dp->d_fvec, dp->d_dist, dp->d_jac, nullptr)) != MMB_OK) return st;
        MMB_CUDA_CHECK(cudaEventRecord(e1, dp->stream));
        step_launches += dp->launches - la;
        // (2) the residual/Jacobian kernel alone (variant tables fresh, marker tracks cold)
        if (flush) MMB_CUDA_CHECK(cudaMemsetAsync(flush, (r + 1) & 0xff, flush_n * sizeof(double), dp->stream));
        if ((st = launch_variants(dp, dp->d_x, true)) != MMB_OK) return st;
        MMB_CUDA_CHECK(cudaEventRecord(ej, dp->stream));
        if ((st = launch_jacobian(dp, dp->d_fvec, dp->d_dist, dp->d_jac, nullptr)) != MMB_OK) return st;
        MMB_CUDA_CHECK(cudaEventRecord(e2, dp->stream));
        MMB_CUDA_CHECK(cudaEventSynchronize(e2));
        float ms = 0.f;
        MMB_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
        total_ms += ms;
        MMB_CUDA_CHECK(cudaEventElapsedTime(&ms, ej, e2));
        jac_ms += ms;
    }
    if (ms_per_eval) *ms_per_eval = total_ms / reps;
    if (ms_jacobian_kernel) *ms_jacobian_kernel = jac_ms / reps;
    if (launches) *launches = (int32_t)step_launches;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(ej); cudaEventDestroy(e2);
    return MMB_OK;
}

int64_t mmb_launch_count(const mmb_problem *p) { return p ? (int64_t)p->dp->launches : 0; }

mmb_status mmb_solve_stats(const mmb_problem *p, int64_t *host_syncs, int64_t *graph_launches) {
    if (!p) return MMB_ERR_INVALID_ARGUMENT;
    if (host_syncs) *host_syncs = p->dp->solve_host_syncs;
    if (graph_launches) *graph_launches = p->dp->solve_graph_launches;
    return MMB_OK;
}

mmb_status mmb_bench_fp64_peaks(int32_t device, double *dfma_tflops, double *dmma_tflops) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMB_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return MMB_ERR_INVALID_ARGUMENT;
    return measure_fp64_peaks(device, dfma_tflops, dmma_tflops);
}

mmb_status mmb_bench_stream_floor(int32_t device, int64_t read_bytes, int64_t write_bytes, int32_t reps, double *ms_median) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMB_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return MMB_ERR_INVALID_ARGUMENT;
    return measure_stream_floor(device, read_bytes, write_bytes, reps, ms_median);
}

static mmb_status lens_args_ok(int32_t device, int32_t model, const double *cam, int32_t direction) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_last_error("no CUDA device: this library has no CPU path");
        return MMB_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev || model < 0 || model > 4 || !cam || (direction != 0 && direction != 1))
        return MMB_ERR_INVALID_ARGUMENT;
    return MMB_OK;
}

mmb_status mmb_lens_apply(int32_t device, int32_t model, const double *coeffs, const double *camera_params,
                          int32_t direction, int64_t n_points, const double *xy_in, double *xy_out) {
    mmb_status st = lens_args_ok(device, model, camera_params, direction);
    if (st != MMB_OK) return st;
    if (n_points < 0 || (n_points > 0 && (!xy_in || !xy_out))) return MMB_ERR_INVALID_ARGUMENT;
    return lens_apply_host(device, model, coeffs, camera_params, direction, n_points, xy_in, xy_out);
}

mmb_status mmb_lens_apply_layered(int32_t device, int32_t n_layers, const int32_t *models, const double *coeffs,
                                  const double *camera_params, int32_t direction, int64_t n_points,
                                  const double *xy_in, double *xy_out) {
    if (!models || n_layers < 1) return MMB_ERR_INVALID_ARGUMENT;
    for (int k = 0; k < n_layers; ++k) {
        mmb_status st = lens_args_ok(device, models[k], camera_params, direction);
        if (st != MMB_OK) return st;
    }
    if (n_points < 0 || (n_points > 0 && (!xy_in || !xy_out))) return MMB_ERR_INVALID_ARGUMENT;
    return lens_apply_layered_host(device, n_layers, models, coeffs, camera_params, direction, n_points, xy_in, xy_out);
}

mmb_status mmb_lens_apply_layered_no_alloc(int32_t device, int32_t n_layers, const int32_t *models, const double *coeffs,
                                           const double *camera_params, int32_t direction, int64_t n_points,
                                           const double *xy_in, double *xy_out) {
    if (!models || n_layers < 1) return MMB_ERR_INVALID_ARGUMENT;
    for (int k = 0; k < n_layers; ++k) {
        mmb_status st = lens_args_ok(device, models[k], camera_params, direction);
        if (st != MMB_OK) return st;
    }
    if (n_points < 0 || (n_points > 0 && (!xy_in || !xy_out))) return MMB_ERR_INVALID_ARGUMENT;
    return lens_apply_layered_no_alloc_host(device, n_layers, models, coeffs, camera_params, direction, n_points, xy_in,
                                            xy_out);
}

mmb_status mmb_bench_lens(int32_t device, int32_t model, const double *coeffs, const double *camera_params,
                          int32_t direction, int64_t n_points, int32_t reps, double *ms_per_pass) {
    mmb_status st = lens_args_ok(device, model, camera_params, direction);
    if (st != MMB_OK) return st;
    return lens_bench(device, model, coeffs, camera_params, direction, n_points, reps, ms_per_pass);
}

mmb_status mmb_debug_chol_trace(uint64_t *out, int32_t count) {
    if (!out || count <= 0) return MMB_ERR_INVALID_ARGUMENT;
    return lm_chol_trace(reinterpret_cast<unsigned long long *>(out), count);
}

mmb_status mmb_release_cached_memory(int32_t device) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MMB_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return MMB_ERR_INVALID_ARGUMENT;
    cudaMemPool_t pool = nullptr;
    MMB_CUDA_CHECK(cudaDeviceGetDefaultMemPool(&pool, device));
    MMB_CUDA_CHECK(cudaMemPoolTrimTo(pool, 0));
    host_block_trim();   // the page-locked host blocks of the per-observation arrays
    return MMB_OK;
}

const char *mmb_status_string(mmb_status s) {
    switch (s) {
        case MMB_OK: return "ok";
        case MMB_ERR_INVALID_ARGUMENT: return "invalid argument";
        case MMB_ERR_CUDA: return "CUDA error";
        case MMB_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU path)";
        case MMB_ERR_NOT_ENOUGH_OBJECTS: return "not enough markers/attributes to solve";
        case MMB_ERR_MORE_PARAMS_THAN_ERRORS: return "more unknown parameters than error measurements";
        case MMB_ERR_UNSUPPORTED: return "unsupported option";
        case MMB_ERR_INTERNAL: return "internal error";
        case MMB_ERR_COMM: return "communicator error (NCCL)";
        default: return "unknown status";
    }
}

const char *mmb_last_error(void) { return get_last_error().c_str(); }

const char *mmb_version(void) { return "mmsolver_b200 0.1.0 (sm_100a)"; }

}  // extern "C"
