// Device-resident problem: owns every HBM buffer of one solve handle.
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <string>
#include <cstring>
#include <vector>

#include "eval_bodies.cuh"
#include "jacobian_tiled.cuh"
#include "problem.h"

namespace mmb {

#define MMB_CUDA_CHECK(expr)                                                      \
    do {                                                                          \
        cudaError_t _e = (expr);                                                  \
        if (_e != cudaSuccess) {                                                  \
            set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e));   \
            return MMB_ERR_CUDA;                                                  \
        }                                                                         \
    } while (0)

void set_last_error(const std::string &msg);
const std::string &get_last_error();

// Arena of device allocations freed together.  Allocations come from the device's
// stream-ordered memory pool (cudaMallocAsync) on the handle's stream; the pool's release
// threshold is raised so that the memory of a destroyed handle is kept for the next one
// (mmSolver is called once per solver step: create / solve / destroy repeats with the same
// shapes).  mmb_release_cached_memory() trims the pool.  Uploads are asynchronous copies
// on the same stream: a pageable source is staged before the call returns (the caller may free
// it right away); the per-observation arrays of the HostProblem live in page-locked blocks
// (NoInitAlloc) and go out as plain DMA -- they stay alive until the handle's destructor has
// synchronised the stream.
struct DeviceArena {
    std::vector<void *> ptrs;
    size_t bytes = 0;
    cudaStream_t stream = nullptr;
    // Batch window (mmb_problem_create): the many small index tables (a few dozen uploads of a few
    // hundred bytes to a few KB) share ONE device slab and ONE page-locked staging block; an upload
    // is a host memcpy into the staging block, and flush() sends what has gathered since the last
    // one in a single copy -- instead of one cudaMallocAsync + one staged pageable copy per table.
    static constexpr size_t kSlabBytes = (size_t)2 << 20, kSlabSmall = (size_t)64 << 10;
    char *slab = nullptr, *stage = nullptr;
    size_t slab_used = 0, slab_flushed = 0;
    bool batching = false;
    ~DeviceArena() { release(); }
    void bind(cudaStream_t s) { stream = s; }
    void release() {
        for (void *p : ptrs) cudaFreeAsync(p, stream);
        ptrs.clear();
        bytes = 0;
        slab = nullptr; slab_used = slab_flushed = 0; batching = false;
        if (stage) { host_block_release(stage); stage = nullptr; }   // (the owner has synchronised the stream)
    }
    template <typename T>
    cudaError_t alloc(T **out, size_t count) {
        *out = nullptr;
        if (count == 0) count = 1;
        void *p = nullptr;
        cudaError_t e = cudaMallocAsync(&p, count * sizeof(T), stream);
        if (e != cudaSuccess) return e;
        ptrs.push_back(p);
        bytes += count * sizeof(T);
        *out = (T *)p;
        return cudaSuccess;
    }
    cudaError_t begin_batch() {
        if (!slab) {
            cudaError_t e = alloc(&slab, kSlabBytes);
            if (e != cudaSuccess) return e;
            stage = static_cast<char *>(host_block_acquire(kSlabBytes));
        }
        batching = true;
        return cudaSuccess;
    }
    // everything uploaded so far is on its way (call before a kernel that reads it is launched)
    cudaError_t flush() {
        if (slab_used == slab_flushed) return cudaSuccess;
        const cudaError_t e = cudaMemcpyAsync(slab + slab_flushed, stage + slab_flushed, slab_used - slab_flushed, cudaMemcpyHostToDevice, stream);
        slab_flushed = slab_used;
        return e;
    }
    cudaError_t end_batch() { batching = false; return flush(); }
    template <typename T, typename Al>
    cudaError_t upload(T **out, const std::vector<T, Al> &v) {
        const size_t nb = (std::max<size_t>(v.size(), 1) * sizeof(T) + 255) & ~(size_t)255;
        if (batching && nb <= kSlabSmall && slab_used + nb <= kSlabBytes) {
            *out = reinterpret_cast<T *>(slab + slab_used);
            if (!v.empty()) std::memcpy(stage + slab_used, v.data(), v.size() * sizeof(T));
            slab_used += nb;
            return cudaSuccess;
        }
        cudaError_t e = alloc(out, v.size());
        if (e != cudaSuccess) return e;
        if (!v.empty()) e = cudaMemcpyAsync(*out, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, stream);
        return e;
    }
};

struct LmWorkspace;  // lm_global.cu
struct Comm;         // comm.h

struct DeviceProblem {
    HostProblem hp;
    DevScene ds;        // device pointers (passed by value to kernels)
    DeviceArena arena;
    cudaStream_t stream = nullptr;
    int device = 0;
    // work vectors
    double *d_x = nullptr;      // [n] parameters (reference order)
    double *d_fvec = nullptr;   // [m]
    double *d_dist = nullptr;   // [nobs]
    double *d_jac = nullptr;    // [kmax * m] block Jacobian (lazy)
    int32_t *d_cols = nullptr;  // [nobs * kmax] (lazy)
    double *d_stat0 = nullptr, *d_anim0 = nullptr;  // pristine attribute store
    double *d_flush = nullptr;  // L2 flush scratch (bench only)
    bool tiled_ok = false;      // register-tiled Jacobian kernel applies
    bool lens_tile_ok = false;  // lens-in-the-loop tiled kernel applies (k_jacobian_lens)
    TiledPlan tiled{};
    // several cameras (cfg5 shape): the rows kernel once per camera over that camera's markers
    std::vector<TiledPlan> cam_plans;
    const CamSlice *d_cam_slices = nullptr;   // the per-camera part of cam_plans, on the device (blockIdx.z of the one launch)
    int rows_mt[2] = {0, 0};    // markers per CTA of k_jacobian_rows (0 = not configured yet)
    LmWorkspace *lm = nullptr;
    Comm *comm = nullptr;       // set by mmb_problem_attach_comm / use_comm: mmb_solve is then the sharded global solve
    bool comm_owned = true;     // false: borrowed from an mmb_comm that outlives the handle
    int frame_begin = 0;        // first frame of this shard in the unsharded scene
    int frames_global = 0;      // frame count of the unsharded scene
    volatile int cancel = 0;
    long long launches = 0;     // kernels launched by this handle
    long long solve_host_syncs = 0;      // host<->device synchronisations inside the last mmb_solve
    long long solve_graph_launches = 0;  // CUDA graph launches of the last mmb_solve (device-resident LM loop)

    ~DeviceProblem();
};

// FP64 ceilings of the device (peaks.cu)
mmb_status measure_fp64_peaks(int device, double *dfma_tflops, double *dmma_tflops);

// bulk lens (un)distortion of point buffers (lens_bulk.cu)
mmb_status lens_apply_host(int device, int model, const double *coeffs, const double *cam, int direction, long long n,
                           const double *xy_in, double *xy_out);
mmb_status lens_apply_layered_host(int device, int n_layers, const int32_t *models, const double *coeffs,
                                   const double *cams, int direction, long long n, const double *xy_in, double *xy_out);
mmb_status measure_stream_floor(int device, long long read_bytes, long long write_bytes, int reps, double *ms_median);
mmb_status lens_apply_layered_no_alloc_host(int device, int n_layers, const int32_t *models, const double *coeffs,
                                            const double *cams, int direction, long long n, const double *xy_in,
                                            double *xy_out);
mmb_status lens_bench(int device, int model, const double *coeffs, const double *cam, int direction, long long n,
                      int reps, double *ms_per_pass);

mmb_status device_problem_create(const mmb_scene_desc *s, const mmb_options *opt, DeviceProblem **out);
mmb_status ensure_jacobian_buffers(DeviceProblem *p);

// kernel launchers (kernels.cu); all asynchronous on p->stream
// d_x == nullptr evaluates the scene at its own attribute values (no parameters applied)
mmb_status launch_variants(DeviceProblem *p, const double *d_x, bool with_perturbed);
mmb_status launch_residuals(DeviceProblem *p, double *d_fvec, double *d_dist);
mmb_status launch_jacobian(DeviceProblem *p, double *d_fvec, double *d_dist, double *d_jac, int32_t *d_cols,
                           double *d_dist_last = nullptr);
// stiffness / smoothness rows at d_x: values into d_row[n_extra] and / or the one Jacobian entry of
// each row into d_jext[n_extra] with its parameter id in d_pid[n_extra] (either may be nullptr)
mmb_status launch_extra_rows(DeviceProblem *p, const double *d_x, double *d_row, double *d_jext, int32_t *d_pid);

}  // namespace mmb
