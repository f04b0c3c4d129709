// Bulk lens (un)distortion of point buffers: the reference's
// LensModel::applyModelUndistort / applyModelDistort loops
// (lib/cppbind/mmlens/src/lens_model_3de_*.cpp, distortion_operations.h:33-96 -- what
// tools/lensdistortion runs per pixel to make ST-maps) as one streaming kernel.
//   direction 0 (undistort): direct polynomial, 32 B per point -> HBM-bound
//   direction 1 (distort):   LDPK fixed-point inverse (<= 22 polynomial evaluations) -> FP64-bound
// Same per-point arithmetic as the solve path (device_math.cuh), so the solver's lens and
// this kernel cannot drift apart.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>

#include "device_problem.h"

namespace mmb {

struct LensBulk {
    double rec[kLensRec];  // prepared record (lens_prepare)
    double after[4];       // undistort: 2x2 applied after the polynomial
    double before[4];      // undistort: 2x2 applied before it
};

// applyModelUndistort: unit -> diagonal-normalised -> [before] -> polynomial -> [after] -> unit
__device__ __forceinline__ void lens_undistort_pt(const LensBulk &L, double px, double py, double &ox, double &oy) {
    const int model = (int)L.rec[0];
    ox = px; oy = py;
    if (model == 0) return;
    const double radius = L.rec[1], w = L.rec[2], h = L.rec[3], lx = L.rec[4], ly = L.rec[5];
    double qx = (((px + 0.5) - 0.5) * w - lx) / radius;
    double qy = (((py + 0.5) - 0.5) * h - ly) / radius;
    {
        const double tx = L.before[0] * qx + L.before[1] * qy, ty = L.before[2] * qx + L.before[3] * qy;
        qx = tx; qy = ty;
    }
    double fx, fy;
    lens_core(model, &L.rec[6], qx, qy, fx, fy);
    const double x = L.after[0] * fx + L.after[1] * fy, y = L.after[2] * fx + L.after[3] * fy;
    double xc = x * radius, yc = y * radius;
    xc += (w / 2) + lx; yc += (h / 2) + ly;
    ox = xc / w - 0.5; oy = yc / h - 0.5;
}

// 4 points per thread per trip, loads issued together (streaming: every byte is touched once)
constexpr int kLensUnroll = 4;
__global__ void __launch_bounds__(256) k_lens_bulk(LensBulk L, int direction, long long n,
                                                   const double2 *__restrict__ in, double2 *__restrict__ out) {
    const long long tile = (long long)blockDim.x * kLensUnroll;
    for (long long base = blockIdx.x * tile; base < n; base += (long long)gridDim.x * tile) {
        double2 p[kLensUnroll];
#pragma unroll
        for (int u = 0; u < kLensUnroll; ++u) {
            const long long i = base + u * blockDim.x + threadIdx.x;
            p[u] = i < n ? __ldcs(&in[i]) : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int u = 0; u < kLensUnroll; ++u) {
            const long long i = base + u * blockDim.x + threadIdx.x;
            double2 o;
            if (direction == 0) lens_undistort_pt(L, p[u].x, p[u].y, o.x, o.y);
            else lens_distort(L.rec, p[u].x, p[u].y, o.x, o.y);
            if (i < n) __stcs(&out[i], o);
        }
    }
}

// Layered lenses (a Maya lens-node chain: each LensModel first applies its inputLensModel,
// lens_model_3de_*.cpp:44-50 / 89-95): layer 0 is the innermost, both directions run
// innermost -> outermost.
constexpr int kLensMaxLayers = 8;
struct LensLayers {
    long long n;
    LensBulk layer[kLensMaxLayers];
};

__global__ void __launch_bounds__(256) k_lens_bulk_layered(const LensLayers *__restrict__ LL, int direction, long long n,
                                                           const double2 *__restrict__ in, double2 *__restrict__ out) {
    __shared__ LensLayers L;
    for (int i = threadIdx.x; i < (int)(sizeof(LensLayers) / sizeof(double)); i += blockDim.x)
        reinterpret_cast<double *>(&L)[i] = reinterpret_cast<const double *>(LL)[i];
    __syncthreads();
    const int nl = (int)L.n;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        double2 p = __ldcs(&in[i]);
        for (int k = 0; k < nl; ++k) {
            double2 o;
            if (direction == 0) lens_undistort_pt(L.layer[k], p.x, p.y, o.x, o.y);
            else lens_distort_raw(L.layer[k].rec, p.x, p.y, o.x, o.y);
            p = o;
        }
        __stcs(&out[i], p);
    }
}

static void make_bulk(int model, const double *coeffs, const double *cam, LensBulk *L) {
    double c[MMB_LENS_MAX_COEFFS];
    for (int i = 0; i < MMB_LENS_MAX_COEFFS; ++i) c[i] = coeffs ? coeffs[i] : 0.0;
    lens_prepare(model, c, cam, L->rec);
    const double ident[4] = {1.0, 0.0, 0.0, 1.0};
    std::memcpy(L->after, ident, sizeof(ident));
    std::memcpy(L->before, ident, sizeof(ident));
    if (model == 2) {
        m2_inv(&L->rec[16], L->after);           // cylindric extender itself (the record holds its inverse)
    } else if (model == 3 || model == 4) {
        m2_inv(&L->rec[16], L->after);           // rotation / squeeze / pixel-aspect product
        m2_inv(&L->rec[20], L->before);          // inverse of the matrix the distort direction applies last
    }
}

static unsigned grid_for(long long n, int device) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const long long want = (n + 256 * kLensUnroll - 1) / (256 * kLensUnroll);
    return (unsigned)std::max<long long>(1, std::min<long long>(want, (long long)sms * 8));
}

mmb_status lens_apply_host(int device, int model, const double *coeffs, const double *cam, int direction,
                           long long n, const double *xy_in, double *xy_out) {
    MMB_CUDA_CHECK(cudaSetDevice(device));
    LensBulk L;
    make_bulk(model, coeffs, cam, &L);
    if (n == 0) return MMB_OK;
    cudaStream_t s = nullptr;
    MMB_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    double2 *d_in = nullptr, *d_out = nullptr;
    cudaError_t e = cudaMallocAsync((void **)&d_in, sizeof(double2) * n, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_out, sizeof(double2) * n, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, xy_in, sizeof(double2) * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        k_lens_bulk<<<grid_for(n, device), 256, 0, s>>>(L, direction, n, d_in, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(xy_out, d_out, sizeof(double2) * n, cudaMemcpyDeviceToHost, s);
    if (d_in) cudaFreeAsync(d_in, s);
    if (d_out) cudaFreeAsync(d_out, s);
    const cudaError_t e2 = cudaStreamSynchronize(s);
    cudaStreamDestroy(s);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) { set_last_error(std::string("lens_apply: ") + cudaGetErrorString(e)); return MMB_ERR_CUDA; }
    return MMB_OK;
}

mmb_status lens_apply_layered_host(int device, int n_layers, const int32_t *models, const double *coeffs,
                                   const double *cams, int direction, long long n, const double *xy_in, double *xy_out) {
    MMB_CUDA_CHECK(cudaSetDevice(device));
    if (n_layers < 1 || n_layers > kLensMaxLayers) {
        set_last_error("lens_apply_layered: 1..8 layers");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    static_assert(sizeof(LensLayers) % sizeof(double) == 0, "copied as doubles");
    LensLayers LL;
    std::memset(&LL, 0, sizeof(LL));
    LL.n = n_layers;
    for (int k = 0; k < n_layers; ++k)
        make_bulk(models[k], coeffs ? coeffs + (size_t)MMB_LENS_MAX_COEFFS * k : nullptr, cams + 6 * (size_t)k, &LL.layer[k]);
    if (n == 0) return MMB_OK;
    cudaStream_t s = nullptr;
    MMB_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    double2 *d_in = nullptr, *d_out = nullptr;
    LensLayers *d_ll = nullptr;
    cudaError_t e = cudaMallocAsync((void **)&d_in, sizeof(double2) * n, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_out, sizeof(double2) * n, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_ll, sizeof(LensLayers), s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ll, &LL, sizeof(LL), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, xy_in, sizeof(double2) * n, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        k_lens_bulk_layered<<<grid_for(n * kLensUnroll, device), 256, 0, s>>>(d_ll, direction, n, d_in, d_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(xy_out, d_out, sizeof(double2) * n, cudaMemcpyDeviceToHost, s);
    if (d_in) cudaFreeAsync(d_in, s);
    if (d_out) cudaFreeAsync(d_out, s);
    if (d_ll) cudaFreeAsync(d_ll, s);
    const cudaError_t e2 = cudaStreamSynchronize(s);
    cudaStreamDestroy(s);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) { set_last_error(std::string("lens_apply_layered: ") + cudaGetErrorString(e)); return MMB_ERR_CUDA; }
    return MMB_OK;
}

// ---- the layered kernel for callers that cannot allocate -------------------------------------
// mmb_lens_apply_layered_no_alloc: same kernel, same host buffers in and out, but every byte of
// device memory it needs -- the layer table the kernel reads and the staging of the points --
// lives in ONE file-scope array instead of cudaMallocAsync'd buffers.  Every GPU that loads the
// library holds these 8 MiB until exit.  All calls share the array and it keeps the last call's
// contents: nothing in it has to start at zero (the layer table is rewritten whole, only the
// first n points of either staging area are read), but two calls that overlap on one GPU race.
constexpr long long kNoAllocMaxPoints = MMB_LENS_NO_ALLOC_MAX_POINTS;
constexpr size_t kNoAllocHeader = 4096;   // LensLayers
constexpr size_t kNoAllocBytes = kNoAllocHeader + 2 * (size_t)kNoAllocMaxPoints * sizeof(double2);
static_assert(sizeof(LensLayers) <= kNoAllocHeader, "layer table must fit the scratch header");
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address is
// asked for once per device, with that device current, and cached.
static mmb_status noalloc_scratch_ptr(int device, unsigned char **out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static unsigned char *ptrs[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) {
        set_last_error("lens_apply_layered_no_alloc: device ordinal out of range");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            set_last_error(std::string("lens_apply_layered_no_alloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
            return MMB_ERR_CUDA;
        }
        ptrs[device] = static_cast<unsigned char *>(p);
    }
    *out = ptrs[device];
    return MMB_OK;
}

mmb_status lens_apply_layered_no_alloc_host(int device, int n_layers, const int32_t *models, const double *coeffs,
                                            const double *cams, int direction, long long n, const double *xy_in,
                                            double *xy_out) {
    if (n_layers < 1 || n_layers > kLensMaxLayers) {
        set_last_error("lens_apply_layered_no_alloc: 1..8 layers");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    if (n > kNoAllocMaxPoints) {
        set_last_error("lens_apply_layered_no_alloc: n_points = " + std::to_string(n) + " exceeds the fixed scratch of " +
                       std::to_string(kNoAllocMaxPoints) + " points (MMB_LENS_NO_ALLOC_MAX_POINTS); split the buffer or "
                       "use mmb_lens_apply_layered");
        return MMB_ERR_INVALID_ARGUMENT;
    }
    MMB_CUDA_CHECK(cudaSetDevice(device));
    unsigned char *base = nullptr;
    const mmb_status st = noalloc_scratch_ptr(device, &base);
    if (st != MMB_OK) return st;
    if (n == 0) return MMB_OK;
    LensLayers LL;
    std::memset(&LL, 0, sizeof(LL));
    LL.n = n_layers;
    for (int k = 0; k < n_layers; ++k)
        make_bulk(models[k], coeffs ? coeffs + (size_t)MMB_LENS_MAX_COEFFS * k : nullptr, cams + 6 * (size_t)k, &LL.layer[k]);
    LensLayers *d_ll = reinterpret_cast<LensLayers *>(base);
    double2 *d_in = reinterpret_cast<double2 *>(base + kNoAllocHeader);
    double2 *d_out = d_in + kNoAllocMaxPoints;
    // legacy default stream, blocking copies: no stream, event or buffer is created here
    MMB_CUDA_CHECK(cudaMemcpy(d_ll, &LL, sizeof(LL), cudaMemcpyHostToDevice));
    MMB_CUDA_CHECK(cudaMemcpy(d_in, xy_in, sizeof(double2) * (size_t)n, cudaMemcpyHostToDevice));
    k_lens_bulk_layered<<<grid_for(n * kLensUnroll, device), 256>>>(d_ll, direction, n, d_in, d_out);
    MMB_CUDA_CHECK(cudaGetLastError());
    MMB_CUDA_CHECK(cudaMemcpy(xy_out, d_out, sizeof(double2) * (size_t)n, cudaMemcpyDeviceToHost));
    return MMB_OK;
}

// device-resident timing on a regular grid of points over the image (an ST-map)
__global__ void k_lens_grid(long long n, int width, double2 *pts) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long y = i / width, x = i - y * width;
        const long long height = (n + width - 1) / width;
        pts[i] = make_double2(((double)x + 0.5) / width - 0.5, ((double)y + 0.5) / (double)height - 0.5);
    }
}

mmb_status lens_bench(int device, int model, const double *coeffs, const double *cam, int direction, long long n,
                      int reps, double *ms_per_pass) {
    MMB_CUDA_CHECK(cudaSetDevice(device));
    if (n <= 0 || reps <= 0) return MMB_ERR_INVALID_ARGUMENT;
    LensBulk L;
    make_bulk(model, coeffs, cam, &L);
    double2 *d_in = nullptr, *d_out = nullptr;
    MMB_CUDA_CHECK(cudaMalloc((void **)&d_in, sizeof(double2) * n));
    cudaError_t e = cudaMalloc((void **)&d_out, sizeof(double2) * n);
    if (e != cudaSuccess) { cudaFree(d_in); set_last_error("lens_bench: out of memory"); return MMB_ERR_CUDA; }
    const unsigned grid = grid_for(n, device);
    k_lens_grid<<<grid, 256>>>(n, 4096, d_in);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_lens_bulk<<<grid, 256>>>(L, direction, n, d_in, d_out);  // warm-up
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) k_lens_bulk<<<grid, 256>>>(L, direction, n, d_in, d_out);
    cudaEventRecord(e1);
    e = cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) { set_last_error(std::string("lens_bench: ") + cudaGetErrorString(e)); return MMB_ERR_CUDA; }
    if (ms_per_pass) *ms_per_pass = ms / reps;
    return MMB_OK;
}

}  // namespace mmb
