// Communicator backends of the sharded solve.  See comm.h.
#include "comm.h"

#include <dlfcn.h>
#include <nccl.h>  // types and prototypes only: the library itself is opened with dlopen

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>

#include "device_problem.h"

namespace mmb {

// ---- NCCL (dlopen) ------------------------------------------------------------------------------
namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

NcclApi *nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *nm : names) {
            api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) { api.error = std::string("libnccl.so.2 not found: ") + dlerror(); return; }
#define MMB_SYM(field, name)                                                        \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name));        \
    if (!api.field) { api.error = std::string("missing NCCL symbol ") + name; return; }
        MMB_SYM(GetUniqueId, "ncclGetUniqueId")
        MMB_SYM(CommInitRank, "ncclCommInitRank")
        MMB_SYM(CommDestroy, "ncclCommDestroy")
        MMB_SYM(AllReduce, "ncclAllReduce")
        MMB_SYM(AllGather, "ncclAllGather")
        MMB_SYM(GetErrorString, "ncclGetErrorString")
#undef MMB_SYM
    });
    return &api;
}

// ---- peer-memory all-reduce ---------------------------------------------------------------------
// Arena of one rank (cudaMalloc, exported with CUDA IPC, mapped by every peer):
//   [0, 4096)          header: arrive[q] (64-byte stride) written by rank q; own epoch / sequence
//                      counters (advanced by the kernels themselves, so a replayed graph works);
//                      error word
//   small slots        [parity 2][rank 8][kPeerSmall doubles]: contributions to small all-reduces
//   big buffer         the in-place operand of the big all-reduce (packed triangle of S), or the ranks'
//                      partial tiles of the reduced system (PeerBig)
//   tile flags         [parity 2][source rank 8][kPeerFlagTiles] words (PeerBig::flags)
constexpr int kPeerMax = 8;
constexpr size_t kPeerSmall = 64 * 1024;   // doubles per slot
constexpr size_t kPeerHeader = 4096;
constexpr size_t kPeerFlagBytes = (size_t)2 * 8 * kPeerFlagTiles * 4;   // tile flags of a sharded factorisation (PeerBig::flags)
struct PeerView {
    char *base[kPeerMax];
    int rank, world;
    size_t small_off, big_off;
};
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// one thread: tell every peer this rank has arrived at barrier `epoch`, wait for all of them
__device__ void peer_barrier(const PeerView &v) {
    unsigned *hdr = reinterpret_cast<unsigned *>(v.base[v.rank]);
    const unsigned e = hdr[256] + 1u;   // own epoch counter (offset 1024)
    hdr[256] = e;
    __threadfence_system();
    for (int q = 0; q < v.world; ++q) st_release_sys(reinterpret_cast<unsigned *>(v.base[q]) + 16 * v.rank, e);
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int q = 0; q < v.world; ++q) {
        const unsigned *flag = hdr + 16 * q;
        while ((int)(ld_acquire_sys(flag) - e) < 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > 20000000000ull) { hdr[272] = 1u; return; }   // 20 s: ranks out of step (error word, offset 1088)
            __nanosleep(40);
        }
    }
}
__global__ void k_peer_barrier(PeerView v) {
    if (threadIdx.x == 0) peer_barrier(v);
}
// small all-reduce: every rank stores its values into its slot of every arena, barrier, every rank
// folds the slots of its own arena in rank order (identical sums on every rank)
__global__ void __launch_bounds__(1024) k_peer_small(PeerView v, double *buf, int count, int is_max) {
    unsigned *hdr = reinterpret_cast<unsigned *>(v.base[v.rank]);
    __shared__ unsigned parity;
    if (threadIdx.x == 0) { parity = hdr[257] & 1u; hdr[257] += 1u; }   // own sequence counter (offset 1028)
    __syncthreads();
    const size_t slot = ((size_t)parity * kPeerMax + v.rank) * kPeerSmall;
    for (int q = 0; q < v.world; ++q) {
        double *dst = reinterpret_cast<double *>(v.base[q] + v.small_off) + slot;
        for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = buf[i];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) peer_barrier(v);
    __syncthreads();
    const double *mine = reinterpret_cast<const double *>(v.base[v.rank] + v.small_off) + (size_t)parity * kPeerMax * kPeerSmall;
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        double a = __ldcg(&mine[i]);
        for (int q = 1; q < v.world; ++q) {
            const double t = __ldcg(&mine[(size_t)q * kPeerSmall + i]);
            a = is_max ? fmax(a, t) : a + t;
        }
        buf[i] = a;
    }
}
// big all-reduce in place on the arenas' big buffers: rank r owns chunk r -- it loads that chunk from
// every arena (rank order), and stores the sum into every arena.  Between two barrier kernels.
__global__ void __launch_bounds__(256) k_peer_big(PeerView v, size_t count) {
    const size_t chunk = ((count + v.world - 1) / v.world + 1) & ~(size_t)1;
    const size_t lo = chunk * v.rank, hi = lo + chunk < count ? lo + chunk : count;
    for (size_t i = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
        double a = __ldcg(reinterpret_cast<const double *>(v.base[0] + v.big_off) + i);
        for (int q = 1; q < v.world; ++q) a += __ldcg(reinterpret_cast<const double *>(v.base[q] + v.big_off) + i);
        for (int q = 0; q < v.world; ++q) reinterpret_cast<double *>(v.base[q] + v.big_off)[i] = a;
    }
    __threadfence_system();
}

struct NcclComm : Comm {
    ncclComm_t comm = nullptr;
    // peer-memory path
    bool peer_ok = false;
    PeerView view{};
    size_t big_doubles = 0;
    void *arena = nullptr;
    ~NcclComm() override {
        if (peer_ok || arena) {
            cudaDeviceSynchronize();
            if (comm && peer_ok) {   // nobody unmaps while a peer may still touch the arenas
                double *d = nullptr;
                if (cudaMalloc((void **)&d, 8) == cudaSuccess) {
                    cudaMemset(d, 0, 8);
                    nccl_api()->AllReduce(d, d, 1, ncclDouble, ncclSum, comm, nullptr);
                    cudaDeviceSynchronize();
                    cudaFree(d);
                }
            }
            for (int q = 0; q < world; ++q)
                if (q != rank && view.base[q]) cudaIpcCloseMemHandle(view.base[q]);
            if (arena) cudaFree(arena);
        }
        if (comm) nccl_api()->CommDestroy(comm);
    }
    // collective: map every rank's arena; all ranks agree on the outcome
    void peer_setup() {
        const char *e = getenv("MMB_PEER");
        if ((e && atoi(e) == 0) || world > kPeerMax || world < 2) return;
        // room for the packed triangle of a mid-size reduced system, or for the partial tiles of a
        // cfg3-size one from 4 ranks (2 buffers x 4 ranks x 300 tiles x 32 KB = 79 MB)
        size_t big_mb = 96;
        if (const char *m = getenv("MMB_PEER_ARENA_MB")) big_mb = (size_t)std::max(1, atoi(m));
        big_doubles = big_mb * 1024 * 1024 / 8;
        const size_t small_bytes = 2 * kPeerMax * kPeerSmall * 8;
        const size_t bytes = kPeerHeader + small_bytes + big_doubles * 8 + kPeerFlagBytes;   // flag region after the big buffer
        NcclApi *api = nccl_api();
        double ok = 1.0;
        cudaIpcMemHandle_t mine{}, all[kPeerMax];
        char *d_send = nullptr, *d_recv = nullptr;
        double *d_ok = nullptr;
        if (cudaMalloc(&arena, bytes) != cudaSuccess) { arena = nullptr; ok = 0.0; }
        if (ok > 0 && (cudaMemset(arena, 0, bytes) != cudaSuccess || cudaIpcGetMemHandle(&mine, arena) != cudaSuccess)) ok = 0.0;
        cudaGetLastError();
        if (cudaMalloc((void **)&d_send, sizeof(mine)) != cudaSuccess || cudaMalloc((void **)&d_recv, sizeof(mine) * world) != cudaSuccess ||
            cudaMalloc((void **)&d_ok, 8) != cudaSuccess) { cudaGetLastError(); return; }   // not even the agreement can run: NCCL only
        cudaMemcpy(d_send, &mine, sizeof(mine), cudaMemcpyHostToDevice);
        bool gathered = api->AllGather(d_send, d_recv, sizeof(mine), ncclChar, comm, nullptr) == ncclSuccess &&
                        cudaDeviceSynchronize() == cudaSuccess &&
                        cudaMemcpy(all, d_recv, sizeof(mine) * world, cudaMemcpyDeviceToHost) == cudaSuccess;
        if (!gathered) ok = 0.0;
        for (int q = 0; q < world; ++q) view.base[q] = nullptr;
        if (ok > 0) {
            view.base[rank] = static_cast<char *>(arena);
            for (int q = 0; q < world && ok > 0; ++q) {
                if (q == rank) continue;
                void *ptr = nullptr;
                if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0.0; }
                view.base[q] = static_cast<char *>(ptr);
            }
        }
        // agreement: the path is on only if it is on everywhere
        cudaMemcpy(d_ok, &ok, 8, cudaMemcpyHostToDevice);
        if (api->AllReduce(d_ok, d_ok, 1, ncclDouble, ncclMin, comm, nullptr) != ncclSuccess || cudaDeviceSynchronize() != cudaSuccess) ok = 0.0;
        else cudaMemcpy(&ok, d_ok, 8, cudaMemcpyDeviceToHost);
        cudaFree(d_send); cudaFree(d_recv); cudaFree(d_ok);
        view.rank = rank; view.world = world;
        view.small_off = kPeerHeader; view.big_off = kPeerHeader + small_bytes;
        peer_ok = ok > 0;
        if (!peer_ok) {
            for (int q = 0; q < world; ++q) if (q != rank && view.base[q]) { cudaIpcCloseMemHandle(view.base[q]); view.base[q] = nullptr; }
            if (arena) { cudaFree(arena); arena = nullptr; }
        }
        if (getenv("MMB_LM_GRAPH_VERBOSE")) fprintf(stderr, "[mmb] rank %d: peer-memory all-reduce %s (arena %zu MB)\n", rank, peer_ok ? "on" : "off", bytes >> 20);
    }
    double *peer_buffer(size_t count) override {
        return (peer_ok && count <= big_doubles) ? reinterpret_cast<double *>(view.base[rank] + view.big_off) : nullptr;
    }
    bool capturable(size_t max_small) const override { return peer_ok && max_small <= kPeerSmall; }
    bool peer_big(PeerBig *out) override {
        if (!peer_ok) return false;
        for (int q = 0; q < 8; ++q) {
            out->base[q] = q < world ? reinterpret_cast<double *>(view.base[q] + view.big_off) : nullptr;
            out->flags[q] = q < world ? reinterpret_cast<unsigned *>(view.base[q] + view.big_off + big_doubles * 8) : nullptr;
        }
        out->rank = rank; out->world = world; out->doubles = big_doubles;
        out->seq = reinterpret_cast<unsigned *>(view.base[rank]) + 258;   // header word at offset 1032
        return true;
    }
    bool peer_error() override {
        if (!peer_ok) return false;
        unsigned err = 0;
        cudaMemcpy(&err, view.base[rank] + 1088, 4, cudaMemcpyDeviceToHost);
        return err != 0;
    }
    mmb_status allreduce(double *buf, size_t count, bool is_max, cudaStream_t stream) override {
        if (peer_ok && !is_max && buf == reinterpret_cast<double *>(view.base[rank] + view.big_off) && count <= big_doubles) {
            k_peer_barrier<<<1, 32, 0, stream>>>(view);          // every rank's operand is in its arena
            const unsigned blocks = (unsigned)std::min<size_t>((count / world + 255) / 256 + 1, 592);
            k_peer_big<<<blocks, 256, 0, stream>>>(view, count);
            k_peer_barrier<<<1, 32, 0, stream>>>(view);          // every chunk is back in every arena
            if (cudaGetLastError() != cudaSuccess) { set_last_error("peer all-reduce launch failed"); return MMB_ERR_CUDA; }
            return MMB_OK;
        }
        if (peer_ok && count <= kPeerSmall) {
            k_peer_small<<<1, 1024, 0, stream>>>(view, buf, (int)count, is_max ? 1 : 0);
            if (cudaGetLastError() != cudaSuccess) { set_last_error("peer all-reduce launch failed"); return MMB_ERR_CUDA; }
            return MMB_OK;
        }
        NcclApi *api = nccl_api();
        const ncclResult_t r = api->AllReduce(buf, buf, count, ncclDouble, is_max ? ncclMax : ncclSum, comm, stream);
        if (r != ncclSuccess) { set_last_error(std::string("ncclAllReduce: ") + api->GetErrorString(r)); return MMB_ERR_COMM; }
        return MMB_OK;
    }
    const char *backend() const override { return "nccl"; }
};

// ---- local: ranks are handles of this process on one device -------------------------------------
constexpr int kLocalMaxWorld = 8;
struct LocalPtrs { double *p[kLocalMaxWorld]; };

__global__ void k_local_allreduce(LocalPtrs ptrs, int world, size_t count, int is_max) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count; i += (size_t)gridDim.x * blockDim.x) {
        double v = ptrs.p[0][i];
        for (int r = 1; r < world; ++r) {
            const double t = ptrs.p[r][i];
            v = is_max ? fmax(v, t) : v + t;
        }
        for (int r = 0; r < world; ++r) ptrs.p[r][i] = v;
    }
}

struct LocalGroup {
    int world = 0;
    std::mutex mu;
    std::condition_variable cv;
    int arrived = 0;
    long long generation = 0;
    LocalPtrs ptrs{};
    cudaError_t status = cudaSuccess;
    void barrier() {
        std::unique_lock<std::mutex> lk(mu);
        const long long gen = generation;
        if (++arrived == world) { arrived = 0; ++generation; cv.notify_all(); }
        else cv.wait(lk, [&] { return generation != gen; });
    }
};

std::mutex g_local_mu;
std::map<std::string, std::weak_ptr<LocalGroup>> g_local_groups;
std::atomic<uint64_t> g_local_counter{1};
const char kLocalMagic[8] = {'M', 'M', 'B', 'L', 'O', 'C', 'A', 'L'};

struct LocalComm : Comm {
    std::shared_ptr<LocalGroup> g;
    mmb_status allreduce(double *buf, size_t count, bool is_max, cudaStream_t stream) override {
        // the rank's earlier work on its stream must be complete before rank 0 reads the buffer
        cudaError_t e = cudaStreamSynchronize(stream);
        g->ptrs.p[rank] = buf;
        g->barrier();
        if (rank == 0) {
            const unsigned blocks = (unsigned)std::min<size_t>((count + 255) / 256, 1184);
            k_local_allreduce<<<blocks, 256, 0, stream>>>(g->ptrs, world, count, is_max ? 1 : 0);
            cudaError_t e2 = cudaStreamSynchronize(stream);
            g->status = e2;
        }
        g->barrier();
        if (e == cudaSuccess) e = g->status;
        if (e != cudaSuccess) { set_last_error(std::string("local all-reduce: ") + cudaGetErrorString(e)); return MMB_ERR_CUDA; }
        return MMB_OK;
    }
    const char *backend() const override { return "local"; }
};

}  // namespace

mmb_status comm_unique_id(uint8_t id[MMB_COMM_ID_BYTES], int local_backend) {
    static_assert(sizeof(ncclUniqueId) <= MMB_COMM_ID_BYTES, "communicator id size");
    std::memset(id, 0, MMB_COMM_ID_BYTES);
    if (local_backend) {
        std::memcpy(id, kLocalMagic, 8);
        const uint64_t c = g_local_counter.fetch_add(1);
        std::memcpy(id + 8, &c, 8);
        return MMB_OK;
    }
    NcclApi *api = nccl_api();
    if (!api->lib || !api->GetUniqueId) { set_last_error(api->error); return MMB_ERR_COMM; }
    ncclUniqueId nid;
    const ncclResult_t r = api->GetUniqueId(&nid);
    if (r != ncclSuccess) { set_last_error(std::string("ncclGetUniqueId: ") + api->GetErrorString(r)); return MMB_ERR_COMM; }
    std::memcpy(id, &nid, sizeof(nid));
    return MMB_OK;
}

mmb_status comm_create(const uint8_t id[MMB_COMM_ID_BYTES], int rank, int world, int device, Comm **out) {
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) { set_last_error("bad rank / world size"); return MMB_ERR_INVALID_ARGUMENT; }
    MMB_CUDA_CHECK(cudaSetDevice(device));
    if (!std::memcmp(id, kLocalMagic, 8)) {
        if (world > kLocalMaxWorld) { set_last_error("local communicator: at most 8 ranks"); return MMB_ERR_INVALID_ARGUMENT; }
        std::shared_ptr<LocalGroup> g;
        {
            std::lock_guard<std::mutex> lk(g_local_mu);
            const std::string key(reinterpret_cast<const char *>(id), 16);
            g = g_local_groups[key].lock();
            if (!g) { g = std::make_shared<LocalGroup>(); g->world = world; g_local_groups[key] = g; }
        }
        if (g->world != world) { set_last_error("local communicator: world size differs between ranks"); return MMB_ERR_INVALID_ARGUMENT; }
        LocalComm *c = new LocalComm();
        c->rank = rank; c->world = world; c->g = g;
        *out = c;
        return MMB_OK;
    }
    NcclApi *api = nccl_api();
    if (!api->lib || !api->CommInitRank) { set_last_error(api->error); return MMB_ERR_COMM; }
    ncclUniqueId nid;
    std::memcpy(&nid, id, sizeof(nid));
    NcclComm *c = new NcclComm();
    c->rank = rank; c->world = world;
    const ncclResult_t r = api->CommInitRank(&c->comm, world, nid, rank);
    if (r != ncclSuccess) {
        set_last_error(std::string("ncclCommInitRank: ") + api->GetErrorString(r));
        c->comm = nullptr;
        delete c;
        return MMB_ERR_COMM;
    }
    c->peer_setup();   // collective: CUDA IPC mapping of the ranks' arenas (falls back to NCCL alone)
    *out = c;
    return MMB_OK;
}

}  // namespace mmb
