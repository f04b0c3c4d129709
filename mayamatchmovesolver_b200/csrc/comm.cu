// Communicator backends of the sharded solve.  See comm.h.
#include "comm.h"

#include <dlfcn.h>
#include <nccl.h>  // types and prototypes only: the library itself is opened with dlopen

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
        MMB_SYM(GetErrorString, "ncclGetErrorString")
#undef MMB_SYM
    });
    return &api;
}

struct NcclComm : Comm {
    ncclComm_t comm = nullptr;
    ~NcclComm() override { if (comm) nccl_api()->CommDestroy(comm); }
    mmb_status allreduce(double *buf, size_t count, bool is_max, cudaStream_t stream) override {
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
    *out = c;
    return MMB_OK;
}

}  // namespace mmb
