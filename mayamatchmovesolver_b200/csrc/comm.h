// Collective plumbing of a sharded global solve: the only
// exchange step of the path is a sum (or max) all-reduce of device buffers.
//
// Two backends behind one interface:
//   * NCCL over NVLink / NVSwitch (one process per GPU).  libnccl.so.2 is
//     opened with dlopen the first time a communicator is asked for, so the
//     single-GPU product does not depend on NCCL being installed.
//   * "local": several handles of ONE process on ONE device (one host thread
//     each).  The reduction is a kernel over the ranks' buffers, ordered by a
//     host barrier.  It exists so the sharded LM logic can be exercised on a
//     one-GPU box; it is not a CPU path (every buffer stays in HBM).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "../../include/mmsolver_b200.h"

namespace mmb {

// every rank's peer-mapped big buffer as one rank sees them (base[rank] is its own), and a sequence
// word in this rank's arena that device code may advance (all ranks advance theirs alike)
struct PeerBig {
    double *base[8];
    unsigned *flags[8];     // a reserved flag region at the end of every arena: [2][8][kPeerFlagTiles] words
    int rank, world;
    size_t doubles;         // usable doubles of a big buffer (the flag region is not part of them)
    unsigned *seq;
};
constexpr int kPeerFlagTiles = 2080;   // lower-triangle tiles of a 64 x 64 block grid

struct Comm {
    int rank = 0, world = 1;
    virtual ~Comm() {}
    // in-place all-reduce of buf[0..count) on `stream`; sum, or max when is_max
    virtual mmb_status allreduce(double *buf, size_t count, bool is_max, cudaStream_t stream) = 0;
    virtual const char *backend() const = 0;
    // Peer-memory path (NVLink / NVSwitch, one process per GPU): the ranks map each other's arena
    // (CUDA IPC) and the all-reduce is a plain kernel that loads the peers' buffers and stores the
    // sum back to them, ordered by flag barriers in the same memory -- no library call, so the
    // collective can sit inside the conditional CUDA graph of the LM loop.
    // peer_buffer: a buffer of >= count doubles inside this rank's arena (the big all-reduce runs
    // in place on it), or nullptr when the path is unavailable / too small.
    virtual double *peer_buffer(size_t count) { (void)count; return nullptr; }
    // can every all-reduce of a solve (big: the peer buffer; small: up to max_small doubles) be
    // captured into a CUDA graph?
    virtual bool capturable(size_t max_small) const { (void)max_small; return false; }
    // the big buffers of all ranks, for kernels that exchange through them directly (false: no peer path)
    virtual bool peer_big(PeerBig *out) { (void)out; return false; }
    // a barrier of the last solve timed out (ranks out of step): results are not to be trusted
    virtual bool peer_error() { return false; }
};

mmb_status comm_unique_id(uint8_t id[MMB_COMM_ID_BYTES], int local_backend);
// collective over all ranks of the id; binds the communicator to `device`
mmb_status comm_create(const uint8_t id[MMB_COMM_ID_BYTES], int rank, int world, int device, Comm **out);

}  // namespace mmb
