// comm.hpp — rank-to-rank plumbing of the sharded state vector.
//
// One Simulator owns one shard (the amplitudes whose top log2(world) index bits equal its rank).
// Ranks are either processes (one per GPU, launched by torchrun/mpirun: NcclComm, NCCL over
// NVLink/NVSwitch) or threads of one process (LocalComm: a rendezvous + device-to-device copies,
// which also works with several ranks on ONE GPU — the "virtual shards" mode the parity tests use
// on a single-GPU box).  The data path needs exactly two collectives:
//   exchange  : personalised all-to-all of packed sub-blocks (global<->local qubit swap)
//   allgather : a few doubles per rank (reduction finishes; summed in rank order on the host, so
//               every rank obtains bit-identical results and takes identical branches)
#pragma once

#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace b200 {

struct Xfer {
    int peer;
    void *ptr;    // device pointer on the calling rank
    size_t bytes;
};

class Comm {
  public:
    virtual ~Comm() = default;
    int rank() const { return rank_; }
    int world() const { return world_; }
    // sends[i] goes to sends[i].peer; recvs[i] is filled by recvs[i].peer.  Every rank calls this
    // with matching lists (peer p's send to me has the size of my recv from p).  Work is enqueued
    // on `stream`; buffers may be reused once the stream reaches the end of the call.
    virtual void exchange(const std::vector<Xfer> &sends, const std::vector<Xfer> &recvs,
                          void *stream) = 0;
    // in: `count` doubles on the device at `send`; out: world*count doubles on the device at `recv`
    virtual void allgather(const double *send, double *recv, size_t count, void *stream) = 0;
    // host-side gather of `count` doubles per rank (recv holds world*count), blocking
    virtual void allgather_host(const double *send, double *recv, size_t count) = 0;
    virtual void barrier() = 0;
    virtual const char *kind() const = 0;
    // Peer-memory view of a device buffer that every rank holds in the same role (the state shard):
    // out[r] = a pointer kernels of THIS rank can dereference to reach rank r's buffer (out[rank] = mine).
    // Collective; every rank gets the same answer.  false = peer memory is not available (no P2P path
    // between the GPUs, IPC refused): the caller falls back to the staged exchange.
    virtual bool map_peers(void *mine, std::vector<void *> &out) = 0;
    // closes this rank's imports of the peers' buffers (call before ANY rank frees its buffer)
    virtual void unmap_peers(void *mine) = 0;
    // stream-ordered rendezvous: work enqueued on `stream` after this call starts only after every
    // rank's work enqueued before its own call has completed
    virtual void stream_barrier(void *stream) = 0;

  protected:
    int rank_ = 0, world_ = 1;
};

// spec: "local:<group-name>" (threads of this process) or "nccl:<256 hex digits of an ncclUniqueId>"
std::unique_ptr<Comm> make_comm(const std::string &spec, int world, int rank, int cuda_device,
                                void *stream);

// 128-byte NCCL unique id (rank 0 creates it, the host program distributes it); 0 on success
int nccl_unique_id(unsigned char out[128], std::string &err);

} // namespace b200
