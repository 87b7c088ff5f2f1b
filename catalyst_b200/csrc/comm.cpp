// comm.cpp — LocalComm (threads of one process, any number of ranks per GPU) and NcclComm
// (one process per GPU; NCCL is resolved with dlopen so that single-GPU use has no NCCL dependency
// and a host program that already loaded an NCCL — e.g. torch's — shares it).
#include "comm.hpp"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <chrono>
#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <sstream>
#include <stdexcept>

namespace b200 {

namespace {

struct CommError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

void cuda_ok(cudaError_t e, const char *what)
{
    if (e != cudaSuccess) {
        std::ostringstream s;
        s << "CUDA error " << (int)e << " (" << cudaGetErrorString(e) << ") in " << what;
        cudaGetLastError();
        throw CommError(s.str());
    }
}

// ---- LocalComm ------------------------------------------------------------------------------------
struct LocalGroup {
    std::mutex mu;
    std::condition_variable cv;
    int world = 0, arrived = 0, attached = 0;
    uint64_t generation = 0;
    std::vector<std::vector<Xfer>> sends; // per rank, valid between the two barriers of a call
    std::vector<const void *> ptrs;       // per rank: published pointer (allgather)
    void barrier()
    {
        std::unique_lock<std::mutex> lk(mu);
        const uint64_t gen = generation;
        if (++arrived == world) {
            arrived = 0;
            generation++;
            cv.notify_all();
        }
        else if (!cv.wait_for(lk, std::chrono::seconds(timeout_s), [&] { return generation != gen; })) {
            arrived--; // a peer never arrived (it failed or diverged): do not hang the caller
            throw CommError("local rank group: a peer did not reach the collective within " +
                            std::to_string(timeout_s) + " s");
        }
    }
    int timeout_s = 120;
};

std::mutex g_groups_mu;
std::map<std::string, std::shared_ptr<LocalGroup>> g_groups;

class LocalComm final : public Comm {
  public:
    LocalComm(const std::string &name, int world, int rank) : name_(name)
    {
        world_ = world;
        rank_ = rank;
        std::lock_guard<std::mutex> lk(g_groups_mu);
        auto &g = g_groups[name];
        if (!g) {
            g = std::make_shared<LocalGroup>();
            g->world = world;
            g->sends.resize(world);
            g->ptrs.resize(world);
        }
        if (g->world != world) throw CommError("local comm group '" + name + "': world size mismatch");
        g->attached++;
        grp_ = g;
    }
    ~LocalComm() override
    {
        std::lock_guard<std::mutex> lk(g_groups_mu);
        if (--grp_->attached == 0) g_groups.erase(name_);
    }
    void exchange(const std::vector<Xfer> &sends, const std::vector<Xfer> &recvs, void *stream) override
    {
        cudaStream_t st = (cudaStream_t)stream;
        cuda_ok(cudaStreamSynchronize(st), "LocalComm::exchange (pre)");
        grp_->sends[rank_] = sends;
        grp_->barrier();
        for (const Xfer &r : recvs) {
            const Xfer *src = nullptr;
            for (const Xfer &s : grp_->sends[r.peer])
                if (s.peer == rank_) src = &s;
            if (!src || src->bytes != r.bytes) throw CommError("LocalComm::exchange: unmatched transfer");
            cuda_ok(cudaMemcpyAsync(r.ptr, src->ptr, r.bytes, cudaMemcpyDefault, st), "LocalComm copy");
        }
        cuda_ok(cudaStreamSynchronize(st), "LocalComm::exchange (post)");
        grp_->barrier();
    }
    void allgather(const double *send, double *recv, size_t count, void *stream) override
    {
        cudaStream_t st = (cudaStream_t)stream;
        cuda_ok(cudaStreamSynchronize(st), "LocalComm::allgather (pre)");
        grp_->ptrs[rank_] = send;
        grp_->barrier();
        for (int r = 0; r < world_; r++)
            cuda_ok(cudaMemcpyAsync(recv + (size_t)r * count, grp_->ptrs[r], count * sizeof(double),
                                    cudaMemcpyDefault, st), "LocalComm allgather copy");
        cuda_ok(cudaStreamSynchronize(st), "LocalComm::allgather (post)");
        grp_->barrier();
    }
    void allgather_host(const double *send, double *recv, size_t count) override
    {
        grp_->ptrs[rank_] = send;
        grp_->barrier();
        for (int r = 0; r < world_; r++)
            std::memcpy(recv + (size_t)r * count, grp_->ptrs[r], count * sizeof(double));
        grp_->barrier();
    }
    void barrier() override { grp_->barrier(); }
    const char *kind() const override { return "local"; }
    bool map_peers(void *mine, std::vector<void *> &out) override
    {
        int my_dev = 0;
        cuda_ok(cudaGetDevice(&my_dev), "LocalComm::map_peers");
        grp_->ptrs[rank_] = mine;
        grp_->barrier();
        out.assign(world_, nullptr);
        bool ok = true;
        for (int r = 0; r < world_; r++) {
            out[r] = const_cast<void *>(grp_->ptrs[r]);
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, out[r]) != cudaSuccess) {
                cudaGetLastError();
                ok = false;
                continue;
            }
            if (at.device != my_dev) { // another GPU of this process: needs peer access
                int can = 0;
                cudaDeviceCanAccessPeer(&can, my_dev, at.device);
                if (!can) ok = false;
                else {
                    const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = false;
                    cudaGetLastError();
                }
            }
        }
        grp_->barrier();
        // agree on the answer
        double flag = ok ? 1.0 : 0.0;
        std::vector<double> all(world_);
        allgather_host(&flag, all.data(), 1);
        for (double f : all) ok = ok && f > 0.5;
        return ok;
    }
    void unmap_peers(void *) override {}
    void stream_barrier(void *stream) override
    {
        cuda_ok(cudaStreamSynchronize((cudaStream_t)stream), "LocalComm::stream_barrier");
        grp_->barrier();
    }

  private:
    std::string name_;
    std::shared_ptr<LocalGroup> grp_;
};

// ---- NcclComm ---------------------------------------------------------------------------------------
// Minimal mirror of the NCCL 2.x C API (stable ABI): only what the data path uses.
typedef struct { char internal[128]; } nccl_uid;
typedef void *nccl_comm;
enum { NCCL_INT8 = 0, NCCL_FLOAT64 = 8 };

struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(nccl_uid *) = nullptr;
    int (*CommInitRank)(nccl_comm *, int, nccl_uid, int) = nullptr;
    int (*CommDestroy)(nccl_comm) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Send)(const void *, size_t, int, int, nccl_comm, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, nccl_comm, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, nccl_comm, cudaStream_t) = nullptr;
    int (*CommGetAsyncError)(nccl_comm, int *) = nullptr; // optional (NCCL >= 2.4)
    std::string error;
};

NcclApi &nccl_api()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) {
            api.error = std::string("cannot load NCCL (libnccl.so.2): ") + dlerror();
            return;
        }
#define B200_SYM(field, name)                                                                      \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name));                    \
    if (!api.field) api.error = std::string("NCCL symbol missing: ") + name;
        B200_SYM(GetUniqueId, "ncclGetUniqueId")
        B200_SYM(CommInitRank, "ncclCommInitRank")
        B200_SYM(CommDestroy, "ncclCommDestroy")
        B200_SYM(GetErrorString, "ncclGetErrorString")
        B200_SYM(GroupStart, "ncclGroupStart")
        B200_SYM(GroupEnd, "ncclGroupEnd")
        B200_SYM(Send, "ncclSend")
        B200_SYM(Recv, "ncclRecv")
        B200_SYM(AllGather, "ncclAllGather")
#undef B200_SYM
        api.CommGetAsyncError = reinterpret_cast<decltype(api.CommGetAsyncError)>(dlsym(api.handle, "ncclCommGetAsyncError"));
    });
    return api;
}

class NcclComm final : public Comm {
  public:
    NcclComm(const std::string &hex, int world, int rank, int device, void *stream)
    {
        world_ = world;
        rank_ = rank;
        NcclApi &a = nccl_api();
        if (!a.error.empty()) throw CommError(a.error);
        if (hex.size() != 256) throw CommError("nccl id must be 256 hex digits");
        nccl_uid id;
        for (int i = 0; i < 128; i++) id.internal[i] = (char)std::stoi(hex.substr(2 * i, 2), nullptr, 16);
        cuda_ok(cudaSetDevice(device), "NcclComm cudaSetDevice");
        ok(a.CommInitRank(&comm_, world, id, rank), "ncclCommInitRank");
        cuda_ok(cudaMalloc(&dbuf_, dbuf_bytes_), "NcclComm scratch");
        (void)stream;
    }
    ~NcclComm() override
    {
        if (comm_) nccl_api().CommDestroy(comm_);
        if (dbuf_) cudaFree(dbuf_);
    }
    void exchange(const std::vector<Xfer> &sends, const std::vector<Xfer> &recvs, void *stream) override
    {
        NcclApi &a = nccl_api();
        ok(a.GroupStart(), "ncclGroupStart");
        for (const Xfer &r : recvs) ok(a.Recv(r.ptr, r.bytes, NCCL_INT8, r.peer, comm_, (cudaStream_t)stream), "ncclRecv");
        for (const Xfer &s : sends) ok(a.Send(s.ptr, s.bytes, NCCL_INT8, s.peer, comm_, (cudaStream_t)stream), "ncclSend");
        ok(a.GroupEnd(), "ncclGroupEnd");
        poll_async("qubit-swap exchange");
    }
    void allgather(const double *send, double *recv, size_t count, void *stream) override
    {
        ok(nccl_api().AllGather(send, recv, count, NCCL_FLOAT64, comm_, (cudaStream_t)stream), "ncclAllGather");
    }
    void allgather_host(const double *send, double *recv, size_t count) override
    {
        const size_t need = (size_t)(world_ + 1) * count * sizeof(double);
        if (need > dbuf_bytes_) {
            cuda_ok(cudaDeviceSynchronize(), "NcclComm resize");
            cudaFree(dbuf_);
            dbuf_bytes_ = need;
            cuda_ok(cudaMalloc(&dbuf_, dbuf_bytes_), "NcclComm scratch");
        }
        double *d = static_cast<double *>(dbuf_);
        cudaStream_t st = 0;
        cuda_ok(cudaMemcpyAsync(d, send, count * sizeof(double), cudaMemcpyHostToDevice, st), "allgather_host h2d");
        ok(nccl_api().AllGather(d, d + count, count, NCCL_FLOAT64, comm_, st), "ncclAllGather");
        cuda_ok(cudaMemcpyAsync(recv, d + count, (size_t)world_ * count * sizeof(double), cudaMemcpyDeviceToHost, st),
                "allgather_host d2h");
        cuda_ok(cudaStreamSynchronize(st), "allgather_host sync");
    }
    void barrier() override
    {
        double x = 0, y[64];
        std::vector<double> big;
        double *r = y;
        if (world_ > 64) {
            big.resize(world_);
            r = big.data();
        }
        allgather_host(&x, r, 1);
    }
    const char *kind() const override { return "nccl"; }
    bool map_peers(void *mine, std::vector<void *> &out) override
    {
        auto it = maps_.find(mine);
        if (it != maps_.end()) {
            out = it->second;
            return true;
        }
        // CUDA IPC: export my allocation, gather every rank's 64-byte handle, import the peers'
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
        double mineh[9] = {0};
        cudaIpcMemHandle_t h;
        bool good = cudaIpcGetMemHandle(&h, mine) == cudaSuccess;
        cudaGetLastError();
        std::memcpy(mineh, &h, 64);
        mineh[8] = good ? 1.0 : 0.0;
        std::vector<double> all((size_t)world_ * 9);
        allgather_host(mineh, all.data(), 9);
        for (int r = 0; r < world_; r++) good = good && all[(size_t)r * 9 + 8] > 0.5;
        std::vector<void *> ptrs(world_, nullptr);
        ptrs[rank_] = mine;
        if (good)
            for (int r = 0; r < world_; r++) {
                if (r == rank_) continue;
                cudaIpcMemHandle_t hr;
                std::memcpy(&hr, &all[(size_t)r * 9], 64);
                if (cudaIpcOpenMemHandle(&ptrs[r], hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    ptrs[r] = nullptr;
                    good = false;
                }
            }
        double flag = good ? 1.0 : 0.0;
        std::vector<double> flags(world_);
        allgather_host(&flag, flags.data(), 1);
        bool all_good = true;
        for (double f : flags) all_good = all_good && f > 0.5;
        if (!all_good) {
            for (int r = 0; r < world_; r++)
                if (r != rank_ && ptrs[r]) cudaIpcCloseMemHandle(ptrs[r]);
            cudaGetLastError();
            return false;
        }
        maps_[mine] = ptrs;
        out = ptrs;
        return true;
    }
    void unmap_peers(void *mine) override
    {
        auto it = maps_.find(mine);
        if (it == maps_.end()) return;
        for (int r = 0; r < world_; r++)
            if (r != rank_ && it->second[r]) cudaIpcCloseMemHandle(it->second[r]);
        cudaGetLastError();
        maps_.erase(it);
    }
    void stream_barrier(void *stream) override
    {
        if (!sync_buf_) cuda_ok(cudaMalloc(&sync_buf_, (size_t)(world_ + 1) * sizeof(double)), "NcclComm sync buffer");
        double *d = static_cast<double *>(sync_buf_);
        ok(nccl_api().AllGather(d, d + 1, 1, NCCL_FLOAT64, comm_, (cudaStream_t)stream), "ncclAllGather (barrier)");
        poll_async("stream barrier");
    }

  private:
    std::map<void *, std::vector<void *>> maps_;
    void *sync_buf_ = nullptr;
    // Errors of work already enqueued (a peer died, a link went down) surface asynchronously on the
    // communicator: poll after every data-path call so that the NEXT device call fails with RT_FAIL instead of
    // the job hanging in a later collective (ncclInProgress = 7 is not an error).
    void poll_async(const char *where)
    {
        NcclApi &a = nccl_api();
        if (!a.CommGetAsyncError) return;
        int st = 0;
        if (a.CommGetAsyncError(comm_, &st) != 0) return;
        if (st != 0 && st != 7) {
            const char *m = a.GetErrorString ? a.GetErrorString(st) : "?";
            throw CommError(std::string("NCCL asynchronous error after ") + where + ": " + m);
        }
    }
    void ok(int rc, const char *what)
    {
        if (rc != 0) {
            const char *m = nccl_api().GetErrorString ? nccl_api().GetErrorString(rc) : "?";
            throw CommError(std::string("NCCL error in ") + what + ": " + m);
        }
    }
    nccl_comm comm_ = nullptr;
    void *dbuf_ = nullptr;
    size_t dbuf_bytes_ = 1 << 16;
};

} // namespace

std::unique_ptr<Comm> make_comm(const std::string &spec, int world, int rank, int cuda_device, void *stream)
{
    if (world < 1 || (world & (world - 1))) throw CommError("world_size must be a power of two");
    if (rank < 0 || rank >= world) throw CommError("rank out of range");
    if (spec.rfind("local:", 0) == 0) return std::make_unique<LocalComm>(spec.substr(6), world, rank);
    if (spec.rfind("nccl:", 0) == 0) return std::make_unique<NcclComm>(spec.substr(5), world, rank, cuda_device, stream);
    throw CommError("unknown comm spec '" + spec + "' (expected local:<name> or nccl:<hex id>)");
}

int nccl_unique_id(unsigned char out[128], std::string &err)
{
    NcclApi &a = nccl_api();
    if (!a.error.empty()) {
        err = a.error;
        return 1;
    }
    nccl_uid id;
    const int rc = a.GetUniqueId(&id);
    if (rc != 0) {
        err = std::string("ncclGetUniqueId: ") + a.GetErrorString(rc);
        return 1;
    }
    std::memcpy(out, id.internal, 128);
    return 0;
}

} // namespace b200
