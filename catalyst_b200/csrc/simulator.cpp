// simulator.cpp — host layer of the B200 state-vector device (see simulator.hpp).
//
// Gate and measurement semantics follow the reference's device contract
// (runtime/include/QuantumDevice.hpp:215-567) as lightning.qubit implements it; conventions that
// the reference pins in-tree are cited where they are used.  The arithmetic runs only on the GPU:
// there is no CPU code path for any amplitude.
#include "simulator.hpp"
#include "comm.hpp"

#include <cuda_runtime.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include <nvtx3/nvToolsExt.h> // header-only; a no-op unless a profiler injects its library

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <functional>
#include <condition_variable>
#include <sstream>
#include <thread>

namespace b200 {

#define B200_FAIL(msg) throw ::b200::Error(std::string(msg))
#define B200_FAIL_IF(cond, msg)                                                                    \
    do {                                                                                           \
        if (cond) B200_FAIL(msg);                                                                  \
    } while (0)
#define CUDA_CHECK(expr)                                                                           \
    do {                                                                                           \
        cudaError_t _e = (cudaError_t)(expr);                                                      \
        if (_e != cudaSuccess) {                                                                   \
            std::ostringstream _s;                                                                 \
            _s << "CUDA error " << (int)_e << " (" << cudaGetErrorString(_e) << ") at "           \
               << __FILE__ << ":" << __LINE__ << " in " #expr;                                     \
            cudaGetLastError();                                                                    \
            throw ::b200::Error(_s.str());                                                         \
        }                                                                                          \
    } while (0)

static const cplx I_(0.0, 1.0);

struct NvtxRange { // ranges per flush / qubit swap / gradient: what nsys or ncu --nvtx groups the launches by
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

class DeviceBuffer {
  public:
    static bool dry; // B200SV_DRYRUN (development): plan + generate + compile only, no device memory
    explicit DeviceBuffer(size_t bytes) : bytes_(bytes)
    {
        if (dry) {
            p_ = std::malloc(64);
            return;
        }
        cudaError_t e = cudaMalloc(&p_, bytes ? bytes : 16);
        if (e == cudaErrorMemoryAllocation) { // buffers kept across ReleaseQubits by ANY device object of the process go first
            cudaGetLastError();
            Simulator::trim_kept_buffers();
            e = cudaMalloc(&p_, bytes ? bytes : 16);
        }
        CUDA_CHECK(e);
    }
    ~DeviceBuffer()
    {
        if (dry) std::free(p_);
        else cudaFree(p_);
    }
    DeviceBuffer(const DeviceBuffer &) = delete;
    DeviceBuffer &operator=(const DeviceBuffer &) = delete;
    double *d() const { return static_cast<double *>(p_); }
    void *p() const { return p_; }
    size_t bytes() const { return bytes_; }

  private:
    void *p_ = nullptr;
    size_t bytes_;
};

bool DeviceBuffer::dry = false;

// ---- allocations kept across ReleaseQubits -------------------------------------------------------------------
// Catalyst's runtime pools device objects: a program is device_init -> AllocateQubits -> ... -> ReleaseQubits ->
// device_release, and the next execution of the same program gets the same object back.  Freeing and re-allocating
// the state (16 GiB at 30 qubits) and the |psi|^2 buffer (8 GiB) costs ~10 ms per execution, so a released device
// KEEPS them for an allocation of the same size (kwarg keep_allocation, default on).  Kept buffers are not lost
// memory: any cudaMalloc of this library that fails first frees every kept buffer of every device object in the
// process (except shards that peers still map) and retries.
static std::mutex g_keep_mu;
static std::vector<Simulator *> g_sims;

void Simulator::trim_kept_buffers()
{
    std::lock_guard<std::mutex> lk(g_keep_mu);
    for (Simulator *sim : g_sims) {
        if (!sim->psi_keep_mapped_) sim->psi_keep_.reset();
        sim->probs_keep_.reset();
        if (!sim->grad_busy_) sim->grad_work_.clear();
    }
}

// -------------------------------------------------------------------------------------------------
Options parse_kwargs(const char *kwargs)
{
    // str(dict) as produced by the frontend, e.g. "{'shots': 0, 'mcmc': False}"
    // (reference parser: runtime/lib/backend/common/Utils.hpp:107-156 — strip braces, quotes and
    // spaces, split on ',' and ':').
    Options o;
    if (!kwargs) return o;
    std::string s;
    for (const char *p = kwargs; *p; p++)
        if (*p != '{' && *p != '}' && *p != '\'' && *p != '"' && *p != ' ') s.push_back(*p);
    std::stringstream ss(s);
    std::string item;
    while (std::getline(ss, item, ',')) {
        const auto pos = item.find(':');
        if (pos == std::string::npos) continue;
        const std::string k = item.substr(0, pos), v = item.substr(pos + 1);
        auto as_int = [&](int64_t dflt) -> int64_t {
            if (v == "True" || v == "true") return 1;
            if (v == "False" || v == "false") return 0;
            if (v == "None" || v.empty()) return dflt;
            try {
                return std::stoll(v);
            }
            catch (...) {
                throw Error("Invalid value for device kwarg '" + k + "': " + v);
            }
        };
        if (k == "shots") o.shots = as_int(0);
        else if (k == "mcmc") o.mcmc = as_int(0) != 0;
        else if (k == "num_burnin") o.num_burnin = as_int(o.num_burnin);
        else if (k == "kernel_name") o.kernel_name = v;
        else if (k == "tile_bits") o.tile_bits = (int)as_int(o.tile_bits);
        else if (k == "min_row_bits") o.min_row_bits = (int)as_int(o.min_row_bits);
        else if (k == "fuse") o.fuse = as_int(1) != 0;
        else if (k == "device") o.device = (int)as_int(-1);
        else if (k == "world_size") o.world = (int)as_int(1);
        else if (k == "rank") o.rank = (int)as_int(0);
        else if (k == "stage_mib") o.stage_mib = as_int(o.stage_mib);
        else if (k == "comm") o.comm = v;
        else if (k == "jit") {
            if (v == "off") o.jit = 0;
            else if (v == "auto") o.jit = 1;
            else if (v == "always") o.jit = 2;
            else o.jit = (int)as_int(o.jit);
        }
        else if (k == "jit_threshold") o.jit_threshold = (int)as_int(o.jit_threshold);
        else if (k == "jit_regbits") o.jit_regbits = (int)as_int(o.jit_regbits);
        else if (k == "jit_min_qubits") o.jit_min_qubits = (int)as_int(o.jit_min_qubits);
        else if (k == "jit_dbuf") o.jit_dbuf = (int)as_int(o.jit_dbuf);
        else if (k == "jit_fuse_diag") o.jit_fuse_diag = (int)as_int(o.jit_fuse_diag);
        else if (k == "jit_factor") o.jit_factor = (int)as_int(o.jit_factor);
        else if (k == "jit_uniform") o.jit_uniform = (int)as_int(o.jit_uniform);
        else if (k == "jit_memhint") o.jit_memhint = (int)as_int(o.jit_memhint);
        else if (k == "jit_direct_store") o.jit_direct_store = (int)as_int(o.jit_direct_store);
        else if (k == "lazy_adjoint") o.lazy_adjoint = as_int(1) != 0;
        else if (k == "p2p_swap") o.p2p_swap = (int)as_int(-1);
        else if (k == "swap_push") o.swap_push = (int)as_int(-1);
        else if (k == "dist_sampler") o.dist_sampler = (int)as_int(-1);
        else if (k == "plan_refine") o.plan_refine = (int)as_int(-1);
        else if (k == "plan_refine_mode") o.plan_refine_mode = (int)as_int(1);
        else if (k == "gather") {
            B200_FAIL_IF(v != "all" && v != "slice", "gather must be 'all' or 'slice'");
            o.gather_slices = v == "slice";
        }
        else if (k == "adaptive_rows") o.adaptive_rows = as_int(1) != 0;
        else if (k == "virtual_swap") o.virtual_swap = as_int(1) != 0;
        else if (k == "batch_overlaps") o.batch_overlaps = as_int(1) != 0;
        else if (k == "keep_allocation") o.keep_allocation = as_int(1) != 0;
        else if (k == "release_entangled") {
            B200_FAIL_IF(v != "error" && v != "project", "release_entangled must be 'error' or 'project'");
            o.release_project = v == "project";
        }
    }
    B200_FAIL_IF(o.world < 1 || (o.world & (o.world - 1)) || o.world > 64, "world_size must be a power of two (<= 64)");
    B200_FAIL_IF(o.rank < 0 || o.rank >= o.world, "rank out of range");
    B200_FAIL_IF(o.world > 1 && o.comm.empty(), "world_size > 1 needs a 'comm' kwarg (nccl:<id> or local:<group>)");
    B200_FAIL_IF(o.stage_mib < 1, "stage_mib out of range");
    B200_FAIL_IF(o.num_burnin < 0, "num_burnin out of range");
    B200_FAIL_IF(o.mcmc && o.kernel_name != "Local" && o.kernel_name != "NonZeroRandom",
                 "kernel_name must be 'Local' or 'NonZeroRandom'");
    B200_FAIL_IF(o.tile_bits < 1 || o.tile_bits > B200K_MAX_TILE_BITS, "tile_bits out of range");
    B200_FAIL_IF(o.min_row_bits < 0, "min_row_bits out of range");
    B200_FAIL_IF(o.jit < 0 || o.jit > 2, "jit must be 0/off, 1/auto or 2/always");
    B200_FAIL_IF(o.jit_regbits < 1 || o.jit_regbits > 5, "jit_regbits out of range (1..5)");
    B200_FAIL_IF(o.jit_threshold < 1, "jit_threshold out of range");
    o.min_row_bits = std::min(o.min_row_bits, o.tile_bits);
    // a tile has at most B200K_MAX_HIGH scattered bits: rows cannot be shorter than the rest of it
    o.min_row_bits = std::max(o.min_row_bits, o.tile_bits - B200K_MAX_HIGH);
    return o;
}

// -------------------------------------------------------------------------------------------------
Simulator::Simulator(const Options &opt) : opt_(opt)
{
    if (const char *e = std::getenv("B200SV_DRYRUN")) dry_ = std::atoi(e) != 0;
    if (const char *e = std::getenv("B200SV_PASS_OP_CAP")) pass_op_cap_ = std::max(8, std::min(460, std::atoi(e))); // experiments
    if (dry_) {
        // Development mode, no GPU needed: gate calls are canonicalised, planned into passes and the
        // specialised kernels are generated and compiled to cubins ($B200SV_JIT_DUMP), nothing is launched.
        // sharded registers too: every rank plans alone (the plan is a function of the op stream and the rank count
        // only), swaps just update the layout, and nothing that needs the other ranks' data can be called
        while ((1 << g_) < opt_.world) g_++;
        DeviceBuffer::dry = true;
        jit_mode_ = opt_.jit;
        if (const char *e = std::getenv("B200SV_JIT")) {
            const std::string v(e);
            jit_mode_ = v == "off" || v == "0" ? 0 : (v == "always" || v == "2" ? 2 : 1);
        }
        if (const char *e = std::getenv("B200SV_JIT_UNIFORM")) opt_.jit_uniform = std::atoi(e);
        if (const char *e = std::getenv("B200SV_JIT_DIRECT")) opt_.jit_direct_store = std::atoi(e);
        if (const char *e = std::getenv("B200SV_JIT_MEMHINT")) opt_.jit_memhint = std::atoi(e);
        if (const char *e = std::getenv("B200SV_PLAN_REFINE")) opt_.plan_refine = std::atoi(e);
        if (const char *e = std::getenv("B200SV_PLAN_REFINE_MODE")) opt_.plan_refine_mode = std::atoi(e);
        if (const char *e = std::getenv("B200SV_JIT_REGBITS")) opt_.jit_regbits = std::max(1, std::min(5, std::atoi(e)));
        if (const char *e = std::getenv("B200SV_MIN_ROW_BITS")) opt_.min_row_bits = std::max(std::max(0, opt_.tile_bits - B200K_MAX_HIGH), std::min(opt_.tile_bits, std::atoi(e)));
        return;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        B200_FAIL("b200sv: no CUDA device available (this device has no CPU fallback): " +
                  std::string(e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e)));
    }
    int dev = opt_.device;
    if (dev < 0) {
        const char *lr = std::getenv("LOCAL_RANK");
        dev = lr ? std::atoi(lr) % count : 0;
    }
    B200_FAIL_IF(dev >= count, "b200sv: CUDA device ordinal out of range");
    cuda_device_ = dev;
    jit_mode_ = opt_.jit;
    if (const char *e = std::getenv("B200SV_JIT")) { // test / benchmark override of the kwarg
        const std::string v(e);
        jit_mode_ = v == "off" || v == "0" ? 0 : (v == "always" || v == "2" ? 2 : 1);
    }
    if (const char *e = std::getenv("B200SV_JIT_DBUF")) opt_.jit_dbuf = std::atoi(e);
    if (const char *e = std::getenv("B200SV_JIT_MEMHINT")) opt_.jit_memhint = std::atoi(e);
    if (const char *e = std::getenv("B200SV_PLAN_REFINE")) opt_.plan_refine = std::atoi(e);
    if (const char *e = std::getenv("B200SV_PLAN_REFINE_MODE")) opt_.plan_refine_mode = std::atoi(e);
    if (const char *e = std::getenv("B200SV_JIT_FACTOR")) opt_.jit_factor = std::atoi(e);
    if (const char *e = std::getenv("B200SV_JIT_UNIFORM")) opt_.jit_uniform = std::atoi(e);
    if (const char *e = std::getenv("B200SV_P2P_SWAP")) opt_.p2p_swap = std::atoi(e);
    if (const char *e = std::getenv("B200SV_SWAP_PUSH")) opt_.swap_push = std::atoi(e);
    if (const char *e = std::getenv("B200SV_MIN_ROW_BITS")) opt_.min_row_bits = std::max(std::max(0, opt_.tile_bits - B200K_MAX_HIGH), std::min(opt_.tile_bits, std::atoi(e)));
    if (const char *e = std::getenv("B200SV_JIT_REGBITS")) opt_.jit_regbits = std::max(1, std::min(5, std::atoi(e)));
    CUDA_CHECK(cudaSetDevice(dev));
    cudaStream_t st;
    CUDA_CHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    stream_ = st;
    cudaEvent_t a, b;
    CUDA_CHECK(cudaEventCreate(&a));
    CUDA_CHECK(cudaEventCreate(&b));
    ev0_ = a;
    ev1_ = b;
    pinned_doubles_ = 1 << 16;
    CUDA_CHECK(cudaMallocHost((void **)&pinned_, pinned_doubles_ * sizeof(double)));
    scratch_ = std::make_unique<DeviceBuffer>((size_t)b200k_reduce_scratch_doubles() * sizeof(double));
    results_ = std::make_unique<DeviceBuffer>(4096 * sizeof(double));
    if (const char *e = std::getenv("B200SV_KEEP_ALLOC")) opt_.keep_allocation = std::atoi(e) != 0;
    while ((1 << g_) < opt_.world) g_++;
    if (opt_.world > 1) {
        try {
            comm_ = make_comm(opt_.comm, opt_.world, opt_.rank, cuda_device_, stream_);
        }
        catch (const std::exception &e) {
            B200_FAIL(std::string("b200sv: cannot join the rank group: ") + e.what());
        }
    }
    // last: a constructor that throws does not run the destructor, so nothing may fail after the registration
    std::lock_guard<std::mutex> lk(g_keep_mu);
    g_sims.push_back(this);
}

Simulator::~Simulator()
{
    if (dry_) return;
    cudaSetDevice(cuda_device_);
    if (stream_) cudaStreamSynchronize((cudaStream_t)stream_);
    drop_psi_peers();
    if (stage_peers_ok_ && comm_) comm_->unmap_peers(stage_peers_ptr_); // no rendezvous in a destructor
    psi_.reset();
    {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        g_sims.erase(std::remove(g_sims.begin(), g_sims.end(), this), g_sims.end());
        psi_keep_.reset();
        probs_keep_.reset();
    }
    tmp_.reset();
    acc_.reset();
    stage_.reset();
    comm_.reset();
    scratch_.reset();
    results_.reset();
    probs_buf_.reset();
    misc_.reset();
    matrix_arena_.reset();
    if (pinned_) cudaFreeHost(pinned_);
    for (int h = 0; h < kStage; h++) {
        if (stage_host_[h]) cudaFreeHost(stage_host_[h]);
        if (stage_ev_[h]) cudaEventDestroy((cudaEvent_t)stage_ev_[h]);
    }
    for (int h = 0; h < 2; h++) {
        if (ev_packed_[h]) cudaEventDestroy((cudaEvent_t)ev_packed_[h]);
        if (ev_moved_[h]) cudaEventDestroy((cudaEvent_t)ev_moved_[h]);
    }
    if (xfer_stream_) cudaStreamDestroy((cudaStream_t)xfer_stream_);
    collect_timing();
    for (auto &p : ev_free_) {
        cudaEventDestroy((cudaEvent_t)p.first);
        cudaEventDestroy((cudaEvent_t)p.second);
    }
    if (ev2_) cudaEventDestroy((cudaEvent_t)ev2_);
    if (ev3_) cudaEventDestroy((cudaEvent_t)ev3_);
    if (ev0_) cudaEventDestroy((cudaEvent_t)ev0_);
    if (ev1_) cudaEventDestroy((cudaEvent_t)ev1_);
    if (stream_) cudaStreamDestroy((cudaStream_t)stream_);
}

void Simulator::ensure_buffer(std::unique_ptr<DeviceBuffer> &buf, size_t bytes)
{
    if (buf && buf->bytes() >= bytes) return;
    if (buf) CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    buf.reset();
    buf = std::make_unique<DeviceBuffer>(bytes);
}

void Simulator::allocate(int64_t n)
{
    if (!dry_) CUDA_CHECK(cudaSetDevice(cuda_device_));
    B200_FAIL_IF(n < 0 || n > 40, "Invalid number of qubits");
    if (n_ > 0 || psi_) release();
    B200_FAIL_IF(sharded() && n < 2 * g_,
                 "AllocateQubits: a register sharded over " + std::to_string(opt_.world) +
                     " ranks needs at least " + std::to_string(2 * g_) + " qubits");
    n_ = n;
    nl_ = n - g_;
    rank_base_ = (uint64_t)opt_.rank << nl_;
    pos_.resize((size_t)n);
    for (int q = 0; q < (int)n; q++) pos_[q] = q;
    wire_bit_.resize((size_t)n);
    for (int64_t w = 0; w < n; w++) wire_bit_[(size_t)w] = (int)(n - 1 - w);
    const size_t bytes = sizeof(cplx) << nl_;
    try {
        // a shard whose peers hold a mapping of it (peer-to-peer qubit swap) is kept across
        // release/allocate of the same size, so the mapping stays valid
        {
            std::lock_guard<std::mutex> lk(g_keep_mu);
            if (psi_keep_ && psi_keep_->bytes() == bytes) psi_ = std::move(psi_keep_);
            psi_keep_mapped_ = false;
        }
        if (!psi_) {
            drop_psi_peers();
            {
                std::lock_guard<std::mutex> lk(g_keep_mu);
                psi_keep_.reset();
            }
            psi_ = std::make_unique<DeviceBuffer>(bytes); // (outside the lock: a failing cudaMalloc trims kept buffers)
        }
    }
    catch (const Error &e) {
        n_ = nl_ = 0;
        throw Error(std::string("AllocateQubits: cannot allocate the state vector in HBM: ") +
                    e.what());
    }
    if (!dry_) CUDA_CHECK(b200k_set_basis(psi_->d(), (int)nl_, opt_.rank == 0 ? 0 : -1, stream_));
    note_launch();
}

void Simulator::allocate_one()
{
    if (!psi_) {
        B200_FAIL_IF(sharded(), "AllocateQubit: the first allocation on a device sharded over " + std::to_string(opt_.world) +
                                    " ranks must be an array of at least " + std::to_string(2 * g_) + " qubits");
        allocate(1);
        return;
    }
    flush();
    // The new wire becomes the new LEAST significant logical bit, physically position 0 of every shard: each
    // rank doubles its own shard (out[2i] = in[i], out[2i+1] = 0), rank bits stay where they are.
    auto nb = std::make_unique<DeviceBuffer>(sizeof(cplx) << (nl_ + 1));
    CUDA_CHECK(b200k_append_qubit(nb->d(), psi_->d(), (int)nl_, stream_));
    note_launch();
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    drop_psi_peers(); // peers mapped the old allocation (collective: every rank is in this call)
    {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        psi_keep_.reset();
    }
    psi_ = std::move(nb);
    tmp_.reset();
    acc_.reset();
    n_ += 1;
    nl_ += 1;
    rank_base_ = (uint64_t)opt_.rank << nl_;
    for (auto &p : pos_) p += 1;      // every existing logical bit moved up by one physical position ...
    pos_.insert(pos_.begin(), 0);     // ... and one logical index: the new bit is logical 0 at position 0
    for (auto &b : wire_bit_) b += 1;
    wire_bit_.push_back(0);
}

// ReleaseQubit (reference: runtime/include/QuantumDevice.hpp:144; id bookkeeping of the caller:
// runtime/lib/backend/common/QubitManager.hpp:124-126 -- the wire leaves, later device indices move down).
// Composed from the paths the rest of the device already uses: marginal of the wire (read pass), collapse
// onto its branch with renormalisation, X if that branch was |1>, a SWAP with the qubit that sits on the top
// LOCAL index position (on a sharded register the executor brings a rank-bit qubit in with one global<->local
// swap), and the lower half of every shard is the (n-1)-qubit state.
void Simulator::release_one(int64_t wire)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    B200_FAIL_IF(wire < 0 || wire >= n_, "Invalid wire index " + std::to_string(wire));
    if (n_ == 1 && !sharded()) {
        release();
        return;
    }
    B200_FAIL_IF(sharded() && n_ - 1 < 2 * g_, "ReleaseQubit: a register sharded over " + std::to_string(opt_.world) +
                                                   " ranks keeps at least " + std::to_string(2 * g_) + " qubits");
    B200_FAIL_IF(recording_, "ReleaseQubit: not supported while the tape is recording");
    double p[2];
    probs({wire}, p, 1); // flushes the queue
    const int keep = p[1] > 0.5 + 1e-9 ? 1 : 0;
    // Deallocation presumes a qubit that is disentangled and in a basis state (what the reference's programs
    // release: frontend/test/pytest/test_dynamic_qubit_allocation.py).  Anything else would be a silent
    // projection of the remaining register, so it is an error unless release_entangled='project' asks for it.
    B200_FAIL_IF(!opt_.release_project && p[keep] < 1.0 - 1e-9,
                 "ReleaseQubit: wire " + std::to_string(wire) + " is not in a basis state (P(|" + std::to_string(keep) +
                     ">) = " + std::to_string(p[keep]) +
                     "); measure or reset it first, or create the device with release_entangled='project'");
    CUDA_CHECK(b200k_collapse(psi_->d(), (int)nl_, rank_base_, pos_[lbit(wire)], keep, 1.0 / std::sqrt(p[keep]), stream_));
    note_launch();
    std::vector<GateOp> ops;
    if (keep) canonicalize(ops, "PauliX", {}, {wire}, false, {}, {}, "");
    const int top_pos = (int)nl_ - 1; // top LOCAL physical position
    auto holder_of_top = [&]() {
        int q_top = -1;
        for (int q = 0; q < (int)n_; q++)
            if (pos_[q] == top_pos) q_top = q;
        return q_top;
    };
    int q_top = holder_of_top();
    if (wire_bit_[(size_t)wire] != q_top) {
        int64_t v = -1;
        for (int64_t u = 0; u < n_; u++)
            if (wire_bit_[(size_t)u] == q_top) v = u;
        canonicalize(ops, "SWAP", {}, {wire, v}, false, {}, {}, ""); // exchanges the contents of the two logical bits
        std::swap(wire_bit_[(size_t)wire], wire_bit_[(size_t)v]);    // ... and the labels
    }
    if (!ops.empty()) enqueue(std::move(ops));
    flush();
    // a global<->local swap during the flush may have moved another logical bit onto the top local position
    q_top = holder_of_top();
    if (wire_bit_[(size_t)wire] != q_top) {
        // (only reachable on a sharded register) the SWAP gate left the wire's logical bit elsewhere: bring it to
        // the top local position physically
        const int lw = wire_bit_[(size_t)wire];
        std::vector<GateOp> more;
        int64_t v = -1;
        for (int64_t u = 0; u < n_; u++)
            if (wire_bit_[(size_t)u] == q_top) v = u;
        make_local({psi_->d()}, pos_, (1ull << lw) | (1ull << q_top));
        canonicalize(more, "SWAP", {}, {wire, v}, false, {}, {}, "");
        std::swap(wire_bit_[(size_t)wire], wire_bit_[(size_t)v]);
        enqueue(std::move(more));
        flush();
        q_top = holder_of_top();
        B200_FAIL_IF(wire_bit_[(size_t)wire] != q_top, "internal: ReleaseQubit could not bring the wire to the top local bit");
    }
    const size_t bytes = sizeof(cplx) << (nl_ - 1);
    auto nb = std::make_unique<DeviceBuffer>(bytes);
    CUDA_CHECK(cudaMemcpyAsync(nb->p(), psi_->p(), bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    drop_psi_peers();
    {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        psi_keep_.reset();
    }
    psi_ = std::move(nb);
    tmp_.reset();
    acc_.reset();
    probs_buf_.reset();
    // logical bit q_top (physical position top_pos) leaves: higher logical bits and higher positions move down
    pos_.erase(pos_.begin() + q_top);
    for (auto &pp : pos_)
        if (pp > top_pos) pp -= 1;
    wire_bit_.erase(wire_bit_.begin() + wire);
    for (auto &bq : wire_bit_)
        if (bq > q_top) bq -= 1;
    n_ -= 1;
    nl_ -= 1;
    rank_base_ = (uint64_t)opt_.rank << nl_;
}

void Simulator::release()
{
    if (stream_ && !dry_) cudaStreamSynchronize((cudaStream_t)stream_);
    queue_.clear();
    {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        if (psi_ && (psi_peers_ok_ || (opt_.keep_allocation && !dry_))) { // peers still map this allocation / pooled device
            psi_keep_ = std::move(psi_);
            psi_keep_mapped_ = psi_peers_ok_;
        }
        if (probs_buf_ && opt_.keep_allocation && !dry_) probs_keep_ = std::move(probs_buf_);
    }
    psi_.reset();
    tmp_.reset();
    acc_.reset();
    probs_buf_.reset();
    if (!opt_.keep_allocation || dry_) {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        grad_work_.clear();
    }
    n_ = nl_ = 0;
    rank_base_ = 0;
    pos_.clear();
    obs_.clear();
    tape_.clear();
    tape_obs_.clear();
    recording_ = false;
}

void Simulator::reset_stats()
{
    collect_timing();
    stats_ = b200sv_stats{};
}

double Simulator::fp64_peak()
{
    B200_FAIL_IF(dry_, "no device in a dry run");
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    double rate = 0.0;
    CUDA_CHECK(b200k_dfma_peak(results_->d(), &rate, stream_));
    return rate;
}

// Counters as JSON, in the spirit of the reference's ResourceTracker report
// (runtime/lib/backend/null_qubit/NullQubit.hpp:104-112, ResourceTracker.hpp): what ran, not how fast.
std::string Simulator::stats_json() const
{
    const b200sv_stats s = stats();
    std::ostringstream o;
    o << "{\"num_qubits\": " << n_ << ", \"local_qubits\": " << nl_ << ", \"world_size\": " << opt_.world
      << ", \"rank\": " << opt_.rank << ", \"num_gates\": " << s.gates_enqueued << ", \"canonical_ops\": " << s.kernel_ops
      << ", \"tile_passes\": " << s.tile_passes << ", \"specialised_passes\": " << s.jit_passes
      << ", \"shared_memory_rounds\": " << s.rounds << ", \"relabelled_swaps\": " << s.relabelled_swaps
      << ", \"other_launches\": " << s.other_launches << ", \"qubit_swaps\": " << s.exchanges
      << ", \"peer_memory_swaps\": " << s.p2p_exchanges << ", \"push_swaps\": " << s.swap_push_exchanges
      << ", \"nvlink_bytes_sent\": " << s.exchange_bytes << ", \"algorithmic_bytes\": " << s.algorithmic_bytes
      << ", \"fp64_instructions\": " << s.fp64_ops << ", \"gate_device_ms\": " << s.gate_device_ms
      << ", \"swap_device_ms\": " << s.swap_device_ms << ", \"kernels_compiled\": " << s.jit_compiled
      << ", \"kernels_from_disk_cache\": " << s.jit_disk_hits << ", \"compilations_pending\": " << s.jit_pending
      << ", \"nvrtc_cpu_ms\": " << s.jit_compile_ms << ", \"h2d_bytes\": " << s.h2d_bytes << ", \"d2h_bytes\": " << s.d2h_bytes
      << "}";
    return o.str();
}

void Simulator::make_current()
{
    if (dry_) return;
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != cuda_device_) CUDA_CHECK(cudaSetDevice(cuda_device_));
}

void *Simulator::device_pointer() { return psi_ ? psi_->p() : nullptr; }

// ---- RNG ----------------------------------------------------------------------------------------
void Simulator::set_prng(b200sv_rng_next_fn fn, void *ctx)
{
    rng_fn_ = fn;
    rng_ctx_ = ctx;
    own_seeded_ = false;
}
void Simulator::seed(uint32_t s)
{
    own_gen_ = std::mt19937(s);
    own_seeded_ = true;
    rng_fn_ = nullptr;
}
uint32_t Simulator::prng_peek()
{
    if (own_seeded_) {
        std::mt19937 c = own_gen_;
        return (uint32_t)c();
    }
    return 0;
}
uint32_t Simulator::rng_next()
{
    if (rng_fn_) return rng_fn_(rng_ctx_);
    return (uint32_t)own_gen_();
}
// One 32-bit seed for a sampling / measurement call: the next output of the runtime's generator when one is
// set (reference stream), else entropy.  On a sharded register every rank must collapse onto the SAME
// outcome, so without a generator rank 0's entropy is distributed to all ranks.
uint32_t Simulator::fresh_seed()
{
    if (rng_fn_ || own_seeded_) return rng_next();
    uint32_t s = (uint32_t)std::random_device{}();
    if (sharded()) {
        double mine = (double)s;
        std::vector<double> all((size_t)opt_.world);
        try {
            comm_->allgather_host(&mine, all.data(), 1);
        }
        catch (const std::exception &e) {
            B200_FAIL(std::string("distributing the measurement seed across ranks failed: ") + e.what());
        }
        s = (uint32_t)all[0];
    }
    return s;
}
namespace {
struct BorrowedEngine { // same min/max as std::mt19937 => identical generate_canonical arithmetic
    using result_type = uint32_t;
    std::function<uint32_t()> next;
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return 0xffffffffu; }
    result_type operator()() { return next(); }
};
} // namespace

// ---- wires ----------------------------------------------------------------------------------------
int Simulator::lbit(int64_t wire) const
{
    B200_FAIL_IF(wire < 0 || wire >= n_, "Invalid wire index " + std::to_string(wire));
    return wire_bit_[(size_t)wire]; // canonically n-1-wire (wire 0 = most significant bit); SWAPs relabel
}

bool Simulator::wire_identity() const
{
    for (int64_t w = 0; w < n_; w++)
        if (wire_bit_[(size_t)w] != (int)(n_ - 1 - w)) return false;
    return true;
}

// Turns the wire relabelling accumulated by SWAP gates back into data movement: after this call wire w
// sits on logical bit n-1-w again (needed where the whole vector leaves in canonical order: State).
void Simulator::restore_wire_order()
{
    for (int64_t w = 0; w < n_; w++) {
        const int want = (int)(n_ - 1 - w);
        if (wire_bit_[(size_t)w] == want) continue;
        int64_t v = -1;
        for (int64_t u = 0; u < n_; u++)
            if (wire_bit_[(size_t)u] == want) v = u;
        std::vector<GateOp> ops;
        canonicalize(ops, "SWAP", {}, {w, v}, false, {}, {}, ""); // exchanges the contents of the two bits
        enqueue(std::move(ops));
        std::swap(wire_bit_[(size_t)w], wire_bit_[(size_t)v]);    // ... and the labels: the logical state is unchanged
    }
}
uint64_t Simulator::bit_of(int64_t wire) const { return 1ull << lbit(wire); }

void Simulator::check_wires(const std::vector<int64_t> &wires, const std::vector<int64_t> &cw) const
{
    uint64_t seen = 0;
    for (auto w : wires) {
        const uint64_t b = bit_of(w);
        B200_FAIL_IF(seen & b, "Repeated wire in operation");
        seen |= b;
    }
    for (auto w : cw) {
        const uint64_t b = bit_of(w);
        B200_FAIL_IF(seen & b, "Control wire overlaps another wire of the operation");
        seen |= b;
    }
}

// ---- gate canonicalisation ------------------------------------------------------------------------
static void dagger2(cplx *m)
{
    const cplx t01 = m[1];
    m[0] = std::conj(m[0]);
    m[1] = std::conj(m[2]);
    m[2] = std::conj(t01);
    m[3] = std::conj(m[3]);
}

void Simulator::canonicalize(std::vector<GateOp> &out, const std::string &name,
                             const std::vector<double> &p, const std::vector<int64_t> &wires,
                             bool inverse, const std::vector<int64_t> &cw,
                             const std::vector<bool> &cv, const std::string &pauli_word) const
{
    B200_FAIL_IF(cw.size() != cv.size(), "Control wires/values size mismatch");
    check_wires(wires, cw);
    uint64_t cmask = 0, cval = 0;
    for (size_t i = 0; i < cw.size(); i++) {
        cmask |= bit_of(cw[i]);
        if (cv[i]) cval |= bit_of(cw[i]);
    }
    auto need = [&](size_t nw, size_t np) {
        B200_FAIL_IF(wires.size() != nw, "Invalid number of wires for " + name);
        B200_FAIL_IF(p.size() != np, "Invalid number of parameters for " + name);
    };
    const size_t first = out.size();
    auto W = [&](size_t i) { return bit_of(wires[i]); };
    auto pair = [&](uint64_t xm, uint64_t a, uint64_t b, cplx m00, cplx m01, cplx m10, cplx m11,
                    uint64_t extra_c = 0) {
        GateOp g;
        g.kind = GateOp::PAIR;
        g.xmask = xm;
        g.a = a;
        g.b = b;
        g.m[0] = m00; g.m[1] = m01; g.m[2] = m10; g.m[3] = m11;
        g.cmask = cmask | extra_c;
        g.cval = cval | extra_c;
        out.push_back(std::move(g));
    };
    auto one = [&](cplx m00, cplx m01, cplx m10, cplx m11, size_t wi = 0, uint64_t extra_c = 0) {
        pair(W(wi), 0, W(wi), m00, m01, m10, m11, extra_c);
    };
    auto phase = [&](cplx e, uint64_t ones) { // e where all `ones` bits are 1 (and controls hold)
        GateOp g;
        g.kind = GateOp::PHASE;
        g.m[0] = e;
        g.cmask = cmask | ones;
        g.cval = cval | ones;
        out.push_back(std::move(g));
    };
    auto parity = [&](uint64_t zm, cplx e0, cplx e1, uint64_t extra_c = 0) {
        GateOp g;
        g.kind = GateOp::PARITY;
        g.zmask = zm;
        g.m[0] = e0;
        g.m[1] = e1;
        g.cmask = cmask | extra_c;
        g.cval = cval | extra_c;
        out.push_back(std::move(g));
    };
    const double th = p.empty() ? 0.0 : p[0];
    const double c = std::cos(th / 2), s = std::sin(th / 2);
    const double isq2 = 1.0 / std::sqrt(2.0);
    const cplx em = std::exp(-I_ * (th / 2)), ep = std::exp(I_ * (th / 2));
    auto rot = [&](double phi, double theta, double omega, cplx *m) {
        const double cc = std::cos(theta / 2), ss = std::sin(theta / 2);
        m[0] = std::exp(-I_ * ((phi + omega) / 2)) * cc;
        m[1] = -std::exp(I_ * ((phi - omega) / 2)) * ss;
        m[2] = std::exp(-I_ * ((phi - omega) / 2)) * ss;
        m[3] = std::exp(I_ * ((phi + omega) / 2)) * cc;
    };

    if (name == "Identity") {
        for (auto w : wires) (void)bit_of(w);
        return;
    }
    else if (name == "PauliX") { need(1, 0); one(0, 1, 1, 0); }
    else if (name == "PauliY") { need(1, 0); one(0, -I_, I_, 0); }
    else if (name == "PauliZ") { need(1, 0); phase(-1.0, W(0)); }
    else if (name == "Hadamard") { need(1, 0); one(isq2, isq2, isq2, -isq2); }
    else if (name == "S") { need(1, 0); phase(I_, W(0)); }
    else if (name == "T") { need(1, 0); phase(std::exp(I_ * (M_PI / 4)), W(0)); }
    else if (name == "PhaseShift") { need(1, 1); phase(std::exp(I_ * th), W(0)); }
    else if (name == "RX") { need(1, 1); one(c, -I_ * s, -I_ * s, c); }
    else if (name == "RY") { need(1, 1); one(c, -s, s, c); }
    else if (name == "RZ") { need(1, 1); parity(W(0), em, ep); }
    else if (name == "Rot") {
        need(1, 3);
        cplx m[4];
        rot(p[0], p[1], p[2], m);
        one(m[0], m[1], m[2], m[3]);
    }
    else if (name == "CNOT") { need(2, 0); one(0, 1, 1, 0, 1, W(0)); }
    else if (name == "CY") { need(2, 0); one(0, -I_, I_, 0, 1, W(0)); }
    else if (name == "CZ") { need(2, 0); phase(-1.0, W(0) | W(1)); }
    else if (name == "SWAP") { need(2, 0); pair(W(0) | W(1), W(1), W(0), 0, 1, 1, 0); }
    else if (name == "ISWAP") { need(2, 0); pair(W(0) | W(1), W(1), W(0), 0, I_, I_, 0); }
    else if (name == "PSWAP") {
        need(2, 1);
        const cplx e = std::exp(I_ * th);
        pair(W(0) | W(1), W(1), W(0), 0, e, e, 0);
    }
    else if (name == "IsingXX") {
        need(2, 1);
        pair(W(0) | W(1), 0, W(0) | W(1), c, -I_ * s, -I_ * s, c);
        pair(W(0) | W(1), W(1), W(0), c, -I_ * s, -I_ * s, c);
    }
    else if (name == "IsingYY") {
        need(2, 1);
        pair(W(0) | W(1), 0, W(0) | W(1), c, I_ * s, I_ * s, c);
        pair(W(0) | W(1), W(1), W(0), c, -I_ * s, -I_ * s, c);
    }
    else if (name == "IsingXY") { need(2, 1); pair(W(0) | W(1), W(1), W(0), c, I_ * s, I_ * s, c); }
    else if (name == "IsingZZ") { need(2, 1); parity(W(0) | W(1), em, ep); }
    else if (name == "ControlledPhaseShift") { need(2, 1); phase(std::exp(I_ * th), W(0) | W(1)); }
    else if (name == "CRX") { need(2, 1); one(c, -I_ * s, -I_ * s, c, 1, W(0)); }
    else if (name == "CRY") { need(2, 1); one(c, -s, s, c, 1, W(0)); }
    else if (name == "CRZ") { need(2, 1); parity(W(1), em, ep, W(0)); }
    else if (name == "CRot") {
        need(2, 3);
        cplx m[4];
        rot(p[0], p[1], p[2], m);
        one(m[0], m[1], m[2], m[3], 1, W(0));
    }
    else if (name == "CSWAP") {
        need(3, 0);
        pair(W(1) | W(2), W(2), W(1), 0, 1, 1, 0, W(0));
    }
    else if (name == "Toffoli") { need(3, 0); one(0, 1, 1, 0, 2, W(0) | W(1)); }
    else if (name == "SingleExcitation") {
        need(2, 1);
        pair(W(0) | W(1), W(1), W(0), c, -s, s, c);
    }
    else if (name == "DoubleExcitation") {
        need(4, 1);
        pair(W(0) | W(1) | W(2) | W(3), W(2) | W(3), W(0) | W(1), c, -s, s, c);
    }
    else if (name == "MultiRZ") {
        B200_FAIL_IF(p.size() != 1, "Invalid number of parameters for MultiRZ");
        B200_FAIL_IF(wires.empty(), "MultiRZ needs at least one wire");
        uint64_t zm = 0;
        for (size_t i = 0; i < wires.size(); i++) zm |= W(i);
        parity(zm, em, ep);
    }
    else if (name == "GlobalPhase") {
        B200_FAIL_IF(p.size() != 1, "Invalid number of parameters for GlobalPhase");
        phase(std::exp(-I_ * th), 0);
    }
    else if (name == "PCPhase") {
        // params = {theta, dim} (reference: runtime/lib/capi/RuntimeCAPI.cpp:470-485): the first
        // `dim` basis states of the wires get e^{i theta}, the rest e^{-i theta}.  Written as a
        // global phase e^{-i theta} times e^{2 i theta} on {val < dim}, the latter split into one
        // bit-pattern predicate per set bit of dim.
        B200_FAIL_IF(p.size() != 2, "Invalid number of parameters for PCPhase");
        const size_t k = wires.size();
        B200_FAIL_IF(k == 0 || k > 62, "Invalid number of wires for PCPhase");
        B200_FAIL_IF(p[1] < 0, "Invalid dimension for PCPhase");
        const uint64_t full = 1ull << k;
        const uint64_t dim = p[1] >= (double)full ? full : (uint64_t)p[1];
        phase(std::exp(-I_ * th), 0);
        const cplx e2 = std::exp(I_ * (2 * th));
        if (dim >= full) phase(e2, 0);
        else
            for (size_t j = 0; j < k; j++) { // j = position from the most significant wire
                const int sh = (int)(k - 1 - j);
                if (!((dim >> sh) & 1)) continue;
                // values with the bits above j equal to dim's and bit j = 0
                GateOp g;
                g.kind = GateOp::PHASE;
                g.m[0] = e2;
                g.cmask = cmask;
                g.cval = cval;
                for (size_t q = 0; q <= j; q++) {
                    g.cmask |= W(q);
                    if (q < j && ((dim >> (k - 1 - q)) & 1)) g.cval |= W(q);
                }
                out.push_back(std::move(g));
            }
    }
    else if (name == "PauliRot") {
        B200_FAIL_IF(p.size() != 1, "Invalid number of parameters for PauliRot");
        B200_FAIL_IF(pauli_word.size() != wires.size(), "Pauli word / wires length mismatch");
        GateOp g;
        g.kind = GateOp::PAULIROT;
        for (size_t i = 0; i < wires.size(); i++) {
            switch (pauli_word[i]) {
            case 'I': break;
            case 'X': g.xmask |= W(i); break;
            case 'Y': g.xmask |= W(i); g.zmask |= W(i); g.ny++; break;
            case 'Z': g.zmask |= W(i); break;
            default: B200_FAIL("Invalid Pauli word");
            }
        }
        if (g.xmask == 0) {
            if (g.zmask == 0) phase(em, 0);
            else parity(g.zmask, em, ep);
        }
        else {
            g.m[0] = cplx(c, s);
            g.cmask = cmask;
            g.cval = cval;
            out.push_back(std::move(g));
        }
    }
    else {
        B200_FAIL("Unsupported gate: " + name);
    }
    if (inverse) {
        for (size_t i = first; i < out.size(); i++) {
            GateOp &g = out[i];
            switch (g.kind) {
            case GateOp::PAIR: dagger2(g.m); break;
            case GateOp::PHASE: g.m[0] = std::conj(g.m[0]); break;
            case GateOp::PARITY: g.m[0] = std::conj(g.m[0]); g.m[1] = std::conj(g.m[1]); break;
            case GateOp::PAULIROT: g.m[0] = cplx(g.m[0].real(), -g.m[0].imag()); break;
            default: break;
            }
        }
        std::reverse(out.begin() + first, out.end()); // all emitted pieces commute; keep it exact
    }
}

void Simulator::canonicalize_matrix(std::vector<GateOp> &out, const CVec &matrix,
                                    const std::vector<int64_t> &wires, bool inverse,
                                    const std::vector<int64_t> &cw, const std::vector<bool> &cv) const
{
    B200_FAIL_IF(cw.size() != cv.size(), "Control wires/values size mismatch");
    B200_FAIL_IF(wires.empty(), "Invalid number of wires for MatrixOperation");
    check_wires(wires, cw);
    const size_t k = wires.size();
    B200_FAIL_IF(k > 10, "MatrixOperation on more than 10 wires is not supported by this device");
    const size_t dim = size_t(1) << k;
    B200_FAIL_IF(matrix.size() != dim * dim, "Invalid given matrix size");
    uint64_t cmask = 0, cval = 0;
    for (size_t i = 0; i < cw.size(); i++) {
        cmask |= bit_of(cw[i]);
        if (cv[i]) cval |= bit_of(cw[i]);
    }
    GateOp g;
    g.cmask = cmask;
    g.cval = cval;
    if (k == 1) {
        g.kind = GateOp::PAIR;
        g.xmask = bit_of(wires[0]);
        g.a = 0;
        g.b = g.xmask;
        for (int i = 0; i < 4; i++) g.m[i] = matrix[i];
        if (inverse) dagger2(g.m);
    }
    else {
        g.kind = GateOp::DENSE;
        auto m = std::make_shared<CVec>(dim * dim);
        for (size_t r = 0; r < dim; r++)
            for (size_t cc = 0; cc < dim; cc++)
                (*m)[r * dim + cc] = inverse ? std::conj(matrix[cc * dim + r]) : matrix[r * dim + cc];
        g.matrix = std::move(m);
        for (auto w : wires) {
            g.targets.push_back(lbit(w));
            g.xmask |= bit_of(w);
        }
    }
    out.push_back(std::move(g));
}

// ---- queue ------------------------------------------------------------------------------------------
static void matmul2(const cplx *a, const cplx *b, cplx *o)
{ // o = a * b
    const cplx r0 = a[0] * b[0] + a[1] * b[2], r1 = a[0] * b[1] + a[1] * b[3];
    const cplx r2 = a[2] * b[0] + a[3] * b[2], r3 = a[2] * b[1] + a[3] * b[3];
    o[0] = r0; o[1] = r1; o[2] = r2; o[3] = r3;
}

void Simulator::enqueue(std::vector<GateOp> &&ops)
{
    for (auto &g : ops) {
        if (opt_.fuse && g.kind == GateOp::PAIR && g.cmask == 0 && g.a == 0 && g.b == g.xmask &&
            !queue_.empty()) {
            GateOp &l = queue_.back();
            if (l.kind == GateOp::PAIR && l.cmask == 0 && l.xmask == g.xmask && l.a == 0 &&
                l.b == l.xmask) {
                matmul2(g.m, l.m, l.m); // new gate acts after the queued one
                continue;
            }
        }
        queue_.push_back(std::move(g));
    }
    if (queue_.size() >= 8192) flush();
}

void Simulator::named_op(const std::string &name, const std::vector<double> &params,
                         const std::vector<int64_t> &wires, bool inverse,
                         const std::vector<int64_t> &cw, const std::vector<bool> &cv,
                         const std::string &pauli_word)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    if (opt_.virtual_swap && !recording_ && cw.empty() && name == "SWAP") {
        // an uncontrolled SWAP moves no data: the two wires trade index bits.  (While a tape is being
        // recorded SWAPs stay real gates, so that the adjoint replay sees the labels it recorded.)
        B200_FAIL_IF(wires.size() != 2, "Invalid number of wires for SWAP");
        check_wires(wires, cw);
        std::swap(wire_bit_[(size_t)wires[0]], wire_bit_[(size_t)wires[1]]);
        stats_.gates_enqueued++;
        stats_.relabelled_swaps++;
        stats_.algorithmic_bytes += 32.0 * std::ldexp(1.0, (int)nl_);
        return;
    }
    std::vector<GateOp> ops;
    canonicalize(ops, name, params, wires, inverse, cw, cv, pauli_word);
    if (name != "Identity") {
        stats_.gates_enqueued++;
        if (name != "GlobalPhase" || !cw.empty())
            stats_.algorithmic_bytes += 32.0 * std::ldexp(1.0, (int)(nl_ - (int64_t)cw.size()));
    }
    enqueue(std::move(ops));
    if (recording_) // reference tape model: runtime/lib/backend/common/CacheManager.hpp:89-103
        tape_.push_back({name, params, wires, inverse, {}, cw, cv, pauli_word});
}

void Simulator::matrix_op(const CVec &matrix, const std::vector<int64_t> &wires, bool inverse,
                          const std::vector<int64_t> &cw, const std::vector<bool> &cv)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    std::vector<GateOp> ops;
    canonicalize_matrix(ops, matrix, wires, inverse, cw, cv);
    stats_.gates_enqueued++;
    stats_.algorithmic_bytes += 32.0 * std::ldexp(1.0, (int)(nl_ - (int64_t)cw.size()));
    enqueue(std::move(ops));
    if (recording_) tape_.push_back({"QubitUnitary", {}, wires, inverse, matrix, cw, cv, ""});
}

void Simulator::flush()
{
    if (queue_.empty()) return;
    if (dry_) {
        std::vector<GateOp> ops;
        ops.swap(queue_);
        execute({psi_->d()}, pos_, ops);
        return;
    }
    CUDA_CHECK(cudaSetDevice(cuda_device_));
    NvtxRange nvtx("b200sv::flush");
    std::vector<GateOp> ops;
    ops.swap(queue_);
    // timing on (benchmarks): one CUDA-event pair per flush, read back LAZILY (stats() / synchronize) so that the
    // host keeps enqueueing the next circuit while this one runs -- a synchronise per flush would serialise the
    // host's gate calls with the device
    std::pair<void *, void *> evp{nullptr, nullptr};
    if (timing_) {
        if (ev_free_.empty()) {
            cudaEvent_t a, b;
            CUDA_CHECK(cudaEventCreate(&a));
            CUDA_CHECK(cudaEventCreate(&b));
            ev_free_.push_back({a, b});
        }
        evp = ev_free_.back();
        ev_free_.pop_back();
        CUDA_CHECK(cudaEventRecord((cudaEvent_t)evp.first, (cudaStream_t)stream_));
    }
    execute({psi_->d()}, pos_, ops);
    if (timing_) {
        CUDA_CHECK(cudaEventRecord((cudaEvent_t)evp.second, (cudaStream_t)stream_));
        ev_pending_.push_back(evp);
        if (ev_pending_.size() > 256) collect_timing();
    }
}

void Simulator::collect_timing()
{
    for (auto &p : ev_pending_) {
        if (cudaEventSynchronize((cudaEvent_t)p.second) == cudaSuccess) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, (cudaEvent_t)p.first, (cudaEvent_t)p.second) == cudaSuccess) stats_.gate_device_ms += ms;
        }
        ev_free_.push_back(p);
    }
    cudaGetLastError();
    ev_pending_.clear();
}

void Simulator::synchronize()
{
    flush();
    if (dry_) return;
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

// ---- planner: queue -> fused tile passes --------------------------------------------------------------
// A pass stages tiles made of the low `c` index bits plus the high bits its ops need as
// non-diagonal targets.  Ops are taken greedily in program order; an op that does not fit is
// deferred and blocks later ops that do not commute with it (two ops commute when, on every
// shared bit, both act diagonally — controls count as diagonal).
static constexpr size_t kArenaBytes = size_t(64) << 20; // device arena of MatrixOperation matrices
static inline size_t arena_aligned(size_t bytes) { return (bytes + 255) & ~size_t(255); }

static int fit_rows(uint64_t req, int t)
{
    if (__builtin_popcountll(req) > t) return -1;
    int c = t;
    while (c > 0 && __builtin_popcountll(req >> c) > t - c) c--;
    return c;
}

void Simulator::run_ops(double *vec, std::vector<GateOp> &ops)
{
    const int t = (int)std::min<int64_t>(opt_.tile_bits, nl_);
    const int cmin = std::min(opt_.min_row_bits, t);
    const size_t LOOKAHEAD = 4096;
    const int refine_iters = opt_.plan_refine >= 0 ? opt_.plan_refine : (nl_ >= 27 ? 6 : 0);
    std::vector<char> done(ops.size(), 0);
    size_t first = 0, remaining = ops.size();
    std::vector<const GateOp *> pass;
    const bool jit_on = jit_mode_ == 2 || (jit_mode_ == 1 && nl_ >= opt_.jit_min_qubits);
    const int jit_k = std::min(opt_.jit_regbits, jit_mode_ == 2 ? t - 1 : t - 5);
    auto jit_dense = [&](const GateOp &op) {
        return jit_on && op.kind == GateOp::DENSE && (int)op.targets.size() <= jit_k &&
               (int64_t)op.matrix->size() * 16 <= 8192;
    };
    // One candidate pass for a given minimum row width: the op indices it would take, the union of
    // their non-diagonal bits and the resulting number of contiguous low bits.
    struct Cand {
        std::vector<size_t> idx;
        uint64_t req = 0;
        int c = 0;
        int cm = 0;          // the minimum row width it was built with
        uint16_t use[64] = {}; // how many of its ops act non-diagonally on each bit
    };
    // `forbid`: bits no op of this candidate may act on non-diagonally (such ops are deferred like any op that does
    // not fit) -- the refinement step below uses it to trade a tile bit that serves one op for one that serves many
    auto build = [&](int cmin_try, bool dense_pass, Cand &out, uint64_t forbid = 0) {
        uint64_t req = 0, blockedX = 0, blockedZ = 0;
        out.cm = cmin_try;
        std::memset(out.use, 0, sizeof(out.use));
        int c_cur = t;
        int64_t pass_params = 0;
        size_t dense_bytes = 0; // matrices of a dense pass live in the device arena until the pass has run
        out.idx.clear();
        const size_t cap = dense_pass ? (size_t)B200K_MAX_OPS : (size_t)pass_op_cap_;
        for (size_t i = first; i < ops.size() && out.idx.size() < cap && i - first < LOOKAHEAD; i++) {
            if (done[i]) continue;
            const GateOp &op = ops[i];
            const uint64_t xs = op.xmask, zs = op.zsupport();
            const bool commutes = !(xs & forbid) && !(xs & (blockedX | blockedZ)) && !(zs & blockedX);
            // a specialised pass carries its matrices in the kernel parameter block (<= 32 KB)
            const int64_t pcost = op.kind == GateOp::DENSE ? 2 * (int64_t)op.matrix->size() : 8;
            const size_t dbytes = op.kind == GateOp::DENSE ? arena_aligned(op.matrix->size() * sizeof(cplx)) : 0;
            const bool room = dense_pass ? (out.idx.empty() || dense_bytes + dbytes <= kArenaBytes)
                                         : (!jit_on || out.idx.empty() || pass_params + pcost <= 3800);
            if (commutes && room && (dense_pass || op.kind != GateOp::DENSE || jit_dense(op))) {
                const uint64_t nreq = req | xs;
                const int c = fit_rows(nreq, t);
                B200_FAIL_IF(c < 0 && out.idx.empty(),
                             "Operation acts non-diagonally on more qubits than one tile holds");
                if (c >= 0 && (c >= cmin_try || out.idx.empty())) {
                    req = nreq;
                    c_cur = c;
                    out.idx.push_back(i);
                    for (uint64_t m = xs; m; m &= m - 1) out.use[__builtin_ctzll(m)]++;
                    pass_params += pcost;
                    dense_bytes += dbytes;
                    if (!opt_.fuse) break;
                    continue;
                }
            }
            if (!opt_.fuse) break;
            blockedX |= xs;
            blockedZ |= zs;
        }
        out.req = req;
        out.c = c_cur;
    };
    // Relative duration of a pass by row width 2^c amplitudes (measured at 30 qubits, specialised
    // passes: 5.9 ms at c >= 7, 6.6 / 7.1 / 7.7 / 8.1 ms at c = 6 / 5 / 4 / 3): shorter rows cost DRAM
    // efficiency but free tile bits for the high qubits, so fewer passes are needed.  64-byte rows (c = 2):
    // 6.8-10.7 ms depending on the rounds they carry, weighted 1.6 (1.45 and 1.8 plan C3 / C5 into 17 / 49 and
    // 18 / 51 passes, 1.6 into 17 / 47); 32-byte rows measured 16-19 ms per pass and are never worth it.
    static const double cost_c2 = [] { const char *e = std::getenv("B200SV_COST_C2"); return e ? std::atof(e) : 1.6; }();
    static const double cost_blend = [] { const char *e = std::getenv("B200SV_COST_BLEND"); return e ? std::atof(e) : 1.0; }();
    auto pass_cost = [](int c) {
        const double k = c >= 7 ? 1.0 : (c == 6 ? 1.12 : (c == 5 ? 1.2 : (c == 4 ? 1.3 : (c == 3 ? 1.38 : (c == 2 ? cost_c2 : 3.0)))));
        return c >= 2 ? 1.0 + cost_blend * (k - 1.0) : k; // experiment knob: < 1 flattens the penalty of short rows
    };
    while (remaining) {
        // a pass that starts with a DENSE op runs on the first-generation kernel (which knows
        // DENSE); every other pass runs on the register-blocked kernel and defers DENSE ops
        // (small DENSE blocks run in registers inside the specialised passes when those are enabled)
        const bool dense_pass = ops[first].kind == GateOp::DENSE && !jit_dense(ops[first]);
        Cand best, cand;
        build(cmin, dense_pass, best);
        if (opt_.adaptive_rows && opt_.fuse && !dense_pass) {
            double best_score = (double)best.idx.size() / pass_cost(best.c);
            for (int cm = cmin + 1; cm <= std::min(7, t); cm++) {
                build(cm, dense_pass, cand);
                const double sc = (double)cand.idx.size() / pass_cost(cand.c);
                if (sc > best_score * 1.0001) {
                    best_score = sc;
                    std::swap(best, cand);
                }
            }
        }
        // Refinement (large states only: it costs about twenty scans of the queue per pass, hidden behind the previous
        // pass on the GPU).  The greedy scan hands tile bits to whichever ops come first; a SCATTERED tile bit (one at
        // or above the row width: bits inside the row are free) that ends up serving a single op may keep out a bit
        // that would serve ten.  So: forbid one of the three least-used scattered bits of the candidate (never one the
        // first op needs) -- whichever rebuild scores best --, keep the better of old and new, and repeat from the
        // rebuilt candidate.  Planner model on the BASELINE circuits (pass count / cost-weighted): C3 17 / 21.2 ->
        // 15 / 18.7, C5 at 30 qubits 47 / 61.3 -> 40 / 51.3, C5 at 33 local qubits 54 / 67.9 -> 43 / 55.1.
        if (refine_iters > 0 && opt_.adaptive_rows && opt_.fuse && !dense_pass && best.idx.size() > 1) {
            auto score = [&](const Cand &k) { return (double)k.idx.size() / pass_cost(k.c); };
            double best_score = score(best);
            const uint64_t keep = ops[best.idx[0]].xmask;
            uint64_t forbid = 0;
            Cand cur = best, trial;
            for (int it = 0; it < refine_iters; it++) {
                // mode 1 (measured): the single least-used bit of the candidate; mode 2: the three least-used SCATTERED
                // bits, each tried
                const int width = opt_.plan_refine_mode >= 2 ? std::min(8, opt_.plan_refine_mode + 1) : 1;
                int victims[8], nv = 0; // ascending by (use, position)
                for (int b = opt_.plan_refine_mode >= 2 ? cur.c : 0; b < 64; b++) {
                    if (!cur.use[b] || ((keep >> b) & 1)) continue;
                    int at = nv;
                    while (at > 0 && cur.use[b] < cur.use[victims[at - 1]]) at--;
                    if (at >= width) continue;
                    for (int q = std::min(nv, width - 1); q > at; q--) victims[q] = victims[q - 1];
                    victims[at] = b;
                    if (nv < width) nv++;
                }
                if (nv == 0) break;
                int chosen = -1;
                double chosen_score = -1.0;
                for (int v = 0; v < nv; v++) {
                    build(cur.cm, dense_pass, trial, forbid | (1ull << victims[v]));
                    const double sc = score(trial);
                    if (sc > chosen_score) { // ties: the lower bit position (victims are visited in ascending order of use)
                        chosen_score = sc;
                        chosen = victims[v];
                        std::swap(cand, trial);
                    }
                }
                forbid |= 1ull << chosen;
                if (chosen_score > best_score * 1.0001) {
                    best_score = chosen_score;
                    best = cand;
                }
                std::swap(cur, cand);
            }
        }
        pass.clear();
        for (size_t i : best.idx) {
            pass.push_back(&ops[i]);
            done[i] = 1;
            remaining--;
        }
        const uint64_t req = best.req;
        const int c_cur = best.c;
        // $B200SV_PLAN_DUMP=<file>: the schedule as the planner made it, one JSON line per pass with its canonical ops
        // in execution order (tests/test_planner_dryrun.py replays it in numpy against the oracle: reordering gates
        // is only legal where they commute, and that is decided here)
        static const char *plan_dump = std::getenv("B200SV_PLAN_DUMP");
        if (plan_dump) {
            std::ostringstream js;
            js.precision(17);
            js << "{\"c\": " << c_cur << ", \"ops\": [";
            for (size_t q = 0; q < pass.size(); q++) {
                const GateOp &g = *pass[q];
                js << (q ? ", " : "") << "{\"kind\": " << (int)g.kind << ", \"x\": " << g.xmask << ", \"cm\": " << g.cmask
                   << ", \"cv\": " << g.cval << ", \"a\": " << g.a << ", \"b\": " << g.b << ", \"z\": " << g.zmask
                   << ", \"ny\": " << g.ny << ", \"m\": [";
                for (int e = 0; e < 4; e++) js << (e ? ", " : "") << g.m[e].real() << ", " << g.m[e].imag();
                js << "]}";
            }
            js << "]}\n";
            if (FILE *f = std::fopen(plan_dump, "a")) {
                std::fputs(js.str().c_str(), f);
                std::fclose(f);
            }
        }
        while (first < ops.size() && done[first]) first++;
        // high tile bits: required positions >= c, padded with the lowest free positions
        std::vector<int> high;
        for (int p = c_cur; p < 64; p++)
            if ((req >> p) & 1) high.push_back(p);
        for (int p = c_cur; (int)high.size() < t - c_cur; p++)
            if (!((req >> p) & 1)) high.push_back(p);
        std::sort(high.begin(), high.end());
        if (dense_pass) launch_pass(vec, pass, high, c_cur);
        else if (launch_pass_jit(vec, pass, high, c_cur)) {}
        else if (std::any_of(pass.begin(), pass.end(), [](const GateOp *g) { return g->kind == GateOp::DENSE; })) {
            // the specialised kernel is not available (yet): DENSE ops one by one on the generic
            // kernel, the runs between them on the interpreter, in program order
            std::vector<const GateOp *> run;
            for (const GateOp *g : pass) {
                if (g->kind != GateOp::DENSE) {
                    run.push_back(g);
                    continue;
                }
                if (!run.empty()) launch_pass3(vec, run, high, c_cur);
                run.clear();
                launch_pass(vec, {g}, high, c_cur);
            }
            if (!run.empty()) launch_pass3(vec, run, high, c_cur);
        }
        else launch_pass3(vec, pass, high, c_cur);
    }
}

// ---- sharded execution ----------------------------------------------------------------
// Physical index = (rank << nl_) | local index.  `pos[q]` is the physical position of logical bit q.
// Diagonal action and control wires on rank bits need no data movement (the kernels evaluate them
// from rank_base); only a NON-diagonal target on a rank bit forces a global<->local qubit swap.
uint64_t Simulator::to_phys(const std::vector<int> &pos, uint64_t m) const
{
    uint64_t r = 0;
    while (m) {
        const int q = __builtin_ctzll(m);
        m &= m - 1;
        r |= 1ull << pos[q];
    }
    return r;
}

GateOp Simulator::phys_op(const std::vector<int> &pos, const GateOp &g) const
{
    GateOp o = g;
    o.xmask = to_phys(pos, g.xmask);
    o.cmask = to_phys(pos, g.cmask);
    o.cval = to_phys(pos, g.cval);
    o.a = to_phys(pos, g.a);
    o.b = to_phys(pos, g.b);
    o.zmask = to_phys(pos, g.zmask);
    for (auto &t : o.targets) t = pos[t];
    return o;
}

void Simulator::execute(const std::vector<double *> &vecs, std::vector<int> &pos, std::vector<GateOp> &ops)
{
    if (!sharded()) { // identity layout: logical == physical
        for (double *v : vecs) run_ops(v, ops);
        return;
    }
    const uint64_t local_mask_all = nl_ >= 64 ? ~0ull : ((1ull << nl_) - 1ull);
    const size_t LOOKAHEAD = 4096;
    std::vector<char> done(ops.size(), 0);
    size_t first = 0, remaining = ops.size();
    std::vector<GateOp> seg;
    while (remaining) {
        // every op that can run under the current layout without overtaking an op it does not
        // commute with (same rule as the tile planner: diagonal action commutes)
        uint64_t blockedX = 0, blockedZ = 0;
        seg.clear();
        for (size_t i = first; i < ops.size() && i - first < LOOKAHEAD; i++) {
            if (done[i]) continue;
            const GateOp &op = ops[i];
            const uint64_t xs = op.xmask, zs = op.zsupport();
            const bool commutes = !(xs & (blockedX | blockedZ)) && !(zs & blockedX);
            if (commutes && !(to_phys(pos, xs) & ~local_mask_all)) {
                seg.push_back(phys_op(pos, op));
                done[i] = 1;
                remaining--;
                continue;
            }
            blockedX |= xs;
            blockedZ |= zs;
        }
        while (first < ops.size() && done[first]) first++;
        if (!seg.empty())
            for (double *v : vecs) run_ops(v, seg);
        if (!remaining) break;
        // ---- choose the swap: the first blocked op must become runnable -------------------------
        const GateOp &head = ops[first];
        B200_FAIL_IF(__builtin_popcountll(head.xmask) > nl_,
                     "Operation acts non-diagonally on more qubits than one shard holds");
        // next non-diagonal use of every logical qubit among the remaining ops
        std::vector<size_t> next_use((size_t)n_, SIZE_MAX);
        for (size_t i = first; i < ops.size() && i - first < LOOKAHEAD; i++) {
            if (done[i]) continue;
            uint64_t m = ops[i].xmask;
            while (m) {
                const int q = __builtin_ctzll(m);
                m &= m - 1;
                if (next_use[q] == SIZE_MAX) next_use[q] = i;
            }
        }
        std::vector<int> inq, victims; // logical qubits: global ones to bring in, local candidates
        for (int q = 0; q < (int)n_; q++) {
            if (pos[q] >= nl_) {
                if (next_use[q] != SIZE_MAX) inq.push_back(q);
            }
            else if (!((head.xmask >> q) & 1)) victims.push_back(q);
        }
        std::sort(inq.begin(), inq.end(), [&](int a, int b) { return next_use[a] < next_use[b]; });
        std::sort(victims.begin(), victims.end(), [&](int a, int b) {
            if (next_use[a] != next_use[b]) return next_use[a] > next_use[b]; // farthest use first
            return pos[a] > pos[b];                                            // then high positions
        });
        std::vector<std::pair<int, int>> pairs;
        size_t vi = 0;
        for (int q : inq) {
            if (vi >= victims.size()) break;
            const int v = victims[vi];
            const bool mandatory = (head.xmask >> q) & 1;
            if (!mandatory && next_use[v] <= next_use[q]) continue; // evicting v would cost more
            pairs.push_back({pos[q], pos[v]});
            vi++;
        }
        B200_FAIL_IF(pairs.empty(), "internal: sharded planner found no swap to make progress");
        swap_bits(vecs, pos, pairs);
        B200_FAIL_IF(to_phys(pos, head.xmask) & ~local_mask_all,
                     "internal: sharded planner could not localise an operation");
    }
}

void Simulator::make_local(const std::vector<double *> &vecs, std::vector<int> &pos, uint64_t logical_mask)
{
    if (!sharded()) return;
    std::vector<int> inq, victims;
    for (int q = 0; q < (int)n_; q++) {
        const bool wanted = (logical_mask >> q) & 1;
        if (wanted && pos[q] >= nl_) inq.push_back(q);
        if (!wanted && pos[q] < nl_) victims.push_back(q);
    }
    if (inq.empty()) return;
    B200_FAIL_IF(inq.size() > victims.size(),
                 "Operation acts non-diagonally on more qubits than one shard holds");
    std::sort(victims.begin(), victims.end(), [&](int a, int b) { return pos[a] > pos[b]; });
    std::vector<std::pair<int, int>> pairs;
    for (size_t i = 0; i < inq.size(); i++) pairs.push_back({pos[inq[i]], pos[victims[i]]});
    swap_bits(vecs, pos, pairs);
}

// Exchange m rank bits with m local bits: a personalised all-to-all inside groups of 2^m ranks.
// Rank pattern a sends, to the rank with pattern b != a, its sub-block whose local bits spell b, and
// receives that rank's sub-block (whose local bits spell a) into the same slots; chunked through a
// bounded staging buffer (pack -> exchange -> unpack on one stream).
void Simulator::swap_bits(const std::vector<double *> &vecs, std::vector<int> &pos,
                          std::vector<std::pair<int, int>> pairs)
{
    if (pairs.empty()) return;
    if (!timing_) {
        swap_bits_impl(vecs, pos, std::move(pairs));
        return;
    }
    // timing on (benchmarks): the swap's share of the step, by CUDA events on the compute stream
    if (!ev2_) {
        cudaEvent_t a, b;
        CUDA_CHECK(cudaEventCreate(&a));
        CUDA_CHECK(cudaEventCreate(&b));
        ev2_ = a;
        ev3_ = b;
    }
    CUDA_CHECK(cudaEventRecord((cudaEvent_t)ev2_, (cudaStream_t)stream_));
    swap_bits_impl(vecs, pos, std::move(pairs));
    CUDA_CHECK(cudaEventRecord((cudaEvent_t)ev3_, (cudaStream_t)stream_));
    CUDA_CHECK(cudaEventSynchronize((cudaEvent_t)ev3_));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, (cudaEvent_t)ev2_, (cudaEvent_t)ev3_));
    stats_.swap_device_ms += ms;
}

void Simulator::swap_bits_impl(const std::vector<double *> &vecs, std::vector<int> &pos,
                               std::vector<std::pair<int, int>> pairs)
{
    const int m = (int)pairs.size();
    if (m == 0) return;
    NvtxRange nvtx("b200sv::qubit_swap");
    B200_FAIL_IF(m > 8 || m > g_, "internal: too many bits in one swap");
    if (dry_) { // planning only: the layout changes, no data moves
        for (auto &pr : pairs) {
            B200_FAIL_IF(pr.first < nl_ || pr.second >= nl_, "internal: malformed swap");
            int qg = -1, ql = -1;
            for (int q = 0; q < (int)n_; q++) {
                if (pos[q] == pr.first) qg = q;
                if (pos[q] == pr.second) ql = q;
            }
            std::swap(pos[qg], pos[ql]);
        }
        stats_.exchanges++;
        stats_.exchange_bytes += (double)(sizeof(cplx) << nl_) * (1.0 - std::ldexp(1.0, -m)) * (double)vecs.size();
        return;
    }
    std::sort(pairs.begin(), pairs.end(), [](auto &x, auto &y) { return x.second < y.second; });
    uint8_t lpos[8];
    int gsel[8];
    for (int i = 0; i < m; i++) {
        B200_FAIL_IF(pairs[i].first < nl_ || pairs[i].second >= nl_, "internal: malformed swap");
        lpos[i] = (uint8_t)pairs[i].second;
        gsel[i] = pairs[i].first - (int)nl_;
    }
    uint32_t mine = 0;
    for (int i = 0; i < m; i++) mine |= (uint32_t)((opt_.rank >> gsel[i]) & 1) << i;
    auto peer_of = [&](uint32_t b) {
        int r = opt_.rank;
        for (int i = 0; i < m; i++) r = (r & ~(1 << gsel[i])) | (int)((b >> i) & 1u) << gsel[i];
        return r;
    };
    const uint32_t npat = 1u << m;
    const uint64_t sub = 1ull << (nl_ - m); // amplitudes per sub-block
    // Peer-to-peer path (the state shard only; gradient work vectors use the staged path): every
    // pairwise sub-block exchange runs as ONE kernel over the peer's mapped shard, bracketed by two
    // stream-ordered rendezvous.  No staging buffer, no pack/unpack pass, no host synchronisation.
    // auto (-1): on between threads of one process (where the staged path is a chain of memcpys), off over
    // NCCL, whose grouped send/recv drives all peers at once and measured faster on 8 GPUs
    const bool want_p2p = opt_.p2p_swap == 1 || (opt_.p2p_swap < 0 && std::string(comm_->kind()) == "local");
    if (want_p2p && vecs.size() == 1 && psi_ && vecs[0] == psi_->d() && sub >= 2 && ensure_psi_peers()) {
        try {
            comm_->stream_barrier(stream_);
            for (uint32_t b = 0; b < npat; b++) {
                if (b == mine) continue;
                const bool lower = mine < b;
                CUDA_CHECK(b200k_swap_p2p(vecs[0], static_cast<double *>(psi_peers_[(size_t)peer_of(b)]), lpos, m, mine,
                                          b, lower ? 0 : sub / 2, lower ? sub / 2 : sub, stream_));
                note_launch();
                stats_.exchange_bytes += (double)sub * sizeof(cplx);
            }
            comm_->stream_barrier(stream_);
        }
        catch (const std::exception &e) {
            B200_FAIL(std::string("peer-to-peer qubit swap failed: ") + e.what());
        }
        stats_.exchanges++;
        stats_.p2p_exchanges++;
        for (int i = 0; i < m; i++) {
            int qg = -1, ql = -1;
            for (int q = 0; q < (int)n_; q++) {
                if (pos[q] == pairs[i].first) qg = q;
                if (pos[q] == pairs[i].second) ql = q;
            }
            std::swap(pos[qg], pos[ql]);
        }
        return;
    }
    // Push path (default over NCCL process groups of TWO ranks): every rank maps its peers' receive staging once; per chunk
    // ONE kernel reads this rank's sub-blocks and writes them straight into all peers' staging over NVLink while
    // it scatters the previously received chunk into the slots that chunk vacated; a stream-ordered rendezvous
    // separates the rounds.  Against the staged path below: no pack pass, no send-side staging, no NCCL transfer.
    // auto (-1), from the A/B on B200s (profiles/README.md r2_v3; 34 qubits): the push kernel wins for pairwise swaps
    // (2 GPUs: 601 ms of swap per circuit against 788 ms staged) and loses when a swap spans 4 or 8 ranks (8 GPUs:
    // 592 ms against 452 ms: its strided gather sits in the same dependency chain as the NVLink writes, while the
    // staged path hides pack / unpack behind the NCCL transfer of the neighbouring chunk)
    // (decided per communicator, not per swap, so that every run uses exactly one of the measured configurations)
    const bool want_push = opt_.swap_push == 1 || (opt_.swap_push < 0 && std::string(comm_->kind()) == "nccl" && opt_.world == 2);
    if (want_push && npat - 1 <= (uint32_t)B200K_PUSH_MAX_PEERS) {
        const uint64_t budget = ((uint64_t)opt_.stage_mib << 20) / sizeof(cplx);
        uint64_t chunk = budget / (2ull * (npat - 1));
        if (chunk > sub) chunk = sub;
        B200_FAIL_IF(chunk == 0, "stage_mib too small for the qubit swap");
        const uint64_t half_amps = (uint64_t)(npat - 1) * chunk;
        if (ensure_stage_peers(2ull * half_amps * sizeof(cplx))) {
            cplx *stage = static_cast<cplx *>(stage_->p());
            struct Job {
                double *vec;
                uint64_t j0, cnt;
            };
            std::vector<Job> jobs;
            for (double *vec : vecs)
                for (uint64_t j0 = 0; j0 < sub; j0 += chunk) jobs.push_back({vec, j0, std::min(chunk, sub - j0)});
            std::vector<uint32_t> pats;
            std::vector<uint64_t> my_slot_there, their_slot_here;
            std::vector<cplx *> peer_stage;
            for (uint32_t b = 0; b < npat; b++) {
                if (b == mine) continue;
                pats.push_back(b);
                my_slot_there.push_back(mine < b ? mine : mine - 1);
                their_slot_here.push_back(b < mine ? b : b - 1);
                peer_stage.push_back(static_cast<cplx *>(stage_peers_[(size_t)peer_of(b)]));
            }
            const int np = (int)pats.size();
            try {
                for (size_t k = 0; k <= jobs.size(); k++) {
                    const bool do_push = k < jobs.size(), do_unpack = k >= 1;
                    std::vector<double *> dst((size_t)np);
                    std::vector<const double *> src((size_t)np);
                    for (int p = 0; p < np; p++) {
                        dst[(size_t)p] = reinterpret_cast<double *>(peer_stage[(size_t)p] + (k & 1) * half_amps + my_slot_there[(size_t)p] * chunk);
                        src[(size_t)p] = reinterpret_cast<const double *>(stage + ((k + 1) & 1) * half_amps + their_slot_here[(size_t)p] * chunk);
                    }
                    const Job *jp = do_push ? &jobs[k] : nullptr, *ju = do_unpack ? &jobs[k - 1] : nullptr;
                    CUDA_CHECK(b200k_swap_push(jp ? jp->vec : vecs[0], ju ? ju->vec : vecs[0], dst.data(), src.data(), pats.data(), np,
                                               lpos, m, jp ? jp->j0 : 0, jp ? jp->cnt : 0, ju ? ju->j0 : 0, ju ? ju->cnt : 0, stream_));
                    note_launch();
                    if (jp) stats_.exchange_bytes += (double)(npat - 1) * (double)jp->cnt * sizeof(cplx);
                    comm_->stream_barrier(stream_); // round k landed everywhere; staging half (k+1)&1 is free again
                }
            }
            catch (const std::exception &e) {
                B200_FAIL(std::string("push-mode qubit swap failed: ") + e.what());
            }
            stats_.exchanges++;
            stats_.swap_push_exchanges++;
            for (int i = 0; i < m; i++) {
                int qg = -1, ql = -1;
                for (int q = 0; q < (int)n_; q++) {
                    if (pos[q] == pairs[i].first) qg = q;
                    if (pos[q] == pairs[i].second) ql = q;
                }
                std::swap(pos[qg], pos[ql]);
            }
            return;
        }
    }
    // Staged path, software-pipelined over two halves of the staging buffer: the transfers run on their own
    // stream, so that packing chunk k+1 and unpacking chunk k-1 (HBM passes on the compute stream) overlap
    // the NVLink transfer of chunk k.  Order on the compute stream: pack 0, pack 1, unpack 0, pack 2,
    // unpack 1, ...; a half is packed again only after the unpack that emptied it, and a transfer starts
    // only after its pack (events), so no buffer is written while it is still being read.
    const uint64_t stage_amps = ((uint64_t)opt_.stage_mib << 20) / sizeof(cplx);
    uint64_t chunk = stage_amps / (4ull * (npat - 1));
    if (chunk > sub) chunk = sub;
    B200_FAIL_IF(chunk == 0, "stage_mib too small for the qubit swap");
    if (stage_peers_ok_ && (!stage_ || stage_->bytes() < 4ull * (npat - 1) * chunk * sizeof(cplx))) drop_stage_peers();
    ensure_buffer(stage_, 4ull * (npat - 1) * chunk * sizeof(cplx));
    cplx *stage = static_cast<cplx *>(stage_->p());
    if (!xfer_stream_) {
        cudaStream_t xs;
        CUDA_CHECK(cudaStreamCreateWithFlags(&xs, cudaStreamNonBlocking));
        xfer_stream_ = xs;
        for (int h = 0; h < 2; h++) {
            cudaEvent_t e0, e1;
            CUDA_CHECK(cudaEventCreateWithFlags(&e0, cudaEventDisableTiming));
            CUDA_CHECK(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming));
            ev_packed_[h] = e0;
            ev_moved_[h] = e1;
        }
    }
    cudaStream_t cs = (cudaStream_t)stream_, xs = (cudaStream_t)xfer_stream_;
    const uint64_t half_amps = 2ull * (npat - 1) * chunk;
    struct Job {
        double *vec;
        uint64_t j0, cnt;
    };
    std::vector<Job> jobs;
    for (double *vec : vecs)
        for (uint64_t j0 = 0; j0 < sub; j0 += chunk) jobs.push_back({vec, j0, std::min(chunk, sub - j0)});
    auto pack_and_send = [&](size_t k) {
        const Job &J = jobs[k];
        const int h = (int)(k & 1);
        cplx *hb = stage + (uint64_t)h * half_amps;
        std::vector<Xfer> sends, recvs;
        uint32_t slot = 0;
        for (uint32_t b = 0; b < npat; b++) {
            if (b == mine) continue;
            cplx *sbuf = hb + (uint64_t)(2 * slot) * chunk, *rbuf = hb + (uint64_t)(2 * slot + 1) * chunk;
            CUDA_CHECK(b200k_pack_bits(reinterpret_cast<double *>(sbuf), J.vec, lpos, m, b, J.j0, J.cnt, stream_));
            note_launch();
            sends.push_back({peer_of(b), sbuf, J.cnt * sizeof(cplx)});
            recvs.push_back({peer_of(b), rbuf, J.cnt * sizeof(cplx)});
            slot++;
        }
        CUDA_CHECK(cudaEventRecord((cudaEvent_t)ev_packed_[h], cs));
        CUDA_CHECK(cudaStreamWaitEvent(xs, (cudaEvent_t)ev_packed_[h], 0));
        try {
            comm_->exchange(sends, recvs, xfer_stream_);
        }
        catch (const std::exception &e) {
            B200_FAIL(std::string("qubit swap failed: ") + e.what());
        }
        CUDA_CHECK(cudaEventRecord((cudaEvent_t)ev_moved_[h], xs));
        stats_.exchange_bytes += (double)(npat - 1) * (double)J.cnt * sizeof(cplx);
    };
    auto unpack = [&](size_t k) {
        const Job &J = jobs[k];
        const int h = (int)(k & 1);
        const cplx *hb = stage + (uint64_t)h * half_amps;
        CUDA_CHECK(cudaStreamWaitEvent(cs, (cudaEvent_t)ev_moved_[h], 0));
        uint32_t slot = 0;
        for (uint32_t b = 0; b < npat; b++) {
            if (b == mine) continue;
            const cplx *rbuf = hb + (uint64_t)(2 * slot + 1) * chunk;
            CUDA_CHECK(b200k_unpack_bits(J.vec, reinterpret_cast<const double *>(rbuf), lpos, m, b, J.j0, J.cnt, stream_));
            note_launch();
            slot++;
        }
    };
    for (size_t k = 0; k < jobs.size(); k++) {
        pack_and_send(k);
        if (k >= 1) unpack(k - 1);
    }
    unpack(jobs.size() - 1);
    stats_.exchanges++;
    // layout: the logical qubits at the exchanged positions trade places
    for (int i = 0; i < m; i++) {
        int qg = -1, ql = -1;
        for (int q = 0; q < (int)n_; q++) {
            if (pos[q] == pairs[i].first) qg = q;
            if (pos[q] == pairs[i].second) ql = q;
        }
        std::swap(pos[qg], pos[ql]);
    }
}

// Collective: (re)allocate the swap staging to at least `bytes` and map every rank's staging into this GPU's
// address space.  Ranks run the same program, so they all come here with the same size at the same point.
bool Simulator::ensure_stage_peers(size_t bytes)
{
    if (stage_ && stage_->bytes() >= bytes && stage_peers_ok_ && stage_peers_ptr_ == stage_->p()) return true;
    if (stage_peers_failed_) return false;
    drop_stage_peers();
    if (!stage_ || stage_->bytes() < bytes) {
        stage_.reset();
        stage_ = std::make_unique<DeviceBuffer>(bytes);
    }
    stage_peers_ptr_ = stage_->p();
    try {
        stage_peers_ok_ = comm_->map_peers(stage_peers_ptr_, stage_peers_);
    }
    catch (const std::exception &) {
        stage_peers_ok_ = false;
    }
    if (!stage_peers_ok_) stage_peers_failed_ = true; // no peer path between these GPUs: the staged exchange serves
    return stage_peers_ok_;
}

// Collective as well: nobody may still be writing into, or keep a mapping of, a staging buffer about to be freed.
void Simulator::drop_stage_peers()
{
    if (!stage_peers_ok_ || !comm_) return;
    cudaStreamSynchronize((cudaStream_t)stream_);
    comm_->unmap_peers(stage_peers_ptr_);
    try {
        comm_->barrier();
    }
    catch (const std::exception &) {
    }
    stage_peers_ok_ = false;
    stage_peers_ptr_ = nullptr;
    stage_peers_.clear();
}

bool Simulator::ensure_psi_peers()
{
    if (!psi_peers_tried_) {
        psi_peers_tried_ = true;
        psi_peers_ptr_ = psi_->p();
        try {
            psi_peers_ok_ = comm_->map_peers(psi_peers_ptr_, psi_peers_);
        }
        catch (const std::exception &) {
            psi_peers_ok_ = false;
        }
    }
    return psi_peers_ok_ && psi_peers_ptr_ == psi_->p();
}

void Simulator::drop_psi_peers()
{
    if (psi_peers_ok_ && comm_) comm_->unmap_peers(psi_peers_ptr_);
    psi_peers_ok_ = psi_peers_tried_ = false;
    psi_peers_ptr_ = nullptr;
    psi_peers_.clear();
}

void Simulator::reduce_results(std::vector<double> &host)
{
    if (!sharded() || host.empty()) return;
    std::vector<double> all(host.size() * (size_t)opt_.world);
    try {
        comm_->allgather_host(host.data(), all.data(), host.size());
    }
    catch (const std::exception &e) {
        B200_FAIL(std::string("reduction across ranks failed: ") + e.what());
    }
    for (size_t i = 0; i < host.size(); i++) {
        double s = 0.0;
        for (int r = 0; r < opt_.world; r++) s += all[(size_t)r * host.size() + i];
        host[i] = s;
    }
}

// ---- interpreted pass: group the pass's ops into rounds --------------------------------------------------
// A register round holds three tile bits in each thread's registers; it takes every op that is diagonal, a
// one-target PAIR on one of those bits, or a two-target PAIR (01<->10, 00<->11) whose two bits are both
// register bits.  Ops are taken greedily in order; an op that does not fit is deferred to a later round and
// blocks the ops that do not commute with it.  Every op carries masks over the thread's 8 register slots.
namespace {
// kind of a PAIR op for the register rounds: 1 = one target (a = 0, b = bit), 2 = two targets
// exchanging 01 <-> 10, 3 = two targets exchanging 00 <-> 11, 0 = none of these
int reg_pair_kind(const GateOp &g)
{
    if (g.kind != GateOp::PAIR) return 0;
    const int k = __builtin_popcountll(g.xmask);
    if (k == 1 && g.a == 0 && g.b == g.xmask) return 1;
    if (k == 2 && (g.a ^ g.b) == g.xmask) {
        if (g.a != 0 && g.b != 0) return 2;
        return 3;
    }
    return 0;
}
} // namespace

void Simulator::launch_pass3(double *vec, const std::vector<const GateOp *> &ops,
                             const std::vector<int> &high, int c)
{
    const int t = c + (int)high.size();
    if (dry_) {
        stats_.tile_passes++;
        return;
    }
    B200_FAIL_IF((int)high.size() > B200K_MAX_HIGH, "too many scattered tile bits");
    int loc[64];
    for (int p = 0; p < 64; p++) loc[p] = p < c ? p : -1;
    for (size_t h = 0; h < high.size(); h++) loc[high[h]] = c + (int)h;
    auto split = [&](uint64_t mask, uint32_t &inside, uint64_t &outside) {
        inside = 0;
        outside = 0;
        while (mask) {
            const int p = __builtin_ctzll(mask);
            mask &= mask - 1;
            if (loc[p] >= 0) inside |= 1u << loc[p];
            else outside |= 1ull << p;
        }
    };
    auto inside_only = [&](uint64_t mask) -> uint32_t {
        uint32_t in;
        uint64_t outm;
        split(mask, in, outm);
        B200_FAIL_IF(outm != 0, "internal: target bit outside the tile");
        return in;
    };

    std::vector<B200k3Round> rounds;
    std::vector<B200k3Op> dops;
    auto emit = [&]() {
        if (rounds.empty()) return;
        B200k2Header H;
        std::memset(&H, 0, sizeof(H));
        H.n_rounds = (int)rounds.size();
        H.n_ops = (int)dops.size();
        H.t = t;
        H.c = c;
        H.nhigh = (int)high.size();
        for (size_t h = 0; h < high.size(); h++) H.high_pos[h] = (uint8_t)high[h];
        H.rank_base = rank_base_;
        H.ntiles = 1ull << (nl_ - t);
        CUDA_CHECK(b200k_tile_program3(vec, (int)nl_, &H, rounds.data(), dops.data(), stream_));
        stats_.h2d_bytes += (double)(sizeof(H) + rounds.size() * sizeof(B200k3Round) + dops.size() * sizeof(B200k3Op));
        stats_.tile_passes++;
        stats_.kernel_ops += (int64_t)dops.size();
        stats_.rounds += (int64_t)rounds.size();
        rounds.clear();
        dops.clear();
    };
    auto base_op = [&](const GateOp &g, B200k3Op &o, uint32_t &ic, uint32_t &iv) {
        std::memset(&o, 0, sizeof(o));
        split(g.cmask, ic, o.co_mask);
        split(g.cval, iv, o.co_val);
        if (o.co_mask) o.flags |= B200K3_F_CO;
    };
    auto put_matrix = [&](B200k3Op &o, const cplx *m) {
        for (int q = 0; q < 4; q++) {
            o.m[2 * q] = m[q].real();
            o.m[2 * q + 1] = m[q].imag();
        }
    };

    std::vector<const GateOp *> rem(ops.begin(), ops.end()), deferred, taken;
    while (!rem.empty()) {
        uint64_t blockedX = 0, blockedZ = 0;
        uint32_t S = 0; // tile-local register bits of this round
        bool smem_round = false;
        deferred.clear();
        taken.clear();
        for (size_t i = 0; i < rem.size(); i++) {
            const GateOp *g = rem[i];
            const uint64_t xs = g->xmask, zs = g->zsupport();
            bool fits = false;
            if (!smem_round && !(xs & (blockedX | blockedZ)) && !(zs & blockedX)) {
                const bool diag = g->kind == GateOp::PHASE || g->kind == GateOp::PARITY;
                const int rk = t >= 3 ? reg_pair_kind(*g) : 0;
                if (t >= 3 && diag) fits = true;
                else if (rk) {
                    const uint32_t b = inside_only(g->xmask);
                    if (__builtin_popcount(S | b) <= 3) {
                        S |= b;
                        fits = true;
                    }
                }
                else if (taken.empty()) { // anything else: a shared-memory round of its own
                    fits = true;
                    smem_round = true;
                }
            }
            if (fits) taken.push_back(g);
            else {
                deferred.push_back(g);
                blockedX |= xs;
                blockedZ |= zs;
            }
        }
        if (dops.size() + taken.size() > (size_t)B200K3_LARGE_OPS || rounds.size() + 1 > (size_t)B200K3_LARGE_ROUNDS)
            emit();
        B200k3Round R;
        std::memset(&R, 0, sizeof(R));
        R.first = (uint16_t)dops.size();
        R.count = (uint16_t)taken.size();
        if (smem_round) {
            R.kind = B200K2_ROUND_SMEM;
            const GateOp &g = *taken[0];
            B200k3Op o;
            uint32_t ic, iv;
            base_op(g, o, ic, iv);
            o.tc_mask = ic;
            o.tc_val = iv;
            switch (g.kind) {
            case GateOp::PAIR: {
                o.kind = B200K3_PAIRG;
                o.tmask = inside_only(g.xmask);
                o.a = inside_only(g.a);
                o.b = inside_only(g.b);
                uint32_t m = o.tmask;
                while (m) {
                    B200_FAIL_IF(o.npos >= 4, "internal: PAIR on more than 4 bits");
                    o.pos[o.npos++] = (uint8_t)__builtin_ctz(m);
                    m &= m - 1;
                }
                put_matrix(o, g.m);
                break;
            }
            case GateOp::PAULIROT:
                o.kind = B200K3_PAULIROT;
                o.tmask = inside_only(g.xmask);
                o.pos[0] = (uint8_t)(31 - __builtin_clz(o.tmask));
                o.npos = 1;
                split(g.zmask, o.tz_mask, o.zo_mask);
                o.ny = (uint8_t)(g.ny & 3);
                o.m[0] = g.m[0].real();
                o.m[1] = g.m[0].imag();
                break;
            case GateOp::PHASE:
                o.kind = B200K3_PHASE;
                o.m[0] = g.m[0].real();
                o.m[1] = g.m[0].imag();
                break;
            case GateOp::PARITY:
                o.kind = B200K3_PARITY;
                split(g.zmask, o.tz_mask, o.zo_mask);
                o.m[0] = g.m[0].real(); o.m[1] = g.m[0].imag();
                o.m[2] = g.m[1].real(); o.m[3] = g.m[1].imag();
                break;
            default: B200_FAIL("internal: DENSE op in a register-blocked pass");
            }
            dops.push_back(o);
        }
        else {
            R.kind = B200K2_ROUND_REG;
            // pad the register bits to three with the highest free tile bits
            for (int b = t - 1; b >= 0 && __builtin_popcount(S) < 3; b--)
                if (!(S & (1u << b))) S |= 1u << b;
            int rb[3], nrb = 0, ridx[32];
            for (int b = 0; b < t; b++)
                if (S & (1u << b)) {
                    ridx[b] = nrb;
                    rb[nrb++] = b;
                }
            R.rb0 = (uint8_t)rb[0];
            R.rb1 = (uint8_t)rb[1];
            R.rb2 = (uint8_t)rb[2];
            for (int a = 0; a < 8; a++) {
                uint32_t idx = 0;
                for (int q = 0; q < 3; q++)
                    if (a & (1 << q)) idx |= 1u << rb[q];
                R.swz[a] = (uint16_t)(idx ^ ((idx >> 3) & 7u));
            }
            auto regpart = [&](uint32_t m) -> uint32_t {
                uint32_t r = 0;
                for (int q = 0; q < 3; q++)
                    if (m & (1u << rb[q])) r |= 1u << q;
                return r;
            };
            for (const GateOp *gp : taken) {
                const GateOp &g = *gp;
                B200k3Op o;
                uint32_t ic, iv, code = 0;
                base_op(g, o, ic, iv);
                const uint32_t rcm = regpart(ic), rcv = regpart(iv);
                o.tc_mask = ic & ~S;
                o.tc_val = iv & ~S;
                if (o.tc_mask) o.flags |= B200K3_F_TC;
                uint32_t slots = 0; // register slots where the control predicate holds
                for (uint32_t a = 0; a < 8; a++)
                    if ((a & rcm) == rcv) slots |= 1u << a;
                if (g.kind == GateOp::PAIR) {
                    o.kind = B200K3_PAIR;
                    const int rk = reg_pair_kind(g);
                    cplx m[4] = {g.m[0], g.m[1], g.m[2], g.m[3]};
                    uint32_t lo[4];
                    int npairs;
                    if (rk == 1) {
                        const int j = ridx[__builtin_ctz(inside_only(g.xmask))];
                        o.j = (uint8_t)j;
                        npairs = 4;
                        static const uint32_t LO[3][4] = {{0, 2, 4, 6}, {0, 1, 4, 5}, {0, 1, 2, 3}};
                        for (int p = 0; p < 4; p++) lo[p] = LO[j][p];
                    }
                    else {
                        uint32_t A = regpart(inside_only(g.a)), B = regpart(inside_only(g.b));
                        const uint32_t tm = A | B;
                        const int pairsel = tm == 3 ? 0 : (tm == 5 ? 1 : 2);
                        if (rk == 2 && A > B) { // kernel convention: first pattern = lower bit set
                            std::swap(A, B);
                            const cplx sw[4] = {m[3], m[2], m[1], m[0]};
                            for (int q = 0; q < 4; q++) m[q] = sw[q];
                        }
                        if (rk == 3 && A != 0) { // first pattern = 00
                            std::swap(A, B);
                            const cplx sw[4] = {m[3], m[2], m[1], m[0]};
                            for (int q = 0; q < 4; q++) m[q] = sw[q];
                        }
                        o.j = (uint8_t)((rk == 2 ? 3 : 6) + pairsel);
                        npairs = 2;
                        const uint32_t freebit = 7u & ~tm;
                        lo[0] = A;
                        lo[1] = A | freebit;
                    }
                    uint32_t pm = 0;
                    for (int p = 0; p < npairs; p++)
                        if (slots & (1u << lo[p])) pm |= 1u << p;
                    o.amask = (uint8_t)pm;
                    bool real = true, rx = true;
                    for (int q = 0; q < 4; q++) real = real && m[q].imag() == 0.0;
                    rx = m[0].imag() == 0.0 && m[3].imag() == 0.0 && m[1].real() == 0.0 && m[2].real() == 0.0;
                    const bool swp = real && m[0].real() == 0.0 && m[3].real() == 0.0 && m[1].real() == 1.0 &&
                                     m[2].real() == 1.0;
                    // permutations (X, CNOT, SWAP) run as REAL: explicit register exchanges made the
                    // compiler keep two copies of the thread's 8 amplitudes (32 moves per op); 8 FP64
                    // ops per pair on an idle FP64 pipe are cheaper
                    (void)swp;
                    o.mclass = real ? B200K3_M_REAL : (rx ? B200K3_M_RX : B200K3_M_GENERAL);
                    put_matrix(o, m);
                    if (pm == 0) continue; // predicate can never hold on these slots
                    const bool full = pm == (npairs == 4 ? 15u : 3u);
                    code = (uint32_t)((o.mclass * 9 + o.j) * 2 + (full ? 1 : 0));
                }
                else if (g.kind == GateOp::PHASE) {
                    o.kind = B200K3_PHASE;
                    o.amask = (uint8_t)slots;
                    o.m[0] = g.m[0].real();
                    o.m[1] = g.m[0].imag();
                    static const uint32_t SPECIAL[10] = {0xff, 0xaa, 0x55, 0xcc, 0x33, 0xf0, 0x0f, 0x88, 0xa0, 0xc0};
                    uint32_t idx = 10;
                    for (uint32_t q = 0; q < 10; q++)
                        if (SPECIAL[q] == slots) idx = q;
                    code = B200K3_CODE_PHASE + idx;
                    if (slots == 0) continue;
                }
                else {
                    o.kind = B200K3_PARITY;
                    o.amask = (uint8_t)slots;
                    uint32_t zi;
                    split(g.zmask, zi, o.zo_mask);
                    const uint32_t rz = regpart(zi);
                    uint32_t rp = 0;
                    for (uint32_t a = 0; a < 8; a++)
                        if (__builtin_popcount(a & rz) & 1) rp |= 1u << a;
                    o.rpar = (uint8_t)rp;
                    o.tz_mask = zi & ~S;
                    if (o.tz_mask || o.zo_mask) o.flags |= B200K3_F_TZ;
                    o.m[0] = g.m[0].real(); o.m[1] = g.m[0].imag();
                    o.m[2] = g.m[1].real(); o.m[3] = g.m[1].imag();
                    static const uint32_t RPARS[8] = {0x00, 0xaa, 0xcc, 0xf0, 0x66, 0x5a, 0x3c, 0x96};
                    uint32_t idx = 0;
                    for (uint32_t q = 0; q < 8; q++)
                        if (RPARS[q] == rp) idx = q;
                    code = B200K3_CODE_PARITY + 2 * idx + (slots == 0xffu ? 1 : 0);
                    if (slots == 0) continue;
                }
                o.ctl = code | (uint32_t)o.amask << 8 | (uint32_t)o.flags << 16 | (uint32_t)o.rpar << 24;
                dops.push_back(o);
            }
            R.count = (uint16_t)(dops.size() - R.first);
        }
        rounds.push_back(R);
        rem.swap(deferred);
    }
    emit();
}

const double *Simulator::upload_matrix(const CVec &m)
{
    const size_t bytes = m.size() * sizeof(cplx);
    const size_t cap = kArenaBytes;
    if (!matrix_arena_) matrix_arena_ = std::make_unique<DeviceBuffer>(cap);
    const size_t aligned = arena_aligned(bytes);
    // launch_pass reserved room for ALL matrices of its pass before the first upload, so the arena never
    // wraps while earlier ops of the same, not yet launched, pass still point into it
    B200_FAIL_IF(arena_used_ + aligned > cap, "internal: device matrix arena overflow inside one pass");
    char *dst = static_cast<char *>(matrix_arena_->p()) + arena_used_;
    CUDA_CHECK(cudaMemcpyAsync(dst, m.data(), bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
    stats_.h2d_bytes += (double)bytes;
    arena_used_ += aligned;
    return reinterpret_cast<const double *>(dst);
}

void Simulator::launch_pass(double *vec, const std::vector<const GateOp *> &ops,
                            const std::vector<int> &high, int c)
{
    if (dry_) {
        stats_.tile_passes++;
        return;
    }
    B200kTileProgram P;
    std::memset(&P, 0, sizeof(P));
    const int t = c + (int)high.size();
    P.t = t;
    P.c = c;
    P.nhigh = (int)high.size();
    B200_FAIL_IF(P.nhigh > B200K_MAX_HIGH, "too many scattered tile bits");
    for (size_t h = 0; h < high.size(); h++) P.high_pos[h] = (uint8_t)high[h];
    P.rank_base = rank_base_;
    int loc[64];
    for (int p = 0; p < 64; p++) loc[p] = p < c ? p : -1;
    for (size_t h = 0; h < high.size(); h++) loc[high[h]] = c + (int)h;
    auto split = [&](uint64_t mask, uint32_t &inside, uint64_t &outside) {
        inside = 0;
        outside = 0;
        while (mask) {
            const int p = __builtin_ctzll(mask);
            mask &= mask - 1;
            if (loc[p] >= 0) inside |= 1u << loc[p];
            else outside |= 1ull << p;
        }
    };
    auto inside_only = [&](uint64_t mask) -> uint32_t {
        uint32_t in;
        uint64_t outm;
        split(mask, in, outm);
        B200_FAIL_IF(outm != 0, "internal: target bit outside the tile");
        return in;
    };
    P.n_ops = (int)ops.size();
    {
        size_t need = 0;
        for (const GateOp *g : ops)
            if (g->kind == GateOp::DENSE) need += arena_aligned(g->matrix->size() * sizeof(cplx));
        B200_FAIL_IF(need > kArenaBytes, "matrix too large for the device matrix arena");
        if (arena_used_ + need > kArenaBytes) { // wrap BEFORE the pass: kernels that read the old contents must be done
            CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
            arena_used_ = 0;
        }
    }
    for (size_t i = 0; i < ops.size(); i++) {
        const GateOp &g = *ops[i];
        B200kTileOp &o = P.ops[i];
        split(g.cmask, o.cmask, o.cmask_o);
        split(g.cval, o.cval, o.cval_o);
        switch (g.kind) {
        case GateOp::PAIR:
            o.kind = B200K_OP_PAIR;
            o.tmask = inside_only(g.xmask);
            o.a = inside_only(g.a);
            o.b = inside_only(g.b);
            for (int q = 0; q < 4; q++) {
                o.u.m[2 * q] = g.m[q].real();
                o.u.m[2 * q + 1] = g.m[q].imag();
            }
            break;
        case GateOp::PHASE:
            o.kind = B200K_OP_PHASE;
            o.u.m[0] = g.m[0].real();
            o.u.m[1] = g.m[0].imag();
            break;
        case GateOp::PARITY:
            o.kind = B200K_OP_PARITY;
            split(g.zmask, o.a, o.zmask_o);
            o.u.m[0] = g.m[0].real();
            o.u.m[1] = g.m[0].imag();
            o.u.m[2] = g.m[1].real();
            o.u.m[3] = g.m[1].imag();
            break;
        case GateOp::PAULIROT:
            o.kind = B200K_OP_PAULIROT;
            o.tmask = inside_only(g.xmask);
            split(g.zmask, o.a, o.zmask_o);
            o.ny = (uint32_t)(g.ny & 3);
            o.u.m[0] = g.m[0].real();
            o.u.m[1] = g.m[0].imag();
            break;
        case GateOp::DENSE:
            o.kind = B200K_OP_DENSE;
            o.k = (uint32_t)g.targets.size();
            o.tmask = inside_only(g.xmask);
            for (size_t j = 0; j < g.targets.size(); j++) o.u.d.tb[j] = (uint8_t)loc[g.targets[j]];
            o.u.d.mat = upload_matrix(*g.matrix);
            break;
        }
    }
    CUDA_CHECK(b200k_tile_program(vec, (int)nl_, &P, stream_));
    stats_.tile_passes++;
    stats_.kernel_ops += (int64_t)ops.size();
}

// ---- state preparation ----------------------------------------------------------------------------
void Simulator::set_basis_state(const std::vector<int8_t> &bits, const std::vector<int64_t> &wires)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    B200_FAIL_IF(bits.size() != wires.size(),
                 "BasisState parameter and wires must be of equal length.");
    check_wires(wires, {});
    queue_.clear(); // the whole register is overwritten
    for (int64_t w = 0; w < n_; w++) wire_bit_[(size_t)w] = (int)(n_ - 1 - w); // canonical wire order again
    uint64_t idx = 0;
    for (size_t j = 0; j < wires.size(); j++) {
        B200_FAIL_IF(bits[j] != 0 && bits[j] != 1, "BasisState parameter must consist of 0 or 1 integers.");
        if (bits[j]) idx |= bit_of(wires[j]);
    }
    for (int q = 0; q < (int)n_; q++) pos_[q] = q; // fresh register: identity layout
    const int64_t local = (idx >> nl_) == (uint64_t)opt_.rank ? (int64_t)(idx & ((1ull << nl_) - 1ull)) : -1;
    if (dry_) return;
    CUDA_CHECK(b200k_set_basis(psi_->d(), (int)nl_, local, stream_));
    note_launch();
}

void Simulator::set_state(const CVec &st, const std::vector<int64_t> &wires)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    check_wires(wires, {});
    const size_t k = wires.size();
    B200_FAIL_IF(st.size() != (size_t(1) << k), "Invalid state size: must be 2^len(wires)");
    queue_.clear();
    for (int q = 0; q < (int)n_; q++) pos_[q] = q; // the register is overwritten: identity layout
    for (int64_t w = 0; w < n_; w++) wire_bit_[(size_t)w] = (int)(n_ - 1 - w); // ... and canonical wire order
    bool natural = (int64_t)k == n_;
    for (size_t j = 0; j < k && natural; j++) natural = wires[j] == (int64_t)j;
    if (natural) { // each rank takes its contiguous slice
        CUDA_CHECK(cudaMemcpyAsync(psi_->p(), st.data() + ((size_t)opt_.rank << nl_), sizeof(cplx) << nl_,
                                   cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        stats_.h2d_bytes += (double)(sizeof(cplx) << nl_);
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
        return;
    }
    // wires on rank bits select a slice of `st`; a rank bit outside `wires` must be 0
    std::vector<uint8_t> tb;
    std::vector<size_t> local_j; // positions (in `wires`) of the wires that live on local bits
    bool empty_shard = false;
    uint64_t covered = 0;
    size_t fixed = 0;
    for (size_t j = 0; j < k; j++) {
        const int b = lbit(wires[j]);
        covered |= 1ull << b;
        if (b < nl_) {
            tb.push_back((uint8_t)b);
            local_j.push_back(j);
        }
        else if ((rank_base_ >> b) & 1) fixed |= size_t(1) << (k - 1 - j);
    }
    for (int b = (int)nl_; b < (int)n_; b++)
        if (!((covered >> b) & 1) && ((rank_base_ >> b) & 1)) empty_shard = true;
    const size_t kl = tb.size();
    CVec part(size_t(1) << kl);
    for (size_t r = 0; r < part.size(); r++) {
        size_t src = fixed;
        for (size_t i = 0; i < kl; i++)
            if ((r >> (kl - 1 - i)) & 1) src |= size_t(1) << (k - 1 - local_j[i]);
        part[r] = empty_shard ? cplx(0.0, 0.0) : st[src];
    }
    ensure_buffer(misc_, part.size() * sizeof(cplx));
    CUDA_CHECK(cudaMemcpyAsync(misc_->p(), part.data(), part.size() * sizeof(cplx), cudaMemcpyHostToDevice,
                               (cudaStream_t)stream_));
    CUDA_CHECK(b200k_scatter_state(psi_->d(), misc_->d(), (int)nl_, tb.data(), (int)kl, stream_));
    note_launch();
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
}

// ---- observables --------------------------------------------------------------------------------------
// Keys are indices into a per-activation table, monotonically increasing (pattern:
// runtime/lib/backend/openqasm/OpenQasmObsManager.hpp:96-189).
int64_t Simulator::obs_named(int id, const CVec &matrix, const std::vector<int64_t> &wires)
{
    B200_FAIL_IF(id < 0 || id > 5, "Invalid observable id");
    for (auto w : wires) (void)bit_of(w);
    ObsNode o;
    if (id == 5) {
        const size_t dim = size_t(1) << wires.size();
        B200_FAIL_IF(wires.empty() || matrix.size() != dim * dim, "Invalid Hermitian matrix size");
        check_wires(wires, {});
        o.kind = ObsKind::Hermitian;
        o.matrix = matrix;
    }
    else {
        B200_FAIL_IF(wires.empty(), "Invalid number of wires for a named observable");
        // the frontend may pass Identity on many wires (frontend/catalyst/jax_tracer.py:1002-1013)
        B200_FAIL_IF(id != 0 && wires.size() != 1, "Named observables act on exactly one wire");
        o.kind = ObsKind::Named;
    }
    o.named_id = id;
    o.wires = wires;
    obs_.push_back(std::move(o));
    return (int64_t)obs_.size() - 1;
}
int64_t Simulator::obs_tensor(const std::vector<int64_t> &keys)
{
    for (auto k : keys) B200_FAIL_IF(k < 0 || (size_t)k >= obs_.size(), "Invalid observable key");
    ObsNode o;
    o.kind = ObsKind::Tensor;
    o.children = keys;
    obs_.push_back(std::move(o));
    return (int64_t)obs_.size() - 1;
}
int64_t Simulator::obs_hamiltonian(const std::vector<double> &coeffs, const std::vector<int64_t> &keys)
{
    B200_FAIL_IF(coeffs.size() != keys.size(), "Invalid size for the list of coefficients and observables");
    for (auto k : keys) B200_FAIL_IF(k < 0 || (size_t)k >= obs_.size(), "Invalid observable key");
    ObsNode o;
    o.kind = ObsKind::Hamiltonian;
    o.coeffs = coeffs;
    o.children = keys;
    obs_.push_back(std::move(o));
    return (int64_t)obs_.size() - 1;
}

void Simulator::flatten(int64_t key, std::vector<ObsTerm> &out) const
{
    B200_FAIL_IF(key < 0 || (size_t)key >= obs_.size(), "Invalid key for cached observables");
    const ObsNode &o = obs_[key];
    switch (o.kind) {
    case ObsKind::Named:
    case ObsKind::Hermitian: {
        ObsTerm t;
        t.factors.push_back({o.named_id, o.wires, o.kind == ObsKind::Hermitian ? &o.matrix : nullptr});
        out.push_back(std::move(t));
        return;
    }
    case ObsKind::Tensor: {
        std::vector<ObsTerm> acc(1);
        for (auto c : o.children) {
            std::vector<ObsTerm> sub;
            flatten(c, sub);
            std::vector<ObsTerm> next;
            next.reserve(acc.size() * sub.size());
            for (auto &a : acc)
                for (auto &s : sub) {
                    ObsTerm t = a;
                    t.coef *= s.coef;
                    t.factors.insert(t.factors.end(), s.factors.begin(), s.factors.end());
                    next.push_back(std::move(t));
                }
            acc.swap(next);
            B200_FAIL_IF(acc.size() > (size_t(1) << 20), "Observable expands to too many terms");
        }
        for (auto &t : acc) out.push_back(std::move(t));
        return;
    }
    case ObsKind::Hamiltonian:
        for (size_t i = 0; i < o.children.size(); i++) {
            std::vector<ObsTerm> sub;
            flatten(o.children[i], sub);
            for (auto &t : sub) {
                t.coef *= o.coeffs[i];
                out.push_back(std::move(t));
            }
        }
        return;
    }
}

bool Simulator::term_to_paulis(const ObsTerm &t, std::vector<PauliTerm> &out) const
{
    uint64_t used = 0;
    int nh = 0;
    for (auto &f : t.factors) {
        if (f.named_id == 5) return false;
        if (f.named_id == 0) continue;
        const uint64_t b = bit_of(f.wires[0]);
        if (used & b) return false; // repeated wire: take the generic (sequential) path
        used |= b;
        if (f.named_id == 4) nh++;
    }
    if (nh > 6) return false;
    std::vector<PauliTerm> acc(1);
    acc[0].coef = t.coef;
    for (auto &f : t.factors) {
        if (f.named_id == 0) continue;
        const uint64_t b = bit_of(f.wires[0]);
        if (f.named_id == 4) { // Hadamard = (X + Z) / sqrt(2)
            std::vector<PauliTerm> next;
            for (auto &a : acc) {
                PauliTerm x = a, z = a;
                x.coef *= 1.0 / std::sqrt(2.0);
                z.coef *= 1.0 / std::sqrt(2.0);
                x.xmask |= b;
                z.zmask |= b;
                next.push_back(x);
                next.push_back(z);
            }
            acc.swap(next);
            continue;
        }
        for (auto &a : acc) {
            if (f.named_id == 1) a.xmask |= b;
            else if (f.named_id == 2) { a.xmask |= b; a.zmask |= b; a.ny++; }
            else a.zmask |= b;
        }
    }
    for (auto &a : acc) out.push_back(a);
    return true;
}

void Simulator::factor_ops(const ObsFactor &f, std::vector<GateOp> &ops) const
{
    static const char *nm[] = {"Identity", "PauliX", "PauliY", "PauliZ", "Hadamard"};
    if (f.named_id == 5) canonicalize_matrix(ops, *f.matrix, f.wires, false, {}, {});
    else if (f.named_id != 0) canonicalize(ops, nm[f.named_id], {}, f.wires, false, {}, {}, "");
}

double *Simulator::result_slot(int slot) { return results_->d() + 2 * (size_t)slot; }

void Simulator::read_results(int nslots, std::vector<double> &host)
{
    host.resize(2 * (size_t)nslots);
    if (nslots == 0) return;
    CUDA_CHECK(cudaMemcpyAsync(host.data(), results_->p(), host.size() * sizeof(double),
                               cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    stats_.d2h_bytes += (double)(host.size() * sizeof(double));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    reduce_results(host);
}

cplx Simulator::inner(const double *a, const double *b)
{
    CUDA_CHECK(b200k_bilinear_pauli(a, b, (int)nl_, rank_base_, 0, 0, 0, 0, 0, scratch_->d(), result_slot(0), stream_));
    note_launch();
    std::vector<double> h;
    read_results(1, h);
    return {h[0], h[1]};
}

uint64_t Simulator::observable_xmask(int64_t key) const
{
    std::vector<ObsTerm> terms;
    flatten(key, terms);
    uint64_t m = 0;
    for (auto &t : terms) {
        std::vector<PauliTerm> pts;
        if (term_to_paulis(t, pts)) {
            for (auto &p : pts) m |= p.xmask;
            continue;
        }
        for (auto &f : t.factors) {
            if (f.named_id == 0 || f.named_id == 3) continue; // Identity, PauliZ: diagonal
            for (auto w : f.wires) m |= bit_of(w);
        }
    }
    return m;
}

// vout = O vin; both vectors share the layout `pos`, under which every non-diagonal target of O
// must already be local (callers run make_local(observable_xmask(key)) first).
void Simulator::apply_observable(int64_t key, const double *vin, double *vout, const std::vector<int> &pos)
{
    std::vector<ObsTerm> terms;
    flatten(key, terms);
    bool first = true;
    const size_t bytes = sizeof(cplx) << nl_;
    // Pauli words are applied B200K_APPLY_WORDS at a time (one read+write of the accumulator per batch)
    std::vector<uint64_t> bx, bz;
    std::vector<int> bny;
    std::vector<double> bc;
    auto flush_words = [&]() {
        if (bx.empty()) return;
        CUDA_CHECK(b200k_pauli_apply_batch(vout, vin, (int)nl_, rank_base_, bx.data(), bz.data(), bny.data(), bc.data(),
                                           (int)bx.size(), first ? 0 : 1, stream_));
        note_launch();
        first = false;
        bx.clear(); bz.clear(); bny.clear(); bc.clear();
    };
    // Pauli words whose X/Y bits fit a tile: tile-staged, up to 96 words per read of vin (k_pauli_words_tile)
    const int fit = (int)std::min<int64_t>(11, nl_) - 3;
    std::vector<TileWordRef> tiled;
    auto tileable = [&](const PauliTerm &p) {
        return opt_.batch_overlaps && nl_ >= 3 && __builtin_popcountll(p.xmask) <= fit;
    };
    for (auto &t : terms) {
        std::vector<PauliTerm> pts;
        if (term_to_paulis(t, pts)) {
            for (auto &p : pts) {
                if (tileable(p)) {
                    tiled.push_back({p.xmask, p.zmask, p.ny, p.coef});
                    continue;
                }
                bx.push_back(to_phys(pos, p.xmask));
                bz.push_back(to_phys(pos, p.zmask));
                bny.push_back(p.ny);
                bc.push_back(p.coef.real());
                bc.push_back(p.coef.imag());
                if (bx.size() == (size_t)B200K_APPLY_WORDS) flush_words();
            }
            continue;
        }
        flush_words();
        ensure_buffer(tmp_, bytes);
        CUDA_CHECK(cudaMemcpyAsync(tmp_->p(), vin, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
        std::vector<GateOp> ops;
        for (auto &f : t.factors) factor_ops(f, ops);
        if (!ops.empty()) {
            std::vector<int> lay = pos; // no swap can happen: targets are local
            execute({tmp_->d()}, lay, ops);
            B200_FAIL_IF(lay != pos, "internal: observable application changed the layout");
        }
        if (first) {
            CUDA_CHECK(b200k_pauli_apply(vout, tmp_->d(), (int)nl_, rank_base_, 0, 0, 0, t.coef, 0.0, 0, stream_));
        }
        else {
            CUDA_CHECK(b200k_axpy(vout, tmp_->d(), t.coef, (int)nl_, stream_));
        }
        note_launch();
        first = false;
    }
    flush_words();
    if (tiled.size() > 1) {
        std::vector<int> lay = pos; // every X/Y bit is local already (precondition): no swap can happen
        pauli_words_binned(std::move(tiled), {}, lay, [&](const B200k2Header &H, const std::vector<B200kTileWord> &ws) {
            CUDA_CHECK(b200k_pauli_words_tile(vout, vin, nullptr, (int)nl_, &H, ws.data(), (int)ws.size(), first ? 0 : 1,
                                              nullptr, nullptr, stream_));
            note_launch();
            first = false;
        });
    }
    else if (tiled.size() == 1) { // a single word: the gather pass moves the same bytes without staging
        const TileWordRef &w = tiled[0];
        CUDA_CHECK(b200k_pauli_apply(vout, vin, (int)nl_, rank_base_, to_phys(pos, w.xmask), to_phys(pos, w.zmask), w.ny,
                                     w.coef.real(), w.coef.imag(), first ? 0 : 1, stream_));
        note_launch();
        first = false;
    }
    if (first) CUDA_CHECK(cudaMemsetAsync(vout, 0, bytes, (cudaStream_t)stream_));
}

double Simulator::expval(int64_t key)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    B200_FAIL_IF(key < 0 || (size_t)key >= obs_.size(), "Invalid key for cached observables");
    if (recording_) tape_obs_.push_back({key, 1}); // CacheManager.hpp:111-114
    flush();
    std::vector<ObsTerm> terms;
    flatten(key, terms);
    if (opt_.shots > 0) { // sampled estimate
        double e = 0.0;
        for (auto &t : terms) e += t.coef * sampled_term_mean(t);
        return e;
    }
    std::vector<PauliTerm> diag, offd;
    std::vector<const ObsTerm *> generic;
    for (auto &t : terms) {
        std::vector<PauliTerm> pts;
        if (term_to_paulis(t, pts)) {
            for (auto &p : pts) (p.xmask ? offd : diag).push_back(p);
        }
        else generic.push_back(&t);
    }
    const int BATCH = 1024;
    const int ndiag_slots = (int)((diag.size() + BATCH - 1) / BATCH);
    const int nslots = ndiag_slots + (int)offd.size() + (int)generic.size();
    ensure_buffer(results_, std::max<size_t>(4096, 2 * (size_t)nslots) * sizeof(double));
    int slot = 0;
    // all-diagonal words: one read pass per batch (rank bits enter through rank_base)
    if (!diag.empty()) {
        const size_t nt = diag.size();
        ensure_buffer(misc_, nt * 16);
        std::vector<uint64_t> zm(nt);
        std::vector<double> cf(nt);
        for (size_t i = 0; i < nt; i++) {
            zm[i] = to_phys(pos_, diag[i].zmask);
            cf[i] = diag[i].coef.real();
        }
        uint64_t *dz = static_cast<uint64_t *>(misc_->p());
        double *dc = reinterpret_cast<double *>(dz + nt);
        CUDA_CHECK(cudaMemcpyAsync(dz, zm.data(), nt * 8, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        CUDA_CHECK(cudaMemcpyAsync(dc, cf.data(), nt * 8, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        for (size_t o = 0; o < nt; o += BATCH) {
            const int cnt = (int)std::min<size_t>(BATCH, nt - o);
            CUDA_CHECK(b200k_expval_diag_batch(psi_->d(), (int)nl_, rank_base_, dz + o, dc + o, cnt,
                                               scratch_->d(), result_slot(slot++), stream_));
            note_launch();
        }
        CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_)); // zm/cf are stack-owned
    }
    // Off-diagonal words: those whose X/Y bits fit one tile are evaluated many per read pass -- sum_w c_w <psi|P_w|psi>
    // per BIN of words (pauli_words_binned -> k_pauli_words_tile in its contraction mode: psi is read once per bin and
    // nothing is written).  A word with more X/Y factors than a tile has free bits takes a read pass of its own.
    // (Round 1 switched batching off for the WHOLE Hamiltonian when a single word did not fit: C4's 1000 terms, one of
    // them with 9 X/Y factors, ran as 950 separate passes.)
    int nbin_slots = 0;
    std::vector<const PauliTerm *> single;
    {
        const int fit = (int)std::min<int64_t>(11, nl_) - 3;
        std::vector<TileWordRef> binned;
        for (auto &p : offd) {
            if (opt_.batch_overlaps && nl_ >= 3 && offd.size() > 1 && __builtin_popcountll(p.xmask) <= fit)
                binned.push_back({p.xmask, p.zmask, p.ny, p.coef});
            else single.push_back(&p);
        }
        if (!binned.empty())
            pauli_words_binned(std::move(binned), {psi_->d()}, pos_, [&](const B200k2Header &H, const std::vector<B200kTileWord> &ws) {
                CUDA_CHECK(b200k_pauli_words_tile(nullptr, psi_->d(), psi_->d(), (int)nl_, &H, ws.data(), (int)ws.size(), 0,
                                                  scratch_->d(), result_slot(slot++), stream_));
                note_launch();
                nbin_slots++;
            });
        // the layout may change between terms (a word with X/Y on a rank bit pulls that qubit in), so
        // masks are translated term by term
        for (auto *pp : single) {
            const PauliTerm &p = *pp;
            make_local({psi_->d()}, pos_, p.xmask);
            CUDA_CHECK(b200k_bilinear_pauli(psi_->d(), psi_->d(), (int)nl_, rank_base_, to_phys(pos_, p.xmask),
                                            to_phys(pos_, p.zmask), p.ny, 0, 0, scratch_->d(),
                                            result_slot(slot++), stream_));
            note_launch();
        }
    }
    const size_t bytes = sizeof(cplx) << nl_;
    for (auto *t : generic) {
        uint64_t xm = 0;
        for (auto &f : t->factors)
            if (f.named_id != 0 && f.named_id != 3)
                for (auto w : f.wires) xm |= bit_of(w);
        make_local({psi_->d()}, pos_, xm);
        ensure_buffer(tmp_, bytes);
        CUDA_CHECK(cudaMemcpyAsync(tmp_->p(), psi_->p(), bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream_));
        std::vector<GateOp> ops;
        for (auto &f : t->factors) factor_ops(f, ops);
        if (!ops.empty()) {
            std::vector<int> lay = pos_;
            execute({tmp_->d()}, lay, ops);
            B200_FAIL_IF(lay != pos_, "internal: observable application changed the layout");
        }
        CUDA_CHECK(b200k_bilinear_pauli(psi_->d(), tmp_->d(), (int)nl_, rank_base_, 0, 0, 0, 0, 0,
                                        scratch_->d(), result_slot(slot++), stream_));
        note_launch();
    }
    std::vector<double> h;
    read_results(nslots, h);
    double e = 0.0;
    slot = 0;
    for (int i = 0; i < ndiag_slots; i++) e += h[2 * slot++];
    for (int i = 0; i < nbin_slots; i++) e += h[2 * slot++]; // coefficients applied on the device
    for (auto *p : single) e += p->coef.real() * h[2 * slot++];
    for (auto *t : generic) e += t->coef * h[2 * slot++];
    return e;
}

double Simulator::var(int64_t key)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    B200_FAIL_IF(key < 0 || (size_t)key >= obs_.size(), "Invalid key for cached observables");
    if (recording_) tape_obs_.push_back({key, 2});
    flush();
    if (opt_.shots > 0) {
        std::vector<ObsTerm> terms;
        flatten(key, terms);
        B200_FAIL_IF(terms.size() != 1, "Var of a sum of observables does not support finite shots");
        double m1, m2; // sampled <O> and <O^2> of the single term (Pauli products: <O^2> = 1)
        sampled_term_moments(terms[0], m1, m2);
        return terms[0].coef * terms[0].coef * (m2 - m1 * m1);
    }
    const size_t bytes = sizeof(cplx) << nl_;
    make_local({psi_->d()}, pos_, observable_xmask(key));
    ensure_buffer(acc_, bytes);
    apply_observable(key, psi_->d(), acc_->d(), pos_);
    CUDA_CHECK(b200k_bilinear_pauli(psi_->d(), acc_->d(), (int)nl_, rank_base_, 0, 0, 0, 0, 0, scratch_->d(),
                                    result_slot(0), stream_));
    CUDA_CHECK(b200k_bilinear_pauli(acc_->d(), acc_->d(), (int)nl_, rank_base_, 0, 0, 0, 0, 0, scratch_->d(),
                                    result_slot(1), stream_));
    note_launch();
    note_launch();
    std::vector<double> h;
    read_results(2, h);
    return h[2] - h[0] * h[0];
}

// ---- State / Probs ----------------------------------------------------------------------------------
// Copy out of a pinned staging buffer into the caller's memory with non-temporal stores: the destination is written
// once and not read here, so ordinary stores would first READ every destination line (read-for-ownership) -- a third of
// the host-memory traffic of the copy, on a memory bus the incoming DMA shares.
#if defined(__x86_64__)
__attribute__((target("avx2"))) static void stream_copy_avx2(char *dst, const char *src, size_t n)
{
    size_t head = (32 - (reinterpret_cast<uintptr_t>(dst) & 31)) & 31;
    if (head > n) head = n;
    std::memcpy(dst, src, head);
    dst += head; src += head; n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 64));
        const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 96), d);
    }
    _mm_sfence();
    std::memcpy(dst + i, src + i, n - i);
}
static void stream_copy(char *dst, const char *src, size_t n)
{
    static const bool avx2 = __builtin_cpu_supports("avx2") && !std::getenv("B200SV_NO_STREAM_COPY");
    if (avx2 && n >= 4096) stream_copy_avx2(dst, src, n);
    else std::memcpy(dst, src, n);
}
#else
static void stream_copy(char *dst, const char *src, size_t n) { std::memcpy(dst, src, n); }
#endif

// A few parked host threads that copy staged chunks into the caller's (pageable) memory: spawning threads per chunk
// costs ~0.4 ms of every ~1.5 ms chunk.  Process-wide, one job at a time, never destroyed (parked threads at exit).
namespace {
class HostCopyPool {
  public:
    static HostCopyPool &get()
    {
        static HostCopyPool *p = new HostCopyPool;
        return *p;
    }
    // runs fn(i0, i1) over [0, n) split into one slice per thread (the caller's thread takes the first slice)
    void run(size_t n, const std::function<void(size_t, size_t)> &fn)
    {
        std::unique_lock<std::mutex> job(job_mu_); // one job at a time
        const size_t parts = workers_.size() + 1, step = (n + parts - 1) / parts;
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn;
            n_ = n;
            step_ = step;
            remaining_ = (int)workers_.size();
            generation_++;
        }
        cv_.notify_all();
        fn(0, std::min(n, step));
        std::unique_lock<std::mutex> lk(mu_);
        done_.wait(lk, [&] { return remaining_ == 0; });
    }

  private:
    HostCopyPool()
    {
        unsigned hw = std::thread::hardware_concurrency();
        unsigned n = hw > 3 ? std::min(15u, hw - 2) : 1;
        if (const char *e = std::getenv("B200SV_COPY_THREADS")) n = (unsigned)std::max(1, std::atoi(e));
        for (unsigned t = 1; t < n; t++) workers_.emplace_back([this, t] { loop(t); });
    }
    void loop(unsigned t)
    {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_.wait(lk, [&] { return generation_ != seen; });
            seen = generation_;
            const std::function<void(size_t, size_t)> *fn = fn_;
            const size_t i0 = std::min(n_, t * step_), i1 = std::min(n_, (t + 1) * step_);
            lk.unlock();
            if (i1 > i0) (*fn)(i0, i1);
            lk.lock();
            if (--remaining_ == 0) done_.notify_all();
        }
    }
    std::mutex job_mu_, mu_;
    std::condition_variable cv_, done_;
    std::vector<std::thread> workers_;
    const std::function<void(size_t, size_t)> *fn_ = nullptr;
    size_t n_ = 0, step_ = 0;
    int remaining_ = 0;
    uint64_t generation_ = 0;
};
} // namespace

// Device -> caller's host buffer (pageable memory, DataView strides: runtime/include/DataView.hpp:124-134).
// Large results (Probs of 30 wires = 8 GiB, State) go through kStage pinned staging buffers: up to kStage - 1 DMAs are
// queued ahead, so the copy engine never waits for the host, while the host-copy pool moves (or scatters, for a strided
// view) the oldest finished chunk into the caller's memory.
void Simulator::d2h_strided(const void *dsrc, size_t elem_bytes, size_t count, void *hdst, int64_t stride)
{
    cudaStream_t st = (cudaStream_t)stream_;
    const size_t total = elem_bytes * count;
    if (stride == 1 && total <= (size_t(1) << 20)) {
        CUDA_CHECK(cudaMemcpyAsync(hdst, dsrc, total, cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaStreamSynchronize(st));
        stats_.d2h_bytes += (double)total;
        return;
    }
    const size_t CH = size_t(64) << 20; // bytes per staging buffer
    const size_t per = CH / elem_bytes; // elements per chunk
    const size_t nchunks = (count + per - 1) / per;
    const int nbuf = (int)std::min<size_t>(kStage, nchunks);
    for (int h = 0; h < nbuf; h++)
        if (!stage_host_[h]) {
            CUDA_CHECK(cudaMallocHost(&stage_host_[h], CH));
            cudaEvent_t e;
            CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            stage_ev_[h] = e;
        }
    auto drain = [&](size_t k) { // chunk k: staging buffer -> caller's memory
        const int h = (int)(k % (size_t)nbuf);
        CUDA_CHECK(cudaEventSynchronize((cudaEvent_t)stage_ev_[h]));
        const size_t o = k * per, cnt = std::min(per, count - o);
        const char *src = static_cast<const char *>(stage_host_[h]);
        auto part = [&](size_t i0, size_t i1) {
            if (stride == 1) {
                stream_copy(static_cast<char *>(hdst) + (o + i0) * elem_bytes, src + i0 * elem_bytes, (i1 - i0) * elem_bytes);
                return;
            }
            for (size_t i = i0; i < i1; i++)
                std::memcpy(static_cast<char *>(hdst) + (int64_t)(o + i) * stride * (int64_t)elem_bytes, src + i * elem_bytes,
                            elem_bytes);
        };
        if (cnt * elem_bytes < (size_t(4) << 20)) part(0, cnt);
        else HostCopyPool::get().run(cnt, part);
    };
    const size_t ahead = (size_t)nbuf - 1 > 0 ? (size_t)nbuf - 1 : 1; // DMAs in flight before the oldest is drained
    for (size_t k = 0; k < nchunks; k++) {
        const int h = (int)(k % (size_t)nbuf);
        const size_t o = k * per, cnt = std::min(per, count - o);
        // buffer h was drained at iteration k - 1 (chunk k - nbuf): free again
        CUDA_CHECK(cudaMemcpyAsync(stage_host_[h], static_cast<const char *>(dsrc) + o * elem_bytes, cnt * elem_bytes,
                                   cudaMemcpyDeviceToHost, st));
        CUDA_CHECK(cudaEventRecord((cudaEvent_t)stage_ev_[h], st));
        if (k + 1 >= ahead && nbuf > 1) drain(k + 1 - ahead);
    }
    for (size_t k = (nchunks >= ahead && nbuf > 1) ? nchunks + 1 - ahead : 0; k < nchunks; k++) drain(k);
    stats_.d2h_bytes += (double)total;
}

void Simulator::state(double *out_ri, int64_t stride)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    if (!wire_identity()) restore_wire_order();
    flush();
    if (!sharded()) {
        d2h_strided(psi_->p(), sizeof(cplx), size_t(1) << n_, out_ri, stride);
        return;
    }
    if (opt_.gather_slices) { // gather='slice': this rank's amplitudes only, at their place in the caller's array
        canonical_local_bits();
        const int64_t off = (int64_t)opt_.rank << nl_;
        d2h_strided(psi_->p(), sizeof(cplx), size_t(1) << nl_, out_ri + 2 * off * stride, stride);
        return;
    }
    // SPMD contract: every rank returns the whole vector.  Shards are gathered through the host and
    // the layout is undone there; meant for the register sizes a host can hold anyway.
    const size_t nloc = size_t(1) << nl_;
    std::vector<double> mine(2 * nloc), all(2 * nloc * (size_t)opt_.world);
    CUDA_CHECK(cudaMemcpyAsync(mine.data(), psi_->p(), nloc * sizeof(cplx), cudaMemcpyDeviceToHost,
                               (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    try {
        comm_->allgather_host(mine.data(), all.data(), mine.size());
    }
    catch (const std::exception &e) {
        B200_FAIL(std::string("State: gather across ranks failed: ") + e.what());
    }
    const size_t N = size_t(1) << n_;
    for (size_t phys = 0; phys < N; phys++) {
        size_t logical = 0;
        for (int q = 0; q < (int)n_; q++) logical |= ((phys >> pos_[q]) & 1) << q;
        out_ri[2 * (int64_t)logical * stride] = all[2 * phys];
        out_ri[2 * (int64_t)logical * stride + 1] = all[2 * phys + 1];
    }
}

// ---- finite shots (reference: QuantumDevice.hpp:476-503) ------------------------------
// Mean eigenvalue of one flattened term (a product of named Paulis / Hadamards) over `shots` samples:
// the state is rotated IN PLACE to the term's eigenbasis (X: H; Y: H S^dagger; Hadamard: RY(-pi/4) —
// the reference rotates a copy, which at 30 qubits would cost a second 16 GiB vector), the term's wires
// are sampled with the alias sampler on the runtime's RNG stream (one seed draw per term), and the
// rotation is undone; rotation and un-rotation of consecutive terms fuse into the same tile passes.
// Eigen-decomposition of a small Hermitian matrix (cyclic complex Jacobi rotations), eigenvalues ascending:
// M = V diag(lam) V^dagger.  Host side of finite-shot Hermitian observables (2^k x 2^k, k <= ~5).
static void hermitian_eig(const CVec &m, size_t dim, std::vector<double> &lam, CVec &V)
{
    CVec A = m;
    V.assign(dim * dim, cplx(0.0));
    for (size_t i = 0; i < dim; i++) V[i * dim + i] = 1.0;
    for (int sweep = 0; sweep < 100; sweep++) {
        double off = 0.0;
        for (size_t p = 0; p < dim; p++)
            for (size_t q = p + 1; q < dim; q++) off += std::norm(A[p * dim + q]);
        if (off < 1e-290) break;
        for (size_t p = 0; p < dim; p++)
            for (size_t q = p + 1; q < dim; q++) {
                const double g = std::abs(A[p * dim + q]);
                if (g == 0.0) continue;
                const cplx e = A[p * dim + q] / g; // phase of the pivot
                const double th = 0.5 * std::atan2(2.0 * g, (A[p * dim + p] - A[q * dim + q]).real());
                const double c = std::cos(th), sn = std::sin(th);
                for (size_t k = 0; k < dim; k++) { // A <- A R
                    const cplx x = A[k * dim + p], y = A[k * dim + q];
                    A[k * dim + p] = c * x + sn * std::conj(e) * y;
                    A[k * dim + q] = -sn * e * x + c * y;
                }
                for (size_t k = 0; k < dim; k++) { // A <- R^dagger A
                    const cplx x = A[p * dim + k], y = A[q * dim + k];
                    A[p * dim + k] = c * x + sn * e * y;
                    A[q * dim + k] = -sn * std::conj(e) * x + c * y;
                }
                for (size_t k = 0; k < dim; k++) {
                    const cplx x = V[k * dim + p], y = V[k * dim + q];
                    V[k * dim + p] = c * x + sn * std::conj(e) * y;
                    V[k * dim + q] = -sn * e * x + c * y;
                }
            }
    }
    std::vector<size_t> order(dim);
    for (size_t i = 0; i < dim; i++) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t x, size_t y) { return A[x * dim + x].real() < A[y * dim + y].real(); });
    lam.resize(dim);
    CVec Vs(dim * dim);
    for (size_t j = 0; j < dim; j++) {
        lam[j] = A[order[j] * dim + order[j]].real();
        for (size_t k = 0; k < dim; k++) Vs[k * dim + j] = V[k * dim + order[j]];
    }
    V.swap(Vs);
}

// First and second moment of one flattened term (a product of named Paulis / Hadamards / Hermitian matrices on
// distinct wires) over `shots` samples.  Hermitian factors are diagonalised on the host (eigenvalues ascending)
// and rotated with V^dagger; a shot's value is the product of its factors' eigenvalues.
void Simulator::sampled_term_moments(const ObsTerm &t, double &m1, double &m2)
{
    std::vector<GateOp> pre, post;
    std::vector<int64_t> wires;
    struct HermPart {
        size_t first, k; // position of its first wire in `wires`, number of wires
        std::vector<double> lam;
    };
    std::vector<HermPart> herm;
    size_t npauli_first = 0; // Pauli-type wires are listed first: their eigenvalue is (-1)^popcount of those bits
    std::vector<const ObsFactor *> hfac;
    for (auto &f : t.factors) {
        if (f.named_id == 0) continue;
        if (f.named_id == 5) {
            hfac.push_back(&f);
            continue;
        }
        B200_FAIL_IF(f.wires.size() != 1, "Named observable acts on one wire");
        for (auto w : wires) B200_FAIL_IF(w == f.wires[0], "Repeated wire in a tensor observable with finite shots");
        wires.push_back(f.wires[0]);
        const std::vector<int64_t> w{f.wires[0]};
        if (f.named_id == 1) {
            canonicalize(pre, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "Hadamard", {}, w, false, {}, {}, "");
        }
        else if (f.named_id == 2) {
            canonicalize(pre, "S", {}, w, true, {}, {}, "");
            canonicalize(pre, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "S", {}, w, false, {}, {}, "");
        }
        else if (f.named_id == 4) {
            canonicalize(pre, "RY", {-M_PI / 4}, w, false, {}, {}, "");
            canonicalize(post, "RY", {M_PI / 4}, w, false, {}, {}, "");
        }
    }
    npauli_first = wires.size();
    for (const ObsFactor *f : hfac) {
        const size_t k = f->wires.size(), dim = size_t(1) << k;
        B200_FAIL_IF(k > 6, "Hermitian observables on more than 6 wires do not support finite shots");
        for (auto w : f->wires) {
            for (auto x : wires) B200_FAIL_IF(x == w, "Repeated wire in a tensor observable with finite shots");
            wires.push_back(w);
        }
        HermPart hp;
        hp.first = wires.size() - k;
        hp.k = k;
        CVec V, Vd(dim * dim);
        hermitian_eig(*f->matrix, dim, hp.lam, V);
        for (size_t i = 0; i < dim; i++)
            for (size_t j = 0; j < dim; j++) Vd[i * dim + j] = std::conj(V[j * dim + i]);
        canonicalize_matrix(pre, Vd, f->wires, false, {}, {});
        canonicalize_matrix(post, V, f->wires, false, {}, {});
        herm.push_back(std::move(hp));
    }
    if (wires.empty()) {
        m1 = m2 = 1.0;
        return;
    }
    enqueue(std::move(pre));
    const auto idx = sample_indices(wires, opt_.shots); // flushes
    enqueue(std::move(post));
    const size_t K = wires.size();
    double a1 = 0.0, a2 = 0.0;
    for (auto i : idx) {
        const uint64_t u = (uint64_t)i;
        double v = (__builtin_popcountll(u >> (K - npauli_first)) & 1) ? -1.0 : 1.0;
        for (auto &hp : herm) {
            const uint64_t sub = (u >> (K - hp.first - hp.k)) & ((1ull << hp.k) - 1ull);
            v *= hp.lam[(size_t)sub];
        }
        a1 += v;
        a2 += v * v;
    }
    m1 = a1 / (double)opt_.shots;
    m2 = a2 / (double)opt_.shots;
}

double Simulator::sampled_term_mean(const ObsTerm &t)
{
    double m1, m2;
    sampled_term_moments(t, m1, m2);
    return m1;
}

// Probs / PartialProbs as the interface sees them: sampled frequencies when shots > 0
void Simulator::probs_measured(const std::vector<int64_t> &wires_in, double *out, int64_t stride)
{
    if (opt_.shots <= 0) return probs(wires_in, out, stride);
    B200_FAIL_IF(!psi_, "No qubits allocated");
    check_wires(wires_in, {});
    const size_t k = wires_in.empty() ? (size_t)n_ : wires_in.size();
    B200_FAIL_IF(k > 30, "Sampled probabilities of more than 30 wires are not supported");
    const auto idx = sample_indices(wires_in, opt_.shots);
    std::vector<double> h(size_t(1) << k, 0.0);
    for (auto i : idx) h[(size_t)i] += 1.0;
    for (size_t i = 0; i < h.size(); i++) out[(int64_t)i * stride] = h[i] / (double)opt_.shots;
}

void Simulator::probs(const std::vector<int64_t> &wires_in, double *out, int64_t stride)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    flush();
    check_wires(wires_in, {});
    bool all = wires_in.empty();
    if ((int64_t)wires_in.size() == n_) {
        all = true;
        for (size_t j = 0; j < wires_in.size(); j++) all = all && wires_in[j] == (int64_t)j;
    }
    const size_t k = wires_in.empty() ? (size_t)n_ : wires_in.size();
    if (sharded() && all && opt_.gather_slices) { // gather='slice': |psi|^2 of this rank's slice only, no collective but the swaps
        canonical_rank_bits();
        ensure_buffer(probs_buf_, sizeof(double) << nl_);
        bool local_identity = true;
        for (int q = 0; q < (int)nl_; q++) local_identity = local_identity && pos_[q] == q;
        if (local_identity) CUDA_CHECK(b200k_probs_all(psi_->d(), probs_buf_->d(), (int)nl_, stream_));
        else {
            uint8_t perm[64];
            for (int q = 0; q < (int)nl_; q++) perm[pos_[q]] = (uint8_t)q;
            CUDA_CHECK(b200k_probs_all_perm(psi_->d(), probs_buf_->d(), (int)nl_, perm, stream_));
        }
        note_launch();
        if (out) d2h_strided(probs_buf_->p(), sizeof(double), size_t(1) << nl_, out + ((int64_t)opt_.rank << nl_) * stride, stride);
        return;
    }
    if (!probs_buf_) {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        if (probs_keep_ && probs_keep_->bytes() >= (sizeof(double) << k)) probs_buf_ = std::move(probs_keep_);
        else probs_keep_.reset();
    }
    ensure_buffer(probs_buf_, sizeof(double) << k);
    if (!sharded()) {
        if (all && wire_identity()) {
            CUDA_CHECK(b200k_probs_all(psi_->d(), probs_buf_->d(), (int)n_, stream_));
        }
        else if (all) { // wires were relabelled by SWAPs: undo the relabelling on the way out
            uint8_t perm[64];
            for (int64_t w = 0; w < n_; w++) perm[wire_bit_[(size_t)w]] = (uint8_t)(n_ - 1 - w);
            CUDA_CHECK(b200k_probs_all_perm(psi_->d(), probs_buf_->d(), (int)n_, perm, stream_));
        }
        else {
            std::vector<uint8_t> tb(k);
            for (size_t j = 0; j < k; j++) tb[j] = (uint8_t)lbit(wires_in[j]);
            const size_t sp = (size_t)b200k_probs_splits((int)n_, (int)k);
            ensure_buffer(misc_, (sizeof(double) << k) * sp);
            CUDA_CHECK(b200k_probs_partial(psi_->d(), (int)n_, tb.data(), (int)k, misc_->d(),
                                           probs_buf_->d(), stream_));
        }
        note_launch();
        if (out) d2h_strided(probs_buf_->p(), sizeof(double), size_t(1) << k, out, stride);
        return;
    }
    // sharded: marginal over the wires that sit on local bits (kernel), wires on rank bits select the
    // output slots; partials are gathered and summed in rank order on every rank
    std::vector<int64_t> wires = wires_in;
    if (wires.empty())
        for (int64_t w = 0; w < n_; w++) wires.push_back(w);
    std::vector<uint8_t> tb;
    std::vector<size_t> local_j;
    for (size_t j = 0; j < k; j++) {
        const int p = pos_[lbit(wires[j])];
        if (p < nl_) {
            tb.push_back((uint8_t)p);
            local_j.push_back(j);
        }
    }
    const size_t kl = tb.size(), nl_bins = size_t(1) << kl;
    B200_FAIL_IF(((double)nl_bins * opt_.world) * 8.0 > 64.0 * 1073741824.0,
                 "Probs: the requested marginal is too large to gather on every rank of a sharded register");
    ensure_buffer(tmp_probs_, sizeof(double) << kl);
    if (kl == 0) { // every requested wire is a rank bit: this shard's total mass
        CUDA_CHECK(b200k_bilinear_pauli(psi_->d(), psi_->d(), (int)nl_, rank_base_, 0, 0, 0, 0, 0,
                                        scratch_->d(), result_slot(0), stream_));
        CUDA_CHECK(cudaMemcpyAsync(tmp_probs_->p(), result_slot(0), sizeof(double), cudaMemcpyDeviceToDevice,
                                   (cudaStream_t)stream_));
    }
    else {
        const size_t sp = (size_t)b200k_probs_splits((int)nl_, (int)kl);
        ensure_buffer(misc_, (sizeof(double) << kl) * sp);
        CUDA_CHECK(b200k_probs_partial(psi_->d(), (int)nl_, tb.data(), (int)kl, misc_->d(), tmp_probs_->d(), stream_));
    }
    note_launch();
    std::vector<double> mine(nl_bins), gathered(nl_bins * (size_t)opt_.world), full(size_t(1) << k, 0.0);
    CUDA_CHECK(cudaMemcpyAsync(mine.data(), tmp_probs_->p(), nl_bins * sizeof(double), cudaMemcpyDeviceToHost,
                               (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    try {
        comm_->allgather_host(mine.data(), gathered.data(), nl_bins);
    }
    catch (const std::exception &e) {
        B200_FAIL(std::string("Probs: gather across ranks failed: ") + e.what());
    }
    for (int r = 0; r < opt_.world; r++) {
        size_t fixed = 0;
        for (size_t j = 0; j < k; j++) {
            const int p = pos_[lbit(wires[j])];
            if (p >= nl_ && ((r >> (p - nl_)) & 1)) fixed |= size_t(1) << (k - 1 - j);
        }
        for (size_t jl = 0; jl < nl_bins; jl++) {
            size_t o = fixed;
            for (size_t i = 0; i < kl; i++)
                if ((jl >> (kl - 1 - i)) & 1) o |= size_t(1) << (k - 1 - local_j[i]);
            full[o] += gathered[(size_t)r * nl_bins + jl];
        }
    }
    CUDA_CHECK(cudaMemcpyAsync(probs_buf_->p(), full.data(), full.size() * sizeof(double), cudaMemcpyHostToDevice,
                               (cudaStream_t)stream_));
    CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream_));
    if (out)
        for (size_t i = 0; i < full.size(); i++) out[(int64_t)i * stride] = full[i];
}

// ---- Sample -------------------------------------------------------------------------------------------
// Reference stream (verified against doc/releases/changelog-0.9.0.md:
// 299-315): one 32-bit output of the runtime generator seeds a fresh mt19937 per call; each shot
// takes two uniform_real_distribution<double> draws from it.
// MCMC sampling (kwargs mcmc / num_burnin / kernel_name, reference shape: frontend/catalyst/from_plxpr/from_plxpr.py:194):
// PARITY UNPINNED -- the algorithm is pennylane-lightning's generate_samples_metropolis, whose transition kernels seed
// themselves from std::random_device; restated with one mt19937 per call (seeded like the alias sampler) so that it
// is reproducible.  The proposal stream is drawn here, the chain runs as one thread on the GPU.
std::vector<int64_t> Simulator::mcmc_indices(const std::vector<int64_t> &wires_in, int64_t shots)
{
    B200_FAIL_IF(sharded(), "MCMC sampling is not supported on a sharded register");
    B200_FAIL_IF(opt_.kernel_name != "Local" && opt_.kernel_name != "NonZeroRandom",
                 "Unsupported MCMC kernel '" + opt_.kernel_name + "' (Local, NonZeroRandom)");
    check_wires(wires_in, {});
    if (!wire_identity()) restore_wire_order(); // the chain walks logical indices: canonical layout
    flush();
    std::vector<int64_t> wires = wires_in;
    if (wires.empty())
        for (int64_t w = 0; w < n_; w++) wires.push_back(w);
    std::mt19937 rng(fresh_seed());
    const bool local = opt_.kernel_name == "Local";
    const int64_t steps = opt_.num_burnin + shots;
    std::vector<int64_t> idx((size_t)shots);
    if (shots == 0) return idx;
    int64_t nnz = 0;
    std::unique_ptr<DeviceBuffer> nzbuf;
    if (!local) {
        nzbuf = std::make_unique<DeviceBuffer>(sizeof(int64_t) << n_);
        CUDA_CHECK(b200k_nonzero_indices(psi_->d(), (int)n_, 1e-24, static_cast<int64_t *>(nzbuf->p()), &nnz, stream_));
        note_launch();
        B200_FAIL_IF(nnz <= 0, "MCMC sampling: the state has no non-zero amplitude");
    }
    std::vector<uint32_t> site(local ? (size_t)steps : 0);
    std::vector<uint64_t> jump(local ? 0 : (size_t)steps);
    std::vector<double> u((size_t)steps);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    for (int64_t s = 0; s < steps; s++) {
        if (local) {
            const uint32_t q = (uint32_t)rng() % (uint32_t)n_, v = (uint32_t)rng() & 1u;
            site[(size_t)s] = q | (v << 31);
        }
        else {
            const uint64_t hi = (uint32_t)rng(), lo = (uint32_t)rng();
            jump[(size_t)s] = (hi << 32) | lo;
        }
        u[(size_t)s] = U(rng);
    }
    const size_t b_site = (site.size() * 4 + 15) & ~size_t(15), b_jump = jump.size() * 8, b_u = u.size() * 8, b_out = idx.size() * 8;
    ensure_buffer(misc_, b_site + b_jump + b_u + b_out + 64);
    char *base = static_cast<char *>(misc_->p());
    uint32_t *dsite = reinterpret_cast<uint32_t *>(base);
    uint64_t *djump = reinterpret_cast<uint64_t *>(base + b_site);
    double *du = reinterpret_cast<double *>(base + b_site + b_jump);
    int64_t *dout = reinterpret_cast<int64_t *>(base + b_site + b_jump + b_u);
    cudaStream_t st = (cudaStream_t)stream_;
    if (local) CUDA_CHECK(cudaMemcpyAsync(dsite, site.data(), site.size() * 4, cudaMemcpyHostToDevice, st));
    else CUDA_CHECK(cudaMemcpyAsync(djump, jump.data(), b_jump, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(cudaMemcpyAsync(du, u.data(), b_u, cudaMemcpyHostToDevice, st));
    CUDA_CHECK(b200k_mcmc_chain(psi_->d(), (int)n_, steps, opt_.num_burnin, local ? dsite : nullptr, local ? nullptr : djump,
                                local ? nullptr : static_cast<const int64_t *>(nzbuf->p()), nnz, du, dout, stream_));
    note_launch();
    CUDA_CHECK(cudaMemcpyAsync(idx.data(), dout, b_out, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    stats_.h2d_bytes += (double)(site.size() * 4 + b_jump + b_u);
    stats_.d2h_bytes += (double)b_out;
    // the chain recorded full-register indices (wire 0 = most significant bit): project onto the requested wires
    const size_t k = wires.size();
    for (auto &x : idx) {
        const uint64_t full = (uint64_t)x;
        uint64_t o = 0;
        for (size_t j = 0; j < k; j++) o = (o << 1) | ((full >> (n_ - 1 - wires[j])) & 1ull);
        x = (int64_t)o;
    }
    return idx;
}

// Sharded register: bring the layout to wire order, so that rank r's shard IS the slice [r << n_local, (r+1) << n_local)
// of the canonical index (wire 0 = most significant bit).  Rank bits first: every wire w < log2(world) whose qubit is
// not on rank bit n-1-w is exchanged into place with qubit swaps (a qubit that sits on ANOTHER rank bit is pulled in
// first); all ranks run the same host logic on the same layout, so the collective calls match.
void Simulator::canonical_rank_bits()
{
    if (!sharded()) return;
    if (!wire_identity()) restore_wire_order(); // virtual SWAP labels back to identity: logical bit of wire w = n-1-w
    flush();
    for (int round = 0; round < 8; round++) {
        std::vector<std::pair<int, int>> pairs;
        uint64_t pull = 0;
        for (int p = (int)nl_; p < (int)n_; p++) { // rank-bit position p must hold logical bit p
            if (pos_[p] == p) continue;
            if (pos_[p] < nl_) pairs.push_back({p, pos_[p]});
            else pull |= 1ull << p;
        }
        if (pairs.empty() && !pull) return;
        if (!pairs.empty()) swap_bits({psi_->d()}, pos_, pairs);
        else make_local({psi_->d()}, pos_, pull); // sits on another rank bit: comes in now, goes to its place next round
    }
    B200_FAIL("internal: rank bits did not reach wire order");
}

// ... then the local bits.  A transposition of two PHYSICAL index bits A, B of the shard is what a content SWAP of the
// two logical bits that sit there does; a sequence of them (each moving one qubit's data to its own position) sorts the
// local bits, after which pos_ is the identity.  All SWAPs run under the layout they were planned in and fuse into
// tile passes; the logical state never changes, only where its qubits are stored.
void Simulator::canonical_local_bits()
{
    if (!sharded()) return;
    canonical_rank_bits();
    std::vector<int> inv0((size_t)n_, -1), cur(pos_.begin(), pos_.end());
    for (int q = 0; q < (int)n_; q++) inv0[(size_t)pos_[q]] = q;
    std::vector<GateOp> ops;
    for (int q = 0; q < (int)nl_; q++) {
        if (cur[q] == q) continue;
        const int A = cur[q], B = q;
        int q2 = -1;
        for (int r = 0; r < (int)nl_; r++)
            if (cur[r] == B) q2 = r;
        B200_FAIL_IF(q2 < 0 || A >= nl_, "internal: local layout is not a permutation of the local bits");
        canonicalize(ops, "SWAP", {}, {n_ - 1 - inv0[(size_t)A], n_ - 1 - inv0[(size_t)B]}, false, {}, {}, "");
        cur[q] = B;
        cur[q2] = A;
    }
    if (ops.empty()) return;
    std::vector<int> work = pos_;
    execute({psi_->d()}, work, ops);
    B200_FAIL_IF(work != pos_, "internal: local SWAPs moved a rank bit");
    pos_ = cur;
}

// Sample / Counts of ALL wires of a sharded register without gathering 2^n probabilities anywhere: every rank keeps
// |psi|^2 of its own slice (canonical order), the chunk sums of the alias restatement are computed per slice and
// gathered (2^n / 1024 x 2 doubles), and every rank draws every shot, reading the few chunks a shot walks from the
// owning rank's slice through peer memory.  Same kernels on the same numbers as on one GPU: bit-identical samples.
// false = not applicable here (peer memory unavailable, slices smaller than a chunk): the caller gathers instead.
bool Simulator::sample_all_distributed(int64_t shots, std::vector<int64_t> &idx)
{
    if (!sharded() || opt_.dist_sampler == 0 || nl_ < 10 || opt_.world > B200K_SAMPLER_MAX_SHARDS) return false;
    canonical_rank_bits();
    cudaStream_t st = (cudaStream_t)stream_;
    const int64_t nloc = int64_t(1) << nl_, ntot = int64_t(1) << n_;
    ensure_buffer(probs_buf_, sizeof(double) << nl_);
    bool local_identity = true;
    for (int q = 0; q < (int)nl_; q++) local_identity = local_identity && pos_[q] == q;
    if (local_identity) CUDA_CHECK(b200k_probs_all(psi_->d(), probs_buf_->d(), (int)nl_, stream_));
    else {
        uint8_t perm[64];
        for (int q = 0; q < (int)nl_; q++) perm[pos_[q]] = (uint8_t)q; // source (physical) bit -> canonical bit
        CUDA_CHECK(b200k_probs_all_perm(psi_->d(), probs_buf_->d(), (int)nl_, perm, stream_));
    }
    note_launch();
    std::vector<void *> peers;
    bool ok = false;
    try {
        ok = comm_->map_peers(probs_buf_->p(), peers); // collective; the same answer on every rank
    }
    catch (const std::exception &) {
        ok = false;
    }
    if (!ok) return false;
    idx.assign((size_t)shots, 0);
    auto done = [&]() { // nobody may still be reading this rank's slice when it is reused
        cudaStreamSynchronize(st);
        comm_->unmap_peers(probs_buf_->p());
        try {
            comm_->barrier();
        }
        catch (const std::exception &) {
        }
    };
    try {
        const int64_t nch_l = nloc / 1024;
        const size_t scr = (size_t)b200k_sampler_scratch_doubles(ntot) * sizeof(double);
        const size_t ubytes = 2 * (size_t)shots * sizeof(double), ibytes = (size_t)shots * sizeof(int64_t);
        const size_t sums = 2 * (size_t)nch_l * sizeof(double);
        auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
        ensure_buffer(misc_, up(scr) + up(ubytes) + up(ibytes) + up(sums) + up(sums * (size_t)opt_.world) + 256);
        char *base = static_cast<char *>(misc_->p());
        double *dscr = reinterpret_cast<double *>(base);
        double *du = reinterpret_cast<double *>(base + up(scr));
        int64_t *di = reinterpret_cast<int64_t *>(base + up(scr) + up(ubytes));
        double *mine = reinterpret_cast<double *>(base + up(scr) + up(ubytes) + up(ibytes));
        double *all = reinterpret_cast<double *>(base + up(scr) + up(ubytes) + up(ibytes) + up(sums));
        CUDA_CHECK(b200k_sampler_chunk_sums(probs_buf_->d(), nloc, ntot, mine, mine + nch_l, stream_));
        comm_->allgather(mine, all, 2 * (size_t)nch_l, stream_);
        int64_t offD = 0, offE = 0;
        b200k_sampler_sum_offsets(ntot, &offD, &offE);
        for (int r = 0; r < opt_.world; r++) { // descending global order: the highest rank's chunks come first
            const int64_t blk = (int64_t)(opt_.world - 1 - r) * nch_l;
            const double *src = all + (size_t)r * 2 * (size_t)nch_l;
            CUDA_CHECK(cudaMemcpyAsync(dscr + offD + blk, src, (size_t)nch_l * 8, cudaMemcpyDeviceToDevice, st));
            CUDA_CHECK(cudaMemcpyAsync(dscr + offE + blk, src + nch_l, (size_t)nch_l * 8, cudaMemcpyDeviceToDevice, st));
        }
        CUDA_CHECK(b200k_sampler_finish(dscr, ntot, stream_));
        std::mt19937 rng(fresh_seed());
        if (shots > 0) {
            std::vector<double> u(2 * (size_t)shots);
            std::uniform_real_distribution<double> U(0.0, 1.0);
            for (auto &x : u) x = U(rng);
            CUDA_CHECK(cudaMemcpyAsync(du, u.data(), ubytes, cudaMemcpyHostToDevice, st));
            std::vector<const double *> sl((size_t)opt_.world);
            for (int r = 0; r < opt_.world; r++) sl[(size_t)r] = static_cast<const double *>(peers[(size_t)r]);
            comm_->stream_barrier(stream_); // every rank's slice is complete before anyone reads it
            CUDA_CHECK(b200k_sampler_draw_sharded(sl.data(), opt_.world, (int)nl_, dscr, du, shots, di, stream_));
            note_launch();
            CUDA_CHECK(cudaMemcpyAsync(idx.data(), di, ibytes, cudaMemcpyDeviceToHost, st));
            CUDA_CHECK(cudaStreamSynchronize(st)); // u is stack-owned
            stats_.h2d_bytes += (double)ubytes;
            stats_.d2h_bytes += (double)ibytes;
        }
    }
    catch (...) {
        done();
        throw;
    }
    done();
    return true;
}

std::vector<int64_t> Simulator::sample_indices(const std::vector<int64_t> &wires, int64_t shots)
{
    B200_FAIL_IF(shots < 0, "Invalid number of shots");
    if (opt_.mcmc) return mcmc_indices(wires, shots);
    if (sharded() && opt_.dist_sampler != 0) {
        B200_FAIL_IF(!psi_, "No qubits allocated");
        check_wires(wires, {});
        bool all = wires.empty();
        if ((int64_t)wires.size() == n_) {
            all = true;
            for (size_t j = 0; j < wires.size(); j++) all = all && wires[j] == (int64_t)j;
        }
        if (all) {
            std::vector<int64_t> idx;
            if (sample_all_distributed(shots, idx)) return idx;
        }
        B200_FAIL_IF(all && opt_.dist_sampler == 1,
                     "dist_sampler=1: the distributed sampler is not available here (no peer memory between the ranks, "
                     "or fewer than 2^10 amplitudes per rank)");
    }
    probs(wires, nullptr, 1); // leaves the distribution in probs_buf_ on the device
    const size_t k = wires.empty() ? (size_t)n_ : wires.size();
    const int64_t nbins = int64_t(1) << k;
    std::mt19937 rng(fresh_seed());
    std::vector<int64_t> idx((size_t)shots);
    if (shots == 0) return idx;
    std::vector<double> u(2 * (size_t)shots);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    for (auto &x : u) x = U(rng);
    const size_t scr = (size_t)b200k_sampler_scratch_doubles(nbins) * sizeof(double);
    const size_t ubytes = u.size() * sizeof(double);
    const size_t ibytes = idx.size() * sizeof(int64_t);
    ensure_buffer(misc_, scr + ubytes + ibytes + 64);
    char *base = static_cast<char *>(misc_->p());
    double *dscr = reinterpret_cast<double *>(base);
    double *du = reinterpret_cast<double *>(base + ((scr + 15) & ~size_t(15)));
    int64_t *di = reinterpret_cast<int64_t *>(reinterpret_cast<char *>(du) + ubytes);
    cudaStream_t st = (cudaStream_t)stream_;
    CUDA_CHECK(cudaMemcpyAsync(du, u.data(), ubytes, cudaMemcpyHostToDevice, st));
    stats_.h2d_bytes += (double)ubytes;
    stats_.d2h_bytes += (double)ibytes;
    CUDA_CHECK(b200k_sampler_prepare(probs_buf_->d(), nbins, dscr, stream_));
    CUDA_CHECK(b200k_sampler_draw(probs_buf_->d(), nbins, dscr, du, shots, di, stream_));
    note_launch();
    CUDA_CHECK(cudaMemcpyAsync(idx.data(), di, ibytes, cudaMemcpyDeviceToHost, st));
    CUDA_CHECK(cudaStreamSynchronize(st));
    return idx;
}

// ---- Measure --------------------------------------------------------------------------------------------
// Draw semantics: runtime/lib/backend/common/Utils.hpp:312-338 (float cast of the draw and one
// extra discarded generator output are part of the stream; verified against
// doc/releases/changelog-0.8.0.md:70-108).
bool Simulator::draw_outcome(const double p[2], int postselect)
{
    // simulateDraw (reference: runtime/lib/backend/common/Utils.hpp:312-338): one float draw from
    // the borrowed mt19937 compared with p0, then one more engine call
    if (postselect != -1) {
        B200_FAIL_IF(postselect < 0 || postselect > 1, "Invalid postselect value");
        B200_FAIL_IF(p[postselect] == 0, "Probability of postselect value is 0");
        return postselect == 1;
    }
    std::uniform_real_distribution<> dis(0., 1.);
    float draw;
    if (rng_fn_ || own_seeded_) {
        BorrowedEngine eng{[this] { return rng_next(); }};
        draw = (float)dis(eng);
        (void)rng_next();
    }
    else {
        std::mt19937 g(fresh_seed());
        draw = (float)dis(g);
    }
    return draw > p[0];
}

bool Simulator::measure(int64_t wire, int postselect)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    double p[2];
    probs({wire}, p, 1);
    const bool res = draw_outcome(p, postselect);
    const double scale = 1.0 / std::sqrt(p[res ? 1 : 0]);
    CUDA_CHECK(b200k_collapse(psi_->d(), (int)nl_, rank_base_, pos_[lbit(wire)], res ? 1 : 0, scale, stream_));
    note_launch();
    return res;
}

// PauliMeasure (reference: runtime/include/QuantumDevice.hpp:531, caller RuntimeCAPI.cpp:962-993):
// projective measurement of the Pauli product `word` on `wires`; result 0 <-> eigenvalue +1.
// The word is rotated to the Z basis (X: H; Y: H S^dagger), the parity expectation is one read pass,
// the projector (1 +- Z..Z)/2 with its renormalisation is ONE diagonal PARITY op, and the rotation
// is undone; rotation, projector and un-rotation fuse into the surrounding tile passes.
bool Simulator::pauli_measure(const std::string &word, const std::vector<int64_t> &wires, int postselect)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    B200_FAIL_IF(word.size() != wires.size(), "The length of the pauli string must be equal to the number of wires.");
    check_wires(wires, {});
    uint64_t zmask = 0;
    std::vector<GateOp> pre, post;
    for (size_t k = 0; k < wires.size(); k++) {
        const std::vector<int64_t> w{wires[k]};
        switch (word[k]) {
        case 'I': break;
        case 'Z': zmask |= bit_of(wires[k]); break;
        case 'X':
            canonicalize(pre, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "Hadamard", {}, w, false, {}, {}, "");
            zmask |= bit_of(wires[k]);
            break;
        case 'Y':
            canonicalize(pre, "S", {}, w, true, {}, {}, "");
            canonicalize(pre, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "Hadamard", {}, w, false, {}, {}, "");
            canonicalize(post, "S", {}, w, false, {}, {}, "");
            zmask |= bit_of(wires[k]);
            break;
        default: B200_FAIL("Invalid Pauli word '" + word + "'");
        }
    }
    double p[2] = {1.0, 0.0};
    if (zmask) {
        enqueue(std::move(pre));
        flush();
        const uint64_t zm = to_phys(pos_, zmask);
        const double one = 1.0;
        ensure_buffer(misc_, 16);
        uint64_t *dz = static_cast<uint64_t *>(misc_->p());
        double *dc = reinterpret_cast<double *>(dz + 1);
        CUDA_CHECK(cudaMemcpyAsync(dz, &zm, 8, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        CUDA_CHECK(cudaMemcpyAsync(dc, &one, 8, cudaMemcpyHostToDevice, (cudaStream_t)stream_));
        CUDA_CHECK(b200k_expval_diag_batch(psi_->d(), (int)nl_, rank_base_, dz, dc, 1, scratch_->d(), result_slot(0),
                                           stream_));
        note_launch();
        std::vector<double> h;
        read_results(1, h); // synchronises; sums over ranks in fixed order
        const double ev = h[0];
        p[0] = std::min(1.0, std::max(0.0, 0.5 * (1.0 + ev)));
        p[1] = std::min(1.0, std::max(0.0, 0.5 * (1.0 - ev)));
    }
    const bool res = draw_outcome(p, postselect);
    if (!zmask) return res; // identity word: eigenvalue +1 with certainty, state untouched
    GateOp proj;
    proj.kind = GateOp::PARITY;
    proj.zmask = zmask;
    const double nrm = 1.0 / std::sqrt(p[res ? 1 : 0]);
    proj.m[0] = res ? cplx(0.0) : cplx(nrm);
    proj.m[1] = res ? cplx(nrm) : cplx(0.0);
    std::vector<GateOp> ops{proj};
    enqueue(std::move(ops));
    enqueue(std::move(post));
    return res;
}

// ---- tape + adjoint Jacobian ------------------------------------------------------------------------
void Simulator::start_tape()
{
    B200_FAIL_IF(recording_, "Cannot re-activate the cache manager");
    recording_ = true;
    tape_.clear();
    tape_obs_.clear();
}
void Simulator::stop_tape()
{
    B200_FAIL_IF(!recording_, "Cannot stop an already stopped cache manager");
    recording_ = false;
}
int64_t Simulator::tape_num_params() const
{
    size_t P = 0;
    for (auto &op : tape_) P += op.params.size();
    return (int64_t)P;
}

namespace {
// U(theta) = exp(i * sigma * theta * G) with G = projector(controls) (x) generator.
struct Generator {
    double sigma = 0;
    bool is_pair = false;
    uint64_t xmask = 0, zmask = 0;
    int ny = 0;               // Pauli form
    uint64_t tmask = 0, a = 0, b = 0;
    double g[8] = {0};        // pair form
    uint64_t extra_cmask = 0; // projector bits that must be 1
};
} // namespace

// <phi| P_t |lam> for a list of Pauli-word terms (logical masks; optional control predicate), written to
// t.out as (re, im).  Terms are independent, so they are packed first-fit into batches of up to
// B200K_BATCH_GENS whose X/Y bits fit one tile, and each batch costs ONE read pass over the two vectors
// (k_bilinear_batch) instead of one pass per term.  `movable` / `lay`: the vectors that share the
// layout of phi and lam (sharded registers: a term with an X/Y factor on a rank bit pulls that qubit in).
void Simulator::overlaps_batched(const double *phi, const double *lam, const std::vector<double *> &movable,
                                 std::vector<int> &lay, const std::vector<OvTerm> &terms, int want)
{
    const int bt = (int)std::min<int64_t>(11, nl_);
    const int cmin = std::min(3, bt);
    struct Bin {
        uint64_t req = 0;
        int rc = 0; // raw component the batch computes (the kernel is uniform in it)
        std::vector<const OvTerm *> t;
    };
    auto rc_of = [&](const OvTerm &t) { return (t.ny & 1) ^ (want == B200K_WANT_IMAG ? 1 : 0); };
    auto launch = [&](const Bin &b) {
        const int c = fit_rows(b.req, bt);
        std::vector<int> high;
        for (int p = c; p < 64; p++)
            if ((b.req >> p) & 1) high.push_back(p);
        for (int p = c; (int)high.size() < bt - c; p++)
            if (!((b.req >> p) & 1)) high.push_back(p);
        std::sort(high.begin(), high.end());
        int loc[64];
        for (int p = 0; p < 64; p++) loc[p] = p < c ? p : -1;
        for (size_t h = 0; h < high.size(); h++) loc[high[h]] = c + (int)h;
        auto split = [&](uint64_t mask, uint32_t &inside, uint64_t &outside) {
            inside = 0;
            outside = 0;
            while (mask) {
                const int p = __builtin_ctzll(mask);
                mask &= mask - 1;
                if (loc[p] >= 0) inside |= 1u << loc[p];
                else outside |= 1ull << p;
            }
        };
        B200k2Header H;
        std::memset(&H, 0, sizeof(H));
        H.t = bt;
        H.c = c;
        H.nhigh = (int)high.size();
        for (size_t h = 0; h < high.size(); h++) H.high_pos[h] = (uint8_t)high[h];
        H.rank_base = rank_base_;
        H.ntiles = 1ull << (nl_ - bt);
        std::vector<B200kBatchGen> gens(b.t.size());
        std::vector<double *> outs(b.t.size());
        for (size_t q = 0; q < b.t.size(); q++) {
            const OvTerm &p = *b.t[q];
            B200kBatchGen &G = gens[q];
            std::memset(&G, 0, sizeof(G));
            uint64_t out_x;
            split(to_phys(lay, p.xmask), G.xin, out_x);
            B200_FAIL_IF(out_x != 0, "internal: generator bit outside the batch tile");
            split(to_phys(lay, p.zmask), G.zin, G.zo);
            split(to_phys(lay, p.cmask), G.cin_m, G.co_m);
            split(to_phys(lay, p.cval), G.cin_v, G.co_v);
            G.ny = p.ny;
            G.want = want;
            outs[q] = p.out;
        }
        ensure_buffer(batch_scratch_, (size_t)b200k_bilinear_batch_scratch_doubles() * sizeof(double));
        CUDA_CHECK(b200k_bilinear_batch(phi, lam, (int)nl_, &H, gens.data(), (int)gens.size(), outs.data(),
                                        batch_scratch_->d(), stream_));
        note_launch();
    };
    // diagonal words (no X/Y factor): a streaming pass per 32 of them, any layout (rank bits enter through rank_base)
    std::vector<const OvTerm *> rest;
    {
        std::vector<const OvTerm *> dg[2];
        for (auto &t : terms) (t.xmask == 0 ? dg[rc_of(t)] : rest).push_back(&t);
        for (auto &grp : dg)
            for (size_t o = 0; o < grp.size(); o += B200K_BATCH_GENS) {
                const int cnt = (int)std::min<size_t>(B200K_BATCH_GENS, grp.size() - o);
                uint64_t z[B200K_BATCH_GENS], cm[B200K_BATCH_GENS], cv[B200K_BATCH_GENS];
                int ny[B200K_BATCH_GENS];
                double *outs[B200K_BATCH_GENS];
                for (int q = 0; q < cnt; q++) {
                    const OvTerm &p = *grp[o + q];
                    z[q] = to_phys(lay, p.zmask);
                    cm[q] = to_phys(lay, p.cmask);
                    cv[q] = to_phys(lay, p.cval);
                    ny[q] = p.ny;
                    outs[q] = p.out;
                }
                ensure_buffer(batch_scratch_, (size_t)b200k_bilinear_batch_scratch_doubles() * sizeof(double));
                CUDA_CHECK(b200k_bilinear_diag_batch(phi, lam, (int)nl_, rank_base_, z, cm, cv, ny, want, cnt, outs,
                                                     batch_scratch_->d(), stream_));
                note_launch();
            }
    }
    while (!rest.empty()) {
        // every term whose X/Y bits are local under the current layout, packed first-fit
        std::vector<Bin> bins;
        std::vector<const OvTerm *> later;
        for (const OvTerm *t : rest) {
            const uint64_t x = to_phys(lay, t->xmask);
            if (x >> nl_) {
                later.push_back(t);
                continue;
            }
            B200_FAIL_IF(fit_rows(x, bt) < 0, "internal: a Pauli word acts non-diagonally on more qubits than a tile holds");
            bool placed = false;
            for (auto &b : bins)
                if (b.rc == rc_of(*t) && b.t.size() < (size_t)B200K_BATCH_GENS && fit_rows(b.req | x, bt) >= cmin) {
                    b.req |= x;
                    b.t.push_back(t);
                    placed = true;
                    break;
                }
            if (!placed) {
                bins.emplace_back();
                bins.back().req = x;
                bins.back().rc = rc_of(*t);
                bins.back().t.push_back(t);
            }
        }
        for (auto &b : bins) launch(b);
        if (!later.empty()) make_local(movable, lay, later[0]->xmask); // sharded: one more qubit comes in
        rest.swap(later);
    }
}

// Weighted Pauli words, many per read pass: words are packed (largest X/Y support first, first fit) into bins whose
// combined X/Y bits fit one tile -- the low c >= 3 index bits plus up to 11 - c scattered higher ones -- and every bin
// is one launch of b200k_pauli_words_tile.  C4 (26 qubits, 1000 random words): 70 bins instead of 950 passes.
void Simulator::pauli_words_binned(std::vector<TileWordRef> words, const std::vector<double *> &movable,
                                   std::vector<int> &lay,
                                   const std::function<void(const B200k2Header &, const std::vector<B200kTileWord> &)> &launch)
{
    const int bt = (int)std::min<int64_t>(11, nl_);
    const int cmin = std::min(3, bt);
    std::stable_sort(words.begin(), words.end(), [](const TileWordRef &a, const TileWordRef &b) {
        return __builtin_popcountll(a.xmask) > __builtin_popcountll(b.xmask);
    });
    struct Bin {
        uint64_t req = 0;
        std::vector<const TileWordRef *> w;
    };
    auto emit = [&](const Bin &b) {
        const int c = fit_rows(b.req, bt);
        std::vector<int> high;
        for (int p = c; p < 64; p++)
            if ((b.req >> p) & 1) high.push_back(p);
        for (int p = c; (int)high.size() < bt - c; p++)
            if (!((b.req >> p) & 1)) high.push_back(p);
        std::sort(high.begin(), high.end());
        int loc[64];
        for (int p = 0; p < 64; p++) loc[p] = p < c ? p : -1;
        for (size_t h = 0; h < high.size(); h++) loc[high[h]] = c + (int)h;
        B200k2Header H;
        std::memset(&H, 0, sizeof(H));
        H.t = bt;
        H.c = c;
        H.nhigh = (int)high.size();
        for (size_t h = 0; h < high.size(); h++) H.high_pos[h] = (uint8_t)high[h];
        H.rank_base = rank_base_;
        H.ntiles = 1ull << (nl_ - bt);
        std::vector<B200kTileWord> ws(b.w.size());
        for (size_t q = 0; q < b.w.size(); q++) {
            const TileWordRef &r = *b.w[q];
            B200kTileWord &W = ws[q];
            std::memset(&W, 0, sizeof(W));
            for (uint64_t m = to_phys(lay, r.xmask); m; m &= m - 1) {
                const int p = __builtin_ctzll(m);
                B200_FAIL_IF(loc[p] < 0, "internal: Pauli word bit outside the tile");
                W.xin |= 1u << loc[p];
            }
            for (uint64_t m = to_phys(lay, r.zmask); m; m &= m - 1) {
                const int p = __builtin_ctzll(m);
                if (loc[p] >= 0) W.zin |= 1u << loc[p];
                else W.zo |= 1ull << p;
            }
            W.cr = r.coef.real();
            W.ci = r.coef.imag();
            W.ny = r.ny;
        }
        launch(H, ws);
    };
    std::vector<const TileWordRef *> rest;
    for (auto &w : words) rest.push_back(&w);
    while (!rest.empty()) {
        std::vector<Bin> bins;
        std::vector<const TileWordRef *> later;
        for (const TileWordRef *w : rest) {
            const uint64_t x = to_phys(lay, w->xmask);
            if (x >> nl_) { // an X/Y factor on a rank bit: that qubit has to come in first
                later.push_back(w);
                continue;
            }
            B200_FAIL_IF(fit_rows(x, bt) < cmin, "internal: a Pauli word acts non-diagonally on more qubits than a tile holds");
            bool placed = false;
            for (auto &b : bins)
                if (b.w.size() < (size_t)B200K_TILE_WORDS && fit_rows(b.req | x, bt) >= cmin) {
                    b.req |= x;
                    b.w.push_back(w);
                    placed = true;
                    break;
                }
            if (!placed) {
                bins.emplace_back();
                bins.back().req = x;
                bins.back().w.push_back(w);
            }
        }
        for (auto &b : bins) emit(b);
        if (!later.empty()) {
            B200_FAIL_IF(movable.empty(), "internal: Pauli word with a non-local X/Y factor");
            make_local(movable, lay, later[0]->xmask);
        }
        rest.swap(later);
    }
}

std::vector<double> Simulator::gradient(const std::vector<int64_t> &train_in)
{
    B200_FAIL_IF(!psi_, "No qubits allocated");
    const size_t P = (size_t)tape_num_params();
    std::vector<size_t> train;
    if (train_in.empty())
        for (size_t i = 0; i < P; i++) train.push_back(i);
    else
        for (auto t : train_in) {
            B200_FAIL_IF(t < 0 || (size_t)t >= P, "Invalid trainable parameter index");
            train.push_back((size_t)t);
        }
    const size_t T = train.size(), K = tape_obs_.size();
    NvtxRange nvtx("b200sv::adjoint_gradient");
    std::vector<double> jac(K * T, 0.0);
    if (K * T == 0) return jac;
    for (auto &o : tape_obs_)
        B200_FAIL_IF(o.second != 1, "Unsupported measurements to compute gradient; Adjoint "
                                    "differentiation method only supports expectation return type");
    B200_FAIL_IF(!std::is_sorted(train.begin(), train.end()), "trainable parameters must be sorted");
    flush();
    const size_t bytes = sizeof(cplx) << nl_;
    cudaStream_t st = (cudaStream_t)stream_;
    // sharded: psi_ keeps its own layout pos_; lambda and the phi_k share the working layout `lay`
    // (they are swapped together when an adjoint gate or a generator needs a rank-bit qubit)
    uint64_t obs_x = 0;
    for (size_t k = 0; k < K; k++) obs_x |= observable_xmask(tape_obs_[k].first);
    make_local({psi_->d()}, pos_, obs_x);
    std::vector<int> lay = pos_;
    // lambda and the phi_k live in buffers kept between calls (see grad_work_)
    struct Lease {
        Simulator *s;
        ~Lease()
        {
            std::lock_guard<std::mutex> lk(g_keep_mu);
            s->grad_busy_ = false;
            if (!s->opt_.keep_allocation) s->grad_work_.clear();
        }
    } lease{this};
    {
        std::lock_guard<std::mutex> lk(g_keep_mu);
        grad_busy_ = true;
        if (!grad_work_.empty() && grad_work_[0]->bytes() != bytes) grad_work_.clear();
    }
    while (grad_work_.size() < K + 1) grad_work_.push_back(std::make_unique<DeviceBuffer>(bytes)); // may trim OTHER devices
    DeviceBuffer &lambda = *grad_work_[0];
    std::vector<DeviceBuffer *> phi;
    CUDA_CHECK(cudaMemcpyAsync(lambda.p(), psi_->p(), bytes, cudaMemcpyDeviceToDevice, st));
    std::vector<double *> vecs{lambda.d()};
    for (size_t k = 0; k < K; k++) {
        phi.push_back(grad_work_[1 + k].get());
        apply_observable(tape_obs_[k].first, psi_->d(), phi[k]->d(), pos_);
        vecs.push_back(phi[k]->d());
    }
    ensure_buffer(results_, std::max<size_t>(4096, 2 * K * T) * sizeof(double));

    auto generator_of = [&](const TapeOp &op) -> Generator {
        Generator g;
        const std::string &nm = op.name;
        auto W = [&](size_t i) { return bit_of(op.wires[i]); };
        auto pauli1 = [&](char c, size_t wi) {
            if (c == 'X') g.xmask |= W(wi);
            else if (c == 'Y') { g.xmask |= W(wi); g.zmask |= W(wi); g.ny++; }
            else if (c == 'Z') g.zmask |= W(wi);
        };
        auto pairg = [&](uint64_t tm, uint64_t a, uint64_t b, cplx g00, cplx g01, cplx g10, cplx g11) {
            g.is_pair = true;
            g.tmask = tm; g.a = a; g.b = b;
            const cplx v[4] = {g00, g01, g10, g11};
            for (int i = 0; i < 4; i++) { g.g[2 * i] = v[i].real(); g.g[2 * i + 1] = v[i].imag(); }
        };
        g.sigma = -0.5;
        if (nm == "RX") pauli1('X', 0);
        else if (nm == "RY") pauli1('Y', 0);
        else if (nm == "RZ") pauli1('Z', 0);
        else if (nm == "PhaseShift") { g.sigma = 1.0; g.extra_cmask = W(0); }
        else if (nm == "IsingXX") { pauli1('X', 0); pauli1('X', 1); }
        else if (nm == "IsingYY") { pauli1('Y', 0); pauli1('Y', 1); }
        else if (nm == "IsingZZ") { pauli1('Z', 0); pauli1('Z', 1); }
        else if (nm == "IsingXY") { g.sigma = 0.5; pairg(W(0) | W(1), W(1), W(0), 0, 1, 1, 0); }
        else if (nm == "ControlledPhaseShift") { g.sigma = 1.0; g.extra_cmask = W(0) | W(1); }
        else if (nm == "CRX") { pauli1('X', 1); g.extra_cmask = W(0); }
        else if (nm == "CRY") { pauli1('Y', 1); g.extra_cmask = W(0); }
        else if (nm == "CRZ") { pauli1('Z', 1); g.extra_cmask = W(0); }
        else if (nm == "SingleExcitation") pairg(W(0) | W(1), W(1), W(0), 0, -I_, I_, 0);
        else if (nm == "DoubleExcitation")
            pairg(W(0) | W(1) | W(2) | W(3), W(2) | W(3), W(0) | W(1), 0, -I_, I_, 0);
        else if (nm == "MultiRZ")
            for (size_t i = 0; i < op.wires.size(); i++) pauli1('Z', i);
        else if (nm == "PauliRot")
            for (size_t i = 0; i < op.wires.size(); i++) pauli1(op.pauli_word[i], i);
        else if (nm == "GlobalPhase") g.sigma = -1.0;
        else if (nm == "PSWAP") { g.sigma = 1.0; pairg(W(0) | W(1), W(1), W(0), 1, 0, 0, 1); }
        else
            B200_FAIL("The operation is not supported using the adjoint differentiation method: " + nm);
        if (op.inverse) g.sigma = -g.sigma;
        return g;
    };

    // Adjoint ops not yet applied to lambda / phi.  They are applied LAZILY: the overlap
    // <phi_j| G |lambda_j> with phi_j = A phi', lambda_j = A lambda' (A = the pending unitaries) equals
    // <phi'| G |lambda'> whenever G commutes with every pending op, i.e. on every shared bit both act
    // diagonally.  A QAOA cost layer (42 ZZ rotations) or a layer of one-qubit rotations on distinct
    // wires then needs ONE fused application instead of one per parameter: per parameter only the read
    // pass of the overlap remains (2 x 16 x 2^n bytes instead of 6 x).
    // Overlaps whose generators commute with the pending gates all refer to the SAME pair of vectors, so
    // they are collected and evaluated in batches: one read pass of (phi_k, lambda) per batch of up to
    // B200K_BATCH_GENS Pauli-word generators whose X/Y bits fit one tile (k_bilinear_batch), instead of
    // one read pass per parameter.
    std::vector<OvTerm> pend_ov;
    auto flush_overlaps = [&]() {
        if (pend_ov.empty()) return;
        for (size_t k = 0; k < K; k++) {
            std::vector<OvTerm> terms = pend_ov;
            for (auto &t : terms) t.out = result_slot((int)(k * T + t.tag));
            overlaps_batched(phi[k]->d(), lambda.d(), vecs, lay, terms, B200K_WANT_IMAG); // jac = -2 sigma Im<phi|G|lambda>
        }
        pend_ov.clear();
    };
    std::vector<GateOp> pending;
    uint64_t pendX = 0, pendZ = 0;
    auto apply_pending = [&]() {
        flush_overlaps();
        if (pending.empty()) return;
        execute(vecs, lay, pending);
        pending.clear();
        pendX = pendZ = 0;
    };

    int64_t tp = (int64_t)T - 1;
    int64_t cur = (int64_t)P - 1;
    std::vector<double> sig(K * T, 0.0);
    for (int64_t j = (int64_t)tape_.size() - 1; j >= 0 && tp >= 0; j--) {
        const TapeOp &op = tape_[j];
        B200_FAIL_IF(op.params.size() > 1,
                     "The operation is not supported using the adjoint differentiation method: " + op.name);
        if (!op.params.empty()) {
            if ((size_t)cur == train[tp]) {
                const Generator g = generator_of(op);
                uint64_t cmask = g.extra_cmask, cval = g.extra_cmask;
                for (size_t i = 0; i < op.ctrl_wires.size(); i++) {
                    cmask |= bit_of(op.ctrl_wires[i]);
                    if (op.ctrl_values[i]) cval |= bit_of(op.ctrl_wires[i]);
                }
                {
                    const uint64_t gx = g.is_pair ? g.tmask : g.xmask;
                    const uint64_t gz = (g.is_pair ? 0 : g.zmask) | cmask;
                    if (!opt_.lazy_adjoint || (gx & (pendX | pendZ)) || (gz & pendX)) apply_pending();
                }
                if (opt_.batch_overlaps && !g.is_pair) { // evaluated together with its neighbours, see flush_overlaps
                    pend_ov.push_back({g.xmask, g.zmask, cmask, cval, g.ny, nullptr, (size_t)tp});
                    for (size_t k = 0; k < K; k++) sig[k * T + (size_t)tp] = g.sigma;
                    tp--;
                    cur--;
                    goto adjoint_of_this_op;
                }
                make_local(vecs, lay, g.is_pair ? g.tmask : g.xmask);
                for (size_t k = 0; k < K; k++) {
                    double *slot = result_slot((int)(k * T + (size_t)tp));
                    if (g.is_pair)
                        CUDA_CHECK(b200k_bilinear_pair(phi[k]->d(), lambda.d(), (int)nl_, rank_base_,
                                                       to_phys(lay, g.tmask), to_phys(lay, g.a),
                                                       to_phys(lay, g.b), g.g, to_phys(lay, cmask),
                                                       to_phys(lay, cval), scratch_->d(), slot, stream_));
                    else
                        CUDA_CHECK(b200k_bilinear_pauli(phi[k]->d(), lambda.d(), (int)nl_, rank_base_,
                                                        to_phys(lay, g.xmask), to_phys(lay, g.zmask), g.ny,
                                                        to_phys(lay, cmask), to_phys(lay, cval),
                                                        scratch_->d(), slot, stream_));
                    note_launch();
                    sig[k * T + (size_t)tp] = g.sigma;
                }
                tp--;
            }
            cur--;
        }
    adjoint_of_this_op:
        const size_t before = pending.size();
        if (op.name == "QubitUnitary")
            canonicalize_matrix(pending, op.matrix, op.wires, !op.inverse, op.ctrl_wires, op.ctrl_values);
        else
            canonicalize(pending, op.name, op.params, op.wires, !op.inverse, op.ctrl_wires,
                         op.ctrl_values, op.pauli_word);
        for (size_t q = before; q < pending.size(); q++) {
            pendX |= pending[q].xmask;
            pendZ |= pending[q].zsupport();
        }
    }
    flush_overlaps();
    std::vector<double> h;
    read_results((int)(K * T), h);
    for (size_t i = 0; i < K * T; i++) jac[i] = -2.0 * sig[i] * h[2 * i + 1];
    return jac;
}

} // namespace b200

// =====================================================================================================
// C ABI
// =====================================================================================================
using b200::Simulator;

struct b200sv_device {
    std::unique_ptr<Simulator> sim;
};
static std::atomic<int> g_live_devices{0};

static thread_local std::string g_err;

#define B200_TRY(...)                                                                              \
    try {                                                                                          \
        d->sim->make_current(); /* the caller's thread may have another CUDA device current */     \
        __VA_ARGS__;                                                                               \
        return 0;                                                                                  \
    }                                                                                              \
    catch (const std::exception &e) {                                                              \
        g_err = e.what();                                                                          \
        return -1;                                                                                 \
    }

static std::vector<int64_t> wv(const int64_t *w, int64_t n) { return std::vector<int64_t>(w, w + (n > 0 ? n : 0)); }
static std::vector<bool> bv(const uint8_t *v, int64_t n)
{
    std::vector<bool> r((size_t)(n > 0 ? n : 0));
    for (int64_t i = 0; i < n; i++) r[(size_t)i] = v[i] != 0;
    return r;
}
static b200::CVec cvv(const double *ri, size_t n)
{
    b200::CVec r(n);
    for (size_t i = 0; i < n; i++) r[i] = b200::cplx(ri[2 * i], ri[2 * i + 1]);
    return r;
}

extern "C" {

const char *b200sv_last_error(void) { return g_err.c_str(); }
const char *b200sv_version(void) { return "b200sv 0.1 (sm_100a)"; }

b200sv_device *b200sv_create(const char *kwargs)
{
    try {
        auto d = std::make_unique<b200sv_device>();
        d->sim = std::make_unique<Simulator>(b200::parse_kwargs(kwargs));
        g_live_devices.fetch_add(1);
        return d.release();
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}
void b200sv_destroy(b200sv_device *d)
{
    if (!d) return;
    // $B200SV_STATS_FILE: one JSON line of counters per device activation, appended when the device goes away
    // (the hook the reference's ResourceTracker plays for null.qubit: NullQubit.hpp:104-112) -- how a host that
    // only sees the plugin (librt_capi -> librtd_b200sv.so) gets at the counters
    if (const char *path = std::getenv("B200SV_STATS_FILE")) {
        try {
            const std::string js = d->sim->stats_json();
            if (FILE *f = fopen(path, "a")) {
                fprintf(f, "%s\n", js.c_str());
                fclose(f);
            }
        }
        catch (...) {
        }
    }
    const bool jit_used = d->sim->uses_jit();
    delete d;
    // the last device of the process: no background compilation may outlive it (see jit::Engine::quiesce)
    if (g_live_devices.fetch_sub(1) == 1 && jit_used) b200::jit_quiesce();
}

int b200sv_allocate_qubits(b200sv_device *d, int64_t n) { B200_TRY(d->sim->allocate(n)) }
int b200sv_allocate_qubit(b200sv_device *d) { B200_TRY(d->sim->allocate_one()) }
int b200sv_release_qubits(b200sv_device *d) { B200_TRY(d->sim->release()) }
int b200sv_release_qubit(b200sv_device *d, int64_t wire) { B200_TRY(d->sim->release_one(wire)) }
int64_t b200sv_num_qubits(b200sv_device *d) { return d->sim->num_qubits(); }
int b200sv_set_shots(b200sv_device *d, int64_t shots) { B200_TRY(d->sim->set_shots(shots)) }
int64_t b200sv_get_shots(b200sv_device *d) { return d->sim->shots(); }
int b200sv_set_prng(b200sv_device *d, b200sv_rng_next_fn next, void *ctx) { B200_TRY(d->sim->set_prng(next, ctx)) }
int b200sv_seed(b200sv_device *d, uint32_t seed) { B200_TRY(d->sim->seed(seed)) }
uint32_t b200sv_prng_peek(b200sv_device *d) { return d->sim->prng_peek(); }

int b200sv_named_operation(b200sv_device *d, const char *name, const double *params, int64_t nparams,
                           const int64_t *wires, int64_t nwires, int inverse,
                           const int64_t *ctrl_wires, const uint8_t *ctrl_values, int64_t nctrl,
                           const char *pauli_word)
{
    B200_TRY(d->sim->named_op(name, std::vector<double>(params, params + (nparams > 0 ? nparams : 0)),
                              wv(wires, nwires), inverse != 0, wv(ctrl_wires, nctrl),
                              bv(ctrl_values, nctrl), pauli_word ? pauli_word : ""))
}
int b200sv_matrix_operation(b200sv_device *d, const double *matrix_ri, const int64_t *wires,
                            int64_t nwires, int inverse, const int64_t *ctrl_wires,
                            const uint8_t *ctrl_values, int64_t nctrl)
{
    B200_TRY(if (nwires < 1 || nwires > 10) throw b200::Error("Invalid number of wires for MatrixOperation");
             d->sim->matrix_op(cvv(matrix_ri, (size_t(1) << nwires) * (size_t(1) << nwires)),
                               wv(wires, nwires), inverse != 0, wv(ctrl_wires, nctrl),
                               bv(ctrl_values, nctrl)))
}
int b200sv_set_state(b200sv_device *d, const double *state_ri, int64_t stride, const int64_t *wires,
                     int64_t nwires)
{
    B200_TRY(if (nwires < 0 || nwires > 40) throw b200::Error("Invalid number of wires");
             b200::CVec st(size_t(1) << nwires);
             for (size_t i = 0; i < st.size(); i++)
                 st[i] = b200::cplx(state_ri[2 * (int64_t)i * stride], state_ri[2 * (int64_t)i * stride + 1]);
             d->sim->set_state(st, wv(wires, nwires)))
}
int b200sv_set_basis_state(b200sv_device *d, const int8_t *bits, int64_t stride, const int64_t *wires,
                           int64_t nwires)
{
    B200_TRY(std::vector<int8_t> b((size_t)(nwires > 0 ? nwires : 0));
             for (int64_t i = 0; i < nwires; i++) b[(size_t)i] = bits[i * stride];
             d->sim->set_basis_state(b, wv(wires, nwires)))
}

int64_t b200sv_observable(b200sv_device *d, int obs_id, const double *matrix_ri, int64_t nmatrix,
                          const int64_t *wires, int64_t nwires)
{
    try {
        d->sim->make_current();
        return d->sim->obs_named(obs_id, cvv(matrix_ri, (size_t)(nmatrix > 0 ? nmatrix : 0)), wv(wires, nwires));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int64_t b200sv_tensor_observable(b200sv_device *d, const int64_t *keys, int64_t nkeys)
{
    try {
        d->sim->make_current();
        return d->sim->obs_tensor(wv(keys, nkeys));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int64_t b200sv_hamiltonian_observable(b200sv_device *d, const double *coeffs, const int64_t *keys,
                                      int64_t nkeys)
{
    try {
        d->sim->make_current();
        return d->sim->obs_hamiltonian(std::vector<double>(coeffs, coeffs + (nkeys > 0 ? nkeys : 0)),
                                       wv(keys, nkeys));
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}

int b200sv_expval(b200sv_device *d, int64_t key, double *out) { B200_TRY(*out = d->sim->expval(key)) }
int b200sv_var(b200sv_device *d, int64_t key, double *out) { B200_TRY(*out = d->sim->var(key)) }
int b200sv_state(b200sv_device *d, double *out_ri, int64_t stride) { B200_TRY(d->sim->state(out_ri, stride)) }
int b200sv_probs(b200sv_device *d, const int64_t *wires, int64_t nwires, double *out, int64_t stride)
{
    B200_TRY(d->sim->probs_measured(wv(wires, nwires), out, stride))
}
int b200sv_sample(b200sv_device *d, const int64_t *wires, int64_t nwires, int64_t shots, double *out,
                  int64_t stride0, int64_t stride1)
{
    // layout: row-major (shot, wire) of 0.0/1.0, first listed wire first
    // (reference: runtime/include/QuantumDevice.hpp:369-396)
    B200_TRY(const auto idx = d->sim->sample_indices(wv(wires, nwires), shots);
             const int64_t k = nwires > 0 ? nwires : d->sim->num_qubits();
             for (int64_t s = 0; s < shots; s++)
                 for (int64_t w = 0; w < k; w++)
                     out[s * stride0 + w * stride1] = (double)((idx[(size_t)s] >> (k - 1 - w)) & 1))
}
int b200sv_counts(b200sv_device *d, const int64_t *wires, int64_t nwires, int64_t shots,
                  double *eigvals, int64_t eig_stride, int64_t *counts, int64_t cnt_stride)
{
    // eigvals[i] = i, counts[i] = #shots whose bitstring (first wire = MSB) is i; every entry
    // is written (reference: runtime/include/QuantumDevice.hpp:409-449)
    B200_TRY(const auto idx = d->sim->sample_indices(wv(wires, nwires), shots);
             const int64_t k = nwires > 0 ? nwires : d->sim->num_qubits();
             const int64_t nb = int64_t(1) << k;
             for (int64_t i = 0; i < nb; i++) {
                 eigvals[i * eig_stride] = (double)i;
                 counts[i * cnt_stride] = 0;
             }
             for (auto i : idx) counts[i * cnt_stride] += 1)
}
int b200sv_pauli_measure(b200sv_device *d, const char *word, const int64_t *wires, int64_t nwires, int postselect,
                         int *out)
{
    B200_TRY(*out = d->sim->pauli_measure(std::string(word ? word : ""), wv(wires, nwires), postselect) ? 1 : 0)
}
int b200sv_measure(b200sv_device *d, int64_t wire, int postselect, int *out)
{
    B200_TRY(*out = d->sim->measure(wire, postselect) ? 1 : 0)
}

int b200sv_start_tape_recording(b200sv_device *d) { B200_TRY(d->sim->start_tape()) }
int b200sv_stop_tape_recording(b200sv_device *d) { B200_TRY(d->sim->stop_tape()) }
int64_t b200sv_tape_num_params(b200sv_device *d) { return d->sim->tape_num_params(); }
int64_t b200sv_tape_num_observables(b200sv_device *d) { return d->sim->tape_num_obs(); }
int b200sv_gradient(b200sv_device *d, double *const *grad_rows, const int64_t *row_sizes,
                    const int64_t *strides, int64_t nrows, const int64_t *train, int64_t ntrain)
{
    B200_TRY(const int64_t K = d->sim->tape_num_obs();
             const int64_t P = d->sim->tape_num_params();
             const int64_t T = ntrain > 0 ? ntrain : P;
             if (K * T == 0) return 0;
             if (nrows != K) throw b200::Error("Invalid number of pre-allocated gradients");
             for (int64_t k = 0; k < K; k++)
                 if (row_sizes[k] < T) throw b200::Error("Invalid size of pre-allocated gradients");
             const auto jac = d->sim->gradient(wv(train, ntrain));
             for (int64_t k = 0; k < K; k++)
                 for (int64_t j = 0; j < T; j++) grad_rows[k][j * strides[k]] = jac[(size_t)(k * T + j)])
}

int b200sv_get_stats(b200sv_device *d, b200sv_stats *out) { B200_TRY(*out = d->sim->stats()) }
int b200sv_reset_stats(b200sv_device *d) { B200_TRY(d->sim->reset_stats()) }
int b200sv_set_timing(b200sv_device *d, int on) { B200_TRY(d->sim->set_timing(on != 0)) }
int b200sv_flush(b200sv_device *d) { B200_TRY(d->sim->flush()) }
int b200sv_fp64_peak(b200sv_device *d, double *dfma_per_second) { B200_TRY(*dfma_per_second = d->sim->fp64_peak()) }
int64_t b200sv_stats_json(b200sv_device *d, char *buf, int64_t cap)
{
    try {
        const std::string js = d->sim->stats_json();
        if ((int64_t)js.size() + 1 > cap) return -1;
        std::memcpy(buf, js.c_str(), js.size() + 1);
        return (int64_t)js.size();
    }
    catch (const std::exception &e) {
        g_err = e.what();
        return -1;
    }
}
int b200sv_synchronize(b200sv_device *d) { B200_TRY(d->sim->synchronize()) }
int b200sv_jit_wait(b200sv_device *d) { B200_TRY(d->sim->jit_wait()) }
void *b200sv_device_pointer(b200sv_device *d) { return d->sim->device_pointer(); }

int b200sv_nccl_unique_id(char hex_out[257])
{
    unsigned char id[128];
    std::string err;
    if (b200::nccl_unique_id(id, err) != 0) {
        g_err = err;
        return 1;
    }
    static const char *H = "0123456789abcdef";
    for (int i = 0; i < 128; i++) {
        hex_out[2 * i] = H[id[i] >> 4];
        hex_out[2 * i + 1] = H[id[i] & 15];
    }
    hex_out[256] = 0;
    return 0;
}
int b200sv_world_size(b200sv_device *d) { return d->sim->world(); }
int b200sv_rank(b200sv_device *d) { return d->sim->rank(); }
int64_t b200sv_layout(b200sv_device *d, int32_t *layout, int64_t cap)
{
    const std::vector<int> l = d->sim->layout();
    int64_t i = 0;
    for (; i < (int64_t)l.size() && i < cap; i++) layout[i] = l[(size_t)i];
    return i;
}

} // extern "C"
