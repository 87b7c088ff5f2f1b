// simulator.hpp — C++ host layer of the B200 state-vector device.
//
// Owns the complex128 state vector in HBM, turns the eager one-gate-at-a-time calls of the
// reference's device interface (runtime/include/QuantumDevice.hpp:215-567) into a lazy queue of
// canonical ops, plans that queue into fused tile passes (one HBM round trip each) and launches the
// sm_100a kernels through the thin C ABI of include/b200sv_kernels.h.  No reference headers and no
// PyTorch here; the vtable adapter for Catalyst's runtime lives in plugin.cpp.
#pragma once

#include <complex>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <optional>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "b200sv.h"
#include "b200sv_kernels.h"

namespace b200 {

using cplx = std::complex<double>;
using CVec = std::vector<cplx>;

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---- canonical op: everything a gate can become, in LOGICAL bit space (bit n-1-w for wire w) ----
struct GateOp {
    enum Kind { PAIR, PHASE, PARITY, PAULIROT, DENSE } kind = PAIR;
    uint64_t xmask = 0;          // non-diagonal target bits
    uint64_t cmask = 0, cval = 0; // control predicate (also selects where a PHASE applies)
    uint64_t a = 0, b = 0;       // PAIR: basis patterns (subsets of xmask)
    uint64_t zmask = 0;          // PARITY / PAULIROT: z part
    int ny = 0;                  // PAULIROT: number of Y factors
    cplx m[4];                   // PAIR: 2x2; PHASE: m[0]; PARITY: m[0] even, m[1] odd;
                                 // PAULIROT: m[0] = (cos, sin)
    std::vector<int> targets;    // DENSE: logical bits, most significant first
    std::shared_ptr<CVec> matrix; // DENSE
    uint64_t zsupport() const { return cmask | (kind == PARITY || kind == PAULIROT ? zmask : 0); }
};

enum class ObsKind { Named, Hermitian, Tensor, Hamiltonian };
struct ObsNode {
    ObsKind kind;
    int named_id = 0;
    std::vector<int64_t> wires;
    CVec matrix;
    std::vector<double> coeffs;
    std::vector<int64_t> children;
};

struct TapeOp {
    std::string name;
    std::vector<double> params;
    std::vector<int64_t> wires;
    bool inverse;
    CVec matrix;
    std::vector<int64_t> ctrl_wires;
    std::vector<bool> ctrl_values;
    std::string pauli_word;
};

// A flattened observable term: coef * product of factors
struct ObsFactor {
    int named_id;  // 0..4 named, 5 hermitian
    std::vector<int64_t> wires;
    const CVec *matrix = nullptr;
};
struct ObsTerm {
    double coef = 1.0;
    std::vector<ObsFactor> factors; // applied first-to-last
};
struct PauliTerm {
    cplx coef;
    uint64_t xmask = 0, zmask = 0;
    int ny = 0;
};

struct Options {
    int64_t shots = 0;
    // lightning.qubit's sampling kwargs (frontend/catalyst/from_plxpr/from_plxpr.py:194): Metropolis chain instead of
    // the alias sampler; transition kernels "Local" and "NonZeroRandom"
    bool mcmc = false;
    int64_t num_burnin = 100;
    std::string kernel_name = "Local";
    int tile_bits = 11;
    int min_row_bits = 2;     // shortest contiguous row of a tile: 2^2 amplitudes = 64 B (two sectors).  128-B rows
                              // were the limit until the end of round 2; 64-B rows cost ~1.6x a long-row pass but
                              // save a pass in C3 (18 -> 17: 132.7 -> 127.8 ms) and five in C5 at 30 qubits (412 -> 406 ms)
    bool adaptive_rows = true; // per pass, pick the row width that maximises gates per unit of pass time
    bool fuse = true;
    int device = -1; // -1: LOCAL_RANK or 0
    // sharding: this object owns the amplitudes whose top log2(world) index bits
    // equal `rank`; `comm` = "local:<group>" (threads of one process) or "nccl:<hex unique id>"
    int world = 1, rank = 0;
    std::string comm;
    int64_t stage_mib = 4096; // staging memory of the global<->local qubit swap (two pipelined halves)
    // run-time specialised tile passes (jit.hpp): 0 = off, 1 = auto (a pass structure seen
    // `jit_threshold` times is compiled in the background, the interpreter runs it until the
    // kernel is ready), 2 = always (compile synchronously at first sight; tests)
    int jit = 1;
    int jit_threshold = 2;
    int jit_regbits = 4;      // a thread owns 2^jit_regbits amplitudes in a specialised pass
    int jit_min_qubits = 22;  // smaller local states stay on the interpreter
    int jit_dbuf = 1;         // cp.async double-buffered tiles in specialised passes
    int jit_fuse_diag = 1;    // fuse runs of diagonal ops inside a register round
    int jit_memhint = 0;      // cache hints in specialised passes (experiment, B200SV_JIT_MEMHINT): 1 = streaming stores
                              // (st.global.cs), 2 = L2::256B prefetch on the tile loads of rows >= 256 B, 4 = L2::128B
    int jit_direct_store = 0; // the last register round of a specialised pass stores straight to global memory
                              // (measured at 30 qubits on a B200: 134.8 ms per circuit with and without, so off)
    int jit_uniform = 1;      // evaluate tile-uniform diagonal factors once per tile (warp-collective), not per thread
    int jit_factor = 1;       // factor u00 out of uncontrolled one-qubit gates, fold the product into one op
    bool virtual_swap = true; // uncontrolled SWAP gates relabel wires instead of moving amplitudes
    int p2p_swap = -1;        // qubit swap over mapped peer memory: 1 on, 0 off, -1 auto (on for 'local:' rank groups)
    int swap_push = -1;       // qubit swap as one push kernel into all peers' staging: 1 on, 0 off, -1 auto (over NCCL
                              // with 2 ranks; larger worlds take the staged NCCL path -- by measurement)
    int plan_refine = -1;     // planner refinement iterations per pass (-1: 6 from 2^27 local amplitudes on, else 0 --
                              // the dozen extra scans of the queue per pass then hide behind the previous pass on the GPU)
    int plan_refine_mode = 2; // 1: trade the least-used bit of the candidate (measured on the B200: C3 119.1 ms); 2: try
                              // the three least-used SCATTERED bits per step (fewer passes; predicted by the per-pass
                              // time model fitted to the measured passes; not measured itself)
    int dist_sampler = -1;    // sharded register, Sample / Counts of ALL wires: every rank draws every shot from the
                              // slices of |psi|^2 the ranks keep, read through peer memory (1 = required, 0 = gather the
                              // whole marginal on every rank, -1 = use it when peer memory is available)
    bool gather_slices = false; // sharded register: State() / Probs() of all wires fill only THIS rank's slice
                                // [rank << n_local, (rank + 1) << n_local) of the caller's array (gather='slice')
    bool batch_overlaps = true; // adjoint gradient: evaluate commuting generators' overlaps in batched read passes
    bool keep_allocation = true;  // a released device keeps its state / probability buffers for the next allocation of the
                                  // same size (the runtime pools device objects); freed process-wide when any cudaMalloc fails
    bool release_project = false; // ReleaseQubit of a wire that is not in a basis state: error (default) or keep
                                  // the more probable branch (release_entangled='project')
    bool lazy_adjoint = true; // adjoint gradient: apply U^dagger only when the next generator does not commute with it
};

class DeviceBuffer;
class Comm;

class Simulator {
  public:
    explicit Simulator(const Options &opt);
    ~Simulator();
    Simulator(const Simulator &) = delete;
    Simulator &operator=(const Simulator &) = delete;

    // lifecycle
    void allocate(int64_t n);
    void allocate_one();
    void release_one(int64_t wire);
    void release();
    int64_t num_qubits() const { return n_; }
    void set_shots(int64_t s) { opt_.shots = s; }
    int64_t shots() const { return opt_.shots; }
    void set_prng(b200sv_rng_next_fn fn, void *ctx);
    void seed(uint32_t s);
    uint32_t prng_peek();

    // gates
    void named_op(const std::string &name, const std::vector<double> &params,
                  const std::vector<int64_t> &wires, bool inverse,
                  const std::vector<int64_t> &cw, const std::vector<bool> &cv,
                  const std::string &pauli_word);
    void matrix_op(const CVec &matrix, const std::vector<int64_t> &wires, bool inverse,
                   const std::vector<int64_t> &cw, const std::vector<bool> &cv);
    void set_state(const CVec &st, const std::vector<int64_t> &wires);
    void set_basis_state(const std::vector<int8_t> &bits, const std::vector<int64_t> &wires);

    // observables
    int64_t obs_named(int id, const CVec &matrix, const std::vector<int64_t> &wires);
    int64_t obs_tensor(const std::vector<int64_t> &keys);
    int64_t obs_hamiltonian(const std::vector<double> &coeffs, const std::vector<int64_t> &keys);

    // measurements
    double expval(int64_t key);
    double var(int64_t key);
    void state(double *out_ri, int64_t stride);
    void probs(const std::vector<int64_t> &wires, double *out, int64_t stride); // analytic
    void probs_measured(const std::vector<int64_t> &wires, double *out, int64_t stride); // sampled if shots > 0
    std::vector<int64_t> sample_indices(const std::vector<int64_t> &wires, int64_t shots);
    std::vector<int64_t> mcmc_indices(const std::vector<int64_t> &wires, int64_t shots);
    bool measure(int64_t wire, int postselect);
    bool pauli_measure(const std::string &word, const std::vector<int64_t> &wires, int postselect);

    // tape / gradient
    void start_tape();
    void stop_tape();
    int64_t tape_num_params() const;
    int64_t tape_num_obs() const { return (int64_t)tape_obs_.size(); }
    std::vector<double> gradient(const std::vector<int64_t> &train); // K x T row-major

    // instrumentation
    b200sv_stats stats() const; // simulator_jit.cpp (adds the process-wide JIT counters)
    void reset_stats();
    double fp64_peak();            // measured thread-level DFMA instructions per second (microbenchmark)
    std::string stats_json() const; // counters as one JSON object
    void make_current(); // every public entry point runs on this object's CUDA device
    static void trim_kept_buffers(); // free every buffer kept across ReleaseQubits, process-wide (a cudaMalloc failed)
    void set_timing(bool on) { timing_ = on; }
    void flush();
    void synchronize();
    void jit_wait(); // block until no specialised pass is being compiled
    bool uses_jit() const { return jit_mode_ != 0 && !dry_; }
    void *device_pointer();
    int world() const { return opt_.world; }
    int rank() const { return opt_.rank; }
    int64_t num_local_bits() const { return nl_; }
    std::vector<int> layout() const { return pos_; } // logical bit -> physical index position

  private:
    // --- helpers -------------------------------------------------------------------------------
    uint64_t bit_of(int64_t wire) const;
    int lbit(int64_t wire) const;
    bool wire_identity() const;
    void restore_wire_order();
    std::vector<int> wire_bit_; // wire -> logical index bit
    void check_wires(const std::vector<int64_t> &wires, const std::vector<int64_t> &cw) const;
    void canonicalize(std::vector<GateOp> &out, const std::string &name,
                      const std::vector<double> &params, const std::vector<int64_t> &wires,
                      bool inverse, const std::vector<int64_t> &cw, const std::vector<bool> &cv,
                      const std::string &pauli_word) const;
    void canonicalize_matrix(std::vector<GateOp> &out, const CVec &matrix,
                             const std::vector<int64_t> &wires, bool inverse,
                             const std::vector<int64_t> &cw, const std::vector<bool> &cv) const;
    void enqueue(std::vector<GateOp> &&ops);
    // sharded execution: `ops` are in LOGICAL bit space, `pos` maps logical bits to physical index
    // positions (>= nl_ means a rank bit).  Runs every op on every vector of `vecs` (which share the
    // layout), inserting global<->local qubit swaps where an op acts non-diagonally on a rank bit.
    void execute(const std::vector<double *> &vecs, std::vector<int> &pos, std::vector<GateOp> &ops);
    void make_local(const std::vector<double *> &vecs, std::vector<int> &pos, uint64_t logical_mask);
    void swap_bits(const std::vector<double *> &vecs, std::vector<int> &pos,
                   std::vector<std::pair<int, int>> pairs); // (global position, local position)
    void swap_bits_impl(const std::vector<double *> &vecs, std::vector<int> &pos, std::vector<std::pair<int, int>> pairs);
    uint64_t to_phys(const std::vector<int> &pos, uint64_t logical_mask) const;
    GateOp phys_op(const std::vector<int> &pos, const GateOp &g) const;
    bool sharded() const { return opt_.world > 1; }
    bool ensure_psi_peers(); // collective: map every rank's state shard into this GPU's address space
    void drop_psi_peers();
    void canonical_rank_bits();                    // sharded: wire w < log2(world) onto rank bit n-1-w (qubit swaps)
    void canonical_local_bits();                   // ... and the local index bits into wire order (SWAP passes)
    bool sample_all_distributed(int64_t shots, std::vector<int64_t> &idx);
    bool ensure_stage_peers(size_t bytes); // collective: swap staging of >= bytes, mapped on every rank
    void drop_stage_peers();
    std::vector<void *> stage_peers_;
    void *stage_peers_ptr_ = nullptr;
    bool stage_peers_ok_ = false, stage_peers_failed_ = false;
    std::vector<void *> psi_peers_;
    void *psi_peers_ptr_ = nullptr;
    bool psi_peers_tried_ = false, psi_peers_ok_ = false;
    std::unique_ptr<DeviceBuffer> psi_keep_, probs_keep_; // guarded by the process-wide keep mutex (simulator.cpp)
    // work vectors of the adjoint method (lambda, phi_k), kept between Gradient calls (and across ReleaseQubits of a
    // pooled device): cudaMalloc + cudaFree of two state-sized buffers cost 10-40 ms per call at 26 qubits, more than
    // some of the gradients they serve.  Freed by trim_kept_buffers() when idle; grad_busy_ guarded by the keep mutex.
    std::vector<std::unique_ptr<DeviceBuffer>> grad_work_;
    bool grad_busy_ = false;
    bool psi_keep_mapped_ = false;                        // peers hold a mapping of psi_keep_: never trimmed
    void reduce_results(std::vector<double> &host); // sum over ranks, fixed order
    // plan + launch a list of canonical ops (PHYSICAL bit space, non-diagonal targets all local) on a
    // device vector holding this rank's 2^nl_ amplitudes
    void run_ops(double *vec, std::vector<GateOp> &ops);
    void launch_pass(double *vec, const std::vector<const GateOp *> &ops,
                     const std::vector<int> &req_high, int c);
    void launch_pass3(double *vec, const std::vector<const GateOp *> &ops,
                      const std::vector<int> &req_high, int c);
    bool launch_pass_jit(double *vec, const std::vector<const GateOp *> &ops,
                         const std::vector<int> &req_high, int c);
    int jit_mode_ = 0;
    bool dry_ = false; // B200SV_DRYRUN: plan, generate and compile only (development, no GPU)
    int pass_op_cap_ = 160;
    const double *upload_matrix(const CVec &m);

    void flatten(int64_t key, std::vector<ObsTerm> &out) const;
    bool term_to_paulis(const ObsTerm &t, std::vector<PauliTerm> &out) const;
    // vec_out = O vec_in  (vec_out != vec_in); uses tmp_ for non-Pauli terms
    void apply_observable(int64_t key, const double *vec_in, double *vec_out, const std::vector<int> &pos);
    uint64_t observable_xmask(int64_t key) const; // logical bits any term touches non-diagonally
    void factor_ops(const ObsFactor &f, std::vector<GateOp> &ops) const;
    cplx inner(const double *a, const double *b);
    void read_results(int nslots, std::vector<double> &host);
    double *result_slot(int slot);

    struct OvTerm { // one Pauli-word bilinear form <phi| P |lam> (logical masks)
        uint64_t xmask, zmask, cmask, cval;
        int ny;
        double *out; // device, (re, im)
        size_t tag;
    };
    void overlaps_batched(const double *phi, const double *lam, const std::vector<double *> &movable,
                          std::vector<int> &lay, const std::vector<OvTerm> &terms, int want);
    struct TileWordRef { // one weighted Pauli word (logical masks)
        uint64_t xmask, zmask;
        int ny;
        std::complex<double> coef;
    };
    // Packs words into bins whose X/Y bits fit one tile and calls `launch` once per bin (b200k_pauli_words_tile).
    void pauli_words_binned(std::vector<TileWordRef> words, const std::vector<double *> &movable, std::vector<int> &lay,
                            const std::function<void(const B200k2Header &, const std::vector<B200kTileWord> &)> &launch);
    bool draw_outcome(const double p[2], int postselect);
    double sampled_term_mean(const ObsTerm &t);
    void sampled_term_moments(const ObsTerm &t, double &m1, double &m2);
    double uniform01();
    uint32_t rng_next();
    uint32_t fresh_seed();

    void ensure_buffer(std::unique_ptr<DeviceBuffer> &buf, size_t bytes);
    void d2h_strided(const void *dsrc, size_t elem_bytes, size_t count, void *hdst, int64_t stride);
    static constexpr int kStage = 4;
    void *stage_host_[kStage] = {nullptr, nullptr, nullptr, nullptr}; // pinned staging of large device -> host results
    void *stage_ev_[kStage] = {nullptr, nullptr, nullptr, nullptr};
    void note_launch() { stats_.other_launches++; }

    Options opt_;
    int64_t n_ = 0;
    int64_t nl_ = 0;         // local index bits = n_ - g_
    int g_ = 0;              // rank bits = log2(world)
    uint64_t rank_base_ = 0; // rank << nl_
    std::vector<int> pos_;   // layout of psi_: logical bit -> physical position
    std::unique_ptr<Comm> comm_;
    std::unique_ptr<DeviceBuffer> stage_, tmp_probs_;
    void *xfer_stream_ = nullptr;                          // transfers of the staged qubit swap
    void *ev_packed_[2] = {nullptr, nullptr}, *ev_moved_[2] = {nullptr, nullptr};
    int cuda_device_ = 0;
    void *stream_ = nullptr;
    std::unique_ptr<DeviceBuffer> psi_, tmp_, acc_, scratch_, results_, probs_buf_, misc_;
    std::unique_ptr<DeviceBuffer> matrix_arena_, batch_scratch_;
    size_t arena_used_ = 0;
    double *pinned_ = nullptr;
    size_t pinned_doubles_ = 0;

    std::vector<GateOp> queue_;
    std::vector<ObsNode> obs_;
    bool recording_ = false;
    std::vector<TapeOp> tape_;
    std::vector<std::pair<int64_t, int>> tape_obs_;

    b200sv_rng_next_fn rng_fn_ = nullptr;
    void *rng_ctx_ = nullptr;
    std::mt19937 own_gen_;
    bool own_seeded_ = false;

    b200sv_stats stats_{};
    bool timing_ = false;
    void *ev0_ = nullptr, *ev1_ = nullptr, *ev2_ = nullptr, *ev3_ = nullptr;
    std::vector<std::pair<void *, void *>> ev_free_, ev_pending_; // event pairs of timed flushes, read back lazily
    void collect_timing();
};

Options parse_kwargs(const char *kwargs);
void jit_quiesce(); // simulator_jit.cpp: cancel queued background compilations, wait for the running ones

} // namespace b200
