/* b200sv.h — C ABI of the B200 state-vector device (libb200sv.so).
 *
 * This is the boundary a binding of the reference's device interface attaches to.  Each entry point
 * names the method of `Catalyst::Runtime::QuantumDevice` it implements (reference:
 * runtime/include/QuantumDevice.hpp, cited per function) and keeps that method's argument meaning
 * and error behaviour; the C++ vtable adapter that Catalyst's runtime dlopen()s
 * (catalyst_b200/csrc/plugin.cpp -> librtd_b200sv.so, factory symbol `B200StateVectorFactory`,
 * reference: QuantumDevice.hpp:30-33) is a thin shim over these calls.  INTEGRATION.md shows both
 * bindings.
 *
 * Conventions
 *   - `wires` are device wire indices 0..n-1; wire 0 is the most significant bit of every
 *     index-like output (reference: QuantumDevice.hpp:409-412,
 *     frontend/test/pytest/test_measurements_results.py:166-184).
 *   - complex128 data crosses as interleaved (re, im) doubles; matrices are row-major with the
 *     first listed wire as the most significant row/column bit (QuantumDevice.hpp:244).
 *   - every int-returning function returns 0 on success, nonzero on error; the message is
 *     retrievable with b200sv_last_error() (the adapter rethrows it as RuntimeException,
 *     reference: runtime/include/Exception.hpp:26-86).
 *   - gate calls only enqueue work; any call that returns data flushes and synchronises.
 *   - there is NO CPU fallback: creating a device without a usable CUDA device fails.
 */
#ifndef B200SV_H
#define B200SV_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200sv_device b200sv_device;

const char *b200sv_last_error(void);
const char *b200sv_version(void);

/* <Name>Factory(kwargs) — QuantumDevice.hpp:30-33; kwargs is the Python str(dict) the frontend
 * bakes into quantum.device (runtime/lib/backend/common/Utils.hpp:107-156).  Recognised keys:
 * shots, mcmc / num_burnin / kernel_name (lightning.qubit's sampling kwargs: Sample, Counts and the finite-shot
 * estimates then draw from a Metropolis chain, kernels 'Local' and 'NonZeroRandom'; parity unpinned, see
 * Simulator::mcmc_indices), tile_bits, min_row_bits, fuse (0/1), device (CUDA ordinal), world_size, rank, comm,
 * stage_mib, swap_push, p2p_swap, dist_sampler, gather ('all' | 'slice') (see "sharding" below), release_entangled
 * ('error' | 'project'), keep_allocation, plan_refine / plan_refine_mode (pass planner: refinement iterations per pass,
 * -1 = automatic from 2^27 local amplitudes; 1 = any-bit variant, >= 2 = scattered-bit variant), jit ('off' |
 * 'auto' | 'always') and its tuning keys.  Unknown keys are ignored.  Returns NULL on error. */
b200sv_device *b200sv_create(const char *kwargs);
void b200sv_destroy(b200sv_device *d);

/* AllocateQubits :88 / AllocateQubit :122 / ReleaseQubits :103 / ReleaseQubit :144 / GetNumQubits :110 */
int b200sv_allocate_qubits(b200sv_device *d, int64_t n);
int b200sv_allocate_qubit(b200sv_device *d); /* appends one wire in |0> */
int b200sv_release_qubits(b200sv_device *d); /* frees the state; object stays reusable */
/* ReleaseQubit :144 -- device wire `wire` leaves the register and every later wire moves down by one
 * (the caller's id map: runtime/lib/backend/common/QubitManager.hpp:124-126).  The qubit is expected in
 * a basis state (deallocation contract); its more probable branch is kept and renormalised. */
int b200sv_release_qubit(b200sv_device *d, int64_t wire);
int64_t b200sv_num_qubits(b200sv_device *d);
/* SetDeviceShots :164 / GetDeviceShots :171 */
int b200sv_set_shots(b200sv_device *d, int64_t shots);
int64_t b200sv_get_shots(b200sv_device *d);
/* SetDevicePRNG :192 — the runtime lends a std::mt19937*; across a C ABI that is a callback that
 * returns the engine's next 32-bit output.  NULL => unseeded (fresh random_device seed per draw,
 * as runtime/lib/backend/common/Utils.hpp:329-334). */
typedef uint32_t (*b200sv_rng_next_fn)(void *ctx);
int b200sv_set_prng(b200sv_device *d, b200sv_rng_next_fn next, void *ctx);
/* convenience for hosts without a runtime generator: own std::mt19937(seed)
 * (reference: runtime/lib/capi/ExecutionContext.hpp:309-321) */
int b200sv_seed(b200sv_device *d, uint32_t seed);
uint32_t b200sv_prng_peek(b200sv_device *d); /* next raw output without consuming (tests) */

/* NamedOperation :215-219 (caller runtime/lib/capi/RuntimeCAPI.cpp:449-814).
 * pauli_word is optional_params[0] of PauliRot (RuntimeCAPI.cpp:812-813), else NULL. */
int b200sv_named_operation(b200sv_device *d, const char *name, const double *params, int64_t nparams,
                           const int64_t *wires, int64_t nwires, int inverse,
                           const int64_t *ctrl_wires, const uint8_t *ctrl_values, int64_t nctrl,
                           const char *pauli_word);
/* MatrixOperation :250-255 (caller RuntimeCAPI.cpp:816-859) */
int b200sv_matrix_operation(b200sv_device *d, const double *matrix_ri, const int64_t *wires,
                            int64_t nwires, int inverse, const int64_t *ctrl_wires,
                            const uint8_t *ctrl_values, int64_t nctrl);
/* SetState :284 / SetBasisState :270 */
int b200sv_set_state(b200sv_device *d, const double *state_ri, int64_t stride, const int64_t *wires,
                     int64_t nwires);
int b200sv_set_basis_state(b200sv_device *d, const int8_t *bits, int64_t stride, const int64_t *wires,
                           int64_t nwires);

/* Observable :317 / TensorObservable :335 / HamiltonianObservable :353 — return key >= 0, -1 on error.
 * obs_id follows enum ObsId of runtime/include/Types.h:36-43 (0 Identity .. 4 Hadamard, 5 Hermitian). */
int64_t b200sv_observable(b200sv_device *d, int obs_id, const double *matrix_ri, int64_t nmatrix,
                          const int64_t *wires, int64_t nwires);
int64_t b200sv_tensor_observable(b200sv_device *d, const int64_t *keys, int64_t nkeys);
int64_t b200sv_hamiltonian_observable(b200sv_device *d, const double *coeffs, const int64_t *keys,
                                      int64_t nkeys);

/* Expval :488 / Var :503 */
int b200sv_expval(b200sv_device *d, int64_t key, double *out);
int b200sv_var(b200sv_device *d, int64_t key, double *out);
/* State :518 — out has 2^n complex entries spaced `stride` complex elements apart (DataView) */
int b200sv_state(b200sv_device *d, double *out_ri, int64_t stride);
/* Probs :462 (nwires == 0) / PartialProbs :472 — out has 2^k doubles spaced `stride` apart */
int b200sv_probs(b200sv_device *d, const int64_t *wires, int64_t nwires, double *out, int64_t stride);
/* Sample :384 / PartialSample :396 — out(shot, wire) = out[shot*stride0 + wire*stride1] in {0.0, 1.0} */
int b200sv_sample(b200sv_device *d, const int64_t *wires, int64_t nwires, int64_t shots, double *out,
                  int64_t stride0, int64_t stride1);
/* Counts :433 / PartialCounts :446 — all 2^k entries of both outputs are written */
int b200sv_counts(b200sv_device *d, const int64_t *wires, int64_t nwires, int64_t shots,
                  double *eigvals, int64_t eig_stride, int64_t *counts, int64_t cnt_stride);
/* Measure :234 — postselect -1 = none (RuntimeCAPI.cpp:950-960); *out = 0/1 */
int b200sv_measure(b200sv_device *d, int64_t wire, int postselect, int *out);
/* PauliMeasure :531 (caller RuntimeCAPI.cpp:962-993) — joint measurement of the Pauli product `word`
 * (one of I/X/Y/Z per wire); *out = 0 for eigenvalue +1, 1 for eigenvalue -1; the state collapses onto
 * that eigenspace.  Same draw as Measure; postselect -1 = none. */
int b200sv_pauli_measure(b200sv_device *d, const char *word, const int64_t *wires, int64_t nwires,
                         int postselect, int *out);

/* StartTapeRecording :574 / StopTapeRecording :581 / Gradient :564-567 (caller RuntimeCAPI.cpp:386-447).
 * grad_rows[k] receives d<O_k>/d theta_j for the recorded expvals k, j over `train` (ntrain == 0 =>
 * all recorded gate parameters in program order); element j of row k is grad_rows[k][j*strides[k]]. */
int b200sv_start_tape_recording(b200sv_device *d);
int b200sv_stop_tape_recording(b200sv_device *d);
int64_t b200sv_tape_num_params(b200sv_device *d);
int64_t b200sv_tape_num_observables(b200sv_device *d);
int b200sv_gradient(b200sv_device *d, double *const *grad_rows, const int64_t *row_sizes,
                    const int64_t *strides, int64_t nrows, const int64_t *train, int64_t ntrain);

/* ---- instrumentation (counters behind the device, not in the reference) ---- */
typedef struct b200sv_stats {
    int64_t gates_enqueued;   /* logical gates received (Identity excluded) */
    int64_t kernel_ops;       /* canonical ops executed inside tile passes */
    int64_t tile_passes;      /* launches of the fused tile kernel = HBM round trips */
    int64_t other_launches;   /* reductions, probs, collapse, sampler, copies */
    double gate_device_ms;    /* CUDA-event time of the tile passes since the last reset */
    double algorithmic_bytes; /* sum over gates of 2*16*2^(n-c) */
    int64_t rounds;           /* shared-memory round trips inside the tile passes */
    int64_t exchanges;        /* global<->local qubit swaps (each = one personalised all-to-all) */
    double exchange_bytes;    /* bytes this rank sent over NVLink in those swaps */
    int64_t jit_passes;       /* tile passes that ran on a run-time specialised kernel */
    int64_t jit_compiled;     /* process-wide: specialised kernels compiled so far */
    int64_t jit_pending;      /* process-wide: compilations still running in the background */
    double jit_compile_ms;    /* process-wide: NVRTC time spent so far (sum over pool threads) */
    int64_t p2p_exchanges;    /* qubit swaps that ran as in-place kernels over mapped peer memory */
    int64_t relabelled_swaps; /* uncontrolled SWAP gates executed as wire relabelling (no data movement) */
    double swap_device_ms;    /* CUDA-event time of the global<->local qubit swaps (timing on), part of gate_device_ms */
    double fp64_ops;          /* thread-level FP64 instructions the specialised passes issued (upper bound:
                                 run-time-conditional op bodies are counted as taken) */
    int64_t jit_disk_hits;    /* process-wide: specialised kernels loaded from the on-disk cubin cache */
    int64_t swap_push_exchanges; /* qubit swaps that ran as ONE push kernel writing into every peer's shard */
    double d2h_bytes;         /* result bytes copied device -> host (State, Probs, Sample indices, scalars excluded) */
    double h2d_bytes;         /* bytes copied host -> device (kernel parameter blocks, matrices, sampler uniforms, SetState) */
} b200sv_stats;
int b200sv_get_stats(b200sv_device *d, b200sv_stats *out);
int b200sv_reset_stats(b200sv_device *d);
int b200sv_set_timing(b200sv_device *d, int on); /* record CUDA events around each flush */
int b200sv_flush(b200sv_device *d);              /* launch everything queued (no host sync) */
/* Thread-level FP64 FMA instructions per second this GPU sustains (microbenchmark, ~20 ms): the second
 * roofline of a fused tile pass next to the HBM bandwidth. */
int b200sv_fp64_peak(b200sv_device *d, double *dfma_per_second);
/* Counters as one JSON object (the role of the reference's ResourceTracker output,
 * runtime/lib/backend/null_qubit/NullQubit.hpp:104-112): gates, passes, swaps, launches, JIT.
 * Returns the length written (excluding the terminating 0), or -1 if `cap` is too small. */
int64_t b200sv_stats_json(b200sv_device *d, char *buf, int64_t cap);
int b200sv_synchronize(b200sv_device *d);        /* flush + wait */
/* Tile passes whose structure recurs are compiled at run time into specialised kernels (kwarg
 * 'jit': 'off' | 'auto' (default) | 'always'; catalyst_b200/csrc/jit.hpp).  In 'auto' mode the
 * compilation runs in the background; this call blocks until none is pending (benchmarks). */
int b200sv_jit_wait(b200sv_device *d);
/* device pointer of the local shard (for benchmarks / zero-copy consumers) */
void *b200sv_device_pointer(b200sv_device *d);

/* ---- sharding across GPUs ------------------------------------------------------
 * The register is sharded by its high-order index bits: rank r of `world_size` (a power of two)
 * owns the amplitudes whose top log2(world_size) bits equal r.  Every rank runs the SAME call
 * sequence (SPMD: one process per GPU under torchrun/mpirun, or one thread per rank); scalar and
 * array results are identical on every rank.  Configured through the factory kwargs:
 *   'world_size': G, 'rank': r, 'comm': 'nccl:<256 hex digits>' | 'local:<group name>'
 * where the hex digits are an ncclUniqueId created once by b200sv_nccl_unique_id() and distributed
 * by the host program (any out-of-band channel: torch.distributed, MPI, a file).  'local:' joins the
 * ranks of one process by name (threads; several ranks may share a GPU — used by the tests).
 * Results on a sharded register:
 *   'gather': 'all' (default) — State / Probs of all wires return the whole vector on every rank (gathered through
 *             the host; small registers); 'slice' — each rank fills ONLY entries [rank << n_local, (rank+1) << n_local)
 *             of the caller's array, nothing is gathered (the layout is brought to wire order first).
 *   'dist_sampler': Sample / Counts of ALL wires draw from the ranks' slices of |psi|^2 through peer memory instead of
 *             gathering 2^n probabilities on every rank; samples are bit-identical to the unsharded device.
 *             -1 (default) when peer memory is available, 0 never, 1 required (error otherwise). */
int b200sv_nccl_unique_id(char hex_out[257]);
int b200sv_world_size(b200sv_device *d);
int b200sv_rank(b200sv_device *d);
/* layout[q] = physical index position of logical bit q (bit n-1-w for wire w); positions >=
 * n - log2(world_size) are rank bits.  Returns the number of entries written (<= cap). */
int64_t b200sv_layout(b200sv_device *d, int32_t *layout, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
