/* b200sv_kernels.h — thin C ABI between the C++ host layer and the hand-written sm_100a kernels.
 *
 * Plain pointers and sizes only: no C++ types, no torch types.  Every entry point enqueues work on
 * the given CUDA stream and returns a cudaError_t-compatible int (0 = success).  Device pointers
 * are complex128 amplitudes stored as interleaved (re, im) doubles (16 bytes per amplitude).
 *
 * Index convention (reference: runtime/include/QuantumDevice.hpp:409-412 and
 * frontend/test/pytest/test_measurements_results.py:166-184 — wire 0 is the most significant bit):
 * this layer only sees *bit positions* of the amplitude index; the host maps wires to positions.
 * On a sharded state the full index is (rank << n_local) | local_index; `rank_base` carries the
 * rank bits so that control predicates and diagonal phases on global qubits need no data movement.
 *
 * What each kernel replaces in the reference stack: the per-gate / per-measurement sweeps that
 * lightning.qubit performs behind QuantumDevice::NamedOperation / MatrixOperation / Expval / Var /
 * Probs / Sample / Measure / Gradient (runtime/include/QuantumDevice.hpp:215-567).
 */
#ifndef B200SV_KERNELS_H
#define B200SV_KERNELS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- fused tile pass --------------------------------------------------------------------------
 * One launch = one HBM round trip over the whole local shard.  Each CTA stages a tile of 2^t
 * amplitudes in shared memory (the low `c` index bits — contiguous 16*2^c-byte rows — plus
 * `nhigh` arbitrary higher bits), runs the whole op list on it, and writes it back. */
enum {
    B200K_OP_PAIR = 0,     /* 2x2 matrix between basis patterns a and b of the target bits      */
    B200K_OP_PHASE = 1,    /* amplitude *= m[0..1] where the control predicate holds            */
    B200K_OP_PARITY = 2,   /* amplitude *= (parity(idx & zmask) ? m[2..3] : m[0..1])            */
    B200K_OP_PAULIROT = 3, /* exp(-i theta/2 P): m[0]=cos, m[1]=sin, x-part = tmask, z-part = a */
    B200K_OP_DENSE = 4     /* 2^k x 2^k row-major matrix in global memory on k tile bits        */
};

#define B200K_MAX_OPS 32
#define B200K_MAX_TILE_BITS 12
#define B200K_MAX_HIGH 10
#define B200K_TILE_THREADS 256

typedef struct B200kTileOp {
    uint32_t kind;
    uint32_t k;       /* DENSE: number of target bits */
    uint32_t tmask;   /* tile-local mask of the non-diagonal target bits */
    uint32_t a, b;    /* PAIR: tile-local patterns; PARITY/PAULIROT: a = tile-local z mask */
    uint32_t cmask, cval; /* control predicate on tile-local bits */
    uint32_t ny;      /* PAULIROT: number of Y factors mod 4 */
    uint64_t cmask_o, cval_o; /* control predicate on the bits outside the tile (incl. rank bits) */
    uint64_t zmask_o;         /* PARITY/PAULIROT: z mask on the bits outside the tile */
    union {
        double m[8];
        struct {
            const double *mat; /* DENSE: device pointer, 2^k * 2^k complex, row-major */
            uint8_t tb[16];    /* DENSE: tile-local target bits, most significant first */
        } d;
    } u;
} B200kTileOp;

typedef struct B200kTileProgram {
    int32_t n_ops;
    int32_t t;     /* tile bits (<= B200K_MAX_TILE_BITS, <= n_local) */
    int32_t c;     /* contiguous low bits inside the tile */
    int32_t nhigh; /* = t - c */
    uint8_t high_pos[16]; /* index positions (>= c, ascending) of tile bits c..t-1 */
    uint64_t rank_base;   /* rank << n_local */
    B200kTileOp ops[B200K_MAX_OPS];
} B200kTileProgram;

int b200k_tile_program(double *psi, int n_local, const B200kTileProgram *prog, void *stream);

/* ---- shared by the register-blocked tile kernels --------------------------------------------------
 * The CTAs are persistent (grid = a multiple of the SM count, striding over the tiles) and the op list is
 * grouped by the host into ROUNDS.  In a register round every thread owns the 8 amplitudes spanned by
 * three tile bits and applies all of the round's ops to them in registers: one shared-memory read+write
 * per round instead of per op.  Shared memory is XOR-swizzled so that any choice of the three bits is
 * bank-conflict free.  Ops a register round cannot express run as single-op shared-memory rounds. */
enum { B200K2_ROUND_REG = 0, B200K2_ROUND_SMEM = 1 };

typedef struct B200k2Header {
    int32_t n_rounds, n_ops;
    int32_t t, c, nhigh, pad;
    uint8_t high_pos[16];
    uint64_t rank_base;
    uint64_t ntiles;
} B200k2Header;

/* ---- fused tile pass, third generation (tile3.cu) ------------------------------------------------
 * What it does follows the ncu reading of its predecessor (profiles/README.md r1_v2: issue-bound on
 * per-op decode, not on HBM or FP64):
 *   - tile load/store offsets are precomputed once per thread (persistent CTAs), not per amplitude;
 *   - a round's eight shared-memory slots are SW(thread base) ^ swz[a] (the XOR swizzle is linear),
 *   - every op carries host-precomputed SLOT MASKS over the thread's 8 registers instead of
 *     per-amplitude predicates, and a matrix class (general / real / real-diagonal+imaginary-off-
 *     diagonal / permutation) that halves or removes the FP64 work of H, RX, RY, X, CNOT;
 *   - two-pattern PAIRs whose two target bits are both register bits (SWAP, ISWAP, PSWAP, IsingXY,
 *     SingleExcitation, CSWAP) run in registers too. */
enum { B200K3_PAIR = 0, B200K3_PHASE = 1, B200K3_PARITY = 2, B200K3_PAIRG = 3, B200K3_PAULIROT = 4 };
enum { B200K3_M_GENERAL = 0, B200K3_M_REAL = 1, B200K3_M_RX = 2 };
enum { B200K3_F_CO = 1, B200K3_F_TC = 2, B200K3_F_TZ = 4 };
enum { B200K3_CODE_PHASE = 54, B200K3_CODE_PARITY = 65, B200K3_NCODES = 81 };

typedef struct B200k3Op { /* 128 bytes; the first 16 are the CONTROL WORD the kernel prefetches */
    uint32_t ctl;             /* register rounds: code | amask << 8 | flags << 16 | rpar << 24, where
                                 code is the dense dispatch code (one straight-line body per code,
                                 reached through a jump table):
                                 PAIR   : ((mclass * 9 + j) * 2 + full), full = all pairs active
                                 PHASE  : B200K3_CODE_PHASE + index of amask in the special-mask list
                                          {ff,aa,55,cc,33,f0,0f,88,a0,c0}, 10 = any other mask
                                 PARITY : B200K3_CODE_PARITY + 2 * index of rpar in
                                          {00,aa,cc,f0,66,5a,3c,96} + (amask == ff)                */
    uint32_t tc_mask, tc_val; /* control predicate over the non-register tile bits (shared-memory
                                 rounds: over all tile bits)                                    */
    uint32_t tz_mask;         /* z mask over the non-register tile bits (shared-memory rounds: all) */
    uint64_t co_mask, co_val; /* control predicate on the bits outside the tile (incl. rank bits) */
    uint64_t zo_mask;         /* z mask outside the tile                                        */
    uint8_t kind;
    uint8_t j;       /* PAIR: 0..2 = one-target on register bit j (pairs a <-> a|bit j);
                        3..5 = two-target 01<->10 on register bits (0,1), (0,2), (1,2) (first
                        pattern = lower bit set); 6..8 = two-target 00<->11 on the same bit pairs */
    uint8_t mclass;  /* PAIR: B200K3_M_*                                                       */
    uint8_t amask;   /* PAIR: mask over the op's 4 (or 2) register pairs; PHASE/PARITY: mask over
                        the 8 register slots (control predicate restricted to the register bits) */
    uint8_t rpar;    /* PARITY: bit a = parity(a & z mask over the register bits)               */
    uint8_t flags;   /* B200K3_F_CO: outside predicate present; _TC: thread predicate present;
                        _TZ: parity has a thread-bit or outside z part                          */
    uint8_t ny, npos; /* shared-memory rounds (PAULIROT / PAIRG)                                 */
    uint32_t tmask;           /* shared-memory rounds: tile-local target mask                   */
    uint32_t a, b;            /* shared-memory rounds: patterns                                 */
    uint8_t pos[4];           /* shared-memory rounds: ascending positions of tmask bits        */
    double m[8];
} B200k3Op;

typedef struct B200k3Round { /* 24 bytes */
    uint8_t kind, rb0, rb1, rb2;
    uint16_t first, count;
    uint16_t swz[8]; /* SW(pattern a spread over rb0..rb2), a = 0..7 */
} B200k3Round;

#define B200K3_SMALL_OPS 24
#define B200K3_SMALL_ROUNDS 16
#define B200K3_LARGE_OPS 215 /* + 1 readable pad entry for the control-word prefetch */
#define B200K3_LARGE_ROUNDS 128

int b200k_tile_program3(double *psi, int n_local, const struct B200k2Header *hdr, const B200k3Round *rounds,
                        const B200k3Op *ops, void *stream);

/* ---- reductions (warp-shuffle tree -> per-block partials -> fixed-order final sum) -------------
 * `scratch` must hold b200k_reduce_scratch_doubles() doubles; results land in `out` (device). */
int64_t b200k_reduce_scratch_doubles(void);

/* out[0..1] = sum_i conj(phi_i) * (P lam)_i over i with ((rank_base|i) & cmask) == cval, where
 * (P v)_i = i^ny * (-1)^popc(j & zmask) * v_j, j = i ^ xmask.  xmask must be local bits. */
int b200k_bilinear_pauli(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                         uint64_t xmask, uint64_t zmask, int ny, uint64_t cmask, uint64_t cval,
                         double *scratch, double *out, void *stream);

/* Allocation-free variant of b200k_bilinear_pauli.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, same result in out[0..1] (device), but the reduction scratch is a file-scope
 * `static __device__` array of B200K_NOALLOC_SCRATCH_DOUBLES doubles (18 944 bytes, held by every GPU that loads
 * the library until exit) instead of a caller buffer; no cudaMalloc / cudaFree / allocating routine is called.
 * Limit: 0 <= n_local <= B200K_NOALLOC_MAX_LOCAL_QUBITS; anything else (or a null pointer, or xmask outside the
 * local index) returns cudaErrorInvalidValue and, when errbuf != NULL, a message in errbuf[0..errcap).
 * All calls share the array: calls that OVERLAP on the GPU race on it -- order them on one stream.  The array
 * keeps the previous call's contents; nothing in it needs to start at zero (every slot read was written by the
 * same call).  The first call on a device resolves the array's address: make it outside graph capture. */
#define B200K_NOALLOC_SCRATCH_DOUBLES 2368 /* 1184 blocks x 2 partials */
#define B200K_NOALLOC_MAX_LOCAL_QUBITS 36
#define B200K_STR_(x) #x
#define B200K_STR(x) B200K_STR_(x)
int b200k_bilinear_pauli_noalloc(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                                 uint64_t xmask, uint64_t zmask, int ny, uint64_t cmask, uint64_t cval,
                                 double *out, void *stream, char *errbuf, int64_t errcap);

/* out[0..1] = sum over base indices i (tmask bits zero, control predicate true) of
 * [phi_{i|a}, phi_{i|b}]^dagger * g(2x2, row-major complex) * [lam_{i|a}, lam_{i|b}]^T. */
int b200k_bilinear_pair(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                        uint64_t tmask, uint64_t a, uint64_t b, const double *g8, uint64_t cmask,
                        uint64_t cval, double *scratch, double *out, void *stream);

/* Batched version of b200k_bilinear_pauli: up to B200K_BATCH_GENS Pauli-word generators evaluated in
 * ONE read pass over phi and lam (tile geometry of the gate passes: header fields t, c, nhigh, high_pos,
 * rank_base, ntiles).  Masks are split like the gate ops: inside the tile (tile-local bit positions) and
 * outside (physical positions, incl. rank bits).  outs[g] receives (re, im) of generator g, of which only the
 * component named by `want` is computed. */
#define B200K_BATCH_GENS 32
typedef struct B200kBatchGen {
    uint32_t xin;            /* X/Y bits (tile-local); all of them must lie inside the tile        */
    uint32_t zin;            /* Z/Y bits inside the tile                                            */
    uint32_t cin_m, cin_v;   /* control predicate inside the tile                                   */
    uint64_t zo;             /* Z bits outside the tile                                             */
    uint64_t co_m, co_v;     /* control predicate outside the tile                                  */
    int32_t ny;              /* number of Y factors (phase i^ny)                                    */
    int32_t want;            /* B200K_WANT_REAL / _IMAG: the part of the result the caller uses; the other
                                is written as 0.  All generators of one call must need the same RAW component
                                ((ny & 1) ^ (want == imag) equal across the batch)                       */
} B200kBatchGen;
enum { B200K_WANT_REAL = 1, B200K_WANT_IMAG = 2 };
int64_t b200k_bilinear_batch_scratch_doubles(void);
int b200k_bilinear_batch(const double *phi, const double *lam, int n_local, const struct B200k2Header *hdr,
                         const B200kBatchGen *gens, int ngen, double *const *outs, double *scratch, void *stream);

/* Sum of up to B200K_TILE_WORDS Pauli words, tile-staged: ONE read of `in` for the whole set instead of one gather
 * per word.  Serves QuantumDevice::Expval of a HamiltonianObservable (runtime/include/QuantumDevice.hpp:353,488;
 * "copy psi, apply O (per-term copy + gates + axpy), inner product: >= 3S traffic per term")
 * and the phi = H psi initialisation of QuantumDevice::Gradient (:564).
 * Tile geometry as in b200k_bilinear_batch (hdr: t <= 11, c, nhigh, high_pos, rank_base, ntiles); every word's
 * X/Y bits lie inside the tile.
 *   out != NULL:  out_i = (accumulate ? out_i : 0) + sum_w coef_w (P_w in)_i            (out must not alias in)
 *   out == NULL:  dot_out[0..1] = sum_i conj(bra_i) * sum_w coef_w (P_w in)_i  (re, im); `scratch` holds
 *                 b200k_reduce_scratch_doubles() doubles; bra may equal in.
 * (P v)_i = i^ny (-1)^popc(j & z) v_j, j = i ^ x, as in b200k_bilinear_pauli. */
#define B200K_TILE_WORDS 96
typedef struct B200kTileWord {
    uint32_t xin;  /* X/Y bits (tile-local positions)             */
    uint32_t zin;  /* Z/Y bits inside the tile (tile-local)       */
    uint64_t zo;   /* Z/Y bits outside the tile (physical, incl. rank bits) */
    double cr, ci; /* coefficient                                 */
    int32_t ny;    /* number of Y factors                         */
    int32_t pad;
} B200kTileWord;
int b200k_pauli_words_tile(double *out, const double *in, const double *bra, int n_local,
                           const struct B200k2Header *hdr, const B200kTileWord *words, int nwords, int accumulate,
                           double *scratch, double *dot_out, void *stream);

/* The same bilinear forms for up to B200K_BATCH_GENS DIAGONAL words (no X/Y factor) in one streaming pass:
 * outs[g] = (re, im) of sum_i [((rank_base|i) & cmasks[g]) == cvals[g]] i^ny (-1)^popc((rank_base|i) & zmasks[g])
 * conj(phi_i) lam_i, of which only the component named by `want` is computed (the other is written as 0).  Masks are
 * physical (rank bits included).  scratch: b200k_bilinear_batch_scratch_doubles() doubles. */
int b200k_bilinear_diag_batch(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                              const uint64_t *zmasks, const uint64_t *cmasks, const uint64_t *cvals, const int *nys,
                              int want, int ngen, double *const *outs, double *scratch, void *stream);

/* out[0] = sum_i |psi_i|^2 * sum_t coef[t] * (-1)^popc((rank_base|i) & zmask[t]). */
int b200k_expval_diag_batch(const double *psi, int n_local, uint64_t rank_base,
                            const uint64_t *zmasks, const double *coefs, int nterms,
                            double *scratch, double *out, void *stream);

/* ---- probabilities ---------------------------------------------------------------------------- */
int b200k_probs_all(const double *psi, double *out, int n_local, void *stream);
/* out[dst(i)] = |psi_i|^2 with source index bit b landing on destination bit perm[b] (a permutation of
 * 0..n-1): undoes the host's wire relabelling (SWAP gates executed as renames) on the way out. */
int b200k_probs_all_perm(const double *psi, double *out, int n, const uint8_t *perm, void *stream);
/* out[o] = sum of |psi_i|^2 over i whose bits at tb[0..k) (most significant first) spell o.
 * scratch: 2^k * b200k_probs_splits(n_local, k) doubles. */
int b200k_probs_splits(int n_local, int k);
int b200k_probs_partial(const double *psi, int n_local, const uint8_t *tb, int k, double *scratch,
                        double *out, void *stream);

/* ---- element-wise passes ---------------------------------------------------------------------- */
/* psi_i = (bit `pos` of (rank_base|i) == keep) ? psi_i * scale : 0 */
int b200k_collapse(double *psi, int n_local, uint64_t rank_base, int pos, int keep, double scale,
                   void *stream);
/* out_i = (accumulate ? out_i : 0) + coef * (P in)_i  (Pauli word as in b200k_bilinear_pauli) */
int b200k_pauli_apply(double *out, const double *in, int n_local, uint64_t rank_base,
                      uint64_t xmask, uint64_t zmask, int ny, double coef_re, double coef_im,
                      int accumulate, void *stream);
/* out_i = (accumulate ? out_i : 0) + sum_t coef[t] * (P_t in)_i for nwords <= B200K_APPLY_WORDS words in one
 * pass (coefs_ri = (re, im) per word) */
#define B200K_APPLY_WORDS 16
int b200k_pauli_apply_batch(double *out, const double *in, int n_local, uint64_t rank_base,
                            const uint64_t *xmasks, const uint64_t *zmasks, const int *nys,
                            const double *coefs_ri, int nwords, int accumulate, void *stream);
/* acc += coef * x */
int b200k_axpy(double *acc, const double *x, double coef, int n_local, void *stream);
/* psi = 0 everywhere, then psi[index] = 1 if index >= 0 */
int b200k_set_basis(double *psi, int n_local, int64_t local_index, void *stream);
/* psi = 0, then psi[index whose bits at tb[0..k) (most significant first) spell r] = vals[r] */
int b200k_scatter_state(double *psi, const double *vals, int n_local, const uint8_t *tb, int k,
                        void *stream);
/* new[(i << 1)] = old[i], new[(i << 1) | 1] = 0  — append one qubit in |0> as the new bit 0 */
int b200k_append_qubit(double *dst, const double *src, int n_local_old, void *stream);

/* ---- global<->local qubit swap (shard.cu) --------------------------------------------------------
 * staged[j] = state[expand(j0 + j)], j in [0, count): expand() inserts, at the m ascending local bit
 * positions lpos[], the bits of `val` (bit i of val -> position lpos[i]).  unpack is the inverse
 * scatter into the same slots.  m <= 8. */
int b200k_pack_bits(double *staged, const double *state, const uint8_t *lpos, int m, uint32_t val,
                    uint64_t j0, uint64_t count, void *stream);
int b200k_unpack_bits(double *state, const double *staged, const uint8_t *lpos, int m, uint32_t val,
                      uint64_t j0, uint64_t count, void *stream);
/* In-place pairwise swap over peer memory (NVLink/NVSwitch; `peer` = the other rank's shard mapped into
 * this GPU's address space): for j in [j_begin, j_end), mine[expand(j) | bits(peer_pat)] <->
 * peer[expand(j) | bits(my_pat)].  Each rank of a pair calls it for its half of the sub-block. */
int b200k_swap_p2p(double *mine, double *peer, const uint8_t *lpos, int m, uint32_t my_pat, uint32_t peer_pat,
                   uint64_t j_begin, uint64_t j_end, void *stream);

/* ---- alias-method sampler on device (reference algorithm: Walker alias table with LIFO pairing,
 * SURVEY.md Appendix A.4; restated as prefix sums so that no sequential pass over 2^n is needed) */
/* MCMC sampling (kernels.cu): one thread walks a Metropolis chain of `steps` proposals over the basis states of a
 * state in canonical layout and writes the visited index after each of the last steps - burnin proposals.
 * Proposal s: nz == NULL ("Local"): qubit site[s] & 0x7fffffff is set to bit 31 of site[s]; else ("NonZeroRandom"):
 * jump to nz[jump[s] % nnz].  Accepted when |psi_new|^2 > u[s] * |psi_old|^2 (or the chain sits on a zero). */
int b200k_mcmc_chain(const double *psi, int n, int64_t steps, int64_t burnin, const uint32_t *site, const uint64_t *jump,
                     const int64_t *nz, int64_t nnz, const double *u, int64_t *out, void *stream);
/* ascending indices i with |psi_i|^2 > eps into out (capacity 2^n), their number into *count; synchronises */
int b200k_nonzero_indices(const double *psi, int n, double eps, int64_t *out, int64_t *count, void *stream);

/* Push-mode qubit swap (shard.cu): one launch writes chunk [push_j0, push_j0 + push_cnt) of this rank's
 * sub-block (of the vector push_state) for every peer p (lpos bits spell peer_pat[p]) into peer_dst[p] (the peer's receive staging, mapped
 * peer memory) and, in the same launch, scatters the previous chunk received from every peer (my_src[p]) into
 * the slots (of the vector unpack_state) that chunk vacated.  Either count may be 0. */
#define B200K_PUSH_MAX_PEERS 15
int b200k_swap_push(double *push_state, double *unpack_state, double *const *peer_dst, const double *const *my_src, const uint32_t *peer_pat,
                    int npeer, const uint8_t *lpos, int m, uint64_t push_j0, uint64_t push_cnt, uint64_t unpack_j0,
                    uint64_t unpack_cnt, void *stream);

/* FP64 pipe microbenchmark: thread-level DFMA instructions per second this GPU sustains (16 independent
 * chains per thread, full occupancy); `scratch` = any device buffer of >= 8 bytes.  Synchronises. */
int b200k_dfma_peak(double *scratch, double *dfma_per_second, void *stream);

int64_t b200k_sampler_scratch_doubles(int64_t nbins);
int b200k_sampler_prepare(const double *probs, int64_t nbins, double *scratch, void *stream);
/* u: 2*nshots uniforms (u1,u2 per shot) on device; out_idx: nshots int64 on device */
int b200k_sampler_draw(const double *probs, int64_t nbins, const double *scratch, const double *u,
                       int64_t nshots, int64_t *out_idx, void *stream);

/* The same sampler over a distribution sharded by its top index bits (Sample / Counts of all wires of a
 * sharded register; QuantumDevice.hpp:384,433).  Every rank: (1) b200k_sampler_chunk_sums over its own slice
 * (nbins_slice = 2^n_local >= 1024 entries, scaled by the bins of the whole distribution) into nbins_slice/1024
 * doubles each of sumD / sumE; (2) the host gathers all ranks' sums into EVERY rank's scratch
 * (b200k_sampler_scratch_doubles(nbins_total) doubles) at the offsets b200k_sampler_sum_offsets returns, in global
 * descending order: rank r's block at chunk index (W-1-r) * nbins_slice/1024; (3) b200k_sampler_finish;
 * (4) b200k_sampler_draw_sharded with slices[r] = a pointer THIS GPU can read rank r's slice through (peer memory).
 * Every rank draws every shot: same arithmetic on the same numbers as on one GPU, bit-identical samples. */
#define B200K_SAMPLER_MAX_SHARDS 64
int b200k_sampler_chunk_sums(const double *probs_slice, int64_t nbins_slice, int64_t nbins_total, double *sumD,
                             double *sumE, void *stream);
void b200k_sampler_sum_offsets(int64_t nbins_total, int64_t *off_sumD, int64_t *off_sumE);
int b200k_sampler_finish(double *scratch, int64_t nbins_total, void *stream);
int b200k_sampler_draw_sharded(const double *const *slices, int nslices, int n_local, const double *scratch,
                               const double *u, int64_t nshots, int64_t *out_idx, void *stream);

#ifdef __cplusplus
}
#endif
#endif
