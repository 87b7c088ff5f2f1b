// tile3.cu — third-generation fused tile pass (see include/b200sv_kernels.h, "third generation").
//
// Evidence that shaped it (profiles/README.md):
//  * ncu --set full of k_tile2 at 30 qubits: DRAM traffic at the algorithmic minimum (34.4 GB per
//    launch) but 33-64 ms per launch; issue slots 57 % busy, FP64 pipe 11 %, ~630 thread-instructions
//    per amplitude per pass of which ~60 FP64.  The cost was the per-op interpreter: field loads,
//    per-amplitude predicate loops whose predicated-off complex multiplies still issue, 64-bit index
//    arithmetic per amplitude in the tile load/store.  Here every decision that is uniform over the
//    CTA is taken once per op on host-precomputed slot masks, everything that does not depend on the
//    tile is computed once per thread.
//  * SASS-level profile of the first k_tile3: explicit register exchanges for permutation gates made
//    the compiler keep two copies of the thread's 8 amplitudes (32 moves per op, 29 % of all issued
//    instructions) -> permutations run as real 2x2 matrices; the compare chain that selected an op
//    body cost ~45 instructions per op -> one dense dispatch code per op (jump table); 30 % of the
//    stall samples sat on the tile's global loads -> the next tile is prefetched with cp.async into
//    a second shared-memory buffer while the current one is processed.
#include <atomic>
#include <mutex>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200sv_kernels.h"

typedef double2 cplx;
#define TT B200K_TILE_THREADS
#define SW(i) ((i) ^ (((i) >> 3) & 7u))

template <int NOPS, int NROUNDS> struct Prog3 {
    B200k2Header h;
    B200k3Round rounds[NROUNDS];
    B200k3Op ops[NOPS + 1]; // + one readable entry for the control-word prefetch
};

__device__ __forceinline__ cplx cmul3(cplx a, cplx b)
{
    // multiplies on a.x first, FMAs on a.y last: both results can be written over their own inputs
    return make_double2(fma(-a.y, b.y, a.x * b.x), fma(a.y, b.x, a.x * b.y));
}
__device__ __forceinline__ uint32_t ins0(uint32_t g, uint32_t b)
{
    return ((g >> b) << (b + 1)) | (g & ((1u << b) - 1u));
}
__device__ __forceinline__ cplx ipow3(cplx a, int k)
{
    switch (k & 3) {
    case 0: return a;
    case 1: return make_double2(-a.y, a.x);
    case 2: return make_double2(-a.x, -a.y);
    default: return make_double2(a.y, -a.x);
    }
}
__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gmem_src)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait()
{
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// register pairs of a PAIR op, by op.j (see B200k3Op): lo slot, hi slot (2 pairs for j >= 3)
__device__ constexpr int PLO[9][4] = {{0, 2, 4, 6}, {0, 1, 4, 5}, {0, 1, 2, 3}, {1, 5, 0, 0}, {1, 3, 0, 0},
                                      {2, 3, 0, 0}, {0, 4, 0, 0}, {0, 2, 0, 0}, {0, 1, 0, 0}};
__device__ constexpr int PHI[9][4] = {{1, 3, 5, 7}, {2, 3, 6, 7}, {4, 5, 6, 7}, {2, 6, 0, 0}, {4, 6, 0, 0},
                                      {4, 5, 0, 0}, {3, 7, 0, 0}, {5, 7, 0, 0}, {6, 7, 0, 0}};

// one 2x2 block on the register pair (x, y); MC selects how much FP64 work the matrix needs
template <int MC> __device__ __forceinline__ void pair2x2(cplx &x, cplx &y, const double (&m)[8])
{
    const cplx a = x, b = y;
    if (MC == B200K3_M_REAL) { // all entries real: H, RY, X, CNOT, SWAP
        x = make_double2(fma(m[2], b.x, m[0] * a.x), fma(m[2], b.y, m[0] * a.y));
        y = make_double2(fma(m[6], b.x, m[4] * a.x), fma(m[6], b.y, m[4] * a.y));
    }
    else if (MC == B200K3_M_RX) { // real diagonal, imaginary off-diagonal: RX, IsingXX blocks
        x = make_double2(fma(-m[3], b.y, m[0] * a.x), fma(m[3], b.x, m[0] * a.y));
        y = make_double2(fma(m[6], b.x, -m[5] * a.y), fma(m[6], b.y, m[5] * a.x));
    }
    else {
        x = make_double2(fma(m[0], a.x, fma(-m[1], a.y, fma(m[2], b.x, -m[3] * b.y))),
                         fma(m[0], a.y, fma(m[1], a.x, fma(m[2], b.y, m[3] * b.x))));
        y = make_double2(fma(m[4], a.x, fma(-m[5], a.y, fma(m[6], b.x, -m[7] * b.y))),
                         fma(m[4], a.y, fma(m[5], a.x, fma(m[6], b.y, m[7] * b.x))));
    }
}

template <int MC, int J, bool FULL>
__device__ __forceinline__ void pair_case(cplx (&v)[8], const B200k3Op &op, uint32_t ctl)
{
    double m[8];
    if (MC == B200K3_M_REAL) {
        m[0] = op.m[0]; m[2] = op.m[2]; m[4] = op.m[4]; m[6] = op.m[6];
        m[1] = m[3] = m[5] = m[7] = 0.0;
    }
    else if (MC == B200K3_M_RX) {
        m[0] = op.m[0]; m[3] = op.m[3]; m[5] = op.m[5]; m[6] = op.m[6];
        m[1] = m[2] = m[4] = m[7] = 0.0;
    }
    else {
#pragma unroll
        for (int q = 0; q < 8; q++) m[q] = op.m[q];
    }
    constexpr int NP = J < 3 ? 4 : 2;
    const uint32_t pm = FULL ? 0xfu : ((ctl >> 8) & 0xffu);
#pragma unroll
    for (int p = 0; p < NP; p++)
        if (FULL || (pm & (1u << p))) pair2x2<MC>(v[PLO[J][p]], v[PHI[J][p]], m);
}

template <int MASK> __device__ __forceinline__ void phase_slots(cplx (&v)[8], const B200k3Op &op)
{
    const cplx e = make_double2(op.m[0], op.m[1]);
#pragma unroll
    for (int a = 0; a < 8; a++)
        if (MASK & (1 << a)) v[a] = cmul3(v[a], e);
}

// v[a] *= ((RPAR >> a) & 1) ? eB : eA on the slots of amask; eA/eB = (e0, e1) ordered by the parity of
// the thread's non-register index bits and of the tile's outside bits
template <int RPAR, bool FULL>
__device__ __forceinline__ void parity_case(cplx (&v)[8], const B200k3Op &op, uint4 ctl, uint32_t ib,
                                            uint64_t obase)
{
    uint32_t p0 = 0;
    if (ctl.x & (B200K3_F_TZ << 16)) p0 = (__popc(ib & ctl.w) + __popcll(obase & op.zo_mask)) & 1u;
    const cplx e0 = make_double2(op.m[0], op.m[1]);
    const cplx e1 = make_double2(op.m[2], op.m[3]);
    const cplx eA = p0 ? e1 : e0, eB = p0 ? e0 : e1;
    const uint32_t am = FULL ? 0xffu : ((ctl.x >> 8) & 0xffu);
#pragma unroll
    for (int a = 0; a < 8; a++)
        if (FULL || (am & (1u << a))) v[a] = cmul3(v[a], ((RPAR >> a) & 1) ? eB : eA);
}

// every register-round op body, selected by the host-computed dense code (jump table)
__device__ __forceinline__ void run_op(cplx (&v)[8], const B200k3Op &op, uint4 ctl, uint32_t ib, uint64_t obase)
{
#define PAIR_J(MC, J)                                                                              \
    case ((MC * 9 + J) * 2 + 1): pair_case<MC, J, true>(v, op, ctl.x); break;                      \
    case ((MC * 9 + J) * 2 + 0): pair_case<MC, J, false>(v, op, ctl.x); break;
#define PAIR_MC(MC)                                                                                \
    PAIR_J(MC, 0) PAIR_J(MC, 1) PAIR_J(MC, 2) PAIR_J(MC, 3) PAIR_J(MC, 4) PAIR_J(MC, 5) PAIR_J(MC, 6)   \
    PAIR_J(MC, 7) PAIR_J(MC, 8)
#define PARITY_R(IDX, RPAR)                                                                        \
    case (B200K3_CODE_PARITY + 2 * IDX + 1): parity_case<RPAR, true>(v, op, ctl, ib, obase); break; \
    case (B200K3_CODE_PARITY + 2 * IDX + 0): parity_case<RPAR, false>(v, op, ctl, ib, obase); break;
    switch (ctl.x & 0xffu) {
        PAIR_MC(B200K3_M_GENERAL)
        PAIR_MC(B200K3_M_REAL)
        PAIR_MC(B200K3_M_RX)
    case B200K3_CODE_PHASE + 0: phase_slots<0xff>(v, op); break;
    case B200K3_CODE_PHASE + 1: phase_slots<0xaa>(v, op); break;
    case B200K3_CODE_PHASE + 2: phase_slots<0x55>(v, op); break;
    case B200K3_CODE_PHASE + 3: phase_slots<0xcc>(v, op); break;
    case B200K3_CODE_PHASE + 4: phase_slots<0x33>(v, op); break;
    case B200K3_CODE_PHASE + 5: phase_slots<0xf0>(v, op); break;
    case B200K3_CODE_PHASE + 6: phase_slots<0x0f>(v, op); break;
    case B200K3_CODE_PHASE + 7: phase_slots<0x88>(v, op); break;
    case B200K3_CODE_PHASE + 8: phase_slots<0xa0>(v, op); break;
    case B200K3_CODE_PHASE + 9: phase_slots<0xc0>(v, op); break;
    case B200K3_CODE_PHASE + 10: {
        const cplx e = make_double2(op.m[0], op.m[1]);
        const uint32_t am = (ctl.x >> 8) & 0xffu;
#pragma unroll
        for (int a = 0; a < 8; a++)
            if (am & (1u << a)) v[a] = cmul3(v[a], e);
        break;
    }
        PARITY_R(0, 0x00)
        PARITY_R(1, 0xaa)
        PARITY_R(2, 0xcc)
        PARITY_R(3, 0xf0)
        PARITY_R(4, 0x66)
        PARITY_R(5, 0x5a)
        PARITY_R(6, 0x3c)
        PARITY_R(7, 0x96)
    default: break;
    }
#undef PAIR_J
#undef PAIR_MC
#undef PARITY_R
}

template <int NOPS, int NROUNDS>
__global__ void __launch_bounds__(TT, 3)
k_tile3(cplx *__restrict__ psi, const __grid_constant__ Prog3<NOPS, NROUNDS> P)
{
    extern __shared__ cplx smem[]; // one tile, or two when the launch double-buffers (P.h.pad != 0)
    __shared__ uint64_t s_offk[32];
    const int t = P.h.t, c = P.h.c, nhigh = P.h.nhigh;
    const uint32_t nelem = 1u << t;
    const uint32_t lowmask = (1u << c) - 1u;
    const uint32_t tid = threadIdx.x;
    const bool dbuf = P.h.pad != 0;

    // tile-independent part of the global offset of element e: off(e) = (e & lowmask) | scatter(e >> c)
    // is an OR over disjoint bit sets of e, so off(tid + 256 k) = off(tid) | off(256 k)
    auto off_of = [&](uint32_t e) -> uint64_t {
        uint64_t o = e & lowmask;
        const uint32_t r = e >> c;
        for (int h = 0; h < nhigh; h++) o |= (uint64_t)((r >> h) & 1u) << P.h.high_pos[h];
        return o;
    };
    auto base_of = [&](uint64_t tile) -> uint64_t {
        uint64_t base = tile << c;
        for (int h = 0; h < nhigh; h++) {
            const int p = P.h.high_pos[h];
            base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
        }
        return base;
    };
    const uint32_t niter = nelem > TT ? nelem / TT : 1u; // <= 16 for t <= 12
    if (tid < niter) s_offk[tid] = off_of(tid * TT);
    const uint64_t off_tid = off_of(tid);
    const uint32_t sw_tid = SW(tid);
    __syncthreads();

    auto prefetch = [&](uint64_t tile, cplx *buf) {
        if (tid < nelem) {
            const cplx *src = psi + (base_of(tile) | off_tid);
#pragma unroll 8
            for (uint32_t k = 0; k < niter; k++) cp_async16(&buf[sw_tid + k * TT], &src[s_offk[k]]);
        }
        cp_async_commit();
    };

    uint32_t cur = 0;
    if (dbuf && blockIdx.x < P.h.ntiles) prefetch(blockIdx.x, smem);
    for (uint64_t tile = blockIdx.x; tile < P.h.ntiles; tile += gridDim.x) {
        const uint64_t base = base_of(tile);
        const uint64_t obase = base | P.h.rank_base;
        cplx *__restrict__ tp = psi + (base | off_tid);
        cplx *__restrict__ s = smem + (dbuf ? cur * nelem : 0u);

        // ---- load ---------------------------------------------------------------------------
        if (dbuf) {
            const uint64_t next = tile + gridDim.x;
            if (next < P.h.ntiles) {
                prefetch(next, smem + (cur ^ 1u) * nelem);
                cp_async_wait<1>(); // the current tile's group is complete, the next one may be in flight
            }
            else cp_async_wait<0>();
        }
        else if (tid < nelem) {
#pragma unroll 8
            for (uint32_t k = 0; k < niter; k++) s[sw_tid + k * TT] = tp[s_offk[k]];
        }
        __syncthreads();

        // ---- rounds -------------------------------------------------------------------------
        for (int ri = 0; ri < P.h.n_rounds; ri++) {
            const B200k3Round &R = P.rounds[ri];
            const int o0 = R.first, o1 = R.first + R.count;
            if (R.kind == B200K2_ROUND_REG) {
                for (uint32_t g = tid; g < (nelem >> 3); g += TT) {
                    const uint32_t ib = ins0(ins0(ins0(g, R.rb0), R.rb1), R.rb2);
                    const uint32_t sb = SW(ib);
                    cplx v[8];
#pragma unroll
                    for (int a = 0; a < 8; a++) v[a] = s[sb ^ R.swz[a]];
                    // the next op's 16-byte control word is fetched while the current op runs (the op
                    // array has one readable entry past the last op)
                    uint4 nxt = *reinterpret_cast<const uint4 *>(&P.ops[o0]);
                    for (int o = o0; o < o1; o++) {
                        const B200k3Op &op = P.ops[o];
                        const uint4 ctl = nxt;
                        nxt = *reinterpret_cast<const uint4 *>(&P.ops[o + 1]);
                        if ((ctl.x & (B200K3_F_CO << 16)) && (obase & op.co_mask) != op.co_val) continue;
                        if ((ctl.x & (B200K3_F_TC << 16)) && (ib & ctl.y) != ctl.z) continue;
                        run_op(v, op, ctl, ib, obase);
                    }
#pragma unroll
                    for (int a = 0; a < 8; a++) s[sb ^ R.swz[a]] = v[a];
                }
            }
            else { // shared-memory round: one op executed on the staged tile
                for (int o = o0; o < o1; o++) {
                    const B200k3Op &op = P.ops[o];
                    const bool active = (obase & op.co_mask) == op.co_val;
                    if (active && op.kind == B200K3_PAIRG) {
                        const cplx m00 = make_double2(op.m[0], op.m[1]), m01 = make_double2(op.m[2], op.m[3]);
                        const cplx m10 = make_double2(op.m[4], op.m[5]), m11 = make_double2(op.m[6], op.m[7]);
                        const uint32_t ng = nelem >> op.npos;
                        for (uint32_t g = tid; g < ng; g += TT) {
                            uint32_t i = g;
                            for (uint32_t q = 0; q < op.npos; q++) i = ins0(i, op.pos[q]);
                            if ((i & op.tc_mask) != op.tc_val) continue;
                            const uint32_t ia = SW(i | op.a), ibb = SW(i | op.b);
                            const cplx va = s[ia], vb = s[ibb];
                            const cplx t0 = cmul3(m00, va), t1 = cmul3(m01, vb);
                            const cplx t2 = cmul3(m10, va), t3 = cmul3(m11, vb);
                            s[ia] = make_double2(t0.x + t1.x, t0.y + t1.y);
                            s[ibb] = make_double2(t2.x + t3.x, t2.y + t3.y);
                        }
                    }
                    else if (active && op.kind == B200K3_PAULIROT) {
                        const double cs = op.m[0], sn = op.m[1];
                        const uint32_t opar = __popcll(obase & op.zo_mask);
                        for (uint32_t g = tid; g < (nelem >> 1); g += TT) {
                            const uint32_t i = ins0(g, op.pos[0]);
                            if ((i & op.tc_mask) != op.tc_val) continue;
                            const uint32_t j = i ^ op.tmask;
                            const cplx vi = s[SW(i)], vj = s[SW(j)];
                            const cplx pi = ipow3(vj, (int)op.ny + 2 * ((__popc(j & op.tz_mask) + opar) & 1));
                            const cplx pj = ipow3(vi, (int)op.ny + 2 * ((__popc(i & op.tz_mask) + opar) & 1));
                            s[SW(i)] = make_double2(cs * vi.x + sn * pi.y, cs * vi.y - sn * pi.x);
                            s[SW(j)] = make_double2(cs * vj.x + sn * pj.y, cs * vj.y - sn * pj.x);
                        }
                    }
                    else if (active && (op.kind == B200K3_PHASE || op.kind == B200K3_PARITY)) { // t < 3 only
                        const uint32_t opar = __popcll(obase & op.zo_mask);
                        for (uint32_t e = tid; e < nelem; e += TT) {
                            if ((e & op.tc_mask) != op.tc_val) continue;
                            cplx ph = make_double2(op.m[0], op.m[1]);
                            if (op.kind == B200K3_PARITY && ((__popc(e & op.tz_mask) + opar) & 1))
                                ph = make_double2(op.m[2], op.m[3]);
                            s[SW(e)] = cmul3(s[SW(e)], ph);
                        }
                    }
                    if (o + 1 < o1) __syncthreads();
                }
            }
            __syncthreads();
        }

        // ---- store --------------------------------------------------------------------------
        if (tid < nelem) {
#pragma unroll 8
            for (uint32_t k = 0; k < niter; k++) tp[s_offk[k]] = s[sw_tid + k * TT];
        }
        __syncthreads(); // the buffer is refilled by the next iteration's prefetch (or load)
        cur ^= 1u;
    }
}

template <int NOPS, int NROUNDS>
static int launch3(double *psi, const B200k2Header *hdr, const B200k3Round *rounds, const B200k3Op *ops,
                   cudaStream_t st)
{
    // Per CUDA device (ranks of one process may sit on different GPUs and launch concurrently): function
    // attributes, SM count and resident CTAs per SM for [double-buffered?][tile bits].  A slot is filled under
    // the mutex and published with a release store; later launches only read it after the acquire load.
    struct PerDevice {
        std::atomic<int> ready{0};
        int sm_count = 0;
        int grid_per_sm[2][B200K_MAX_TILE_BITS + 1];
    };
    static PerDevice per_device[64];
    static std::mutex init_mu;
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    PerDevice &D = per_device[cur_dev & 63];
    auto kern = k_tile3<NOPS, NROUNDS>;
    if (!D.ready.load(std::memory_order_acquire)) {
        std::lock_guard<std::mutex> lk(init_mu);
        if (!D.ready.load(std::memory_order_relaxed)) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)(2 * sizeof(cplx) << B200K_MAX_TILE_BITS));
            if (e != cudaSuccess) return (int)e;
            cudaDeviceGetAttribute(&D.sm_count, cudaDevAttrMultiProcessorCount, cur_dev);
            for (int d = 0; d < 2; d++)
                for (int t = 0; t <= B200K_MAX_TILE_BITS; t++) {
                    int nb = 0;
                    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, TT, (size_t)(d + 1) * (sizeof(cplx) << t));
                    D.grid_per_sm[d][t] = nb;
                }
            D.ready.store(1, std::memory_order_release);
        }
    }
    const int sm_count = D.sm_count;
    const auto &grid_per_sm = D.grid_per_sm;
    Prog3<NOPS, NROUNDS> P;
    P.h = *hdr;
    for (int i = 0; i < hdr->n_rounds; i++) P.rounds[i] = rounds[i];
    for (int i = 0; i < hdr->n_ops; i++) P.ops[i] = ops[i];
    // double-buffer when it costs no residency and every CTA handles several tiles
    const int t = hdr->t;
    int dbuf = grid_per_sm[1][t] >= grid_per_sm[0][t] && grid_per_sm[1][t] > 0 &&
               hdr->ntiles >= 2ull * (uint64_t)sm_count * (uint64_t)grid_per_sm[1][t] && t >= 8;
    if (hdr->pad == 2) dbuf = 0; // caller override (benchmarking the plain path)
    const int per_sm = grid_per_sm[dbuf][t] > 0 ? grid_per_sm[dbuf][t] : 1;
    P.h.pad = dbuf;
    uint64_t grid = (uint64_t)sm_count * per_sm;
    if (grid > hdr->ntiles) grid = hdr->ntiles;
    kern<<<(unsigned)grid, TT, (size_t)(dbuf + 1) * (sizeof(cplx) << t), st>>>(reinterpret_cast<cplx *>(psi), P);
    return (int)cudaGetLastError();
}

extern "C" int b200k_tile_program3(double *psi, int n_local, const B200k2Header *hdr,
                                   const B200k3Round *rounds, const B200k3Op *ops, void *stream)
{
    if (hdr->t > B200K_MAX_TILE_BITS || hdr->t > n_local || hdr->nhigh > B200K_MAX_HIGH ||
        hdr->c + hdr->nhigh != hdr->t || hdr->ntiles != (1ull << (n_local - hdr->t)) ||
        hdr->n_ops > B200K3_LARGE_OPS || hdr->n_rounds > B200K3_LARGE_ROUNDS)
        return (int)cudaErrorInvalidValue;
    if (hdr->n_ops <= B200K3_SMALL_OPS && hdr->n_rounds <= B200K3_SMALL_ROUNDS)
        return launch3<B200K3_SMALL_OPS, B200K3_SMALL_ROUNDS>(psi, hdr, rounds, ops, (cudaStream_t)stream);
    return launch3<B200K3_LARGE_OPS, B200K3_LARGE_ROUNDS>(psi, hdr, rounds, ops, (cudaStream_t)stream);
}
