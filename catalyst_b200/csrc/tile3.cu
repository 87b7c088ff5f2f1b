// tile3.cu — third-generation fused tile pass (see include/b200sv_kernels.h, "third generation").
//
// Evidence that shaped it (profiles/README.md, ncu --set full of k_tile2 at 30 qubits): DRAM traffic
// at the algorithmic minimum (34.4 GB per launch) but 33-64 ms per launch; issue slots 57 % busy,
// FP64 pipe 11 %, ~630 thread-instructions per amplitude per pass of which ~60 FP64.  The cost was
// the per-op interpreter: field loads, per-amplitude predicate loops whose predicated-off complex
// multiplies still issue, 64-bit index arithmetic per amplitude in the tile load/store.  Here every
// decision that is uniform over the CTA is taken once per op on host-precomputed slot masks, and
// everything that does not depend on the tile is computed once per thread.
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200sv_kernels.h"

typedef double2 cplx;
#define TT B200K_TILE_THREADS
#define SW(i) ((i) ^ (((i) >> 3) & 7u))

template <int NOPS, int NROUNDS> struct Prog3 {
    B200k2Header h;
    B200k3Round rounds[NROUNDS];
    B200k3Op ops[NOPS];
};

__device__ __forceinline__ cplx cmul3(cplx a, cplx b)
{
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
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

// one 2x2 block on the register pair (x, y); MC selects how much FP64 work the matrix needs
template <int MC> __device__ __forceinline__ void pair2x2(cplx &x, cplx &y, const double (&m)[8])
{
    const cplx a = x, b = y;
    if (MC == B200K3_M_SWAP) {
        x = b;
        y = a;
    }
    else if (MC == B200K3_M_REAL) { // all entries real: H, RY, X-like with signs
        x = make_double2(fma(m[0], a.x, m[2] * b.x), fma(m[0], a.y, m[2] * b.y));
        y = make_double2(fma(m[4], a.x, m[6] * b.x), fma(m[4], a.y, m[6] * b.y));
    }
    else if (MC == B200K3_M_RX) { // real diagonal, imaginary off-diagonal: RX, IsingXX blocks
        x = make_double2(fma(m[0], a.x, -m[3] * b.y), fma(m[0], a.y, m[3] * b.x));
        y = make_double2(fma(m[6], b.x, -m[5] * a.y), fma(m[6], b.y, m[5] * a.x));
    }
    else {
        x = make_double2(fma(m[0], a.x, fma(-m[1], a.y, fma(m[2], b.x, -m[3] * b.y))),
                         fma(m[0], a.y, fma(m[1], a.x, fma(m[2], b.y, m[3] * b.x))));
        y = make_double2(fma(m[4], a.x, fma(-m[5], a.y, fma(m[6], b.x, -m[7] * b.y))),
                         fma(m[4], a.y, fma(m[5], a.x, fma(m[6], b.y, m[7] * b.x))));
    }
}

template <int MC, int L0, int H0, int L1, int H1, int L2, int H2, int L3, int H3>
__device__ __forceinline__ void apply4(cplx (&v)[8], const double (&m)[8], uint32_t pm)
{
    if (pm == 15u) {
        pair2x2<MC>(v[L0], v[H0], m);
        pair2x2<MC>(v[L1], v[H1], m);
        pair2x2<MC>(v[L2], v[H2], m);
        pair2x2<MC>(v[L3], v[H3], m);
    }
    else { // controls on register bits: CTA-uniform branches
        if (pm & 1u) pair2x2<MC>(v[L0], v[H0], m);
        if (pm & 2u) pair2x2<MC>(v[L1], v[H1], m);
        if (pm & 4u) pair2x2<MC>(v[L2], v[H2], m);
        if (pm & 8u) pair2x2<MC>(v[L3], v[H3], m);
    }
}
template <int MC, int L0, int H0, int L1, int H1>
__device__ __forceinline__ void apply2(cplx (&v)[8], const double (&m)[8], uint32_t pm)
{
    if (pm & 1u) pair2x2<MC>(v[L0], v[H0], m);
    if (pm & 2u) pair2x2<MC>(v[L1], v[H1], m);
}

template <int MC> __device__ __forceinline__ void pair_op(cplx (&v)[8], const B200k3Op &op)
{
    double m[8];
#pragma unroll
    for (int q = 0; q < 8; q++) m[q] = op.m[q];
    const uint32_t pm = op.amask;
    switch (op.j) {
    case 0: apply4<MC, 0, 1, 2, 3, 4, 5, 6, 7>(v, m, pm); break;
    case 1: apply4<MC, 0, 2, 1, 3, 4, 6, 5, 7>(v, m, pm); break;
    case 2: apply4<MC, 0, 4, 1, 5, 2, 6, 3, 7>(v, m, pm); break;
    // two-target 01 <-> 10: first pattern = lower register bit set, second = higher bit set
    case 3: apply2<MC, 1, 2, 5, 6>(v, m, pm); break; // bits (0,1), free bit 2
    case 4: apply2<MC, 1, 4, 3, 6>(v, m, pm); break; // bits (0,2), free bit 1
    case 5: apply2<MC, 2, 4, 3, 5>(v, m, pm); break; // bits (1,2), free bit 0
    // two-target 00 <-> 11
    case 6: apply2<MC, 0, 3, 4, 7>(v, m, pm); break; // bits (0,1)
    case 7: apply2<MC, 0, 5, 2, 7>(v, m, pm); break; // bits (0,2)
    default: apply2<MC, 0, 6, 1, 7>(v, m, pm); break; // bits (1,2)
    }
}

template <int MASK> __device__ __forceinline__ void phase_slots(cplx (&v)[8], cplx e)
{
#pragma unroll
    for (int a = 0; a < 8; a++)
        if (MASK & (1 << a)) v[a] = cmul3(v[a], e);
}

// v[a] *= ((RPAR >> a) & 1) ? eB : eA on the slots of `am`
template <int RPAR> __device__ __forceinline__ void parity_slots(cplx (&v)[8], cplx eA, cplx eB, uint32_t am)
{
    if (am == 0xffu) {
#pragma unroll
        for (int a = 0; a < 8; a++) v[a] = cmul3(v[a], ((RPAR >> a) & 1) ? eB : eA);
    }
    else {
#pragma unroll
        for (int a = 0; a < 8; a++)
            if (am & (1u << a)) v[a] = cmul3(v[a], ((RPAR >> a) & 1) ? eB : eA);
    }
}

template <int NOPS, int NROUNDS>
__global__ void __launch_bounds__(TT, 3)
k_tile3(cplx *__restrict__ psi, const __grid_constant__ Prog3<NOPS, NROUNDS> P)
{
    extern __shared__ cplx s[];
    __shared__ uint64_t s_offk[32];
    const int t = P.h.t, c = P.h.c, nhigh = P.h.nhigh;
    const uint32_t nelem = 1u << t;
    const uint32_t lowmask = (1u << c) - 1u;
    const uint32_t tid = threadIdx.x;

    // tile-independent part of the global offset of element e: off(e) = (e & lowmask) | scatter(e >> c)
    // is an OR over disjoint bit sets of e, so off(tid + 256 k) = off(tid) | off(256 k)
    auto off_of = [&](uint32_t e) -> uint64_t {
        uint64_t o = e & lowmask;
        const uint32_t r = e >> c;
        for (int h = 0; h < nhigh; h++) o |= (uint64_t)((r >> h) & 1u) << P.h.high_pos[h];
        return o;
    };
    const uint32_t niter = nelem > TT ? nelem / TT : 1u; // <= 16 for t <= 12
    if (tid < niter) s_offk[tid] = off_of(tid * TT);
    const uint64_t off_tid = off_of(tid);
    const uint32_t sw_tid = SW(tid);
    __syncthreads();

    for (uint64_t tile = blockIdx.x; tile < P.h.ntiles; tile += gridDim.x) {
        uint64_t base = tile << c;
        for (int h = 0; h < nhigh; h++) {
            const int p = P.h.high_pos[h];
            base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
        }
        const uint64_t obase = base | P.h.rank_base;
        cplx *__restrict__ tp = psi + (base | off_tid);

        // ---- load ---------------------------------------------------------------------------
        if (tid < nelem) {
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
                    for (int o = o0; o < o1; o++) {
                        const B200k3Op &op = P.ops[o];
                        const uint32_t fl = op.flags;
                        if ((fl & B200K3_F_CO) && (obase & op.co_mask) != op.co_val) continue;
                        if ((fl & B200K3_F_TC) && (ib & op.tc_mask) != op.tc_val) continue;
                        if (op.kind == B200K3_PAIR) {
                            switch (op.mclass) {
                            case B200K3_M_SWAP: pair_op<B200K3_M_SWAP>(v, op); break;
                            case B200K3_M_REAL: pair_op<B200K3_M_REAL>(v, op); break;
                            case B200K3_M_RX: pair_op<B200K3_M_RX>(v, op); break;
                            default: pair_op<B200K3_M_GENERAL>(v, op); break;
                            }
                        }
                        else if (op.kind == B200K3_PHASE) {
                            const cplx e = make_double2(op.m[0], op.m[1]);
                            switch (op.amask) {
                            case 0xff: phase_slots<0xff>(v, e); break;
                            case 0xaa: phase_slots<0xaa>(v, e); break;
                            case 0x55: phase_slots<0x55>(v, e); break;
                            case 0xcc: phase_slots<0xcc>(v, e); break;
                            case 0x33: phase_slots<0x33>(v, e); break;
                            case 0xf0: phase_slots<0xf0>(v, e); break;
                            case 0x0f: phase_slots<0x0f>(v, e); break;
                            case 0x88: phase_slots<0x88>(v, e); break;
                            case 0xa0: phase_slots<0xa0>(v, e); break;
                            case 0xc0: phase_slots<0xc0>(v, e); break;
                            default: {
                                const uint32_t am = op.amask;
#pragma unroll
                                for (int a = 0; a < 8; a++)
                                    if (am & (1u << a)) v[a] = cmul3(v[a], e);
                            }
                            }
                        }
                        else { // PARITY
                            uint32_t p0 = 0;
                            if (fl & B200K3_F_TZ) p0 = (__popc(ib & op.tz_mask) + __popcll(obase & op.zo_mask)) & 1u;
                            const cplx e0 = make_double2(op.m[0], op.m[1]);
                            const cplx e1 = make_double2(op.m[2], op.m[3]);
                            const cplx eA = p0 ? e1 : e0, eB = p0 ? e0 : e1;
                            const uint32_t am = op.amask;
                            switch (op.rpar) {
                            case 0x00: parity_slots<0x00>(v, eA, eB, am); break;
                            case 0xaa: parity_slots<0xaa>(v, eA, eB, am); break;
                            case 0xcc: parity_slots<0xcc>(v, eA, eB, am); break;
                            case 0xf0: parity_slots<0xf0>(v, eA, eB, am); break;
                            case 0x66: parity_slots<0x66>(v, eA, eB, am); break;
                            case 0x5a: parity_slots<0x5a>(v, eA, eB, am); break;
                            case 0x3c: parity_slots<0x3c>(v, eA, eB, am); break;
                            default: parity_slots<0x96>(v, eA, eB, am); break;
                            }
                        }
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
        __syncthreads();
    }
}

template <int NOPS, int NROUNDS>
static int launch3(double *psi, const B200k2Header *hdr, const B200k3Round *rounds, const B200k3Op *ops,
                   cudaStream_t st)
{
    static int grid_per_sm[B200K_MAX_TILE_BITS + 1];
    static int sm_count = 0;
    static bool attr_set = false;
    auto kern = k_tile3<NOPS, NROUNDS>;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(sizeof(cplx) << B200K_MAX_TILE_BITS));
        if (e != cudaSuccess) return (int)e;
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
        for (int t = 0; t <= B200K_MAX_TILE_BITS; t++) {
            int nb = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, TT, sizeof(cplx) << t);
            grid_per_sm[t] = nb > 0 ? nb : 1;
        }
        attr_set = true;
    }
    Prog3<NOPS, NROUNDS> P;
    P.h = *hdr;
    for (int i = 0; i < hdr->n_rounds; i++) P.rounds[i] = rounds[i];
    for (int i = 0; i < hdr->n_ops; i++) P.ops[i] = ops[i];
    uint64_t grid = (uint64_t)sm_count * grid_per_sm[hdr->t];
    if (grid > hdr->ntiles) grid = hdr->ntiles;
    kern<<<(unsigned)grid, TT, sizeof(cplx) << hdr->t, st>>>(reinterpret_cast<cplx *>(psi), P);
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
