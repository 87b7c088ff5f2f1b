// tile2.cu — second-generation fused tile pass (see include/b200sv_kernels.h, "second generation").
//
// Evidence that shaped it (profiles/r1_v1_*: ncu --set full of the first kernel at 30 qubits):
// DRAM traffic was already the algorithmic minimum (34.3 GB per pass) but the pass took 58 ms,
// not 5.3 ms: the XU pipe (POPC / FLO / BREV from per-amplitude index arithmetic) sat at its
// ceiling, FP64 at 4 %, and 120 registers capped occupancy at 2 CTAs/SM.  Hence here:
//   * every bit position is precomputed on the host, no POPC/FFS/CLZ per amplitude;
//   * a thread keeps 8 amplitudes (three tile bits) in registers for a whole ROUND of ops —
//     one shared-memory read+write per round, per-op control/parity decisions once per thread;
//   * XOR-swizzled shared memory: conflict-free for any choice of the three register bits;
//   * persistent CTAs (grid = SMs x resident CTAs) striding over the tiles;
//   * 2x2 matrices tagged real / swap so that CNOT, X, H, RY cost no or half the FP64 work.
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200sv_kernels.h"

typedef double2 cplx;
#define TT B200K_TILE_THREADS
#define SW(i) ((i) ^ (((i) >> 3) & 7u))

template <int NOPS, int NROUNDS> struct Prog2 {
    B200k2Header h;
    B200k2Round rounds[NROUNDS];
    B200k2Op ops[NOPS];
};

__device__ __forceinline__ cplx cmul2(cplx a, cplx b)
{
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ uint32_t ins0(uint32_t g, uint32_t b)
{
    return ((g >> b) << (b + 1)) | (g & ((1u << b) - 1u));
}
__device__ __forceinline__ cplx ipow2(cplx a, int k)
{
    switch (k & 3) {
    case 0: return a;
    case 1: return make_double2(-a.y, a.x);
    case 2: return make_double2(-a.x, -a.y);
    default: return make_double2(a.y, -a.x);
    }
}

template <int LO, int HI>
__device__ __forceinline__ void pair_apply(cplx (&v)[8], const B200k2Op &op)
{
    if ((LO & op.rc_mask) != op.rc_val) return; // uniform across the CTA
    const cplx a = v[LO], b = v[HI];
    if (op.mflag == B200K2_M_SWAP) {
        v[LO] = b;
        v[HI] = a;
    }
    else if (op.mflag == B200K2_M_REAL) {
        const double m00 = op.m[0], m01 = op.m[2], m10 = op.m[4], m11 = op.m[6];
        v[LO] = make_double2(fma(m00, a.x, m01 * b.x), fma(m00, a.y, m01 * b.y));
        v[HI] = make_double2(fma(m10, a.x, m11 * b.x), fma(m10, a.y, m11 * b.y));
    }
    else {
        const double r00 = op.m[0], i00 = op.m[1], r01 = op.m[2], i01 = op.m[3];
        const double r10 = op.m[4], i10 = op.m[5], r11 = op.m[6], i11 = op.m[7];
        v[LO] = make_double2(fma(r00, a.x, fma(-i00, a.y, fma(r01, b.x, -i01 * b.y))),
                             fma(r00, a.y, fma(i00, a.x, fma(r01, b.y, i01 * b.x))));
        v[HI] = make_double2(fma(r10, a.x, fma(-i10, a.y, fma(r11, b.x, -i11 * b.y))),
                             fma(r10, a.y, fma(i10, a.x, fma(r11, b.y, i11 * b.x))));
    }
}

template <int NOPS, int NROUNDS>
__global__ void __launch_bounds__(TT, 3)
k_tile2(cplx *__restrict__ psi, const __grid_constant__ Prog2<NOPS, NROUNDS> P)
{
    extern __shared__ cplx s[];
    const int t = P.h.t, c = P.h.c, nhigh = P.h.nhigh;
    const uint32_t nelem = 1u << t;
    const uint32_t lowmask = (1u << c) - 1u;
    const uint32_t tid = threadIdx.x;

    for (uint64_t tile = blockIdx.x; tile < P.h.ntiles; tile += gridDim.x) {
        uint64_t base = tile << c;
        for (int h = 0; h < nhigh; h++) {
            const int p = P.h.high_pos[h];
            base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
        }
        const uint64_t obase = base | P.h.rank_base;

        // ---- load ---------------------------------------------------------------------------
#pragma unroll 4
        for (uint32_t e = tid; e < nelem; e += TT) {
            uint64_t gi = base | (e & lowmask);
            const uint32_t r = e >> c;
            for (int h = 0; h < nhigh; h++) gi |= (uint64_t)((r >> h) & 1u) << P.h.high_pos[h];
            s[SW(e)] = psi[gi];
        }
        __syncthreads();

        // ---- rounds -------------------------------------------------------------------------
        for (int ri = 0; ri < P.h.n_rounds; ri++) {
            const B200k2Round R = P.rounds[ri];
            const int o0 = R.first, o1 = R.first + R.count;
            if (R.kind == B200K2_ROUND_REG) {
                const uint32_t p1 = 1u << R.rb0, p2 = 1u << R.rb1, p4 = 1u << R.rb2;
                for (uint32_t g = tid; g < (nelem >> 3); g += TT) {
                    const uint32_t ib = ins0(ins0(ins0(g, R.rb0), R.rb1), R.rb2);
                    cplx v[8];
#pragma unroll
                    for (int a = 0; a < 8; a++) {
                        const uint32_t idx = ib | ((a & 1) ? p1 : 0u) | ((a & 2) ? p2 : 0u) | ((a & 4) ? p4 : 0u);
                        v[a] = s[SW(idx)];
                    }
                    for (int o = o0; o < o1; o++) {
                        const B200k2Op &op = P.ops[o];
                        if ((obase & op.co_mask) != op.co_val) continue;
                        if ((ib & op.tc_mask) != op.tc_val) continue;
                        if (op.kind == B200K2_PAIR1) {
                            if (op.j == 0) {
                                pair_apply<0, 1>(v, op); pair_apply<2, 3>(v, op);
                                pair_apply<4, 5>(v, op); pair_apply<6, 7>(v, op);
                            }
                            else if (op.j == 1) {
                                pair_apply<0, 2>(v, op); pair_apply<1, 3>(v, op);
                                pair_apply<4, 6>(v, op); pair_apply<5, 7>(v, op);
                            }
                            else {
                                pair_apply<0, 4>(v, op); pair_apply<1, 5>(v, op);
                                pair_apply<2, 6>(v, op); pair_apply<3, 7>(v, op);
                            }
                        }
                        else if (op.kind == B200K2_PHASE) {
                            const cplx e = make_double2(op.m[0], op.m[1]);
#pragma unroll
                            for (int a = 0; a < 8; a++)
                                if ((a & op.rc_mask) == op.rc_val) v[a] = cmul2(v[a], e);
                        }
                        else { // PARITY
                            const uint32_t par0 = (__popc(ib & op.tz_mask) + __popcll(obase & op.zo_mask)) & 1u;
                            const cplx e0 = make_double2(op.m[0], op.m[1]);
                            const cplx e1 = make_double2(op.m[2], op.m[3]);
#pragma unroll
                            for (int a = 0; a < 8; a++) {
                                if ((a & op.rc_mask) != op.rc_val) continue;
                                const uint32_t pa = par0 ^ ((0x96u >> (a & op.rz_mask)) & 1u);
                                v[a] = cmul2(v[a], pa ? e1 : e0);
                            }
                        }
                    }
#pragma unroll
                    for (int a = 0; a < 8; a++) {
                        const uint32_t idx = ib | ((a & 1) ? p1 : 0u) | ((a & 2) ? p2 : 0u) | ((a & 4) ? p4 : 0u);
                        s[SW(idx)] = v[a];
                    }
                }
            }
            else { // shared-memory round: ops executed one after the other on the staged tile
                for (int o = o0; o < o1; o++) {
                    const B200k2Op &op = P.ops[o];
                    const bool active = (obase & op.co_mask) == op.co_val;
                    if (active && op.kind == B200K2_PAIRG) {
                        const cplx m00 = make_double2(op.m[0], op.m[1]), m01 = make_double2(op.m[2], op.m[3]);
                        const cplx m10 = make_double2(op.m[4], op.m[5]), m11 = make_double2(op.m[6], op.m[7]);
                        const uint32_t ng = nelem >> op.npos;
                        for (uint32_t g = tid; g < ng; g += TT) {
                            uint32_t i = g;
                            for (uint32_t q = 0; q < op.npos; q++) i = ins0(i, op.pos[q]);
                            if ((i & op.tc_mask) != op.tc_val) continue;
                            const uint32_t ia = SW(i | op.a), ibb = SW(i | op.b);
                            const cplx va = s[ia], vb = s[ibb];
                            const cplx t0 = cmul2(m00, va), t1 = cmul2(m01, vb);
                            const cplx t2 = cmul2(m10, va), t3 = cmul2(m11, vb);
                            s[ia] = make_double2(t0.x + t1.x, t0.y + t1.y);
                            s[ibb] = make_double2(t2.x + t3.x, t2.y + t3.y);
                        }
                    }
                    else if (active && op.kind == B200K2_PAULIROT) {
                        const double cs = op.m[0], sn = op.m[1];
                        const uint32_t opar = __popcll(obase & op.zo_mask);
                        for (uint32_t g = tid; g < (nelem >> 1); g += TT) {
                            const uint32_t i = ins0(g, op.pos[0]);
                            if ((i & op.tc_mask) != op.tc_val) continue;
                            const uint32_t j = i ^ op.tmask;
                            const cplx vi = s[SW(i)], vj = s[SW(j)];
                            const cplx pi = ipow2(vj, (int)op.ny + 2 * ((__popc(j & op.tz_mask) + opar) & 1));
                            const cplx pj = ipow2(vi, (int)op.ny + 2 * ((__popc(i & op.tz_mask) + opar) & 1));
                            s[SW(i)] = make_double2(cs * vi.x + sn * pi.y, cs * vi.y - sn * pi.x);
                            s[SW(j)] = make_double2(cs * vj.x + sn * pj.y, cs * vj.y - sn * pj.x);
                        }
                    }
                    else if (active && (op.kind == B200K2_PHASE || op.kind == B200K2_PARITY)) {
                        const uint32_t opar = __popcll(obase & op.zo_mask);
                        for (uint32_t e = tid; e < nelem; e += TT) {
                            if ((e & op.tc_mask) != op.tc_val) continue;
                            cplx ph = make_double2(op.m[0], op.m[1]);
                            if (op.kind == B200K2_PARITY && ((__popc(e & op.tz_mask) + opar) & 1))
                                ph = make_double2(op.m[2], op.m[3]);
                            s[SW(e)] = cmul2(s[SW(e)], ph);
                        }
                    }
                    if (o + 1 < o1) __syncthreads();
                }
            }
            __syncthreads();
        }

        // ---- store --------------------------------------------------------------------------
#pragma unroll 4
        for (uint32_t e = tid; e < nelem; e += TT) {
            uint64_t gi = base | (e & lowmask);
            const uint32_t r = e >> c;
            for (int h = 0; h < nhigh; h++) gi |= (uint64_t)((r >> h) & 1u) << P.h.high_pos[h];
            psi[gi] = s[SW(e)];
        }
        __syncthreads();
    }
}

template <int NOPS, int NROUNDS>
static int launch2(double *psi, const B200k2Header *hdr, const B200k2Round *rounds, const B200k2Op *ops,
                   cudaStream_t st)
{
    static int grid_per_sm[B200K_MAX_TILE_BITS + 1];
    static int sm_count = 0;
    // function attributes are per device: one bit per CUDA ordinal (ranks of one process may sit on
    // different GPUs)
    static std::atomic<unsigned long long> attr_mask{0}; // threads of one process may launch concurrently
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    const bool attr_set = (attr_mask.load(std::memory_order_acquire) >> (cur_dev & 63)) & 1ull;
    auto kern = k_tile2<NOPS, NROUNDS>;
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
        attr_mask.fetch_or(1ull << (cur_dev & 63), std::memory_order_release);
    }
    Prog2<NOPS, NROUNDS> P;
    P.h = *hdr;
    for (int i = 0; i < hdr->n_rounds; i++) P.rounds[i] = rounds[i];
    for (int i = 0; i < hdr->n_ops; i++) P.ops[i] = ops[i];
    uint64_t grid = (uint64_t)sm_count * grid_per_sm[hdr->t];
    if (grid > hdr->ntiles) grid = hdr->ntiles;
    kern<<<(unsigned)grid, TT, sizeof(cplx) << hdr->t, st>>>(reinterpret_cast<cplx *>(psi), P);
    return (int)cudaGetLastError();
}

extern "C" int b200k_tile_program2(double *psi, int n_local, const B200k2Header *hdr,
                                   const B200k2Round *rounds, const B200k2Op *ops, void *stream)
{
    if (hdr->t > B200K_MAX_TILE_BITS || hdr->t > n_local || hdr->nhigh > B200K_MAX_HIGH ||
        hdr->c + hdr->nhigh != hdr->t || hdr->ntiles != (1ull << (n_local - hdr->t)) ||
        hdr->n_ops > B200K2_LARGE_OPS || hdr->n_rounds > B200K2_LARGE_ROUNDS)
        return (int)cudaErrorInvalidValue;
    if (hdr->n_ops <= B200K2_SMALL_OPS && hdr->n_rounds <= B200K2_SMALL_ROUNDS)
        return launch2<B200K2_SMALL_OPS, B200K2_SMALL_ROUNDS>(psi, hdr, rounds, ops, (cudaStream_t)stream);
    return launch2<B200K2_LARGE_OPS, B200K2_LARGE_ROUNDS>(psi, hdr, rounds, ops, (cudaStream_t)stream);
}
