// kernels.cu — hand-written sm_100a kernels of the B200 state-vector device, behind the C ABI of
// include/b200sv_kernels.h.  No C++ standard library, no reference headers: this translation unit
// only knows amplitudes, bit positions and masks.
//
// Every kernel here is HBM-bound complex128 streaming work (<= 1 flop/byte), so the design rules
// are: 128-bit (double2) coalesced accesses, >= 2 KiB contiguous rows per tile, as few HBM round
// trips as possible (many gates per staged tile), deterministic warp-shuffle tree reductions.
//
// What the kernels replace in the reference stack (runtime/include/QuantumDevice.hpp):
//   k_tile_program ........ NamedOperation :215 / MatrixOperation :250 (gate sweeps of lightning.qubit)
//   k_bilinear_* .......... Expval :488, Var :503, Gradient :564 (inner products <phi|G|lambda>)
//   k_probs_* ............. Probs :462, PartialProbs :472, first half of Measure :234
//   k_collapse ............ second half of Measure :234
//   k_sampler_* ........... Sample :384 / Counts :433 (alias sampler of SURVEY.md Appendix A.4)
#include <atomic>
#include <cuda_runtime.h>
#include <thrust/copy.h>
#include <thrust/execution_policy.h>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/system/cuda/execution_policy.h>
#include <stdint.h>
#include <string.h>

#include "b200sv_kernels.h"

typedef double2 cplx;

#define TT B200K_TILE_THREADS

__device__ __forceinline__ cplx cmul(cplx a, cplx b)
{
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ cplx cfma(cplx a, cplx b, cplx c)
{ // a*b + c
    return make_double2(fma(a.x, b.x, fma(-a.y, b.y, c.x)), fma(a.x, b.y, fma(a.y, b.x, c.y)));
}
__device__ __forceinline__ cplx cadd(cplx a, cplx b) { return make_double2(a.x + b.x, a.y + b.y); }
// conj(a) * b
__device__ __forceinline__ cplx cconjmul(cplx a, cplx b)
{
    return make_double2(a.x * b.x + a.y * b.y, a.x * b.y - a.y * b.x);
}
// multiply by i^k
__device__ __forceinline__ cplx cmul_ipow(cplx a, int k)
{
    switch (k & 3) {
    case 0: return a;
    case 1: return make_double2(-a.y, a.x);
    case 2: return make_double2(-a.x, -a.y);
    default: return make_double2(a.y, -a.x);
    }
}

// insert zero bits into g at every set position of mask (ascending)
__device__ __forceinline__ uint32_t insert_zeros32(uint32_t g, uint32_t mask)
{
    while (mask) {
        const int b = __ffs(mask) - 1;
        g = ((g >> b) << (b + 1)) | (g & ((1u << b) - 1u));
        mask &= mask - 1;
    }
    return g;
}
__device__ __forceinline__ uint64_t insert_zeros64(uint64_t g, uint64_t mask)
{
    while (mask) {
        const int b = __ffsll((long long)mask) - 1;
        g = ((g >> b) << (b + 1)) | (g & ((1ull << b) - 1ull));
        mask &= mask - 1;
    }
    return g;
}

// =================================================================================================
// Fused tile pass
// =================================================================================================
__global__ void __launch_bounds__(TT)
k_tile_program(cplx *__restrict__ psi, const __grid_constant__ B200kTileProgram P)
{
    extern __shared__ cplx s[];
    const int t = P.t, c = P.c, nhigh = P.nhigh;
    const uint32_t nelem = 1u << t;
    const uint32_t lowmask = (1u << c) - 1u;
    const uint32_t tid = threadIdx.x;

    // tile id -> base index: spread the tile id over the bits that are not tile bits
    uint64_t base = (uint64_t)blockIdx.x << c;
    for (int h = 0; h < nhigh; h++) {
        const int p = P.high_pos[h];
        base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
    }
    const uint64_t obase = base | P.rank_base;

    // ---- load: rows of 2^c contiguous amplitudes --------------------------------------------
#pragma unroll 4
    for (uint32_t e = tid; e < nelem; e += TT) {
        uint64_t gi = base | (e & lowmask);
        const uint32_t r = e >> c;
        for (int h = 0; h < nhigh; h++) gi |= (uint64_t)((r >> h) & 1u) << P.high_pos[h];
        s[e] = psi[gi];
    }
    __syncthreads();

    // ---- op list ----------------------------------------------------------------------------
    int o = 0;
    while (o < P.n_ops) {
        const uint32_t kind = P.ops[o].kind;
        if (kind == B200K_OP_PHASE || kind == B200K_OP_PARITY) {
            // run of diagonal ops: element-local, one pass over the tile for the whole run
            int o1 = o + 1;
            while (o1 < P.n_ops &&
                   (P.ops[o1].kind == B200K_OP_PHASE || P.ops[o1].kind == B200K_OP_PARITY))
                o1++;
            for (uint32_t e = tid; e < nelem; e += TT) {
                cplx v = s[e];
                for (int q = o; q < o1; q++) {
                    const B200kTileOp &op = P.ops[q];
                    if ((obase & op.cmask_o) != op.cval_o) continue;
                    if ((e & op.cmask) != op.cval) continue;
                    if (op.kind == B200K_OP_PHASE) {
                        v = cmul(v, make_double2(op.u.m[0], op.u.m[1]));
                    }
                    else {
                        const int par = (__popc(e & op.a) + __popcll(obase & op.zmask_o)) & 1;
                        v = cmul(v, par ? make_double2(op.u.m[2], op.u.m[3])
                                        : make_double2(op.u.m[0], op.u.m[1]));
                    }
                }
                s[e] = v;
            }
            __syncthreads();
            o = o1;
            continue;
        }
        const B200kTileOp &op = P.ops[o];
        const bool active = (obase & op.cmask_o) == op.cval_o;
        if (kind == B200K_OP_PAIR) {
            if (active) {
                const cplx m00 = make_double2(op.u.m[0], op.u.m[1]);
                const cplx m01 = make_double2(op.u.m[2], op.u.m[3]);
                const cplx m10 = make_double2(op.u.m[4], op.u.m[5]);
                const cplx m11 = make_double2(op.u.m[6], op.u.m[7]);
                const uint32_t ngroups = nelem >> __popc(op.tmask);
                for (uint32_t g = tid; g < ngroups; g += TT) {
                    const uint32_t i = insert_zeros32(g, op.tmask);
                    if ((i & op.cmask) != op.cval) continue;
                    const cplx va = s[i | op.a], vb = s[i | op.b];
                    s[i | op.a] = cfma(m00, va, cmul(m01, vb));
                    s[i | op.b] = cfma(m10, va, cmul(m11, vb));
                }
            }
            __syncthreads();
        }
        else if (kind == B200K_OP_PAULIROT) {
            if (active) {
                const double cs = op.u.m[0], sn = op.u.m[1];
                const uint32_t xm = op.tmask;
                const uint32_t top = 1u << (31 - __clz(xm));
                const int opar = __popcll(obase & op.zmask_o);
                for (uint32_t g = tid; g < (nelem >> 1); g += TT) {
                    const uint32_t i = insert_zeros32(g, top);
                    if ((i & op.cmask) != op.cval) continue;
                    const uint32_t j = i ^ xm;
                    const cplx vi = s[i], vj = s[j];
                    // (P v)_i = i^ny (-1)^par(j) v_j ;  new = cos * v - i sin * (P v)
                    cplx pi = cmul_ipow(vj, (int)op.ny + 2 * ((__popc(j & op.a) + opar) & 1));
                    cplx pj = cmul_ipow(vi, (int)op.ny + 2 * ((__popc(i & op.a) + opar) & 1));
                    s[i] = make_double2(cs * vi.x + sn * pi.y, cs * vi.y - sn * pi.x);
                    s[j] = make_double2(cs * vj.x + sn * pj.y, cs * vj.y - sn * pj.x);
                }
            }
            __syncthreads();
        }
        else { // B200K_OP_DENSE
            constexpr int MAXPER = (1 << B200K_MAX_TILE_BITS) / TT;
            cplx outv[MAXPER];
            if (active) {
                const int k = (int)op.k;
                const uint32_t dim = 1u << k;
                const cplx *__restrict__ M = reinterpret_cast<const cplx *>(op.u.d.mat);
#pragma unroll
                for (int it = 0; it < MAXPER; it++) {
                    const uint32_t e = tid + it * TT;
                    if (e >= nelem) break;
                    if ((e & op.cmask) != op.cval) {
                        outv[it] = s[e];
                        continue;
                    }
                    uint32_t row = 0;
                    for (int j = 0; j < k; j++) row = (row << 1) | ((e >> op.u.d.tb[j]) & 1u);
                    const uint32_t ebase = e & ~op.tmask;
                    cplx acc = make_double2(0.0, 0.0);
                    for (uint32_t col = 0; col < dim; col++) {
                        uint32_t src = ebase;
                        for (int j = 0; j < k; j++)
                            src |= ((col >> (k - 1 - j)) & 1u) << op.u.d.tb[j];
                        acc = cfma(__ldg(&M[row * dim + col]), s[src], acc);
                    }
                    outv[it] = acc;
                }
            }
            __syncthreads();
            if (active) {
#pragma unroll
                for (int it = 0; it < MAXPER; it++) {
                    const uint32_t e = tid + it * TT;
                    if (e >= nelem) break;
                    s[e] = outv[it];
                }
            }
            __syncthreads();
        }
        o++;
    }

    // ---- store --------------------------------------------------------------------------------
#pragma unroll 4
    for (uint32_t e = tid; e < nelem; e += TT) {
        uint64_t gi = base | (e & lowmask);
        const uint32_t r = e >> c;
        for (int h = 0; h < nhigh; h++) gi |= (uint64_t)((r >> h) & 1u) << P.high_pos[h];
        psi[gi] = s[e];
    }
}

extern "C" int b200k_tile_program(double *psi, int n_local, const B200kTileProgram *prog,
                                  void *stream)
{
    if (prog->t > B200K_MAX_TILE_BITS || prog->t > n_local || prog->n_ops > B200K_MAX_OPS ||
        prog->nhigh > B200K_MAX_HIGH || prog->c + prog->nhigh != prog->t)
        return (int)cudaErrorInvalidValue;
    // function attributes are per device: one bit per CUDA ordinal (ranks of one process may sit on
    // different GPUs)
    static std::atomic<unsigned long long> attr_mask{0}; // threads of one process may launch concurrently
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    const bool attr_set = (attr_mask.load(std::memory_order_acquire) >> (cur_dev & 63)) & 1ull;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_tile_program,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)(sizeof(cplx) << B200K_MAX_TILE_BITS));
        if (e != cudaSuccess) return (int)e;
        attr_mask.fetch_or(1ull << (cur_dev & 63), std::memory_order_release);
    }
    const uint64_t ntiles = 1ull << (n_local - prog->t);
    if (ntiles > 0x7fffffffull) return (int)cudaErrorInvalidValue;
    const size_t smem = sizeof(cplx) << prog->t;
    k_tile_program<<<(unsigned)ntiles, TT, smem, (cudaStream_t)stream>>>(
        reinterpret_cast<cplx *>(psi), *prog);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Reductions: per-thread grid-stride accumulate -> warp shuffle tree -> per-block partials in
// `scratch` -> single-block fixed-order final sum.  Deterministic run to run.
// =================================================================================================
#define RED_BLOCKS 1184 /* 148 SMs x 8 */
#define RED_THREADS 256
#define RED_MAXVALS 2

extern "C" int64_t b200k_reduce_scratch_doubles(void) { return (int64_t)RED_BLOCKS * RED_MAXVALS; }

template <int NV> __device__ __forceinline__ void block_reduce_store(double (&v)[NV], double *partial)
{
    __shared__ double sh[NV][RED_THREADS / 32];
#pragma unroll
    for (int q = 0; q < NV; q++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v[q] += __shfl_down_sync(0xffffffffu, v[q], off);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) sh[q][w] = v[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double a = 0.0;
            for (int i = 0; i < RED_THREADS / 32; i++) a += sh[q][i];
            partial[(size_t)blockIdx.x * NV + q] = a;
        }
    }
}

template <int NV> __global__ void __launch_bounds__(RED_THREADS)
k_reduce_final(const double *__restrict__ partial, int nblocks, double *__restrict__ out)
{
    double v[NV];
#pragma unroll
    for (int q = 0; q < NV; q++) v[q] = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += RED_THREADS) {
#pragma unroll
        for (int q = 0; q < NV; q++) v[q] += partial[(size_t)i * NV + q];
    }
    __shared__ double sh[NV][RED_THREADS / 32];
#pragma unroll
    for (int q = 0; q < NV; q++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v[q] += __shfl_down_sync(0xffffffffu, v[q], off);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) sh[q][w] = v[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double a = 0.0;
            for (int i = 0; i < RED_THREADS / 32; i++) a += sh[q][i];
            out[q] = a;
        }
    }
}

static inline int red_blocks_for(uint64_t work)
{
    uint64_t b = (work + RED_THREADS - 1) / RED_THREADS;
    if (b < 1) b = 1;
    if (b > RED_BLOCKS) b = RED_BLOCKS;
    return (int)b;
}

// ---- <phi| P |lam> for a Pauli word --------------------------------------------------------------
__global__ void __launch_bounds__(RED_THREADS)
k_bilinear_pauli(const cplx *__restrict__ phi, const cplx *__restrict__ lam, int n, uint64_t rank_base,
                 uint64_t xm, uint64_t zm, int ny, uint64_t cmask, uint64_t cval,
                 double *__restrict__ partial)
{
    double acc[2] = {0.0, 0.0};
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t start = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (xm == 0) {
        const uint64_t N = 1ull << n;
        for (uint64_t i = start; i < N; i += stride) {
            const uint64_t gi = i | rank_base;
            if ((gi & cmask) != cval) continue;
            cplx tq = cconjmul(phi[i], lam[i]);
            tq = cmul_ipow(tq, ny + 2 * (__popcll(gi & zm) & 1));
            acc[0] += tq.x;
            acc[1] += tq.y;
        }
    }
    else {
        const uint64_t top = 1ull << (63 - __clzll((long long)xm));
        const uint64_t NP = 1ull << (n - 1);
        const bool same = (phi == lam);
        for (uint64_t g = start; g < NP; g += stride) {
            const uint64_t i = insert_zeros64(g, top);
            const uint64_t gi = i | rank_base;
            if ((gi & cmask) != cval) continue;
            const uint64_t j = i ^ xm, gj = gi ^ xm;
            const cplx li = lam[i], lj = lam[j];
            const cplx pi = same ? li : phi[i];
            const cplx pj = same ? lj : phi[j];
            // conj(phi_i) * ph(j) * lam_j + conj(phi_j) * ph(i) * lam_i
            cplx t1 = cmul_ipow(cconjmul(pi, lj), ny + 2 * (__popcll(gj & zm) & 1));
            cplx t2 = cmul_ipow(cconjmul(pj, li), ny + 2 * (__popcll(gi & zm) & 1));
            acc[0] += t1.x + t2.x;
            acc[1] += t1.y + t2.y;
        }
    }
    block_reduce_store<2>(acc, partial);
}

extern "C" int b200k_bilinear_pauli(const double *phi, const double *lam, int n_local,
                                    uint64_t rank_base, uint64_t xmask, uint64_t zmask, int ny,
                                    uint64_t cmask, uint64_t cval, double *scratch, double *out,
                                    void *stream)
{
    if (xmask >> n_local) return (int)cudaErrorInvalidValue;
    const uint64_t work = xmask ? (1ull << (n_local - 1)) : (1ull << n_local);
    const int nb = red_blocks_for(work);
    k_bilinear_pauli<<<nb, RED_THREADS, 0, (cudaStream_t)stream>>>(
        (const cplx *)phi, (const cplx *)lam, n_local, rank_base, xmask, zmask, ny, cmask, cval,
        scratch);
    k_reduce_final<2><<<1, RED_THREADS, 0, (cudaStream_t)stream>>>(scratch, nb, out);
    return (int)cudaGetLastError();
}

// ---- <phi| (proj_ctrl (x) g on span{a,b}) |lam> ---------------------------------------------------
struct G8 {
    double v[8];
};
__global__ void __launch_bounds__(RED_THREADS)
k_bilinear_pair(const cplx *__restrict__ phi, const cplx *__restrict__ lam, int n, uint64_t rank_base,
                uint64_t tmask, uint64_t a, uint64_t b, G8 g, uint64_t cmask, uint64_t cval,
                double *__restrict__ partial)
{
    double acc[2] = {0.0, 0.0};
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t NG = 1ull << (n - __popcll(tmask));
    const cplx g00 = make_double2(g.v[0], g.v[1]), g01 = make_double2(g.v[2], g.v[3]);
    const cplx g10 = make_double2(g.v[4], g.v[5]), g11 = make_double2(g.v[6], g.v[7]);
    for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < NG; q += stride) {
        const uint64_t i = insert_zeros64(q, tmask);
        if (((i | rank_base) & cmask) != cval) continue;
        const cplx la = lam[i | a], lb = lam[i | b];
        const cplx pa = phi[i | a], pb = phi[i | b];
        const cplx ga = cfma(g00, la, cmul(g01, lb));
        const cplx gb = cfma(g10, la, cmul(g11, lb));
        const cplx t1 = cconjmul(pa, ga), t2 = cconjmul(pb, gb);
        acc[0] += t1.x + t2.x;
        acc[1] += t1.y + t2.y;
    }
    block_reduce_store<2>(acc, partial);
}

extern "C" int b200k_bilinear_pair(const double *phi, const double *lam, int n_local,
                                   uint64_t rank_base, uint64_t tmask, uint64_t a, uint64_t b,
                                   const double *g8, uint64_t cmask, uint64_t cval, double *scratch,
                                   double *out, void *stream)
{
    if (tmask == 0 || (tmask >> n_local)) return (int)cudaErrorInvalidValue;
    G8 g;
    for (int i = 0; i < 8; i++) g.v[i] = g8[i];
    const uint64_t work = 1ull << (n_local - __builtin_popcountll(tmask));
    const int nb = red_blocks_for(work);
    k_bilinear_pair<<<nb, RED_THREADS, 0, (cudaStream_t)stream>>>(
        (const cplx *)phi, (const cplx *)lam, n_local, rank_base, tmask, a, b, g, cmask, cval,
        scratch);
    k_reduce_final<2><<<1, RED_THREADS, 0, (cudaStream_t)stream>>>(scratch, nb, out);
    return (int)cudaGetLastError();
}

// ---- sum_t coef_t <psi| Z-string_t |psi> in one read pass -----------------------------------------
__global__ void __launch_bounds__(RED_THREADS)
k_expval_diag_batch(const cplx *__restrict__ psi, int n, uint64_t rank_base,
                    const uint64_t *__restrict__ zmasks, const double *__restrict__ coefs, int nterms,
                    double *__restrict__ partial)
{
    extern __shared__ unsigned char smraw[];
    uint64_t *sz = reinterpret_cast<uint64_t *>(smraw);
    double *sc = reinterpret_cast<double *>(smraw + sizeof(uint64_t) * nterms);
    for (int q = threadIdx.x; q < nterms; q += blockDim.x) {
        sz[q] = zmasks[q];
        sc[q] = coefs[q];
    }
    __syncthreads();
    double acc[1] = {0.0};
    const uint64_t N = 1ull << n;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        const cplx v = psi[i];
        const double p = v.x * v.x + v.y * v.y;
        const uint64_t gi = i | rank_base;
        double w = 0.0;
        for (int q = 0; q < nterms; q++) w += (__popcll(gi & sz[q]) & 1) ? -sc[q] : sc[q];
        acc[0] = fma(p, w, acc[0]);
    }
    block_reduce_store<1>(acc, partial);
}

extern "C" int b200k_expval_diag_batch(const double *psi, int n_local, uint64_t rank_base,
                                       const uint64_t *zmasks, const double *coefs, int nterms,
                                       double *scratch, double *out, void *stream)
{
    if (nterms < 1 || nterms > 2048) return (int)cudaErrorInvalidValue;
    const int nb = red_blocks_for(1ull << n_local);
    k_expval_diag_batch<<<nb, RED_THREADS, 16 * (size_t)nterms, (cudaStream_t)stream>>>(
        (const cplx *)psi, n_local, rank_base, zmasks, coefs, nterms, scratch);
    k_reduce_final<1><<<1, RED_THREADS, 0, (cudaStream_t)stream>>>(scratch, nb, out);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Probabilities
// =================================================================================================
__global__ void __launch_bounds__(256)
k_probs_all(const cplx *__restrict__ psi, double *__restrict__ out, uint64_t N)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x * 2;
    for (uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < N; i += stride) {
        const cplx a = psi[i], b = psi[i + 1];
        *reinterpret_cast<double2 *>(out + i) =
            make_double2(a.x * a.x + a.y * a.y, b.x * b.x + b.y * b.y);
    }
}
__global__ void k_probs_all1(const cplx *psi, double *out)
{
    out[0] = psi[0].x * psi[0].x + psi[0].y * psi[0].y;
}

extern "C" int b200k_probs_all(const double *psi, double *out, int n_local, void *stream)
{
    const uint64_t N = 1ull << n_local;
    if (N == 1) {
        k_probs_all1<<<1, 1, 0, (cudaStream_t)stream>>>((const cplx *)psi, out);
        return (int)cudaGetLastError();
    }
    uint64_t nb = (N / 2 + 255) / 256;
    if (nb > 148 * 32) nb = 148 * 32;
    k_probs_all<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>((const cplx *)psi, out, N);
    return (int)cudaGetLastError();
}

// |psi|^2 written through a BIT PERMUTATION of the index: out[dst(i)] = |psi_i|^2 where source bit b
// lands on destination bit perm[b] (the host's wire relabelling — SWAP gates executed as renames — is
// undone on the way out, so Probs()/Sample() of all wires see canonical wire order at the cost of the
// plain pass).  A tile = the low C source bits plus the source bits that become the low C destination
// bits: loads are contiguous in the source, stores contiguous in the destination, the transposition
// happens in shared memory.
#define PP_C 6
struct PermTile {
    uint8_t nt;          // tile bits (C .. 2C)
    uint8_t n;
    uint8_t src_pos[2 * PP_C];   // tile bit j (source order, ascending) -> source position
    uint8_t dst_rank[2 * PP_C];  // tile bit in destination order j' -> its index j in source order
    uint8_t dst_pos[2 * PP_C];   // tile bit in destination order j' -> destination position
    uint8_t rest_src[64];        // non-tile source positions (ascending)
    uint8_t rest_dst[64];        // their destination positions
    uint8_t nrest;
    // shared-memory swizzle: index ^= bit(index, swz_from[q]) << swz_to[q].  The host picks it so that BOTH phases are
    // bank-conflict free for any permutation: a half-warp writes 16 consecutive source-order slots and reads the 16
    // slots spanned by the source-order bits dst_rank[0..3]; each of those that is not itself one of the low four
    // bits is folded onto a low bit none of the others uses (a plain 64 x 64 transposition read with stride 64:
    // 16-way conflicts, 12.1 ms at 30 qubits instead of the ~5 ms the 25.8 GB of traffic cost).
    uint8_t swz_from[4], swz_to[4], nswz;
};
__device__ __forceinline__ uint32_t perm_swz(uint32_t i, const PermTile &T)
{
#pragma unroll
    for (int q = 0; q < 4; q++)
        if (q < T.nswz) i ^= ((i >> T.swz_from[q]) & 1u) << T.swz_to[q];
    return i;
}
__global__ void __launch_bounds__(256)
k_probs_all_perm(const cplx *__restrict__ psi, double *__restrict__ out, const __grid_constant__ PermTile T)
{
    __shared__ double sm[1 << (2 * PP_C)];
    const uint32_t nel = 1u << T.nt;
    const uint64_t ntiles = 1ull << T.nrest;
    const uint32_t tid = threadIdx.x;
    const int nlow = T.nt < 8 ? T.nt : 8; // tile bits covered by the thread index; the rest by the slice counter k
    const uint32_t nk = nel > 256u ? nel / 256u : 1u;
    // per-thread parts of the offsets (hoisted out of the tile loop)
    uint64_t s_tid = 0, d_tid = 0;
    uint32_t se_tid = 0;
    for (int j = 0; j < nlow; j++) {
        const uint32_t bit = (tid >> j) & 1u;
        s_tid |= (uint64_t)bit << T.src_pos[j];
        d_tid |= (uint64_t)bit << T.dst_pos[j];
        se_tid |= bit << T.dst_rank[j];
    }
    for (uint64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        uint64_t sbase = 0, dbase = 0;
        for (int q = 0; q < T.nrest; q++) {
            const uint64_t bit = (tile >> q) & 1ull;
            sbase |= bit << T.rest_src[q];
            dbase |= bit << T.rest_dst[q];
        }
        cplx a[16];
#pragma unroll
        for (uint32_t k = 0; k < 16; k++) { // all loads of the thread in flight before the first use
            if (k < nk && tid + k * 256u < nel) {
                uint64_t s_k = 0;
                for (int j = 8; j < T.nt; j++) s_k |= (uint64_t)((k >> (j - 8)) & 1u) << T.src_pos[j];
                a[k] = psi[sbase | s_tid | s_k];
            }
        }
#pragma unroll
        for (uint32_t k = 0; k < 16; k++)
            if (k < nk && tid + k * 256u < nel) sm[perm_swz(tid + k * 256u, T)] = a[k].x * a[k].x + a[k].y * a[k].y;
        __syncthreads();
#pragma unroll
        for (uint32_t k = 0; k < 16; k++) { // slot tid + 256 k now enumerates in DESTINATION order
            if (k < nk && tid + k * 256u < nel) {
                uint64_t d_k = 0;
                uint32_t se_k = 0;
                for (int j = 8; j < T.nt; j++) {
                    const uint32_t bit = (k >> (j - 8)) & 1u;
                    d_k |= (uint64_t)bit << T.dst_pos[j];
                    se_k |= bit << T.dst_rank[j];
                }
                out[dbase | d_tid | d_k] = sm[perm_swz(se_tid | se_k, T)];
            }
        }
        __syncthreads();
    }
}

extern "C" int b200k_probs_all_perm(const double *psi, double *out, int n, const uint8_t *perm, void *stream)
{
    if (n < 1 || n > 62) return (int)cudaErrorInvalidValue;
    PermTile T;
    memset(&T, 0, sizeof(T));
    T.n = (uint8_t)n;
    const int C = n < PP_C ? n : PP_C;
    bool in_tile[64] = {false};
    for (int b = 0; b < C; b++) in_tile[b] = true;                 // low source bits
    for (int b = 0; b < n; b++)
        if (perm[b] < C) in_tile[b] = true;                        // sources of the low destination bits
    int nt = 0;
    for (int b = 0; b < n; b++)
        if (in_tile[b]) T.src_pos[nt++] = (uint8_t)b;
    T.nt = (uint8_t)nt;
    // destination order: tile bits sorted by destination position
    int order[2 * PP_C];
    for (int j = 0; j < nt; j++) order[j] = j;
    for (int i = 0; i < nt; i++)
        for (int j = i + 1; j < nt; j++)
            if (perm[T.src_pos[order[j]]] < perm[T.src_pos[order[i]]]) {
                const int tmp = order[i];
                order[i] = order[j];
                order[j] = tmp;
            }
    for (int j = 0; j < nt; j++) {
        T.dst_rank[j] = (uint8_t)order[j];
        T.dst_pos[j] = perm[T.src_pos[order[j]]];
    }
    // swizzle (see PermTile): the source-order bits behind the four lowest destination-order bits
    {
        bool used[4] = {false, false, false, false};
        const int nd = nt < 4 ? nt : 4;
        for (int j = 0; j < nd; j++)
            if (T.dst_rank[j] < 4) used[T.dst_rank[j]] = true;
        int free_low = 0;
        for (int j = 0; j < nd; j++) {
            if (T.dst_rank[j] < 4) continue;
            while (free_low < 4 && used[free_low]) free_low++;
            if (free_low >= 4) break;
            T.swz_from[T.nswz] = T.dst_rank[j];
            T.swz_to[T.nswz] = (uint8_t)free_low;
            used[free_low] = true;
            T.nswz++;
        }
    }
    int nr = 0;
    for (int b = 0; b < n; b++)
        if (!in_tile[b]) {
            T.rest_src[nr] = (uint8_t)b;
            T.rest_dst[nr] = perm[b];
            nr++;
        }
    T.nrest = (uint8_t)nr;
    uint64_t nb = 1ull << nr;
    if (nb > 148 * 7) nb = 148 * 7; // 32 KiB of shared memory per CTA: 7 resident CTAs per SM
    k_probs_all_perm<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>((const cplx *)psi, out, T);
    return (int)cudaGetLastError();
}

// Marginal on k target bits.  A group of GS threads (1, 32 or 256) owns one (bin, split) work item
// and sums |psi|^2 over its slice of the 2^(n-k) complementary indices.
struct TB64 {
    uint8_t tb[64];
};
static inline void probs_shape(int n, int k, int *gs, int *splits)
{
    const int rest = n - k; // log2 of elements per bin
    int g = rest >= 8 ? 256 : (rest >= 5 ? 32 : 1);
    // want ~2^18 threads in flight: bins * splits * gs >= 2^18 where possible
    int s = 0; // log2 splits
    while ((k + s) + (g == 256 ? 8 : g == 32 ? 5 : 0) < 18 && rest - s > (g == 256 ? 10 : g == 32 ? 7 : 2))
        s++;
    *gs = g;
    *splits = 1 << s;
}
extern "C" int b200k_probs_splits(int n_local, int k)
{
    int gs, sp;
    probs_shape(n_local, k, &gs, &sp);
    return sp;
}

template <int GS> __global__ void __launch_bounds__(256)
k_probs_partial(const cplx *__restrict__ psi, int n, int k, TB64 T, uint64_t tmask, int log_splits,
                double *__restrict__ partial)
{
    const uint64_t nitems = 1ull << (k + log_splits);
    const uint64_t gid = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / GS;
    const int lane = threadIdx.x % GS;
    if (gid >= nitems) return; // whole group exits together (GS divides blockDim)
    const uint64_t bin = gid >> log_splits, split = gid & ((1ull << log_splits) - 1);
    uint64_t binbits = 0;
    for (int j = 0; j < k; j++) binbits |= ((bin >> (k - 1 - j)) & 1ull) << T.tb[j];
    const uint64_t per = 1ull << (n - k - log_splits);
    const uint64_t r0 = split * per;
    double acc = 0.0;
    for (uint64_t r = lane; r < per; r += GS) {
        const uint64_t i = insert_zeros64(r0 + r, tmask) | binbits;
        const cplx v = psi[i];
        acc += v.x * v.x + v.y * v.y;
    }
    if (GS >= 32) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, off);
    }
    if (GS == 256) {
        __shared__ double sh[8];
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0;
            for (int i = 0; i < 8; i++) a += sh[i];
            partial[gid] = a;
        }
    }
    else if (lane == 0) {
        partial[gid] = acc;
    }
}
__global__ void k_probs_finish(const double *__restrict__ partial, int log_splits, uint64_t nbins,
                               double *__restrict__ out)
{
    const uint64_t o = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= nbins) return;
    double a = 0.0;
    const uint64_t ns = 1ull << log_splits;
    for (uint64_t q = 0; q < ns; q++) a += partial[(o << log_splits) + q];
    out[o] = a;
}

extern "C" int b200k_probs_partial(const double *psi, int n_local, const uint8_t *tb, int k,
                                   double *scratch, double *out, void *stream)
{
    if (k < 0 || k > n_local || k > 62) return (int)cudaErrorInvalidValue;
    TB64 T;
    uint64_t tmask = 0;
    for (int j = 0; j < k; j++) {
        T.tb[j] = tb[j];
        tmask |= 1ull << tb[j];
    }
    int gs, sp;
    probs_shape(n_local, k, &gs, &sp);
    int ls = 0;
    while ((1 << ls) < sp) ls++;
    const uint64_t nitems = 1ull << (k + ls);
    const uint64_t nthreads = nitems * (uint64_t)gs;
    const uint64_t nb = (nthreads + 255) / 256;
    if (nb > 0x7fffffffull) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    const cplx *p = (const cplx *)psi;
    if (gs == 256) k_probs_partial<256><<<(unsigned)nb, 256, 0, st>>>(p, n_local, k, T, tmask, ls, scratch);
    else if (gs == 32) k_probs_partial<32><<<(unsigned)nb, 256, 0, st>>>(p, n_local, k, T, tmask, ls, scratch);
    else k_probs_partial<1><<<(unsigned)nb, 256, 0, st>>>(p, n_local, k, T, tmask, ls, scratch);
    const uint64_t nbins = 1ull << k;
    k_probs_finish<<<(unsigned)((nbins + 255) / 256), 256, 0, st>>>(scratch, ls, nbins, out);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Element-wise passes
// =================================================================================================
__global__ void __launch_bounds__(256)
k_collapse(cplx *__restrict__ psi, uint64_t N, uint64_t rank_base, uint64_t bit, int keep, double scale)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        const bool b = ((i | rank_base) & bit) != 0;
        if (b == (keep != 0)) {
            cplx v = psi[i];
            psi[i] = make_double2(v.x * scale, v.y * scale);
        }
        else {
            psi[i] = make_double2(0.0, 0.0);
        }
    }
}
static inline unsigned ew_blocks(uint64_t N)
{
    uint64_t nb = (N + 255) / 256;
    if (nb > 148 * 32) nb = 148 * 32;
    return (unsigned)nb;
}
extern "C" int b200k_collapse(double *psi, int n_local, uint64_t rank_base, int pos, int keep,
                              double scale, void *stream)
{
    const uint64_t N = 1ull << n_local;
    k_collapse<<<ew_blocks(N), 256, 0, (cudaStream_t)stream>>>((cplx *)psi, N, rank_base, 1ull << pos,
                                                              keep, scale);
    return (int)cudaGetLastError();
}

__global__ void __launch_bounds__(256)
k_pauli_apply(cplx *__restrict__ out, const cplx *__restrict__ in, uint64_t N, uint64_t rank_base,
              uint64_t xm, uint64_t zm, int ny, cplx coef, int accumulate)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        const uint64_t j = i ^ xm;
        cplx v = cmul_ipow(in[j], ny + 2 * (__popcll((j | rank_base) & zm) & 1));
        v = cmul(v, coef);
        if (accumulate) v = cadd(v, out[i]);
        out[i] = v;
    }
}
extern "C" int b200k_pauli_apply(double *out, const double *in, int n_local, uint64_t rank_base,
                                 uint64_t xmask, uint64_t zmask, int ny, double coef_re,
                                 double coef_im, int accumulate, void *stream)
{
    if (xmask >> n_local) return (int)cudaErrorInvalidValue;
    const uint64_t N = 1ull << n_local;
    k_pauli_apply<<<ew_blocks(N), 256, 0, (cudaStream_t)stream>>>(
        (cplx *)out, (const cplx *)in, N, rank_base, xmask, zmask, ny,
        make_double2(coef_re, coef_im), accumulate);
    return (int)cudaGetLastError();
}

// out_i (+)= sum_t coef_t * (P_t in)_i for up to B200K_APPLY_WORDS Pauli words per pass: the words'
// permuted reads of `in` are independent gathers (aligned blocks map to aligned blocks under i ^ x), the
// accumulator is read and written once per pass instead of once per word: (K + 2) instead of 3 K vector
// transfers for K words (Hamiltonian observables of the adjoint gradient: phi = H psi).
struct ApplyWords {
    uint64_t xm[B200K_APPLY_WORDS], zm[B200K_APPLY_WORDS];
    double cr[B200K_APPLY_WORDS], ci[B200K_APPLY_WORDS];
    int ny[B200K_APPLY_WORDS];
    int n;
};
__global__ void __launch_bounds__(256)
k_pauli_apply_batch(cplx *__restrict__ out, const cplx *__restrict__ in, uint64_t N, uint64_t rank_base,
                    const __grid_constant__ ApplyWords W, int accumulate)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        cplx acc = accumulate ? out[i] : make_double2(0.0, 0.0);
#pragma unroll 4
        for (int t = 0; t < W.n; t++) {
            const uint64_t j = i ^ W.xm[t];
            cplx v = cmul_ipow(in[j], W.ny[t] + 2 * (__popcll((j | rank_base) & W.zm[t]) & 1));
            acc = cfma(make_double2(W.cr[t], W.ci[t]), v, acc);
        }
        out[i] = acc;
    }
}
extern "C" int b200k_pauli_apply_batch(double *out, const double *in, int n_local, uint64_t rank_base,
                                       const uint64_t *xmasks, const uint64_t *zmasks, const int *nys,
                                       const double *coefs_ri, int nwords, int accumulate, void *stream)
{
    if (nwords < 1 || nwords > B200K_APPLY_WORDS) return (int)cudaErrorInvalidValue;
    ApplyWords W;
    memset(&W, 0, sizeof(W));
    W.n = nwords;
    for (int t = 0; t < nwords; t++) {
        if (xmasks[t] >> n_local) return (int)cudaErrorInvalidValue;
        W.xm[t] = xmasks[t];
        W.zm[t] = zmasks[t];
        W.ny[t] = nys[t];
        W.cr[t] = coefs_ri[2 * t];
        W.ci[t] = coefs_ri[2 * t + 1];
    }
    const uint64_t N = 1ull << n_local;
    k_pauli_apply_batch<<<ew_blocks(N), 256, 0, (cudaStream_t)stream>>>((cplx *)out, (const cplx *)in, N, rank_base, W,
                                                                      accumulate);
    return (int)cudaGetLastError();
}

__global__ void __launch_bounds__(256)
k_axpy(cplx *__restrict__ acc, const cplx *__restrict__ x, double coef, uint64_t N)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
        const cplx a = acc[i], b = x[i];
        acc[i] = make_double2(fma(coef, b.x, a.x), fma(coef, b.y, a.y));
    }
}
extern "C" int b200k_axpy(double *acc, const double *x, double coef, int n_local, void *stream)
{
    const uint64_t N = 1ull << n_local;
    k_axpy<<<ew_blocks(N), 256, 0, (cudaStream_t)stream>>>((cplx *)acc, (const cplx *)x, coef, N);
    return (int)cudaGetLastError();
}

__global__ void k_set_one(cplx *psi, int64_t idx) { psi[idx] = make_double2(1.0, 0.0); }
extern "C" int b200k_set_basis(double *psi, int n_local, int64_t local_index, void *stream)
{
    cudaError_t e = cudaMemsetAsync(psi, 0, sizeof(cplx) << n_local, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    if (local_index >= 0) k_set_one<<<1, 1, 0, (cudaStream_t)stream>>>((cplx *)psi, local_index);
    return (int)cudaGetLastError();
}

__global__ void __launch_bounds__(256)
k_scatter_state(cplx *__restrict__ psi, const cplx *__restrict__ vals, int k, TB64 T)
{
    const uint64_t K = 1ull << k;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < K; r += stride) {
        uint64_t idx = 0;
        for (int j = 0; j < k; j++) idx |= ((r >> (k - 1 - j)) & 1ull) << T.tb[j];
        psi[idx] = vals[r];
    }
}
extern "C" int b200k_scatter_state(double *psi, const double *vals, int n_local, const uint8_t *tb,
                                   int k, void *stream)
{
    if (k < 0 || k > n_local || k > 62) return (int)cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(psi, 0, sizeof(cplx) << n_local, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    TB64 T;
    for (int j = 0; j < k; j++) T.tb[j] = tb[j];
    k_scatter_state<<<ew_blocks(1ull << k), 256, 0, (cudaStream_t)stream>>>((cplx *)psi,
                                                                          (const cplx *)vals, k, T);
    return (int)cudaGetLastError();
}

__global__ void __launch_bounds__(256)
k_append_qubit(cplx *__restrict__ dst, const cplx *__restrict__ src, uint64_t Nold)
{
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < Nold; i += stride) {
        dst[2 * i] = src[i];
        dst[2 * i + 1] = make_double2(0.0, 0.0);
    }
}
extern "C" int b200k_append_qubit(double *dst, const double *src, int n_local_old, void *stream)
{
    const uint64_t N = 1ull << n_local_old;
    k_append_qubit<<<ew_blocks(N), 256, 0, (cudaStream_t)stream>>>((cplx *)dst, (const cplx *)src, N);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Alias sampler without the sequential table build.
//
// Reference algorithm (verified against doc/releases/changelog-0.9.0.md:
// 299-315): first[i] = N p[i]; indices with first < 1 go on an `under` stack, the others on an
// `over` stack (both filled in ascending order, so popped in DESCENDING index order); the top
// `over` item donates to successive `under` items until it drops below 1, becomes the top `under`
// item itself and is served first by the next donor.  A shot draws idx = floor(u1 N) and moves to
// partner[idx] iff u2 >= first[idx].
//
// Restated with prefix sums in descending index order: D = inclusive prefix of deficits 1-first
// over `under` items, E = inclusive prefix of excesses first-1 over `over` items.
//   under item u : first unchanged; partner = first over item m with E_m >= D_u - deficit_u
//   over item m  : let u* = first under item with D_u* > E_m.  None -> never drops below 1 ->
//                  always accepted.  Else first = 1 - (D_u* - E_m), partner = next over item.
// Only the buckets the shots actually hit are evaluated; chunk sums (descending chunks of CH
// entries, summed sequentially in a fixed order) give each query an O(CH + log) walk.
// =================================================================================================
#define SAMP_CH 1024

// scratch layout (doubles): [cumD: nch+1][cumE: nch+1][sumD: nch+1][sumE: nch+1][segD: 1024][segE: 1024]
// (cum*: exclusive prefix over descending chunks)
extern "C" int64_t b200k_sampler_scratch_doubles(int64_t nbins)
{
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    return 4 * (nch + 1) + 2048;
}

// One thread per chunk sums its SAMP_CH entries sequentially in descending index order (the order is part of the
// arithmetic the sampled stream depends on); the loads are made coalesced by a warp-level transposition: a warp owns
// 32 consecutive chunks, its lanes fetch 32 consecutive entries of one chunk at a time (256-byte segments) into shared
// memory, then every lane walks its own chunk's row.  (One uncoalesced 8-byte load per thread and step: 3.9 ms at
// 2^30 entries; now bound by the 8 GiB read.)
#define SAMP_WARPS 4
__global__ void __launch_bounds__(32 * SAMP_WARPS)
k_samp_chunk_sums(const double *__restrict__ p, int64_t N, double dN, int64_t nch, double *__restrict__ sumD,
                  double *__restrict__ sumE)
{ // N: entries of p (this rank's slice when the distribution is sharded); dN: bins of the WHOLE distribution
    __shared__ double sm[SAMP_WARPS][32][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t q0 = ((int64_t)blockIdx.x * SAMP_WARPS + w) * 32; // first chunk of this warp
    if (q0 >= nch) return;                                           // whole warp
    const int64_t q = q0 + lane; // the chunk this lane sums
    double d = 0.0, e = 0.0;
    for (int t = 0; t < SAMP_CH / 32; t++) {
#pragma unroll 8
        for (int row = 0; row < 32; row++) {
            const int64_t qr = q0 + row;
            const int64_t i = N - 1 - qr * SAMP_CH - 32 * (int64_t)t - lane; // descending within the chunk
            sm[w][row][lane] = (qr < nch && i >= 0) ? dN * p[i] : -1.0;        // -1: past the end of the vector
        }
        __syncwarp();
        if (q < nch) {
#pragma unroll 8
            for (int c = 0; c < 32; c++) {
                const double f = sm[w][lane][c];
                if (f >= 0.0) {
                    if (f < 1.0) d += 1.0 - f;
                    else e += f - 1.0;
                }
            }
        }
        __syncwarp();
    }
    if (q < nch) {
        sumD[q] = d;
        sumE[q] = e;
    }
}

// Exclusive scan of the chunk sums in a fixed sequential order: 1024 segments of `per` consecutive chunks; segment
// totals and the running values inside a segment are plain sequential sums (again part of the arithmetic), the 1024
// totals are scanned by one thread.  Segments are walked by one lane each with the same warp transposition, 32 segments
// per warp, one warp per CTA (a single CTA of 1024 threads with strided loads took 4.2 ms at 2^20 chunks).
__global__ void __launch_bounds__(32)
k_samp_seg_totals(const double *__restrict__ sumD, const double *__restrict__ sumE, int64_t nch, int64_t per,
                  double *__restrict__ segD, double *__restrict__ segE)
{
    __shared__ double smD[32][33], smE[32][33];
    const int lane = threadIdx.x;
    const int64_t seg0 = (int64_t)blockIdx.x * 32;
    int64_t q0 = (seg0 + lane) * per, q1 = q0 + per;
    if (q0 > nch) q0 = nch;
    if (q1 > nch) q1 = nch;
    double d = 0.0, e = 0.0;
    for (int64_t t = 0; t < per; t += 32) {
        for (int row = 0; row < 32; row++) {
            int64_t r0 = (seg0 + row) * per, r1 = r0 + per;
            if (r1 > nch) r1 = nch;
            const int64_t i = r0 + t + lane;
            const bool ok = i < r1;
            smD[row][lane] = ok ? sumD[i] : 0.0;
            smE[row][lane] = ok ? sumE[i] : 0.0;
        }
        __syncwarp();
        for (int c = 0; c < 32; c++)
            if (q0 + t + c < q1) {
                d += smD[lane][c];
                e += smE[lane][c];
            }
        __syncwarp();
    }
    segD[seg0 + lane] = d;
    segE[seg0 + lane] = e;
}

__global__ void k_samp_seg_scan(double *__restrict__ segD, double *__restrict__ segE, double *__restrict__ totD,
                                double *__restrict__ totE)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    double rd = 0.0, re = 0.0;
    for (int i = 0; i < 1024; i++) {
        const double td = segD[i], te = segE[i];
        segD[i] = rd;
        segE[i] = re;
        rd += td;
        re += te;
    }
    *totD = rd;
    *totE = re;
}

__global__ void __launch_bounds__(32)
k_samp_seg_write(const double *__restrict__ sumD, const double *__restrict__ sumE, int64_t nch, int64_t per,
                 const double *__restrict__ segD, const double *__restrict__ segE, double *__restrict__ cumD,
                 double *__restrict__ cumE)
{
    __shared__ double smD[32][33], smE[32][33];
    const int lane = threadIdx.x;
    const int64_t seg0 = (int64_t)blockIdx.x * 32;
    int64_t q0 = (seg0 + lane) * per, q1 = q0 + per;
    if (q0 > nch) q0 = nch;
    if (q1 > nch) q1 = nch;
    const double bd = segD[seg0 + lane], be = segE[seg0 + lane];
    double d = 0.0, e = 0.0;
    for (int64_t t = 0; t < per; t += 32) {
        for (int row = 0; row < 32; row++) {
            int64_t r0 = (seg0 + row) * per, r1 = r0 + per;
            if (r1 > nch) r1 = nch;
            const int64_t i = r0 + t + lane;
            const bool ok = i < r1;
            smD[row][lane] = ok ? sumD[i] : 0.0;
            smE[row][lane] = ok ? sumE[i] : 0.0;
        }
        __syncwarp();
        for (int c = 0; c < 32; c++)
            if (q0 + t + c < q1) { // exclusive: the value BEFORE adding this chunk
                const double xd = smD[lane][c], xe = smE[lane][c];
                smD[lane][c] = bd + d;
                smE[lane][c] = be + e;
                d += xd;
                e += xe;
            }
        __syncwarp();
        for (int row = 0; row < 32; row++) {
            int64_t r0 = (seg0 + row) * per, r1 = r0 + per;
            if (r1 > nch) r1 = nch;
            const int64_t i = r0 + t + lane;
            if (i < r1) {
                cumD[i] = smD[row][lane];
                cumE[i] = smE[row][lane];
            }
        }
        __syncwarp();
    }
}

static int sampler_finish(double *scratch, int64_t nbins, cudaStream_t st)
{
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    double *cumD = scratch, *cumE = scratch + (nch + 1);
    double *sumD = scratch + 2 * (nch + 1), *sumE = scratch + 3 * (nch + 1);
    double *segD = scratch + 4 * (nch + 1), *segE = segD + 1024;
    const int64_t per = (nch + 1023) / 1024;
    k_samp_seg_totals<<<32, 32, 0, st>>>(sumD, sumE, nch, per, segD, segE);
    k_samp_seg_scan<<<1, 32, 0, st>>>(segD, segE, cumD + nch, cumE + nch);
    k_samp_seg_write<<<32, 32, 0, st>>>(sumD, sumE, nch, per, segD, segE, cumD, cumE);
    return (int)cudaGetLastError();
}

extern "C" int b200k_sampler_prepare(const double *probs, int64_t nbins, double *scratch,
                                     void *stream)
{
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    double *sumD = scratch + 2 * (nch + 1), *sumE = scratch + 3 * (nch + 1);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t warps = (nch + 31) / 32;
    k_samp_chunk_sums<<<(unsigned)((warps + SAMP_WARPS - 1) / SAMP_WARPS), 32 * SAMP_WARPS, 0, st>>>(probs, nbins, (double)nbins, nch, sumD, sumE);
    return sampler_finish(scratch, nbins, st);
}

// ---- the same sampler over a distribution SHARDED by its top index bits ------------------------------------------
// Chunk sums are independent per chunk, so each rank sums the chunks of its own slice (scaled by the bins of the WHOLE
// distribution); the host gathers them into the scratch of every rank in global descending order (the slice of rank
// r holds global chunks (W-1-r) * nch_local ...), after which the segment totals / scan / exclusive prefixes are the
// very same kernels on the very same numbers as on one GPU: the sampled stream is bit-identical.
extern "C" int b200k_sampler_chunk_sums(const double *probs_slice, int64_t nbins_slice, int64_t nbins_total, double *sumD,
                                        double *sumE, void *stream)
{
    if (nbins_slice < SAMP_CH || nbins_slice % SAMP_CH) return (int)cudaErrorInvalidValue; // slices hold whole chunks
    const int64_t nch = nbins_slice / SAMP_CH;
    const int64_t warps = (nch + 31) / 32;
    k_samp_chunk_sums<<<(unsigned)((warps + SAMP_WARPS - 1) / SAMP_WARPS), 32 * SAMP_WARPS, 0, (cudaStream_t)stream>>>(
        probs_slice, nbins_slice, (double)nbins_total, nch, sumD, sumE);
    return (int)cudaGetLastError();
}
extern "C" void b200k_sampler_sum_offsets(int64_t nbins, int64_t *off_sumD, int64_t *off_sumE)
{
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    *off_sumD = 2 * (nch + 1);
    *off_sumE = 3 * (nch + 1);
}
extern "C" int b200k_sampler_finish(double *scratch, int64_t nbins_total, void *stream)
{
    return sampler_finish(scratch, nbins_total, (cudaStream_t)stream);
}

// first chunk q in [0, nch) with cum[q+1] >= x (strict=false) or cum[q+1] > x (strict=true);
// returns nch if none.
__device__ __forceinline__ int64_t samp_find_chunk(const double *cum, int64_t nch, double x, bool strict)
{
    int64_t lo = 0, hi = nch; // answer in [lo, hi]
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        const double v = cum[mid + 1];
        const bool ok = strict ? (v > x) : (v >= x);
        if (ok) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}

struct SampFlat { // the whole distribution in one buffer
    const double *p;
    __device__ __forceinline__ double operator[](int64_t i) const { return p[i]; }
};
struct SampSharded { // slice r = entries [r << nl, (r + 1) << nl), reachable through peer memory
    const double *s[B200K_SAMPLER_MAX_SHARDS];
    int nl;
    __device__ __forceinline__ double operator[](int64_t i) const { return s[i >> nl][i & ((int64_t(1) << nl) - 1)]; }
};
template <class Acc> __global__ void __launch_bounds__(128)
k_samp_draw(const Acc p, int64_t N, int64_t nch, const double *__restrict__ cumD,
            const double *__restrict__ cumE, const double *__restrict__ u, int64_t nshots,
            int64_t *__restrict__ out)
{
    const int64_t sidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (sidx >= nshots) return;
    const double dN = (double)N;
    const double u1 = u[2 * sidx], u2 = u[2 * sidx + 1];
    int64_t idx = (int64_t)(u1 * dN);
    if (idx >= N) idx = N - 1;
    const double f = dN * p[idx];
    const int64_t q = (N - 1 - idx) / SAMP_CH;
    const int64_t chunk_hi = N - 1 - q * SAMP_CH;
    int64_t result = idx;
    if (f < 1.0) {
        // under item: accepted iff u2 < first = f
        if (u2 >= f) {
            // X = exclusive prefix of deficits before idx
            double x = 0.0;
            for (int64_t i = chunk_hi; i > idx; i--) {
                const double fi = dN * p[i];
                if (fi < 1.0) x += 1.0 - fi;
            }
            x = cumD[q] + x;
            // partner = first over item m (descending) with E_m >= x
            int64_t qq = samp_find_chunk(cumE, nch, x, false);
            bool found = false;
            while (qq < nch && !found) {
                const int64_t hi = N - 1 - qq * SAMP_CH;
                int64_t lo = hi - SAMP_CH + 1;
                if (lo < 0) lo = 0;
                double e = 0.0;
                for (int64_t i = hi; i >= lo; i--) {
                    const double fi = dN * p[i];
                    if (fi >= 1.0) {
                        e += fi - 1.0;
                        if (cumE[qq] + e >= x) {
                            result = i;
                            found = true;
                            break;
                        }
                    }
                }
                qq++;
            }
        }
    }
    else {
        // over item: E_m inclusive
        double e = 0.0;
        for (int64_t i = chunk_hi; i >= idx; i--) {
            const double fi = dN * p[i];
            if (fi >= 1.0) e += fi - 1.0;
        }
        const double Em = cumE[q] + e;
        // first under item with D > Em
        int64_t qq = samp_find_chunk(cumD, nch, Em, true);
        bool found = false;
        double Dstar = 0.0;
        while (qq < nch && !found) {
            const int64_t hi = N - 1 - qq * SAMP_CH;
            int64_t lo = hi - SAMP_CH + 1;
            if (lo < 0) lo = 0;
            double d = 0.0;
            for (int64_t i = hi; i >= lo; i--) {
                const double fi = dN * p[i];
                if (fi < 1.0) {
                    d += 1.0 - fi;
                    if (cumD[qq] + d > Em) {
                        Dstar = cumD[qq] + d;
                        found = true;
                        break;
                    }
                }
            }
            qq++;
        }
        if (found) {
            const double first = 1.0 - (Dstar - Em);
            if (u2 >= first) {
                // partner = next over item below idx
                for (int64_t i = idx - 1; i >= 0; i--) {
                    if (dN * p[i] >= 1.0) {
                        result = i;
                        break;
                    }
                }
            }
        }
    }
    out[sidx] = result;
}

extern "C" int b200k_sampler_draw(const double *probs, int64_t nbins, const double *scratch,
                                  const double *u, int64_t nshots, int64_t *out_idx, void *stream)
{
    if (nshots <= 0) return 0;
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    const double *cumD = scratch, *cumE = scratch + (nch + 1);
    k_samp_draw<SampFlat><<<(unsigned)((nshots + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        SampFlat{probs}, nbins, nch, cumD, cumE, u, nshots, out_idx);
    return (int)cudaGetLastError();
}

extern "C" int b200k_sampler_draw_sharded(const double *const *slices, int nslices, int n_local, const double *scratch,
                                          const double *u, int64_t nshots, int64_t *out_idx, void *stream)
{
    if (nshots <= 0) return 0;
    if (nslices < 1 || nslices > B200K_SAMPLER_MAX_SHARDS || (nslices & (nslices - 1)) || n_local < 10 || n_local > 40)
        return (int)cudaErrorInvalidValue;
    const int64_t nbins = (int64_t)nslices << n_local;
    const int64_t nch = (nbins + SAMP_CH - 1) / SAMP_CH;
    const double *cumD = scratch, *cumE = scratch + (nch + 1);
    SampSharded A;
    memset(&A, 0, sizeof(A));
    for (int r = 0; r < nslices; r++) A.s[r] = slices[r];
    A.nl = n_local;
    k_samp_draw<SampSharded><<<(unsigned)((nshots + 127) / 128), 128, 0, (cudaStream_t)stream>>>(A, nbins, nch, cumD, cumE, u,
                                                                                              nshots, out_idx);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Batched bilinear forms <phi| P_g |lam> for up to B200K_BATCH_GENS Pauli-word generators in ONE read
// pass over the two vectors (adjoint gradient: every parameter whose generator commutes with the
// not-yet-applied adjoint gates is evaluated on the same pair of vectors).  Tile geometry as in the
// gate passes: a tile of phi and the same tile of lam are staged in shared memory; a generator whose
// X/Y bits all lie inside the tile pairs amplitude i with i ^ x inside the tile; Z factors, controls
// and rank bits outside the tile enter through the tile's base index.  Per (generator, tile) every
// warp reduces with shuffles into its own shared-memory slot; per-CTA partials go to `partial`, a
// second kernel sums them in fixed order (run-to-run reproducible) and applies the i^ny factor.
// =================================================================================================
__device__ __forceinline__ void kw_cp_async16(void *smem_dst, const void *gmem_src)
{
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void kw_cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void kw_cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

struct BatchParams {
    B200k2Header h;
    B200kBatchGen g[B200K_BATCH_GENS];
    int ngen;
};

// What each generator costs is issue + shared-memory bandwidth, not HBM (2^26 amplitudes x up to 32 generators per
// 1 GiB read), so the kernel is built around the instruction count per (amplitude, generator):
//  * every caller needs ONE real number per generator (Expval the real part of <phi|P|lam>, the adjoint gradient the
//    imaginary part), which after the i^ny phase is either Re or Im of the raw sum: a batch is uniform in that choice
//    (template RC), tr = a . b, ti = (i a) . b, two DFMA per pair;
//  * the thread's own amplitudes of phi stay in registers across all generators of the tile, only the partner
//    lam[e ^ x] comes from shared memory (one LDS.128);
//  * one register accumulator per generator lives across ALL tiles of the CTA: the warp-shuffle tree runs once per
//    kernel instead of once per (generator, tile);
//  * the Z-parity sign costs no POPC per pair (POPC issues at 16 / clk / SM): parity((e ^ x) & z) splits into a part
//    shared by the thread's 8 elements (its tid bits, the tile's outside bits, parity(x & z)) and a host-made 8-bit
//    table over the elements' bits 8..10 (upper half of the zin field), XORed into the sign bit of phi's amplitude;
//  * tiles are double-buffered with cp.async (lam: two buffers; phi: one, copied to registers first), so the next tile
//    is in flight while the generators of the current one are summed.  A C2 RX layer packs only 7 generators per
//    launch: without the prefetch such a launch took 5.6 ms at 28 qubits against 1.5 ms of HBM time.
// (Round 1's version reloaded phi per generator, computed both components and reduced per generator and tile:
// ~0.2 ms per word at 26 qubits, the cost of a separate read pass per word.)
// GENERIC = 0: full 2^11 tiles and no generator with a control inside the tile (the common case: no bounds or control
// predicate per element); GENERIC = 1: anything else.
template <int RC, int GENERIC> __global__ void __launch_bounds__(256, 2)
k_bilinear_batch(const cplx *__restrict__ phi, const cplx *__restrict__ lam, const __grid_constant__ BatchParams P,
                 double *__restrict__ partial)
{
    extern __shared__ cplx sT[]; // lam tile x 2, then the phi tile (when phi != lam)
    __shared__ double wacc[B200K_BATCH_GENS][8];
    __shared__ unsigned long long koff[8];
    const int t = P.h.t, c = P.h.c;
    const uint32_t nelem = 1u << t, tid = threadIdx.x;
    // global offset of tile-local index e = tid + 256 k: the tid part is hoisted, the k part (tile bits 8..10) tabulated
    auto pos_of = [&](int bpos) -> int { return bpos < c ? bpos : (int)P.h.high_pos[bpos - c]; };
    uint64_t off_tid = 0;
    for (int bpos = 0; bpos < 8 && bpos < t; bpos++) off_tid |= (uint64_t)((tid >> bpos) & 1u) << pos_of(bpos);
    if (tid < 8) {
        uint64_t o = 0;
        for (int j = 0; j < 3; j++)
            if (8 + j < t && ((tid >> j) & 1u)) o |= 1ull << pos_of(8 + j);
        koff[tid] = o;
    }
    double acc[B200K_BATCH_GENS];
#pragma unroll
    for (int g = 0; g < B200K_BATCH_GENS; g++) acc[g] = 0.0;
    const bool same = phi == lam;
    cplx *sPhi = sT + ((size_t)2 << t);
    auto base_of = [&](uint64_t tile) {
        uint64_t base = tile << c;
        for (int hh = 0; hh < P.h.nhigh; hh++) {
            const int p = P.h.high_pos[hh];
            base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
        }
        return base;
    };
    auto fetch = [&](uint64_t tile, cplx *lbuf) {
        const uint64_t b = base_of(tile) | off_tid;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t e = tid + 256u * k;
            if (!GENERIC || e < nelem) {
                kw_cp_async16(&lbuf[e], &lam[b | koff[k]]);
                if (!same) kw_cp_async16(&sPhi[e], &phi[b | koff[k]]);
            }
        }
    };
    __syncthreads(); // koff
    if (blockIdx.x < P.h.ntiles) fetch(blockIdx.x, sT);
    kw_cp_async_commit();
    int cur = 0;
    for (uint64_t tile = blockIdx.x; tile < P.h.ntiles; tile += gridDim.x, cur ^= 1) {
        cplx *sB = sT + ((size_t)cur << t);
        kw_cp_async_wait<0>();
        __syncthreads(); // this tile has landed; every thread is done with the previous tile's buffers
        const uint64_t obase = base_of(tile) | P.h.rank_base;
        cplx av[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t e = tid + 256u * k;
            av[k] = make_double2(0.0, 0.0);
            if (!GENERIC || e < nelem) {
                const cplx a = same ? sB[e] : sPhi[e];
                av[k] = RC ? make_double2(-a.y, a.x) : a; // (i a) . b = Im(conj(a) b);  a . b = Re(conj(a) b)
            }
        }
        __syncthreads(); // phi's buffer is free again
        const uint64_t next = tile + gridDim.x;
        if (next < P.h.ntiles) fetch(next, sT + ((size_t)(cur ^ 1) << t));
        kw_cp_async_commit();
#pragma unroll
        for (int g = 0; g < B200K_BATCH_GENS; g++) {
            if (g < P.ngen) {
                const B200kBatchGen &G = P.g[g];
                if ((obase & G.co_m) == G.co_v) { // uniform per tile
                    // bit k of M = parity of the Z factors at the partner of element k
                    const uint32_t common = (__popc(tid & G.zin) + (G.zin >> 24) + __popcll(obase & G.zo)) & 1u;
                    const uint32_t M = (0u - common) ^ ((G.zin >> 16) & 0xffu);
                    // partner of e = tid | k << 8 is (tid ^ x) ^ (k << 8): one XOR per thread and generator, then a
                    // constant XOR per element; the sign goes onto the product (in place, no register-pair copies)
                    const uint32_t A = tid ^ G.xin;
                    double s0 = 0.0, s1 = 0.0;
#pragma unroll
                    for (int k = 0; k < 8; k++) {
                        const uint32_t e = tid + 256u * k;
                        if (!GENERIC || (e < nelem && (e & G.cin_m) == G.cin_v)) {
                            const cplx b = sB[A ^ ((uint32_t)k << 8)];
                            double v = fma(av[k].y, b.y, av[k].x * b.x);
                            v = __hiloint2double(__double2hiint(v) ^ (int)((M << (31 - k)) & 0x80000000u), __double2loint(v));
                            if (k & 1) s1 += v;
                            else s0 += v;
                        }
                    }
                    acc[g] += s0 + s1;
                }
            }
        }
    }
    kw_cp_async_wait<0>();
    // one reduction per kernel: warp tree, then the eight warps in fixed order
#pragma unroll
    for (int g = 0; g < B200K_BATCH_GENS; g++) {
        if (g < P.ngen) {
            double v = acc[g];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
            if ((tid & 31u) == 0) wacc[g][tid >> 5] = v;
        }
    }
    __syncthreads();
    for (int g = tid; g < P.ngen; g += blockDim.x) {
        double s = 0.0;
        for (int w = 0; w < 8; w++) s += wacc[g][w];
        partial[(size_t)blockIdx.x * B200K_BATCH_GENS + g] = s;
    }
}

struct BatchOut {
    double *out[B200K_BATCH_GENS];
    int8_t slot[B200K_BATCH_GENS]; // 0: the sum is the real part of the result, 1: the imaginary part
    int8_t neg[B200K_BATCH_GENS];  // result = -sum
};
__global__ void __launch_bounds__(256)
k_bilinear_batch_final(const double *__restrict__ partial, int nblocks, BatchOut O)
{
    const int g = blockIdx.x;
    double v = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 256) v += partial[(size_t)i * B200K_BATCH_GENS + g];
    __shared__ double sh[8];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int i = 0; i < 8; i++) a += sh[i];
        if (O.neg[g]) a = -a;
        O.out[g][O.slot[g]] = a;
        O.out[g][O.slot[g] ^ 1] = 0.0; // the component the caller did not ask for is not computed
    }
}

#define BATCH_MAX_BLOCKS (148 * 4)
extern "C" int64_t b200k_bilinear_batch_scratch_doubles(void) { return (int64_t)BATCH_MAX_BLOCKS * B200K_BATCH_GENS; }

// Diagonal generators (Z words: RZ, MultiRZ, IsingZZ, CRZ's Z factor, PhaseShift-type projectors): <phi|P_g|lam> =
// sum_i sign_g(i) conj(phi_i) lam_i, so the one real number r_i every caller needs (Re or Im of conj(phi_i) lam_i) is
// computed ONCE per amplitude and each generator costs a parity, a sign flip and an add -- no tile, no shared memory,
// a plain streaming pass over the two vectors.  (Through k_bilinear_batch a diagonal generator cost the same ~0.1 ms
// at 26 qubits as an X/Y one: half of C4's and 60 % of C2's parameters.)
// The parity of (i & z_g) for all 32 generators at once is a GF(2) matrix-vector product: S(i) = XOR over the set bits b
// of i of the column word C_b (bit g of C_b = bit b of z_g).  POPC runs at 16 / clk / SM, and 2 x 32 of them per amplitude
// made the first version of this kernel POPC-bound (5.4 ms per launch at 28 qubits against 1.5 ms of HBM time); here
// S(i) = S(tid) ^ T0[bits 8..17] ^ T1[bits 18..27] ^ T2[bits 28..37] ^ S(rank bits) with three 1024-entry tables built
// per CTA, and generator g's sign is bit g of S(i) shifted into the sign bit of r.
#define DIAG_COLS 38
struct DiagBatchParams {
    uint64_t cm[B200K_BATCH_GENS], cv[B200K_BATCH_GENS];
    uint32_t col[DIAG_COLS]; // column words of the local index bits
    uint32_t s_const;        // contribution of the rank bits (and of any bit >= DIAG_COLS)
    uint64_t rank_base;
    int ngen;
};
template <int RC, int HASCTL> __global__ void __launch_bounds__(256, 3)
k_bilinear_diag_batch(const cplx *__restrict__ phi, const cplx *__restrict__ lam, uint64_t N,
                      const __grid_constant__ DiagBatchParams P, double *__restrict__ partial)
{
    __shared__ double wacc[B200K_BATCH_GENS][8];
    __shared__ uint32_t T[3][1024];
    const uint32_t tid = threadIdx.x;
    for (uint32_t idx = tid; idx < 3 * 1024; idx += 256) {
        const uint32_t tb = idx >> 10, v = idx & 1023u;
        uint32_t sgn = 0;
        for (int b = 0; b < 10; b++)
            if ((v >> b) & 1u) sgn ^= P.col[8 + 10 * tb + b];
        T[tb][v] = sgn;
    }
    uint32_t s_thread = P.s_const;
    for (int b = 0; b < 8; b++)
        if ((tid >> b) & 1u) s_thread ^= P.col[b];
    __syncthreads();
    double acc[B200K_BATCH_GENS];
#pragma unroll
    for (int g = 0; g < B200K_BATCH_GENS; g++) acc[g] = 0.0;
    const bool same = phi == lam;
    const uint64_t stride = (uint64_t)gridDim.x * 256u; // a multiple of 256: the low 8 index bits of a thread are its tid
    for (uint64_t i = (uint64_t)blockIdx.x * 256u + tid; i < N; i += stride) {
        const cplx b = lam[i];
        const cplx a = same ? b : phi[i];
        const double r = RC ? fma(a.x, b.y, -(a.y * b.x)) : fma(a.x, b.x, a.y * b.y); // Im / Re of conj(a) b
        const uint64_t q = i >> 8;
        const uint32_t S = s_thread ^ T[0][q & 1023u] ^ T[1][(q >> 10) & 1023u] ^ T[2][(q >> 20) & 1023u];
        const uint64_t gi = i | P.rank_base;
#pragma unroll
        for (int g = 0; g < B200K_BATCH_GENS; g++) {
            if (g < P.ngen && (!HASCTL || (gi & P.cm[g]) == P.cv[g])) {
                const long long sign = (long long)((S << (31 - g)) & 0x80000000u) << 32;
                acc[g] += __longlong_as_double(__double_as_longlong(r) ^ sign);
            }
        }
    }
#pragma unroll
    for (int g = 0; g < B200K_BATCH_GENS; g++) {
        if (g < P.ngen) {
            double v = acc[g];
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
            if ((tid & 31u) == 0) wacc[g][tid >> 5] = v;
        }
    }
    __syncthreads();
    for (int g = tid; g < P.ngen; g += blockDim.x) {
        double s = 0.0;
        for (int w = 0; w < 8; w++) s += wacc[g][w];
        partial[(size_t)blockIdx.x * B200K_BATCH_GENS + g] = s;
    }
}

extern "C" int b200k_bilinear_diag_batch(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                                         const uint64_t *zmasks, const uint64_t *cmasks, const uint64_t *cvals,
                                         const int *nys, int want, int ngen, double *const *outs, double *scratch,
                                         void *stream)
{
    if (ngen < 1 || ngen > B200K_BATCH_GENS || (want != B200K_WANT_REAL && want != B200K_WANT_IMAG))
        return (int)cudaErrorInvalidValue;
    const bool imag = want == B200K_WANT_IMAG;
    int rc = -1;
    DiagBatchParams P;
    memset(&P, 0, sizeof(P));
    P.rank_base = rank_base;
    P.ngen = ngen;
    BatchOut O;
    memset(&O, 0, sizeof(O));
    for (int g = 0; g < ngen; g++) {
        const int ny = nys[g] & 3; // a diagonal word has no Y factor; the phase argument is honoured all the same
        const int my_rc = (ny & 1) ^ (imag ? 1 : 0);
        if (rc < 0) rc = my_rc;
        if (rc != my_rc) return (int)cudaErrorInvalidValue;
        for (int b = 0; b < 64; b++) {
            if (!((zmasks[g] >> b) & 1ull)) continue;
            if (b < DIAG_COLS && b < n_local) P.col[b] |= 1u << g;
            else if ((rank_base >> b) & 1ull) P.s_const ^= 1u << g; // a bit the local index does not reach
        }
        P.cm[g] = cmasks[g];
        P.cv[g] = cvals[g];
        O.out[g] = outs[g];
        O.slot[g] = imag ? 1 : 0;
        O.neg[g] = imag ? (ny == 2 || ny == 3) : (ny == 1 || ny == 2);
    }
    if (n_local > DIAG_COLS) return (int)cudaErrorInvalidValue;
    const uint64_t N = 1ull << n_local;
    uint64_t grid = (N + 255) / 256;
    if (grid > 148 * 3) grid = 148 * 3; // three resident CTAs per SM (registers): one wave
    bool ctl = false;
    for (int g = 0; g < ngen; g++) ctl = ctl || cmasks[g] != 0;
    const cplx *ph = (const cplx *)phi, *la = (const cplx *)lam;
    cudaStream_t st = (cudaStream_t)stream;
    if (rc && ctl) k_bilinear_diag_batch<1, 1><<<(unsigned)grid, 256, 0, st>>>(ph, la, N, P, scratch);
    else if (rc) k_bilinear_diag_batch<1, 0><<<(unsigned)grid, 256, 0, st>>>(ph, la, N, P, scratch);
    else if (ctl) k_bilinear_diag_batch<0, 1><<<(unsigned)grid, 256, 0, st>>>(ph, la, N, P, scratch);
    else k_bilinear_diag_batch<0, 0><<<(unsigned)grid, 256, 0, st>>>(ph, la, N, P, scratch);
    k_bilinear_batch_final<<<ngen, 256, 0, (cudaStream_t)stream>>>(scratch, (int)grid, O);
    return (int)cudaGetLastError();
}


extern "C" int b200k_bilinear_batch(const double *phi, const double *lam, int n_local, const B200k2Header *hdr,
                                    const B200kBatchGen *gens, int ngen, double *const *outs, double *scratch,
                                    void *stream)
{
    if (ngen < 1 || ngen > B200K_BATCH_GENS || hdr->t > 11 || hdr->t > n_local || hdr->c + hdr->nhigh != hdr->t ||
        hdr->ntiles != (1ull << (n_local - hdr->t)))
        return (int)cudaErrorInvalidValue;
    // result = i^ny (tr + i ti): which raw component carries the wanted part, and with which sign
    //   real part: ny 0: tr, 1: -ti, 2: -tr, 3: ti        imaginary part: ny 0: ti, 1: tr, 2: -ti, 3: -tr
    int rc = -1;
    bool ctl = false;
    BatchParams P;
    memset(&P, 0, sizeof(P));
    P.h = *hdr;
    P.ngen = ngen;
    BatchOut O;
    memset(&O, 0, sizeof(O));
    for (int g = 0; g < ngen; g++) {
        const int ny = gens[g].ny & 3, want = gens[g].want;
        if (want != B200K_WANT_REAL && want != B200K_WANT_IMAG) return (int)cudaErrorInvalidValue;
        const bool imag = want == B200K_WANT_IMAG;
        const int my_rc = (ny & 1) ^ (imag ? 1 : 0);
        if (rc < 0) rc = my_rc;
        if (rc != my_rc) return (int)cudaErrorInvalidValue; // a batch is uniform in the raw component it needs
        P.g[g] = gens[g];
        if ((gens[g].xin >> hdr->t) || (gens[g].zin >> hdr->t)) return (int)cudaErrorInvalidValue;
        // sign table of the elements' bits 8..10 (bits 16..23 of zin) and parity(xin & zin) (bit 24): see the kernel
        uint32_t kpar = 0;
        for (uint32_t kk = 0; kk < 8; kk++) kpar |= (uint32_t)(__builtin_popcount(kk & (gens[g].zin >> 8)) & 1) << kk;
        P.g[g].zin = gens[g].zin | (kpar << 16) | ((uint32_t)(__builtin_popcount(gens[g].xin & gens[g].zin) & 1) << 24);
        ctl = ctl || gens[g].cin_m != 0;
        O.out[g] = outs[g];
        O.slot[g] = imag ? 1 : 0;
        O.neg[g] = imag ? (ny == 2 || ny == 3) : (ny == 1 || ny == 2);
    }
    const bool same = phi == lam;
    const size_t smem = (same ? 2 : 3) * (sizeof(cplx) << hdr->t);
    static std::atomic<unsigned long long> attr_mask{0}; // threads of one process may launch concurrently
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    if (!((attr_mask.load(std::memory_order_acquire) >> (cur_dev & 63)) & 1ull)) {
        const int cap = 3 * 16 * 2048;
        cudaError_t e = cudaFuncSetAttribute(k_bilinear_batch<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bilinear_batch<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bilinear_batch<1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bilinear_batch<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
        if (e != cudaSuccess) return (int)e;
        attr_mask.fetch_or(1ull << (cur_dev & 63), std::memory_order_release);
    }
    uint64_t grid = 148 * 2;
    if (grid > hdr->ntiles) grid = hdr->ntiles;
    const cplx *ph = (const cplx *)phi, *la = (const cplx *)lam;
    cudaStream_t st = (cudaStream_t)stream;
    const bool generic = ctl || hdr->t != 11;
    if (rc && generic) k_bilinear_batch<1, 1><<<(unsigned)grid, 256, smem, st>>>(ph, la, P, scratch);
    else if (rc) k_bilinear_batch<1, 0><<<(unsigned)grid, 256, smem, st>>>(ph, la, P, scratch);
    else if (generic) k_bilinear_batch<0, 1><<<(unsigned)grid, 256, smem, st>>>(ph, la, P, scratch);
    else k_bilinear_batch<0, 0><<<(unsigned)grid, 256, smem, st>>>(ph, la, P, scratch);
    k_bilinear_batch_final<<<ngen, 256, 0, (cudaStream_t)stream>>>(scratch, (int)grid, O);
    return (int)cudaGetLastError();
}

// =================================================================================================
// Sum of Pauli words on tiles: sigma = sum_w coef_w P_w in, either written out (phi = H psi of the adjoint method)
// or contracted on the fly with a bra (Expval of a Hamiltonian: no vector is written at all).
// k_pauli_apply_batch gathers in[i ^ x_w] from HBM/L2 once per WORD (C4, 26 qubits, 1000 words: 63 launches of
// 2.2 ms); here the tile is staged once per launch and every word whose X/Y bits lie inside it is a shared-memory
// permutation, so HBM sees 1 read (+ the accumulator) per BIN of up to 96 words.
// Per (amplitude, word): one LDS.128, the Z-parity sign folded into the coefficient's sign bit, 2 DFMA for the usual
// real-coefficient word (i^ny makes the factor purely real or purely imaginary), 4 for a general one.
// =================================================================================================
struct TileWordK { // 32 bytes, coefficient with i^ny and (-1)^popc(xin & zin) folded in
    uint32_t xin, zin; // zin: bits 0..10 the mask, bits 16..23 the sign table of the elements' bits 8..10
    uint64_t zo;
    double cr, ci;
};
struct TileWordsParams {
    B200k2Header h;
    int n, accumulate;
    TileWordK w[B200K_TILE_WORDS];
};

// Persistent CTAs, two tile buffers: the next tile's amplitudes are in flight (cp.async, no registers) while the words
// of the current tile are summed.  (The first version loaded tile after tile with nothing in flight and spent ~1.2 ms
// per launch of ~14 words at 26 qubits, against 0.17 ms of HBM time and ~0.4 ms of shared-memory time.)
template <int DOT, int FULL> __global__ void __launch_bounds__(256, 2)
k_pauli_words_tile(cplx *__restrict__ out, const cplx *__restrict__ in, const cplx *__restrict__ bra,
                   const __grid_constant__ TileWordsParams P, double *__restrict__ partial)
{
    extern __shared__ cplx sT[]; // 2 x 2^t amplitudes
    __shared__ unsigned long long koff[8];
    const int t = P.h.t, c = P.h.c;
    const uint32_t nelem = 1u << t, tid = threadIdx.x;
    auto pos_of = [&](int bpos) -> int { return bpos < c ? bpos : (int)P.h.high_pos[bpos - c]; };
    uint64_t off_tid = 0;
    for (int bpos = 0; bpos < 8 && bpos < t; bpos++) off_tid |= (uint64_t)((tid >> bpos) & 1u) << pos_of(bpos);
    if (tid < 8) {
        uint64_t o = 0;
        for (int j = 0; j < 3; j++)
            if (8 + j < t && ((tid >> j) & 1u)) o |= 1ull << pos_of(8 + j);
        koff[tid] = o;
    }
    auto base_of = [&](uint64_t tile) {
        uint64_t base = tile << c;
        for (int hh = 0; hh < P.h.nhigh; hh++) {
            const int p = P.h.high_pos[hh];
            base = ((base >> p) << (p + 1)) | (base & ((1ull << p) - 1ull));
        }
        return base;
    };
    auto fetch = [&](uint64_t tile, cplx *buf) {
        const uint64_t b = base_of(tile) | off_tid;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint32_t e = tid + 256u * k;
            if (FULL || e < nelem) kw_cp_async16(&buf[e], &in[b | koff[k]]);
        }
    };
    double dot[2] = {0.0, 0.0};
    const bool same = bra == in;
    __syncthreads(); // koff
    if (blockIdx.x < P.h.ntiles) fetch(blockIdx.x, sT);
    kw_cp_async_commit();
    int cur = 0;
    for (uint64_t tile = blockIdx.x; tile < P.h.ntiles; tile += gridDim.x, cur ^= 1) {
        cplx *sB = sT + ((size_t)cur << t);
        __syncthreads(); // every thread is done with the buffer the prefetch below overwrites
        const uint64_t next = tile + gridDim.x;
        if (next < P.h.ntiles) fetch(next, sT + ((size_t)(cur ^ 1) << t));
        kw_cp_async_commit();
        kw_cp_async_wait<1>(); // this tile's group has landed; the next one may still be in flight
        __syncthreads();
        const uint64_t base = base_of(tile);
        const uint64_t obase = base | P.h.rank_base;
        cplx acc[8];
#pragma unroll
        for (int k = 0; k < 8; k++) { // accumulating into out: its old values are loaded first, behind the word loop
            const uint32_t e = tid + 256u * k;
            acc[k] = (!DOT && P.accumulate && (FULL || e < nelem)) ? out[base | off_tid | koff[k]] : make_double2(0.0, 0.0);
        }
        for (int w = 0; w < P.n; w++) {
            const TileWordK &W = P.w[w];
            // sign of element e = tid + 256 k: parity(e & zin) ^ parity(obase & zo).  The tid part and the tile part are
            // shared by the thread's 8 elements; the k part is a host-made 8-bit table (bit k = parity(k & (zin >> 8)))
            // in the upper half of the zin field.  M bit k = sign of element k.
            const uint32_t common = (__popc(tid & W.zin) + __popcll(obase & W.zo)) & 1u;
            const uint32_t M = (0u - common) ^ (W.zin >> 16);
            const uint32_t x16 = W.xin;
            if (W.ci == 0.0) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t e = tid + 256u * k;
                    if (FULL || e < nelem) {
                        const cplx b = sB[e ^ x16];
                        const long long sg = (long long)((M << (31 - k)) & 0x80000000u) << 32;
                        const double cr = __longlong_as_double(__double_as_longlong(W.cr) ^ sg);
                        acc[k].x = fma(cr, b.x, acc[k].x);
                        acc[k].y = fma(cr, b.y, acc[k].y);
                    }
                }
            }
            else if (W.cr == 0.0) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t e = tid + 256u * k;
                    if (FULL || e < nelem) {
                        const cplx b = sB[e ^ x16];
                        const long long sg = (long long)((M << (31 - k)) & 0x80000000u) << 32;
                        const double ci = __longlong_as_double(__double_as_longlong(W.ci) ^ sg);
                        acc[k].x = fma(-ci, b.y, acc[k].x);
                        acc[k].y = fma(ci, b.x, acc[k].y);
                    }
                }
            }
            else {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const uint32_t e = tid + 256u * k;
                    if (FULL || e < nelem) {
                        const cplx b = sB[e ^ x16];
                        const long long sg = (long long)((M << (31 - k)) & 0x80000000u) << 32;
                        const double cr = __longlong_as_double(__double_as_longlong(W.cr) ^ sg);
                        const double ci = __longlong_as_double(__double_as_longlong(W.ci) ^ sg);
                        acc[k].x = fma(cr, b.x, fma(-ci, b.y, acc[k].x));
                        acc[k].y = fma(cr, b.y, fma(ci, b.x, acc[k].y));
                    }
                }
            }
        }
        if (DOT) {
#pragma unroll
            for (int k = 0; k < 8; k++) { // conj(bra) * acc
                const uint32_t e = tid + 256u * k;
                if (FULL || e < nelem) {
                    const cplx a = same ? sB[e] : bra[base | off_tid | koff[k]];
                    dot[0] += fma(a.x, acc[k].x, a.y * acc[k].y);
                    dot[1] += fma(a.x, acc[k].y, -(a.y * acc[k].x));
                }
            }
        }
        else {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const uint32_t e = tid + 256u * k;
                if (FULL || e < nelem) {
                    out[base | off_tid | koff[k]] = acc[k];
                }
            }
        }
    }
    kw_cp_async_wait<0>();
    if (DOT) block_reduce_store<2>(dot, partial);
}

extern "C" int b200k_pauli_words_tile(double *out, const double *in, const double *bra, int n_local,
                                      const B200k2Header *hdr, const B200kTileWord *words, int nwords, int accumulate,
                                      double *scratch, double *dot_out, void *stream)
{
    if (nwords < 1 || nwords > B200K_TILE_WORDS || hdr->t > 11 || hdr->t > n_local || hdr->c + hdr->nhigh != hdr->t ||
        hdr->ntiles != (1ull << (n_local - hdr->t)) || !in || out == in)
        return (int)cudaErrorInvalidValue;
    if (!out && (!bra || !scratch || !dot_out)) return (int)cudaErrorInvalidValue;
    static_assert(sizeof(TileWordsParams) <= 4000, "kernel parameter space");
    TileWordsParams P;
    memset(&P, 0, sizeof(P));
    P.h = *hdr;
    P.n = nwords;
    P.accumulate = accumulate;
    for (int w = 0; w < nwords; w++) {
        const B200kTileWord &W = words[w];
        if ((W.xin >> hdr->t) || (W.zin >> hdr->t)) return (int)cudaErrorInvalidValue;
        // coef * i^ny * (-1)^popc(xin & zin): the Z factors act at the partner index j = e ^ xin
        double cr = W.cr, ci = W.ci;
        switch (W.ny & 3) {
        case 1: { const double r = -ci; ci = cr; cr = r; break; }
        case 2: cr = -cr; ci = -ci; break;
        case 3: { const double r = ci; ci = -cr; cr = r; break; }
        default: break;
        }
        if (__builtin_popcount(W.xin & W.zin) & 1) { cr = -cr; ci = -ci; }
        if (cr == 0.0) cr = 0.0; // no negative zeros: the kernel classifies words by exact zero tests
        if (ci == 0.0) ci = 0.0;
        P.w[w].xin = W.xin;
        uint32_t kpar = 0; // bit k = parity(k & (zin >> 8)), k = the element's tile bits 8..10
        for (uint32_t kk = 0; kk < 8; kk++) kpar |= (uint32_t)(__builtin_popcount(kk & (W.zin >> 8)) & 1) << kk;
        P.w[w].zin = W.zin | (kpar << 16);
        P.w[w].zo = W.zo;
        P.w[w].cr = cr;
        P.w[w].ci = ci;
    }
    static std::atomic<unsigned long long> attr_mask{0};
    int cur_dev = 0;
    cudaGetDevice(&cur_dev);
    if (!((attr_mask.load(std::memory_order_acquire) >> (cur_dev & 63)) & 1ull)) {
        cudaError_t e = cudaFuncSetAttribute(k_pauli_words_tile<0, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 16 * 2048);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_pauli_words_tile<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * 16 * 2048);
        if (e != cudaSuccess) return (int)e;
        attr_mask.fetch_or(1ull << (cur_dev & 63), std::memory_order_release);
    }
    const size_t smem = 2 * (sizeof(cplx) << hdr->t); // two tile buffers
    const bool full = hdr->t == 11;
    // persistent CTAs: as many as are resident at once (2 or 3 per SM, by registers and the two tile buffers)
    static std::atomic<int> resident[2] = {{0}, {0}};
    int per_sm = resident[out ? 0 : 1].load(std::memory_order_relaxed);
    if (per_sm == 0 || !full) {
        per_sm = 2;
        if (full) {
            int nb = 0;
            cudaError_t e = out ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_pauli_words_tile<0, 1>, 256, smem)
                                : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_pauli_words_tile<1, 1>, 256, smem);
            if (e == cudaSuccess && nb >= 1) per_sm = nb;
            resident[out ? 0 : 1].store(per_sm, std::memory_order_relaxed);
        }
    }
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cur_dev);
    uint64_t grid = (uint64_t)sms * per_sm;
    if (grid > hdr->ntiles) grid = hdr->ntiles;
    if (grid > RED_BLOCKS) grid = RED_BLOCKS; // the contraction's partials live in the shared reduction scratch
    cudaStream_t st = (cudaStream_t)stream;
    if (out) {
        if (full) k_pauli_words_tile<0, 1><<<(unsigned)grid, 256, smem, st>>>((cplx *)out, (const cplx *)in, nullptr, P, nullptr);
        else k_pauli_words_tile<0, 0><<<(unsigned)grid, 256, smem, st>>>((cplx *)out, (const cplx *)in, nullptr, P, nullptr);
    }
    else {
        if (full) k_pauli_words_tile<1, 1><<<(unsigned)grid, 256, smem, st>>>(nullptr, (const cplx *)in, (const cplx *)bra, P, scratch);
        else k_pauli_words_tile<1, 0><<<(unsigned)grid, 256, smem, st>>>(nullptr, (const cplx *)in, (const cplx *)bra, P, scratch);
        k_reduce_final<2><<<1, RED_THREADS, 0, (cudaStream_t)stream>>>(scratch, (int)grid, dot_out);
    }
    return (int)cudaGetLastError();
}

// =================================================================================================
// FP64 pipe microbenchmark: the second roofline of a fused tile pass.  A pass that fuses ~50 gates issues
// ~10^2 DFMA per amplitude, so it can be bound by the FP64 pipe rather than by HBM; bench.py reports each
// pass time against max(HBM time, DFMA time) with this measured rate as the DFMA denominator.
// 16 independent FMA chains per thread (latency fully hidden), 8 resident CTAs of 256 threads per SM.
// =================================================================================================
__global__ void __launch_bounds__(256)
k_dfma_peak(double *__restrict__ out, int iters, double a, double b)
{
    double v[16];
#pragma unroll
    for (int k = 0; k < 16; k++) v[k] = (double)(threadIdx.x + k);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) v[k] = fma(v[k], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += v[k];
    if (s == 123.456) out[0] = s; // keeps the chains alive; never true for the arguments used
}

extern "C" int b200k_dfma_peak(double *scratch, double *dfma_per_second, void *stream)
{
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t e0, e1;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return (int)cudaGetLastError();
    const int iters = 1 << 14, grid = sms * 8;
    k_dfma_peak<<<grid, 256, 0, st>>>(scratch, 256, 0.999999, 1e-9); // warm-up
    double best = 0.0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0, st);
        k_dfma_peak<<<grid, 256, 0, st>>>(scratch, iters, 0.999999, 1e-9);
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double rate = (double)grid * 256.0 * 16.0 * (double)iters / (ms * 1e-3);
        if (rate > best) best = rate;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *dfma_per_second = best;
    return (int)cudaGetLastError();
}

// =================================================================================================
// MCMC sampling (device kwargs mcmc / num_burnin / kernel_name of lightning.qubit).
// A Metropolis chain over basis states is sequential by nature: ONE thread walks it, reading two amplitudes
// per step from HBM; the proposal stream (site / value or a random rank among the non-zero amplitudes, plus
// one uniform per step) is drawn on the host from the same mt19937 the alias sampler uses, so the chain is
// reproducible and comparable with the CPU restatement.  The state must be in canonical layout (index =
// logical index).  "NonZeroRandom" needs the ascending list of basis states with |amplitude|^2 > eps:
// an ordered stream compaction (thrust::copy_if over a counting iterator).
// =================================================================================================
__global__ void k_mcmc_chain(const cplx *__restrict__ psi, int n, int64_t steps, int64_t burnin, const uint32_t *__restrict__ site,
                             const uint64_t *__restrict__ jump, const int64_t *__restrict__ nz, int64_t nnz,
                             const double *__restrict__ u, int64_t *__restrict__ out)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    uint64_t cur = 0;
    cplx a = psi[0];
    double pc = a.x * a.x + a.y * a.y;
    for (int64_t s = 0; s < steps; s++) {
        uint64_t cand;
        if (nz) cand = (uint64_t)nz[jump[s] % (uint64_t)nnz];
        else {
            const uint32_t w = site[s]; // bit 31: value, low bits: qubit site
            const uint32_t q = w & 0x7fffffffu;
            cand = (cur & ~(1ull << q)) | ((uint64_t)(w >> 31) << q);
        }
        const cplx b = psi[cand];
        const double pn = b.x * b.x + b.y * b.y;
        if (pc == 0.0 || pn > u[s] * pc) {
            cur = cand;
            pc = pn;
        }
        if (s >= burnin) out[s - burnin] = (int64_t)cur;
    }
}

extern "C" int b200k_mcmc_chain(const double *psi, int n, int64_t steps, int64_t burnin, const uint32_t *site,
                                const uint64_t *jump, const int64_t *nz, int64_t nnz, const double *u, int64_t *out,
                                void *stream)
{
    if (steps <= burnin) return 0;
    if (nz && nnz <= 0) return (int)cudaErrorInvalidValue;
    k_mcmc_chain<<<1, 32, 0, (cudaStream_t)stream>>>((const cplx *)psi, n, steps, burnin, site, jump, nz, nnz, u, out);
    return (int)cudaGetLastError();
}

struct NonZeroAmp {
    const cplx *psi;
    double eps;
    __device__ bool operator()(int64_t i) const
    {
        const cplx a = psi[i];
        return a.x * a.x + a.y * a.y > eps;
    }
};

// out (capacity 2^n int64) <- ascending indices with |psi_i|^2 > eps; *count receives how many.  Synchronises.
extern "C" int b200k_nonzero_indices(const double *psi, int n, double eps, int64_t *out, int64_t *count, void *stream)
{
    const int64_t N = int64_t(1) << n;
    NonZeroAmp pred{(const cplx *)psi, eps};
    try {
        auto end = thrust::copy_if(thrust::cuda::par.on((cudaStream_t)stream), thrust::counting_iterator<int64_t>(0),
                                   thrust::counting_iterator<int64_t>(N), out, pred);
        *count = (int64_t)(end - out);
    }
    catch (...) {
        cudaGetLastError();
        return (int)cudaErrorUnknown;
    }
    return (int)cudaStreamSynchronize((cudaStream_t)stream);
}

// =================================================================================================
// Allocation-free variant of b200k_bilinear_pauli (no counterpart in the reference).
// Some callers cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it.  The reduction scratch (the per-block partials the plain entry point takes
// from the caller) lives in a file-scope device array instead: B200K_NOALLOC_SCRATCH_DOUBLES doubles =
// 18 944 bytes, held by every GPU that loads this library until the process exits.
//   * In HOST code the array's name is a host-side stand-in, not a device address.  Its device address is
//     resolved with cudaGetSymbolAddress once per device, with that device current, and cached per ordinal.
//   * Nothing in the array must start at zero: every block of k_bilinear_pauli stores its own partial slot
//     before k_reduce_final (next on the same stream) reads exactly those slots, so the previous call's
//     leftovers are never read.
//   * ALL calls share the array: calls that overlap on the GPU (different streams, or different host threads
//     on the same device without stream ordering between them) race on it.  Order them on one stream.
//   * This function calls no cudaMalloc / cudaFree / allocating library routine.  The first call on a device
//     resolves the symbol (which may load the module lazily): make one call outside graph capture first.
// =================================================================================================
static __device__ __align__(256) double g_noalloc_scratch[B200K_NOALLOC_SCRATCH_DOUBLES];

static void noalloc_err(char *buf, int64_t cap, const char *msg)
{
    if (!buf || cap <= 0) return;
    int64_t i = 0;
    for (; msg[i] && i < cap - 1; i++) buf[i] = msg[i];
    buf[i] = 0;
}

extern "C" int b200k_bilinear_pauli_noalloc(const double *phi, const double *lam, int n_local, uint64_t rank_base,
                                            uint64_t xmask, uint64_t zmask, int ny, uint64_t cmask, uint64_t cval,
                                            double *out, void *stream, char *errbuf, int64_t errcap)
{
    static_assert(B200K_NOALLOC_SCRATCH_DOUBLES >= RED_BLOCKS * RED_MAXVALS, "no-alloc scratch smaller than the reduction grid");
    noalloc_err(errbuf, errcap, "");
    if (!phi || !lam || !out) {
        noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: null input / output pointer");
        return (int)cudaErrorInvalidValue;
    }
    if (n_local < 0 || n_local > B200K_NOALLOC_MAX_LOCAL_QUBITS) {
        noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: n_local outside 0.." B200K_STR(B200K_NOALLOC_MAX_LOCAL_QUBITS)
                                    " (the static scratch serves vectors of at most 2^" B200K_STR(B200K_NOALLOC_MAX_LOCAL_QUBITS)
                                    " amplitudes)");
        return (int)cudaErrorInvalidValue;
    }
    if (xmask >> n_local) {
        noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: xmask has bits outside the local index");
        return (int)cudaErrorInvalidValue;
    }
    const uint64_t work = xmask ? (1ull << (n_local > 0 ? n_local - 1 : 0)) : (1ull << n_local);
    const int nb = red_blocks_for(work);
    if ((int64_t)nb * RED_MAXVALS > (int64_t)B200K_NOALLOC_SCRATCH_DOUBLES) {
        noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: reduction needs more partial slots than the static scratch holds");
        return (int)cudaErrorInvalidValue;
    }
    // device address of the array: once per device, with that device current
    static std::atomic<double *> dev_ptr[64];
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: no current CUDA device (or ordinal >= 64)");
        return (int)(e != cudaSuccess ? e : cudaErrorInvalidDevice);
    }
    double *scratch = dev_ptr[dev].load(std::memory_order_acquire);
    if (!scratch) {
        e = cudaGetSymbolAddress((void **)&scratch, g_noalloc_scratch);
        if (e != cudaSuccess || !scratch) {
            cudaGetLastError();
            noalloc_err(errbuf, errcap, "b200k_bilinear_pauli_noalloc: cudaGetSymbolAddress of the static scratch failed");
            return (int)(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
        }
        dev_ptr[dev].store(scratch, std::memory_order_release); // the same value whichever thread stores it
    }
    k_bilinear_pauli<<<nb, RED_THREADS, 0, (cudaStream_t)stream>>>((const cplx *)phi, (const cplx *)lam, n_local, rank_base,
                                                                   xmask, zmask, ny, cmask, cval, scratch);
    k_reduce_final<2><<<1, RED_THREADS, 0, (cudaStream_t)stream>>>(scratch, nb, out);
    e = cudaGetLastError();
    if (e != cudaSuccess) noalloc_err(errbuf, errcap, cudaGetErrorString(e));
    return (int)e;
}
