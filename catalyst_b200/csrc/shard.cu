// shard.cu — pack / unpack kernels of the global<->local qubit swap.
//
// A swap of m global index bits with m local bits at positions lpos[0..m) sends, to the peer whose
// global-bit pattern is `val`, the sub-block of the shard whose bits at lpos spell `val`, and
// receives that peer's sub-block for the same slots.  Pack gathers such a sub-block (2^(n_local-m)
// amplitudes, addressed by the index with the lpos bits squeezed out) into a contiguous staging
// buffer; unpack scatters a received buffer back into the same slots.  Pure HBM streaming: one
// 128-bit load and store per amplitude; contiguous runs of 2^lpos[0] amplitudes, so the host prefers
// to evict high local bits.
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200sv_kernels.h"

typedef double2 cplx;

struct PackPos {
    uint8_t pos[8];
};

template <bool PACK>
__global__ void __launch_bounds__(256)
k_pack_bits(cplx *__restrict__ staged, cplx *__restrict__ state, PackPos P, int m, uint32_t val,
            uint64_t j0, uint64_t count)
{
    uint64_t deposit = 0;
    for (int i = 0; i < m; i++) deposit |= (uint64_t)((val >> i) & 1u) << P.pos[i];
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < count; j += stride) {
        uint64_t g = j0 + j;
        for (int i = 0; i < m; i++) {
            const int p = P.pos[i];
            g = ((g >> p) << (p + 1)) | (g & ((1ull << p) - 1ull));
        }
        g |= deposit;
        if (PACK) staged[j] = state[g];
        else state[g] = staged[j];
    }
}

static int pack_launch(bool pack, double *staged, double *state, const uint8_t *lpos, int m, uint32_t val,
                       uint64_t j0, uint64_t count, void *stream)
{
    if (m < 0 || m > 8) return (int)cudaErrorInvalidValue;
    PackPos P;
    for (int i = 0; i < 8; i++) P.pos[i] = i < m ? lpos[i] : 0;
    for (int i = 1; i < m; i++)
        if (P.pos[i] <= P.pos[i - 1]) return (int)cudaErrorInvalidValue;
    if (count == 0) return 0;
    uint64_t blocks = (count + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    if (pack)
        k_pack_bits<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((cplx *)staged, (cplx *)state, P, m, val, j0, count);
    else
        k_pack_bits<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((cplx *)staged, (cplx *)state, P, m, val, j0, count);
    return (int)cudaGetLastError();
}

extern "C" int b200k_pack_bits(double *staged, const double *state, const uint8_t *lpos, int m, uint32_t val,
                               uint64_t j0, uint64_t count, void *stream)
{
    return pack_launch(true, staged, const_cast<double *>(state), lpos, m, val, j0, count, stream);
}

extern "C" int b200k_unpack_bits(double *state, const double *staged, const uint8_t *lpos, int m, uint32_t val,
                                 uint64_t j0, uint64_t count, void *stream)
{
    return pack_launch(false, const_cast<double *>(staged), state, lpos, m, val, j0, count, stream);
}

// ---- in-place peer-to-peer swap over NVLink / NVSwitch ---------------------------------------------
// The m-bit global<->local swap is a set of PAIRWISE exchanges: rank pattern a trades its sub-block
// whose lpos bits spell b with rank b's sub-block whose lpos bits spell a.  With the peer's shard
// mapped into this GPU's address space (CUDA IPC between processes, plain peer access between
// threads) the exchange needs no staging buffer, no pack/unpack pass and no NCCL transfer: each
// element pair is swapped by exactly ONE thread of ONE of the two GPUs (the lower pattern owns the
// lower half of the sub-block, the higher pattern the upper half), which reads the local and the
// remote amplitude and writes them crosswise — 128-bit loads/stores, both NVLink directions busy at
// once, four independent pairs in flight per thread to cover the remote-read latency.
__global__ void __launch_bounds__(256)
k_swap_p2p(cplx *__restrict__ mine, cplx *__restrict__ peer, PackPos P, int m, uint32_t my_pat, uint32_t peer_pat,
           uint64_t j_begin, uint64_t j_end)
{
    uint64_t dep_mine = 0, dep_peer = 0;
    for (int i = 0; i < m; i++) {
        dep_mine |= (uint64_t)((peer_pat >> i) & 1u) << P.pos[i];
        dep_peer |= (uint64_t)((my_pat >> i) & 1u) << P.pos[i];
    }
    auto expand = [&](uint64_t g) {
        for (int i = 0; i < m; i++) {
            const int p = P.pos[i];
            g = ((g >> p) << (p + 1)) | (g & ((1ull << p) - 1ull));
        }
        return g;
    };
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t j = j_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < j_end; j += 4 * stride) {
        uint64_t g[4];
        cplx x[4], y[4];
        bool on[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            on[u] = j + u * stride < j_end;
            g[u] = expand(j + u * stride);
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (on[u]) y[u] = peer[g[u] | dep_peer];
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (on[u]) x[u] = mine[g[u] | dep_mine];
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (on[u]) {
                peer[g[u] | dep_peer] = x[u];
                mine[g[u] | dep_mine] = y[u];
            }
    }
}

extern "C" int b200k_swap_p2p(double *mine, double *peer, const uint8_t *lpos, int m, uint32_t my_pat,
                              uint32_t peer_pat, uint64_t j_begin, uint64_t j_end, void *stream)
{
    if (m < 1 || m > 8 || my_pat == peer_pat) return (int)cudaErrorInvalidValue;
    PackPos P;
    for (int i = 0; i < 8; i++) P.pos[i] = i < m ? lpos[i] : 0;
    for (int i = 1; i < m; i++)
        if (P.pos[i] <= P.pos[i - 1]) return (int)cudaErrorInvalidValue;
    if (j_end <= j_begin) return 0;
    uint64_t blocks = (j_end - j_begin + 1023) / 1024;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_swap_p2p<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((cplx *)mine, (cplx *)peer, P, m, my_pat, peer_pat,
                                                                   j_begin, j_end);
    return (int)cudaGetLastError();
}

// ---- push-mode swap: ONE launch drives all peers at once, write-only over NVLink ------------------------
// Why (profiles/README.md, round 1: 34 qubits on 8 GPUs): the staged swap costs pack (HBM read + write of 7/8
// of the shard) -> ncclSend/ncclRecv -> unpack (another HBM read + write); the in-place pairwise kernel above
// avoids both but PULLS (remote reads, one peer after the other) and loses on 8 GPUs.  Here the sender reads
// its strided sub-blocks once and WRITES them straight into the receive staging of every peer (posted NVLink
// writes to all 2^m - 1 peers concurrently, no remote read, no send-side staging, no pack pass); the unpack of
// the previous chunk (staging -> the slots that chunk vacated, HBM-bound) runs in the SAME launch, so the NVLink
// transfer of chunk k overlaps the HBM traffic of chunk k-1.  blockIdx.y = peer, blockIdx.z = role.
struct PushArgs {
    cplx *dst[B200K_PUSH_MAX_PEERS];       // push: peer's receive staging for the chunk being sent (slot of THIS rank)
    const cplx *src[B200K_PUSH_MAX_PEERS]; // unpack: my receive staging of the previous chunk (slot of that peer)
    uint32_t pat[B200K_PUSH_MAX_PEERS];    // peer's pattern over the m exchanged bits
    PackPos P;
    int m, npeer;
    uint64_t push_j0, push_cnt;            // sub-block range sent now (cnt = 0: nothing to send)
    uint64_t unpack_j0, unpack_cnt;        // sub-block range received in the previous round (cnt = 0: nothing)
};

__global__ void __launch_bounds__(256)
k_swap_push(cplx *__restrict__ push_state, cplx *__restrict__ unpack_state, const __grid_constant__ PushArgs A)
{
    const int p = blockIdx.y;
    const bool unpack = blockIdx.z == 1;
    const uint64_t cnt = unpack ? A.unpack_cnt : A.push_cnt, j0 = unpack ? A.unpack_j0 : A.push_j0;
    if (cnt == 0) return;
    uint64_t deposit = 0;
    for (int i = 0; i < A.m; i++) deposit |= (uint64_t)((A.pat[p] >> i) & 1u) << A.P.pos[i];
    auto expand = [&](uint64_t g) {
        for (int i = 0; i < A.m; i++) {
            const int q = A.P.pos[i];
            g = ((g >> q) << (q + 1)) | (g & ((1ull << q) - 1ull));
        }
        return g | deposit;
    };
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (unpack) {
        const cplx *__restrict__ s = A.src[p];
        for (uint64_t j = t0; j < cnt; j += 4 * stride) {
            cplx v[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (j + u * stride < cnt) v[u] = __ldcs(&s[j + u * stride]);
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (j + u * stride < cnt) unpack_state[expand(j0 + j + u * stride)] = v[u];
        }
    }
    else {
        cplx *__restrict__ d = A.dst[p];
        for (uint64_t j = t0; j < cnt; j += 4 * stride) {
            cplx v[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (j + u * stride < cnt) v[u] = push_state[expand(j0 + j + u * stride)];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (j + u * stride < cnt) d[j + u * stride] = v[u]; // posted write into the peer's HBM
        }
    }
}

extern "C" int b200k_swap_push(double *push_state, double *unpack_state, double *const *peer_dst, const double *const *my_src, const uint32_t *peer_pat,
                               int npeer, const uint8_t *lpos, int m, uint64_t push_j0, uint64_t push_cnt,
                               uint64_t unpack_j0, uint64_t unpack_cnt, void *stream)
{
    if (m < 1 || m > 8 || npeer < 1 || npeer > B200K_PUSH_MAX_PEERS) return (int)cudaErrorInvalidValue;
    PushArgs A;
    for (int i = 0; i < 8; i++) A.P.pos[i] = i < m ? lpos[i] : 0;
    for (int i = 1; i < m; i++)
        if (A.P.pos[i] <= A.P.pos[i - 1]) return (int)cudaErrorInvalidValue;
    for (int p = 0; p < B200K_PUSH_MAX_PEERS; p++) {
        A.dst[p] = p < npeer ? (cplx *)peer_dst[p] : nullptr;
        A.src[p] = p < npeer ? (const cplx *)my_src[p] : nullptr;
        A.pat[p] = p < npeer ? peer_pat[p] : 0;
    }
    A.m = m;
    A.npeer = npeer;
    A.push_j0 = push_j0;
    A.push_cnt = push_cnt;
    A.unpack_j0 = unpack_j0;
    A.unpack_cnt = unpack_cnt;
    const uint64_t big = push_cnt > unpack_cnt ? push_cnt : unpack_cnt;
    if (big == 0) return 0;
    uint64_t bx = (big + 1023) / 1024;
    const uint64_t cap = (148ull * 16ull) / (uint64_t)(2 * npeer) + 1; // ~16 CTAs per SM over all (peer, role) slices
    if (bx > cap) bx = cap;
    dim3 grid((unsigned)bx, (unsigned)npeer, 2);
    k_swap_push<<<grid, 256, 0, (cudaStream_t)stream>>>((cplx *)push_state, (cplx *)unpack_state, A);
    return (int)cudaGetLastError();
}
