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
