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
