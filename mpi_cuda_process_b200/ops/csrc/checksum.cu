// Position-dependent 64-bit checksum of the owned cells of a slab, computed on the device: the wrapping sum over cells
// of mix64(bit pattern ^ golden * global cell index).  The sum is order independent, so the checksums of the slabs of
// N ranks add up (mod 2^64) to the checksum of the whole field whatever N is -- a field of 16384^2 cells produced on 8
// GPUs can be compared bit for bit with the same solve on 1 GPU or with another backend without gathering it.
// No counterpart in the reference (it never looks at its result, MDF_kernel.cu:185,218).
#include <algorithm>

#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

template <class B, int N>
struct alignas(sizeof(B) * N) SumVec {
    B v[N];
};

template <class B>  // B = unsigned (4-byte cells) or unsigned long long (fp64)
__global__ void __launch_bounds__(256) checksum2d_kernel(const B* __restrict__ u, long long pitch, int w, int row_lo,
                                                         int nrows, long long grow0, unsigned long long* out) {
    constexpr int VPT = 16 / (int)sizeof(B);
    unsigned long long acc = 0;
    const bool vec = ((uintptr_t)u % 16 == 0) && (pitch % VPT == 0);
    const int nvec = vec ? w / VPT : 0;
    const int tx = blockIdx.x * blockDim.x + threadIdx.x, sx = gridDim.x * blockDim.x;
    for (int row = blockIdx.y; row < nrows; row += gridDim.y) {
        const B* s = u + (long long)(row_lo + row) * pitch;
        const unsigned long long base = (unsigned long long)(grow0 + row) * (unsigned long long)w;
        for (int v = tx; v < nvec; v += sx) {
            const SumVec<B, VPT> x = reinterpret_cast<const SumVec<B, VPT>*>(s)[v];  // one 16-byte load
#pragma unroll
            for (int k = 0; k < VPT; ++k)
                acc += mix64((unsigned long long)x.v[k] ^ ((base + (unsigned)(v * VPT + k)) * 0x9e3779b97f4a7c15ull));
        }
        for (int c = nvec * VPT + tx; c < w; c += sx)
            acc += mix64((unsigned long long)s[c] ^ ((base + c) * 0x9e3779b97f4a7c15ull));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc != 0) atomicAdd(out, acc);
}

}  // namespace mcp

using namespace mcp;

// out_dev: one 64-bit word, zeroed by this call; rows [row_lo, row_hi) of the buffer are global rows grow0 + (r - row_lo).
extern "C" int mcp_checksum2d(int dtype, const void* u_dev, int64_t pitch, int w, int row_lo, int row_hi, int64_t grow0,
                              void* out_dev, void* stream) {
    MCP_REQUIRE(u_dev && out_dev && w >= 1 && row_lo <= row_hi, "mcp_checksum2d: bad arguments");
    MCP_REQUIRE(dtype == MCP_DTYPE_F32 || dtype == MCP_DTYPE_F64 || dtype == MCP_DTYPE_I32,
                "mcp_checksum2d: dtype must be f32, f64 or i32");
    cudaStream_t s = (cudaStream_t)stream;
    MCP_CUDA_TRY(cudaMemsetAsync(out_dev, 0, 8, s));
    const int nrows = row_hi - row_lo;
    if (nrows == 0) return MCP_OK;
    dim3 grid(std::max(1, std::min(4, (w / 4 + 255) / 256)), std::min(nrows, 148 * 8));
    if (dtype == MCP_DTYPE_F64)
        checksum2d_kernel<unsigned long long><<<grid, 256, 0, s>>>((const unsigned long long*)u_dev, pitch, w, row_lo,
                                                                   nrows, grow0, (unsigned long long*)out_dev);
    else
        checksum2d_kernel<unsigned><<<grid, 256, 0, s>>>((const unsigned*)u_dev, pitch, w, row_lo, nrows, grow0,
                                                         (unsigned long long*)out_dev);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}
