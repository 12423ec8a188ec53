// Unfused residual of two fields: (sum (a-b)^2, max |a-b|) over a row range; a-b is formed in the field's own
// precision (like the fused kernel), squares and sums are accumulated in fp64.
// Two passes through a scratch array of per-block partials, combined in a fixed order, so the result is
// bit-reproducible run to run.  The reference has no reduction of any kind (no residual, no norm, no
// convergence test); this is the baseline for the residual fused into the sweep kernel.
#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

constexpr int kResBlocks = 592;  // 148 SMs x 4
constexpr int kResThreads = 256;

constexpr int kResGridX = 4;                       // blocks along a row
constexpr int kResGridY = kResBlocks / kResGridX;  // blocks over the rows (grid-stride)

template <class T, int N>
struct alignas(sizeof(T) * N) ResVec {
    T v[N];
};

// blockIdx.y walks the rows, threads walk a row in 16-byte vectors (scalar tail / fallback when w or the base
// address does not allow it): no per-element divide, 16-byte loads.
template <class T>
__device__ __forceinline__ void block_partial(const T* __restrict__ a, const T* __restrict__ b, long long pitch,
                                              int w, int row_lo, int row_hi, double* part) {
    constexpr int VPT = 16 / (int)sizeof(T);
    double ss = 0.0, mx = 0.0;
    const bool vec = ((uintptr_t)a % 16 == 0) && ((uintptr_t)b % 16 == 0) && (pitch % VPT == 0);
    const int nvec = vec ? w / VPT : 0;
    const int tx = blockIdx.x * blockDim.x + threadIdx.x, sx = gridDim.x * blockDim.x;
    for (int r = row_lo + blockIdx.y; r < row_hi; r += gridDim.y) {
        const T* ra = a + (long long)r * pitch;
        const T* rb = b + (long long)r * pitch;
        for (int v = tx; v < nvec; v += sx) {
            const ResVec<T, VPT> va = reinterpret_cast<const ResVec<T, VPT>*>(ra)[v];
            const ResVec<T, VPT> vb = reinterpret_cast<const ResVec<T, VPT>*>(rb)[v];
#pragma unroll
            for (int k = 0; k < VPT; ++k) {
                const double d = (double)(T)(va.v[k] - vb.v[k]);  // difference in the field's own precision
                ss = fma(d, d, ss);
                mx = fmax(mx, fabs(d));
            }
        }
        for (int c = nvec * VPT + tx; c < w; c += sx) {
            const double d = (double)(T)(ra[c] - rb[c]);
            ss = fma(d, d, ss);
            mx = fmax(mx, fabs(d));
        }
    }
    __shared__ double s_ss[kResThreads / 32], s_mx[kResThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        s_ss[threadIdx.x >> 5] = ss;
        s_mx[threadIdx.x >> 5] = mx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a2 = 0.0, m2 = 0.0;
        for (int k = 0; k < kResThreads / 32; ++k) {
            a2 += s_ss[k];
            m2 = fmax(m2, s_mx[k]);
        }
        const int blk = blockIdx.y * gridDim.x + blockIdx.x;
        part[2 * blk] = a2;
        part[2 * blk + 1] = m2;
    }
}

template <class T>
__global__ void __launch_bounds__(kResThreads)
residual_pass1(const T* __restrict__ a, const T* __restrict__ b, long long pitch, int w, int row_lo, int row_hi,
               double* __restrict__ part) {
    block_partial<T>(a, b, pitch, w, row_lo, row_hi, part);
}

__global__ void residual_pass2(const double* __restrict__ part, int nblocks, double* __restrict__ out) {
    // one warp, fixed order: lane l sums blocks l, l+32, ...; then a fixed butterfly
    double ss = 0.0, mx = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 32) {
        ss += part[2 * i];
        mx = fmax(mx, part[2 * i + 1]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (threadIdx.x == 0) {
        out[0] = ss;
        out[1] = mx;
    }
}

}  // namespace mcp

using namespace mcp;

extern "C" {

size_t mcp_residual2d_scratch_bytes(void) { return (size_t)kResBlocks * 2 * sizeof(double); }

int mcp_residual2d(int dtype, const void* a, const void* b, int64_t pitch, int w, int row_lo, int row_hi,
                   void* scratch, double* out, void* stream) {
    MCP_REQUIRE(a && b && scratch && out, "mcp_residual2d: null argument");
    MCP_REQUIRE(row_lo <= row_hi && w >= 1 && pitch >= w, "mcp_residual2d: bad geometry");
    cudaStream_t s = (cudaStream_t)stream;
    double* part = (double*)scratch;
    if (dtype == MCP_DTYPE_F32)
        residual_pass1<float><<<dim3(kResGridX, kResGridY), kResThreads, 0, s>>>((const float*)a, (const float*)b, pitch, w, row_lo,
                                                                 row_hi, part);
    else if (dtype == MCP_DTYPE_F64)
        residual_pass1<double><<<dim3(kResGridX, kResGridY), kResThreads, 0, s>>>((const double*)a, (const double*)b, pitch, w,
                                                                  row_lo, row_hi, part);
    else if (dtype == MCP_DTYPE_I32)
        residual_pass1<int><<<dim3(kResGridX, kResGridY), kResThreads, 0, s>>>((const int*)a, (const int*)b, pitch, w, row_lo, row_hi,
                                                               part);
    else
        MCP_REQUIRE(false, "mcp_residual2d: bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    residual_pass2<<<1, 32, 0, s>>>(part, kResBlocks, out);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // extern "C"
