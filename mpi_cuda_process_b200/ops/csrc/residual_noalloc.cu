// Residual of two fields for callers that must not allocate device memory.
//
// Who this is for: hosts where every byte of device memory is owned by somebody else -- a framework's
// caching allocator that has already claimed the whole GPU, a service running under a fixed memory
// budget, a plug-in that is forbidden to allocate after start-up -- and latency-critical loops that cannot
// afford the device-wide synchronisation a cudaMalloc/cudaFree may imply.  Such a caller has nowhere to get
// the scratch buffer mcp_residual2d() asks for, so this variant brings its own: a fixed array that lives in
// the library image itself.
//
// The price of that convenience:
//   * the array has a fixed capacity (MCP_NOALLOC_MAX_ROWS rows per call, 1 MiB of device memory); it is
//     reserved on every GPU that loads the library and stays reserved until the process exits;
//   * there is ONE array per GPU, shared by every call: two calls that overlap on the same GPU (different
//     streams, different threads) race on it and both results are garbage -- serialise them;
//   * the array keeps whatever the previous call left behind.  Nothing here needs to start from zero: pass 1
//     overwrites every entry that pass 2 reads, and pass 2 reads only the entries of the current call.
//
// This entry point has no counterpart in the reference (which computes no residual at all).
#include <mutex>

#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

constexpr int kNoAllocMaxRows = MCP_NOALLOC_MAX_ROWS;
constexpr int kRowThreads = 128;
constexpr int kMaxDevices = 64;

// One (sum of squares, max abs) pair per row.  File-scope device storage: part of the module image, so no
// allocation call is ever made for it.  In HOST code this name is only a stand-in for the symbol -- it is
// not a device address; the address a kernel can use comes from cudaGetSymbolAddress (see scratch_for_device).
static __device__ __align__(256) double g_row_partials[2 * kNoAllocMaxRows];

template <class T>
__global__ void __launch_bounds__(kRowThreads)
noalloc_row_pass(const T* __restrict__ a, const T* __restrict__ b, long long pitch, int w, int row_lo,
                 double* __restrict__ part) {
    // one block per row: fixed assignment of columns to threads, fixed combination order
    const int r = blockIdx.x;
    const long long base = (long long)(row_lo + r) * pitch;
    double ss = 0.0, mx = 0.0;
    for (int c = threadIdx.x; c < w; c += kRowThreads) {
        const double d = (double)(T)(a[base + c] - b[base + c]);  // difference in the field's own precision
        ss = fma(d, d, ss);
        mx = fmax(mx, fabs(d));
    }
    __shared__ double s_ss[kRowThreads / 32], s_mx[kRowThreads / 32];
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
        for (int k = 0; k < kRowThreads / 32; ++k) {
            a2 += s_ss[k];
            m2 = fmax(m2, s_mx[k]);
        }
        part[2 * r] = a2;
        part[2 * r + 1] = m2;
    }
}

__global__ void __launch_bounds__(256) noalloc_final_pass(const double* __restrict__ part, int rows,
                                                          double* __restrict__ out) {
    double ss = 0.0, mx = 0.0;
    for (int i = threadIdx.x; i < rows; i += 256) {
        ss += part[2 * i];
        mx = fmax(mx, part[2 * i + 1]);
    }
    __shared__ double s_ss[8], s_mx[8];
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
        for (int k = 0; k < 8; ++k) {
            a2 += s_ss[k];
            m2 = fmax(m2, s_mx[k]);
        }
        out[0] = a2;
        out[1] = m2;
    }
}

// Device address of the scratch array on the CURRENT device, looked up once per device.
// cudaGetSymbolAddress only reads the module's symbol table; it reserves nothing.
static int scratch_for_device(double** out) {
    static std::mutex mu;
    static double* table[kMaxDevices] = {nullptr};
    int dev = -1;
    MCP_CUDA_TRY(cudaGetDevice(&dev));
    MCP_REQUIRE(dev >= 0 && dev < kMaxDevices, "mcp_residual2d_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (table[dev] == nullptr) {
        double* p = nullptr;
        MCP_CUDA_TRY(cudaGetSymbolAddress((void**)&p, g_row_partials));
        MCP_REQUIRE(p != nullptr, "mcp_residual2d_noalloc: built-in scratch has no device address on device %d", dev);
        table[dev] = p;
    }
    *out = table[dev];
    return MCP_OK;
}

}  // namespace mcp

using namespace mcp;

extern "C" {

int mcp_residual2d_noalloc_max_rows(void) { return kNoAllocMaxRows; }

int mcp_residual2d_noalloc(int dtype, const void* a, const void* b, int64_t pitch, int w, int row_lo, int row_hi,
                           double* out, void* stream) {
    MCP_REQUIRE(a && b && out, "mcp_residual2d_noalloc: null argument");
    MCP_REQUIRE(row_lo >= 0 && row_lo <= row_hi && w >= 1 && pitch >= w, "mcp_residual2d_noalloc: bad geometry");
    const long long rows = (long long)row_hi - row_lo;
    if (rows > kNoAllocMaxRows) {
        set_error("mcp_residual2d_noalloc: %lld rows exceed the built-in scratch capacity of %d rows per call; "
                  "split the row range, or use mcp_residual2d with scratch you provide",
                  rows, kNoAllocMaxRows);
        return MCP_ERR_ARG;
    }
    double* part = nullptr;
    int rc = scratch_for_device(&part);   // first call on a device resolves the symbol address
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (rows == 0) {
        noalloc_final_pass<<<1, 256, 0, s>>>(part, 0, out);
        MCP_CUDA_TRY(cudaGetLastError());
        return MCP_OK;
    }
    if (dtype == MCP_DTYPE_F32)
        noalloc_row_pass<float><<<(unsigned)rows, kRowThreads, 0, s>>>((const float*)a, (const float*)b, pitch, w,
                                                                      row_lo, part);
    else if (dtype == MCP_DTYPE_F64)
        noalloc_row_pass<double><<<(unsigned)rows, kRowThreads, 0, s>>>((const double*)a, (const double*)b, pitch, w,
                                                                       row_lo, part);
    else if (dtype == MCP_DTYPE_I32)
        noalloc_row_pass<int><<<(unsigned)rows, kRowThreads, 0, s>>>((const int*)a, (const int*)b, pitch, w, row_lo,
                                                                    part);
    else
        MCP_REQUIRE(false, "mcp_residual2d_noalloc: bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    noalloc_final_pass<<<1, 256, 0, s>>>(part, (int)rows, out);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // extern "C"
