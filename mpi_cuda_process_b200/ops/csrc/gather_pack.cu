// Fused gather + cast/pack: every rank's kernel writes its owned rows straight into the final output buffer
// (local, or rank 0's buffer mapped over NVLink), converting on the fly.  Reference: print_array
// (kernel.cu:115-129) defines the only output format ('0' for a live cell, ' ' otherwise, '\n' per row) but is
// commented out at every call site and the reference has no gather at all (every rank holds the full grid).
#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

template <class TS, class TD>
__global__ void pack_cast_kernel(const TS* __restrict__ src, long long pitch, int w, int row_lo, int nrows,
                                 TD* __restrict__ dst, long long dst_row0) {
    const long long n = (long long)nrows * w;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / w), c = (int)(i - (long long)r * w);
        dst[(dst_row0 + r) * w + c] = (TD)src[(long long)(row_lo + r) * pitch + c];
    }
}

__global__ void pack_ascii_kernel(const int* __restrict__ src, long long pitch, int w, int row_lo, int nrows,
                                  char* __restrict__ dst, long long dst_row0) {
    const long long wl = (long long)w + 1;
    const long long n = (long long)nrows * wl;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / wl), c = (int)(i - (long long)r * wl);
        char ch = '\n';
        if (c < w) ch = src[(long long)(row_lo + r) * pitch + c] == 1 ? '0' : ' ';
        dst[(dst_row0 + r) * wl + c] = ch;
    }
}

// Block version for pencils: owned planes/rows of a (planes, rows, pitch) buffer -> dense (D, H, W) destination.
template <class TS, class TD>
__global__ void pack_block_kernel(const TS* __restrict__ src, long long pitch, long long plane_pitch, int z_lo, int nz,
                                  int y_lo, int ny, int w, TD* __restrict__ dst, long long H, long long dst_z0,
                                  long long dst_y0) {
    const long long n = (long long)nz * ny * w;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % w);
        const long long t = i / w;
        const int y = (int)(t % ny), z = (int)(t / ny);
        dst[((dst_z0 + z) * H + (dst_y0 + y)) * w + c] =
            (TD)src[(long long)(z_lo + z) * plane_pitch + (long long)(y_lo + y) * pitch + c];
    }
}

}  // namespace mcp

using namespace mcp;

extern "C" int mcp_pack_block(int dtype, int mode, const void* src, int64_t pitch, int64_t plane_pitch, int z_lo, int nz,
                              int y_lo, int ny, int w, void* dst, int64_t dst_h, int64_t dst_z0, int64_t dst_y0,
                              void* stream) {
    MCP_REQUIRE(src && dst && nz >= 0 && ny >= 0 && w >= 1, "mcp_pack_block: bad arguments");
    if (nz == 0 || ny == 0) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const long long n = (long long)nz * ny * w;
    const int grid = (int)std::min<long long>(148 * 8, (n + 255) / 256);
    if (mode == 0 && dtype == MCP_DTYPE_F32)
        pack_block_kernel<float, float><<<grid, 256, 0, s>>>((const float*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w,
                                                             (float*)dst, dst_h, dst_z0, dst_y0);
    else if (mode == 0 && dtype == MCP_DTYPE_F64)
        pack_block_kernel<double, double><<<grid, 256, 0, s>>>((const double*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny,
                                                               w, (double*)dst, dst_h, dst_z0, dst_y0);
    else if (mode == 0 && dtype == MCP_DTYPE_I32)
        pack_block_kernel<int, int><<<grid, 256, 0, s>>>((const int*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w,
                                                         (int*)dst, dst_h, dst_z0, dst_y0);
    else if (mode == 1 && dtype == MCP_DTYPE_F64)
        pack_block_kernel<double, float><<<grid, 256, 0, s>>>((const double*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny,
                                                              w, (float*)dst, dst_h, dst_z0, dst_y0);
    else {
        set_error("mcp_pack_block: unsupported mode %d for dtype %d", mode, dtype);
        return MCP_ERR_UNSUPPORTED;
    }
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

extern "C" int mcp_pack_rows(int dtype, int mode, const void* src, int64_t pitch, int w, int row_lo, int row_hi,
                             void* dst, int64_t dst_row0, void* stream) {
    MCP_REQUIRE(src && dst && row_lo <= row_hi && w >= 1, "mcp_pack_rows: bad arguments");
    const int nrows = row_hi - row_lo;
    if (nrows == 0) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const long long n = (long long)nrows * (w + 1);
    const int grid = (int)std::min<long long>(148 * 8, (n + 255) / 256);
    if (mode == 0 && dtype == MCP_DTYPE_F32)
        pack_cast_kernel<float, float><<<grid, 256, 0, s>>>((const float*)src, pitch, w, row_lo, nrows, (float*)dst, dst_row0);
    else if (mode == 0 && dtype == MCP_DTYPE_F64)
        pack_cast_kernel<double, double><<<grid, 256, 0, s>>>((const double*)src, pitch, w, row_lo, nrows, (double*)dst, dst_row0);
    else if (mode == 0 && dtype == MCP_DTYPE_I32)
        pack_cast_kernel<int, int><<<grid, 256, 0, s>>>((const int*)src, pitch, w, row_lo, nrows, (int*)dst, dst_row0);
    else if (mode == 1 && dtype == MCP_DTYPE_F64)
        pack_cast_kernel<double, float><<<grid, 256, 0, s>>>((const double*)src, pitch, w, row_lo, nrows, (float*)dst, dst_row0);
    else if (mode == 2 && dtype == MCP_DTYPE_I32)
        pack_ascii_kernel<<<grid, 256, 0, s>>>((const int*)src, pitch, w, row_lo, nrows, (char*)dst, dst_row0);
    else {
        set_error("mcp_pack_rows: unsupported mode %d for dtype %d", mode, dtype);
        return MCP_ERR_UNSUPPORTED;
    }
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}
