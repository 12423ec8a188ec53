// Fused gather + cast/pack: every rank's kernel writes its owned rows straight into the final output buffer
// (local, or rank 0's buffer mapped over NVLink), converting on the fly.  Reference: print_array
// (kernel.cu:115-129) defines the only output format ('0' for a live cell, ' ' otherwise, '\n' per row) but is
// commented out at every call site and the reference has no gather at all (every rank holds the full grid).
//
// Shape of the kernels: blockIdx.y walks the rows (grid-stride), threads walk a row in 16-byte vectors, so there is no
// per-element divide and every NVLink store is a full 16-byte (fp64: 2 x 16-byte) transaction; a scalar tail handles
// widths that are not a multiple of the vector.
#include <algorithm>

#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

template <class T, int N>
struct alignas(sizeof(T) * N) Vec {
    T v[N];
};

// Owned planes/rows of a (planes, rows, pitch) buffer -> dense (D, H, W) destination; 4 elements per thread-step.
// VECTOR: w % 4 == 0 and the destination base is 16-byte aligned, so every row start is aligned too.
template <class TS, class TD, bool VECTOR>
__global__ void __launch_bounds__(256) pack_block_kernel(const TS* __restrict__ src, long long pitch,
                                                         long long plane_pitch, int z_lo, int nz, int y_lo, int ny,
                                                         int w, TD* __restrict__ dst, long long H, long long dst_z0,
                                                         long long dst_y0) {
    const int nrows = nz * ny;
    for (int row = blockIdx.y; row < nrows; row += gridDim.y) {
        const int z = row / ny, y = row - z * ny;  // one divide per ROW
        const TS* s = src + (long long)(z_lo + z) * plane_pitch + (long long)(y_lo + y) * pitch;
        TD* d = dst + ((dst_z0 + z) * H + (dst_y0 + y)) * (long long)w;
        if (VECTOR) {
            const int nvec = w >> 2;
            for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
                const Vec<TS, 4> in = reinterpret_cast<const Vec<TS, 4>*>(s)[v];
                Vec<TD, 4> o;
#pragma unroll
                for (int k = 0; k < 4; ++k) o.v[k] = (TD)in.v[k];
                reinterpret_cast<Vec<TD, 4>*>(d)[v] = o;
            }
        } else {
            for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < w; c += gridDim.x * blockDim.x) d[c] = (TD)s[c];
        }
    }
}

// Bit-packed life (32 cells per word, cell x = bit x % 32 of word x / 32) -> one int32 per cell.  One word per
// thread-step = 32 cells = eight 16-byte stores.
__global__ void __launch_bounds__(256) unpack_bits_kernel(const unsigned* __restrict__ src, long long pitch_words,
                                                          int row_lo, int nrows, int w, int* __restrict__ dst,
                                                          long long dst_row0, int vector) {
    const int nwords = (w + 31) >> 5;
    for (int row = blockIdx.y; row < nrows; row += gridDim.y) {
        const unsigned* s = src + (long long)(row_lo + row) * pitch_words;
        int* d = dst + (dst_row0 + row) * (long long)w;
        for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nwords; k += gridDim.x * blockDim.x) {
            const unsigned bits = s[k];
            const int c0 = k << 5;
            if (vector && c0 + 32 <= w) {
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    int4 o;
                    o.x = (bits >> (4 * q)) & 1u;
                    o.y = (bits >> (4 * q + 1)) & 1u;
                    o.z = (bits >> (4 * q + 2)) & 1u;
                    o.w = (bits >> (4 * q + 3)) & 1u;
                    reinterpret_cast<int4*>(d + c0)[q] = o;
                }
            } else {
                for (int b = 0; b < 32 && c0 + b < w; ++b) d[c0 + b] = (bits >> b) & 1u;
            }
        }
    }
}

// print_array bytes: '0' for a live cell, ' ' otherwise, '\n' after every row (kernel.cu:115-129).
__global__ void __launch_bounds__(256) pack_ascii_kernel(const int* __restrict__ src, long long pitch, int w, int row_lo,
                                                         int nrows, char* __restrict__ dst, long long dst_row0) {
    const long long wl = (long long)w + 1;
    for (int row = blockIdx.y; row < nrows; row += gridDim.y) {
        const int* s = src + (long long)(row_lo + row) * pitch;
        char* d = dst + (dst_row0 + row) * wl;
        for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= w; c += gridDim.x * blockDim.x)
            d[c] = c == w ? '\n' : (s[c] == 1 ? '0' : ' ');
    }
}

template <class TS, class TD>
int launch_pack_block(const void* src, int64_t pitch, int64_t plane_pitch, int z_lo, int nz, int y_lo, int ny, int w,
                      void* dst, int64_t dst_h, int64_t dst_z0, int64_t dst_y0, cudaStream_t s) {
    const bool vec = (w % 4 == 0) && ((uintptr_t)dst % (4 * sizeof(TD)) == 0) && ((uintptr_t)src % (4 * sizeof(TS)) == 0) &&
                     (pitch % 4 == 0) && (plane_pitch % 4 == 0);
    const int per_row = vec ? w / 4 : w;
    const int gx = std::max(1, std::min(8, (per_row + 255) / 256));
    const int gy = (int)std::min<long long>((long long)nz * ny, 148 * 16 / gx);
    dim3 grid(gx, std::max(1, gy));
    if (vec)
        pack_block_kernel<TS, TD, true><<<grid, 256, 0, s>>>((const TS*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w,
                                                             (TD*)dst, dst_h, dst_z0, dst_y0);
    else
        pack_block_kernel<TS, TD, false><<<grid, 256, 0, s>>>((const TS*)src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w,
                                                              (TD*)dst, dst_h, dst_z0, dst_y0);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // namespace mcp

using namespace mcp;

extern "C" int mcp_pack_block(int dtype, int mode, const void* src, int64_t pitch, int64_t plane_pitch, int z_lo, int nz,
                              int y_lo, int ny, int w, void* dst, int64_t dst_h, int64_t dst_z0, int64_t dst_y0,
                              void* stream) {
    MCP_REQUIRE(src && dst && nz >= 0 && ny >= 0 && w >= 1, "mcp_pack_block: bad arguments");
    if (nz == 0 || ny == 0) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == 0 && dtype == MCP_DTYPE_F32)
        return launch_pack_block<float, float>(src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w, dst, dst_h, dst_z0, dst_y0, s);
    if (mode == 0 && dtype == MCP_DTYPE_F64)
        return launch_pack_block<double, double>(src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w, dst, dst_h, dst_z0, dst_y0, s);
    if (mode == 0 && dtype == MCP_DTYPE_I32)
        return launch_pack_block<int, int>(src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w, dst, dst_h, dst_z0, dst_y0, s);
    if (mode == 1 && dtype == MCP_DTYPE_F64)
        return launch_pack_block<double, float>(src, pitch, plane_pitch, z_lo, nz, y_lo, ny, w, dst, dst_h, dst_z0, dst_y0, s);
    set_error("mcp_pack_block: unsupported mode %d for dtype %d", mode, dtype);
    return MCP_ERR_UNSUPPORTED;
}

extern "C" int mcp_pack_rows(int dtype, int mode, const void* src, int64_t pitch, int w, int row_lo, int row_hi,
                             void* dst, int64_t dst_row0, void* stream) {
    MCP_REQUIRE(src && dst && row_lo <= row_hi && w >= 1, "mcp_pack_rows: bad arguments");
    const int nrows = row_hi - row_lo;
    if (nrows == 0) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == 2 && dtype == MCP_DTYPE_I32) {
        dim3 grid(std::max(1, std::min(8, (w + 256) / 256)), std::min(nrows, 148 * 4));
        pack_ascii_kernel<<<grid, 256, 0, s>>>((const int*)src, pitch, w, row_lo, nrows, (char*)dst, dst_row0);
        MCP_CUDA_TRY(cudaGetLastError());
        return MCP_OK;
    }
    if (mode == 3 && dtype == MCP_DTYPE_B1) {  // bit-packed words -> one int32 per cell
        const int nwords = (w + 31) / 32;
        dim3 grid(std::max(1, std::min(8, (nwords + 255) / 256)), std::min(nrows, 148 * 8));
        const int vector = (w % 4 == 0) && ((uintptr_t)dst % 16 == 0);
        unpack_bits_kernel<<<grid, 256, 0, s>>>((const unsigned*)src, pitch, row_lo, nrows, w, (int*)dst, dst_row0, vector);
        MCP_CUDA_TRY(cudaGetLastError());
        return MCP_OK;
    }
    if ((mode == 0 && (dtype == MCP_DTYPE_F32 || dtype == MCP_DTYPE_F64 || dtype == MCP_DTYPE_I32)) ||
        (mode == 1 && dtype == MCP_DTYPE_F64))
        // rows of a 2-D slab are the single-plane case of the block kernel: destination (1, H, W) with H unused
        return mcp_pack_block(dtype, mode, src, pitch, 0, 0, 1, row_lo, nrows, w, dst, 0, 0, dst_row0, stream);
    set_error("mcp_pack_rows: unsupported mode %d for dtype %d", mode, dtype);
    return MCP_ERR_UNSUPPORTED;
}
