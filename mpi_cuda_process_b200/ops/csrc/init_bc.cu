// Initial-condition and boundary-condition kernels (sm_100a), global-coordinate aware so that every rank
// fills only its own rows (plus ghost rows) of the distributed field.
// Reference: create_universe (MDF_kernel.cu:88-99: ring = 100, interior = 0; kernel.cu:131-146: ring dead,
// interior alive with p = 0.15) and the "else -> 100 / 0" branches of the kernels
// (MDF_kernel.cu:35,43,59,67; kernel.cu:86,102,110).  The reference builds the field on the host and
// ships the whole grid over PCIe every generation (MDF_kernel.cu:161); here it is born on the device.
#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

template <class T>
__global__ void init_ring2d_kernel(T* __restrict__ u, long long pitch, int rows_alloc, int row0, int h, int w, T ring,
                                   T interior) {
    const long long n = (long long)rows_alloc * pitch;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / pitch), c = (int)(i - (long long)r * pitch);
        const int gy = row0 + r;
        T v = T(0);
        if (gy >= 0 && gy < h && c < w) v = (gy == 0 || gy == h - 1 || c == 0 || c == w - 1) ? ring : interior;
        u[i] = v;
    }
}

__device__ __forceinline__ unsigned hash32(unsigned x, unsigned seed) {
    x ^= seed * 0x9E3779B9u;
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}

__global__ void init_life_hash_kernel(int* __restrict__ u, long long pitch, int rows_alloc, int row0, int h, int w,
                                      unsigned thr, unsigned seed) {
    const long long n = (long long)rows_alloc * pitch;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / pitch), c = (int)(i - (long long)r * pitch);
        const int gy = row0 + r;
        int v = 0;
        if (gy > 0 && gy < h - 1 && c > 0 && c < w - 1)
            v = hash32((unsigned)((long long)gy * w + c), seed) < thr ? 1 : 0;
        u[i] = v;
    }
}

// Bit-packed variant of init_life_hash_kernel: one thread builds one 32-cell word (cell x = bit x%32 of word x/32).
__global__ void init_life_hash_bits_kernel(unsigned* __restrict__ u, long long pitch, int rows_alloc, int row0, int h,
                                           int w, unsigned thr, unsigned seed) {
    const long long n = (long long)rows_alloc * pitch;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / pitch), e = (int)(i - (long long)r * pitch);
        const int gy = row0 + r;
        unsigned word = 0u;
        if (gy > 0 && gy < h - 1) {
#pragma unroll 8
            for (int b = 0; b < 32; ++b) {
                const int c = e * 32 + b;
                if (c > 0 && c < w - 1 && hash32((unsigned)((long long)gy * w + c), seed) < thr) word |= 1u << b;
            }
        }
        u[i] = word;
    }
}

// Synthetic initial fields beyond the reference's ring (all exactly reproducible on the host, oracle.py):
// kind 1: interior cells uniform in [interior, ring) from hash32(i*w+j, seed) (24 random bits);
// kind 2: a hot box (value ring) over the central third of the domain, interior elsewhere.  Ring = ring either way.
template <class T>
__global__ void init_shape2d_kernel(T* __restrict__ u, long long pitch, int rows_alloc, int row0, int h, int w,
                                    double ring, double interior, int kind, unsigned seed) {
    const long long n = (long long)rows_alloc * pitch;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / pitch), c = (int)(i - (long long)r * pitch);
        const int gy = row0 + r;
        double v = 0.0;
        if (gy >= 0 && gy < h && c < w) {
            if (gy == 0 || gy == h - 1 || c == 0 || c == w - 1) {
                v = ring;
            } else if (kind == 1) {
                const unsigned hsh = hash32((unsigned)((long long)gy * w + c), seed);
                v = interior + (ring - interior) * ((double)(hsh >> 8) * (1.0 / 16777216.0));
            } else {
                const bool hot = gy >= h / 3 && gy < h - h / 3 && c >= w / 3 && c < w - w / 3;
                v = hot ? ring : interior;
            }
        }
        u[i] = (T)v;
    }
}

template <class T>
__global__ void apply_ring2d_kernel(T* __restrict__ u, long long pitch, int rows_alloc, int row0, int h, int w,
                                    T ring) {
    // Only ring cells are touched: two full rows (if owned) + two columns.
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = tid; i < (long long)rows_alloc * 2; i += stride) {
        const int r = (int)(i >> 1);
        const int gy = row0 + r;
        if (gy < 0 || gy >= h) continue;
        u[(long long)r * pitch + ((i & 1) ? (w - 1) : 0)] = ring;
    }
    for (int edge = 0; edge < 2; ++edge) {
        const int gy = edge ? h - 1 : 0;
        const int r = gy - row0;
        if (r < 0 || r >= rows_alloc) continue;
        for (long long c = tid; c < w; c += stride) u[(long long)r * pitch + c] = ring;
    }
}

template <class T>
__global__ void init_shell3d_kernel(T* __restrict__ u, long long pitch, long long plane_pitch, int planes_alloc,
                                    int rows_alloc, int z0, int y0, int d, int h, int w, T ring, T interior) {
    const long long n = (long long)planes_alloc * plane_pitch;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int pz = (int)(i / plane_pitch);
        const long long rem = i - (long long)pz * plane_pitch;
        const int r = (int)(rem / pitch), c = (int)(rem - (long long)r * pitch);
        const int gz = z0 + pz, gy = y0 + r;
        T v = T(0);
        if (r < rows_alloc && gz >= 0 && gz < d && gy >= 0 && gy < h && c < w)
            v = (gz == 0 || gz == d - 1 || gy == 0 || gy == h - 1 || c == 0 || c == w - 1) ? ring : interior;
        u[i] = v;
    }
}

static int grid_for(long long n) {
    long long b = (n + 255) / 256;
    return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

}  // namespace mcp

using namespace mcp;

extern "C" {

int mcp_init_ring2d(int dtype, void* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                    double interior, void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_ring2d: bad arguments");
    const long long n = (long long)rows_alloc * pitch;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        init_ring2d_kernel<float><<<grid_for(n), 256, 0, s>>>((float*)u, pitch, rows_alloc, row0, h, w, (float)ring,
                                                              (float)interior);
    else if (dtype == MCP_DTYPE_F64)
        init_ring2d_kernel<double><<<grid_for(n), 256, 0, s>>>((double*)u, pitch, rows_alloc, row0, h, w, ring,
                                                               interior);
    else if (dtype == MCP_DTYPE_I32)
        init_ring2d_kernel<int><<<grid_for(n), 256, 0, s>>>((int*)u, pitch, rows_alloc, row0, h, w, (int)ring,
                                                            (int)interior);
    else
        MCP_REQUIRE(false, "bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

int mcp_init_shape2d(int dtype, int kind, void* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                     double interior, uint32_t seed, void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_shape2d: bad arguments");
    MCP_REQUIRE(kind == 1 || kind == 2, "mcp_init_shape2d: kind must be 1 (hash) or 2 (box)");
    const long long n = (long long)rows_alloc * pitch;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        init_shape2d_kernel<float><<<grid_for(n), 256, 0, s>>>((float*)u, pitch, rows_alloc, row0, h, w, ring, interior,
                                                               kind, seed);
    else if (dtype == MCP_DTYPE_F64)
        init_shape2d_kernel<double><<<grid_for(n), 256, 0, s>>>((double*)u, pitch, rows_alloc, row0, h, w, ring,
                                                                interior, kind, seed);
    else
        MCP_REQUIRE(false, "mcp_init_shape2d: bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

int mcp_init_life_hash(int32_t* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double prob, uint32_t seed,
                       void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_life_hash: bad arguments");
    double t = prob * 4294967296.0;
    const unsigned thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (t <= 0 ? 0u : (unsigned)t);
    const long long n = (long long)rows_alloc * pitch;
    init_life_hash_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(u, pitch, rows_alloc, row0, h, w, thr, seed);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

int mcp_init_life_hash_bits(uint32_t* u, int64_t pitch_words, int rows_alloc, int row0, int h, int w, double prob,
                            uint32_t seed, void* stream) {
    MCP_REQUIRE(u && pitch_words * 32 >= w && rows_alloc > 0, "mcp_init_life_hash_bits: bad arguments");
    double t = prob * 4294967296.0;
    const unsigned thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (t <= 0 ? 0u : (unsigned)t);
    const long long n = (long long)rows_alloc * pitch_words;
    init_life_hash_bits_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(u, pitch_words, rows_alloc, row0, h, w,
                                                                              thr, seed);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

int mcp_apply_ring2d(int dtype, void* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                     void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_apply_ring2d: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    const int g = grid_for(std::max<long long>(w, 2LL * rows_alloc));
    if (dtype == MCP_DTYPE_F32)
        apply_ring2d_kernel<float><<<g, 256, 0, s>>>((float*)u, pitch, rows_alloc, row0, h, w, (float)ring);
    else if (dtype == MCP_DTYPE_F64)
        apply_ring2d_kernel<double><<<g, 256, 0, s>>>((double*)u, pitch, rows_alloc, row0, h, w, ring);
    else if (dtype == MCP_DTYPE_I32)
        apply_ring2d_kernel<int><<<g, 256, 0, s>>>((int*)u, pitch, rows_alloc, row0, h, w, (int)ring);
    else
        MCP_REQUIRE(false, "bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

int mcp_init_shell3d(int dtype, void* u, int64_t pitch, int64_t plane_pitch, int planes_alloc, int rows_alloc, int z0,
                     int y0, int d, int h, int w, double ring, double interior, void* stream) {
    MCP_REQUIRE(u && pitch >= w && plane_pitch >= pitch * rows_alloc && planes_alloc > 0, "mcp_init_shell3d: bad arguments");
    const long long n = (long long)planes_alloc * plane_pitch;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        init_shell3d_kernel<float><<<grid_for(n), 256, 0, s>>>((float*)u, pitch, plane_pitch, planes_alloc, rows_alloc,
                                                               z0, y0, d, h, w, (float)ring, (float)interior);
    else if (dtype == MCP_DTYPE_F64)
        init_shell3d_kernel<double><<<grid_for(n), 256, 0, s>>>((double*)u, pitch, plane_pitch, planes_alloc,
                                                                rows_alloc, z0, y0, d, h, w, ring, interior);
    else
        MCP_REQUIRE(false, "bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // extern "C"
