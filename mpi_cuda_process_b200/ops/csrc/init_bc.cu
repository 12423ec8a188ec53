// Initial-condition and boundary-condition kernels (sm_100a), global-coordinate aware so that every rank
// fills only its own rows (plus ghost rows) of the distributed field.
// Reference: create_universe (MDF_kernel.cu:88-99: ring = 100, interior = 0; kernel.cu:131-146: ring dead,
// interior alive with p = 0.15) and the "else -> 100 / 0" branches of the kernels
// (MDF_kernel.cu:35,43,59,67; kernel.cu:86,102,110).  The reference builds the field on the host and
// ships the whole grid over PCIe every generation (MDF_kernel.cu:161); here it is born on the device.
//
// Shape of the fill kernels: blockIdx.y walks the rows (grid-stride), threads walk a row in 16-byte vectors; the value
// of a cell is a functor of (global plane, global row, column).  No per-element divide, every store is 16 bytes.
#include <algorithm>

#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

template <class T, int N>
struct alignas(sizeof(T) * N) FillVec {
    T v[N];
};

// Rows are `nplanes` x `rows_alloc` (a 2-D buffer is one plane); F::operator()(pz, r, c) -> T for local plane pz,
// local row r, column c.  VECTOR: the buffer is 16-byte aligned and pitch % (16 / sizeof(T)) == 0.
template <class T, class F, bool VECTOR>
__global__ void __launch_bounds__(256) fill_rows_kernel(T* __restrict__ u, long long pitch, long long plane_pitch,
                                                        int nplanes, int rows_alloc, F f) {
    constexpr int VPT = 16 / (int)sizeof(T);
    const int nrows = nplanes * rows_alloc;
    for (int row = blockIdx.y; row < nrows; row += gridDim.y) {
        const int pz = row / rows_alloc, r = row - pz * rows_alloc;  // one divide per ROW
        T* d = u + (long long)pz * plane_pitch + (long long)r * pitch;
        if (VECTOR) {
            const int nvec = (int)(pitch / VPT);
            for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
                FillVec<T, VPT> o;
#pragma unroll
                for (int k = 0; k < VPT; ++k) o.v[k] = f(pz, r, v * VPT + k);
                reinterpret_cast<FillVec<T, VPT>*>(d)[v] = o;
            }
        } else {
            for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < (int)pitch; c += gridDim.x * blockDim.x)
                d[c] = f(pz, r, c);
        }
    }
}

template <class T, class F>
int launch_fill(T* u, long long pitch, long long plane_pitch, int nplanes, int rows_alloc, const F& f, cudaStream_t s) {
    constexpr int VPT = 16 / (int)sizeof(T);
    const bool vec = ((uintptr_t)u % 16 == 0) && (pitch % VPT == 0) && (plane_pitch % VPT == 0);
    const long long per_row = vec ? pitch / VPT : pitch;
    const int gx = (int)std::max<long long>(1, std::min<long long>(8, (per_row + 255) / 256));
    const int gy = (int)std::max<long long>(1, std::min<long long>((long long)nplanes * rows_alloc, 148 * 16 / gx));
    if (vec)
        fill_rows_kernel<T, F, true><<<dim3(gx, gy), 256, 0, s>>>(u, pitch, plane_pitch, nplanes, rows_alloc, f);
    else
        fill_rows_kernel<T, F, false><<<dim3(gx, gy), 256, 0, s>>>(u, pitch, plane_pitch, nplanes, rows_alloc, f);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

__host__ __device__ __forceinline__ unsigned hash32(unsigned x, unsigned seed) {
    x ^= seed * 0x9E3779B9u;
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}

// ring = `ring`, interior = `interior`, cells outside the h x w domain (ghost rows of edge ranks, pad columns) = 0
template <class T>
struct RingFill {
    int row0, h, w;
    T ring, interior;
    __device__ __forceinline__ T operator()(int, int r, int c) const {
        const int gy = row0 + r;
        if (gy < 0 || gy >= h || c >= w) return T(0);
        return (gy == 0 || gy == h - 1 || c == 0 || c == w - 1) ? ring : interior;
    }
};

struct LifeHashFill {
    int row0, h, w;
    unsigned thr, seed;
    __device__ __forceinline__ int operator()(int, int r, int c) const {
        const int gy = row0 + r;
        if (gy <= 0 || gy >= h - 1 || c <= 0 || c >= w - 1) return 0;
        return hash32((unsigned)((long long)gy * w + c), seed) < thr ? 1 : 0;
    }
};

// Bit-packed variant: element e of a row is the 32-cell word of cells 32e .. 32e+31 (cell x = bit x % 32).
struct LifeHashBitsFill {
    int row0, h, w;
    unsigned thr, seed;
    __device__ __forceinline__ unsigned operator()(int, int r, int e) const {
        const int gy = row0 + r;
        unsigned word = 0u;
        if (gy > 0 && gy < h - 1) {
#pragma unroll 8
            for (int b = 0; b < 32; ++b) {
                const int c = e * 32 + b;
                if (c > 0 && c < w - 1 && hash32((unsigned)((long long)gy * w + c), seed) < thr) word |= 1u << b;
            }
        }
        return word;
    }
};

// Synthetic initial fields beyond the reference's ring (all exactly reproducible on the host, oracle.py):
// kind 1: interior cells uniform in [interior, ring) from hash32(i*w+j, seed) (24 random bits);
// kind 2: a hot box (value ring) over the central third of the domain, interior elsewhere.  Ring = ring either way.
template <class T>
struct ShapeFill {
    int row0, h, w, kind;
    unsigned seed;
    double ring, interior;
    __device__ __forceinline__ T operator()(int, int r, int c) const {
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
        return (T)v;
    }
};

template <class T>
struct ShellFill {
    int z0, y0, d, h, w;
    T ring, interior;
    __device__ __forceinline__ T operator()(int pz, int r, int c) const {
        const int gz = z0 + pz, gy = y0 + r;
        if (gz < 0 || gz >= d || gy < 0 || gy >= h || c >= w) return T(0);
        return (gz == 0 || gz == d - 1 || gy == 0 || gy == h - 1 || c == 0 || c == w - 1) ? ring : interior;
    }
};

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

static int grid_for(long long n) {
    long long b = (n + 255) / 256;
    return (int)(b < 1 ? 1 : (b > 148 * 16 ? 148 * 16 : b));
}

static unsigned prob_threshold(double prob) {
    const double t = prob * 4294967296.0;
    return t >= 4294967295.0 ? 0xFFFFFFFFu : (t <= 0 ? 0u : (unsigned)t);
}

}  // namespace mcp

using namespace mcp;

extern "C" {

int mcp_init_ring2d(int dtype, void* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                    double interior, void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_ring2d: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        return launch_fill((float*)u, pitch, 0, 1, rows_alloc, RingFill<float>{row0, h, w, (float)ring, (float)interior}, s);
    if (dtype == MCP_DTYPE_F64)
        return launch_fill((double*)u, pitch, 0, 1, rows_alloc, RingFill<double>{row0, h, w, ring, interior}, s);
    if (dtype == MCP_DTYPE_I32)
        return launch_fill((int*)u, pitch, 0, 1, rows_alloc, RingFill<int>{row0, h, w, (int)ring, (int)interior}, s);
    MCP_REQUIRE(false, "bad dtype %d", dtype);
}

int mcp_init_shape2d(int dtype, int kind, void* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                     double interior, uint32_t seed, void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_shape2d: bad arguments");
    MCP_REQUIRE(kind == 1 || kind == 2, "mcp_init_shape2d: kind must be 1 (hash) or 2 (box)");
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        return launch_fill((float*)u, pitch, 0, 1, rows_alloc, ShapeFill<float>{row0, h, w, kind, seed, ring, interior}, s);
    if (dtype == MCP_DTYPE_F64)
        return launch_fill((double*)u, pitch, 0, 1, rows_alloc, ShapeFill<double>{row0, h, w, kind, seed, ring, interior}, s);
    MCP_REQUIRE(false, "mcp_init_shape2d: bad dtype %d", dtype);
}

int mcp_init_life_hash(int32_t* u, int64_t pitch, int rows_alloc, int row0, int h, int w, double prob, uint32_t seed,
                       void* stream) {
    MCP_REQUIRE(u && pitch >= w && rows_alloc > 0, "mcp_init_life_hash: bad arguments");
    return launch_fill((int*)u, pitch, 0, 1, rows_alloc, LifeHashFill{row0, h, w, prob_threshold(prob), seed},
                       (cudaStream_t)stream);
}

int mcp_init_life_hash_bits(uint32_t* u, int64_t pitch_words, int rows_alloc, int row0, int h, int w, double prob,
                            uint32_t seed, void* stream) {
    MCP_REQUIRE(u && pitch_words * 32 >= w && rows_alloc > 0, "mcp_init_life_hash_bits: bad arguments");
    return launch_fill((unsigned*)u, pitch_words, 0, 1, rows_alloc,
                       LifeHashBitsFill{row0, h, w, prob_threshold(prob), seed}, (cudaStream_t)stream);
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
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == MCP_DTYPE_F32)
        return launch_fill((float*)u, pitch, plane_pitch, planes_alloc, rows_alloc,
                           ShellFill<float>{z0, y0, d, h, w, (float)ring, (float)interior}, s);
    if (dtype == MCP_DTYPE_F64)
        return launch_fill((double*)u, pitch, plane_pitch, planes_alloc, rows_alloc,
                           ShellFill<double>{z0, y0, d, h, w, ring, interior}, s);
    MCP_REQUIRE(false, "bad dtype %d", dtype);
}

}  // extern "C"
