// Plain one-generation kernels: one launch = one generation, no communication inside.
// These are the baseline path (NCCL-halo and host-staged builds, and the 3-D workload's first version):
// the function of the reference's middle_kernel/border_kernel (MDF_kernel.cu:24-70, kernel.cu:70-113)
// without its defects (out-of-bounds reads D7, write-write race D9, floor-divided grid D6).
// 16-byte vector accesses, each thread marches a few rows and keeps a 3-row window in registers.
#include "../../../include/mcp_b200.h"
#include "common.cuh"
#include "stencil_ops.cuh"

namespace mcp {

constexpr int kRowsPerThread = 8;

template <class T, int VEC>
struct Vec;
template <>
struct Vec<float, 4> {
    using type = float4;
};
template <>
struct Vec<double, 2> {
    using type = double2;
};
template <>
struct Vec<int, 4> {
    using type = int4;
};

template <class T, int VEC>
struct RowRegs {
    T c[VEC + 2];  // [0] = left neighbour, [1..VEC] = cells, [VEC+1] = right neighbour
};

template <class T, int VEC>
__device__ __forceinline__ RowRegs<T, VEC> load_row(const T* __restrict__ in, long long pitch, int rows_alloc, int r,
                                                    int cx) {
    RowRegs<T, VEC> o;
#pragma unroll
    for (int k = 0; k < VEC + 2; ++k) o.c[k] = T(0);
    if (r < 0 || r >= rows_alloc) return o;
    const T* row = in + (long long)r * pitch;
    using V = typename Vec<T, VEC>::type;
    const V v = *reinterpret_cast<const V*>(row + cx);
    const T* pv = reinterpret_cast<const T*>(&v);
#pragma unroll
    for (int k = 0; k < VEC; ++k) o.c[k + 1] = pv[k];
    if (cx > 0) o.c[0] = row[cx - 1];
    if (cx + VEC < pitch) o.c[VEC + 1] = row[cx + VEC];
    return o;
}

struct PlainMdfF32 {
    using T = float;
    static constexpr int VEC = 4;
    static __device__ __forceinline__ T cell(const RowRegs<T, VEC>& n, const RowRegs<T, VEC>& c,
                                             const RowRegs<T, VEC>& s, int k) {
        return mdf_cell(c.c[k], n.c[k], s.c[k], c.c[k - 1], c.c[k + 1]);
    }
};
struct PlainMdfF32Strict {
    using T = float;
    static constexpr int VEC = 4;
    static __device__ __forceinline__ T cell(const RowRegs<T, VEC>& n, const RowRegs<T, VEC>& c,
                                             const RowRegs<T, VEC>& s, int k) {
        return mdf_cell_strict(c.c[k], n.c[k], s.c[k], c.c[k - 1], c.c[k + 1]);
    }
};
struct PlainMdfF64 {
    using T = double;
    static constexpr int VEC = 2;
    static __device__ __forceinline__ T cell(const RowRegs<T, VEC>& n, const RowRegs<T, VEC>& c,
                                             const RowRegs<T, VEC>& s, int k) {
        return mdf_cell(c.c[k], n.c[k], s.c[k], c.c[k - 1], c.c[k + 1]);
    }
};
struct PlainLife {
    using T = int;
    static constexpr int VEC = 4;
    static __device__ __forceinline__ T cell(const RowRegs<T, VEC>& n, const RowRegs<T, VEC>& c,
                                             const RowRegs<T, VEC>& s, int k) {
        const int cnt = n.c[k - 1] + n.c[k] + n.c[k + 1] + c.c[k - 1] + c.c[k + 1] + s.c[k - 1] + s.c[k] + s.c[k + 1];
        return life_rule(cnt, c.c[k]);
    }
};

template <class Op>
__global__ void __launch_bounds__(256)
step2d_kernel(const typename Op::T* __restrict__ in, typename Op::T* __restrict__ out, long long pitch,
              int rows_alloc, int row0, int row_lo, int row_hi, int ylo, int yhi, int xlo, int xhi) {
    using T = typename Op::T;
    constexpr int VEC = Op::VEC;
    using V = typename Vec<T, VEC>::type;
    const int cx = (blockIdx.x * blockDim.x + threadIdx.x) * VEC;
    if (cx >= pitch) return;
    const int r_begin = row_lo + blockIdx.y * kRowsPerThread;
    const int r_end = min(row_hi, r_begin + kRowsPerThread);
    if (r_begin >= r_end) return;
    RowRegs<T, VEC> n = load_row<T, VEC>(in, pitch, rows_alloc, r_begin - 1, cx);
    RowRegs<T, VEC> c = load_row<T, VEC>(in, pitch, rows_alloc, r_begin, cx);
    for (int r = r_begin; r < r_end; ++r) {
        const RowRegs<T, VEC> s = load_row<T, VEC>(in, pitch, rows_alloc, r + 1, cx);
        const int gy = row0 + r;
        const bool rowfix = gy < ylo || gy > yhi;
        V o;
        T* po = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const int gx = cx + k;
            const bool fix = rowfix || gx < xlo || gx > xhi;
            po[k] = fix ? c.c[k + 1] : Op::cell(n, c, s, k + 1);
        }
        *reinterpret_cast<V*>(out + (long long)r * pitch + cx) = o;
        n = c;
        c = s;
    }
}

// 3-D 7-point: thread = VEC cells in x at one (y); marches along z keeping U/C/D in registers.
template <class T, int VEC>
__global__ void __launch_bounds__(256)
step3d_kernel(const T* __restrict__ in, T* __restrict__ out, long long pitch, long long plane_pitch, int planes_alloc,
              int rows_alloc, int z0, int y0, int d, int h, int w, int z_lo, int z_hi, int y_lo, int y_hi, T coeff,
              int z_per_block) {
    using V = typename Vec<T, VEC>::type;
    const int cx = (blockIdx.x * blockDim.x + threadIdx.x) * VEC;
    const int r = y_lo + blockIdx.y * blockDim.y + threadIdx.y;
    if (cx >= pitch || r >= y_hi) return;
    const int zb = z_lo + blockIdx.z * z_per_block;
    const int ze = min(z_hi, zb + z_per_block);
    if (zb >= ze) return;
    const int gy = y0 + r;
    auto ldv = [&](int pz) -> V {
        if (pz < 0 || pz >= planes_alloc) {
            V z;
            T* pz_ = reinterpret_cast<T*>(&z);
#pragma unroll
            for (int k = 0; k < VEC; ++k) pz_[k] = T(0);
            return z;
        }
        return *reinterpret_cast<const V*>(in + (long long)pz * plane_pitch + (long long)r * pitch + cx);
    };
    V U = ldv(zb - 1), C = ldv(zb);
    for (int pz = zb; pz < ze; ++pz) {
        const V D = ldv(pz + 1);
        const T* plane = in + (long long)pz * plane_pitch;
        const RowRegs<T, VEC> cr = load_row<T, VEC>(plane, pitch, rows_alloc, r, cx);
        V N, S;
        {
            T* pn = reinterpret_cast<T*>(&N);
            T* ps = reinterpret_cast<T*>(&S);
#pragma unroll
            for (int k = 0; k < VEC; ++k) pn[k] = ps[k] = T(0);
            if (r - 1 >= 0) N = *reinterpret_cast<const V*>(plane + (long long)(r - 1) * pitch + cx);
            if (r + 1 < rows_alloc) S = *reinterpret_cast<const V*>(plane + (long long)(r + 1) * pitch + cx);
        }
        const int gz = z0 + pz;
        const bool planefix = gz <= 0 || gz >= d - 1 || gy <= 0 || gy >= h - 1;
        V o;
        T* po = reinterpret_cast<T*>(&o);
        const T* pU = reinterpret_cast<const T*>(&U);
        const T* pD = reinterpret_cast<const T*>(&D);
        const T* pN = reinterpret_cast<const T*>(&N);
        const T* pS = reinterpret_cast<const T*>(&S);
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const int gx = cx + k;
            const bool fix = planefix || gx <= 0 || gx >= w - 1;
            po[k] = fix ? cr.c[k + 1]
                        : heat3d_cell(cr.c[k + 1], pN[k], pS[k], cr.c[k], cr.c[k + 2], pU[k], pD[k], coeff);
        }
        *reinterpret_cast<V*>(out + (long long)pz * plane_pitch + (long long)r * pitch + cx) = o;
        U = C;
        C = D;
    }
}

template <class Op>
int launch_step2d(const void* in, void* out, int64_t pitch, int rows_alloc, int row0, int row_lo, int row_hi, int ylo,
                  int yhi, int xlo, int xhi, cudaStream_t s) {
    using T = typename Op::T;
    const int nvec = (int)(pitch / Op::VEC);
    dim3 block(256), grid((unsigned)ceil_div(nvec, 256), (unsigned)ceil_div(row_hi - row_lo, kRowsPerThread));
    step2d_kernel<Op><<<grid, block, 0, s>>>((const T*)in, (T*)out, pitch, rows_alloc, row0, row_lo, row_hi, ylo, yhi,
                                             xlo, xhi);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // namespace mcp

using namespace mcp;

extern "C" {

int mcp_step2d(int op, int dtype, const void* in, void* out, int64_t pitch, int rows_alloc, int row0, int h, int w,
               int row_lo, int row_hi, int ylo, int yhi, int xlo, int xhi, void* stream) {
    MCP_REQUIRE(in && out && in != out, "mcp_step2d: null or aliased buffers");
    MCP_REQUIRE(pitch % 4 == 0 && pitch >= w, "mcp_step2d: pitch must be a multiple of 4 and >= w");
    MCP_REQUIRE(row_lo >= 0 && row_hi <= rows_alloc && row_lo <= row_hi, "mcp_step2d: bad row range");
    (void)h;
    if (row_lo == row_hi) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int strict = op & 0x100;
    op &= 0xff;
    if (op == MCP_OP_MDF && dtype == MCP_DTYPE_F32)
        return strict ? launch_step2d<PlainMdfF32Strict>(in, out, pitch, rows_alloc, row0, row_lo, row_hi, ylo, yhi,
                                                         xlo, xhi, s)
                      : launch_step2d<PlainMdfF32>(in, out, pitch, rows_alloc, row0, row_lo, row_hi, ylo, yhi, xlo,
                                                   xhi, s);
    if (op == MCP_OP_MDF && dtype == MCP_DTYPE_F64)
        return launch_step2d<PlainMdfF64>(in, out, pitch, rows_alloc, row0, row_lo, row_hi, ylo, yhi, xlo, xhi, s);
    if (op == MCP_OP_LIFE && dtype == MCP_DTYPE_I32)
        return launch_step2d<PlainLife>(in, out, pitch, rows_alloc, row0, row_lo, row_hi, ylo, yhi, xlo, xhi, s);
    set_error("mcp_step2d: unsupported op/dtype %d/%d", op, dtype);
    return MCP_ERR_UNSUPPORTED;
}

int mcp_step3d(int dtype, const void* in, void* out, int64_t pitch, int64_t plane_pitch, int planes_alloc,
               int rows_alloc, int z0, int y0, int d, int h, int w, int z_lo, int z_hi, int y_lo, int y_hi, double coeff,
               void* stream) {
    MCP_REQUIRE(in && out && in != out, "mcp_step3d: null or aliased buffers");
    MCP_REQUIRE(pitch % 4 == 0 && pitch >= w && plane_pitch >= pitch * rows_alloc, "mcp_step3d: bad pitches");
    MCP_REQUIRE(z_lo >= 0 && z_hi <= planes_alloc && y_lo >= 0 && y_hi <= rows_alloc, "mcp_step3d: bad ranges");
    if (z_lo >= z_hi || y_lo >= y_hi) return MCP_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int vec = dtype == MCP_DTYPE_F64 ? 2 : 4;
    const int nvec = (int)(pitch / vec);
    dim3 block(64, 4);
    const int nz = z_hi - z_lo;
    // enough z-blocks to fill the machine, each marching >= 8 planes
    int zpb = std::max(8, (int)ceil_div(nz, 64));
    dim3 grid((unsigned)ceil_div(nvec, 64), (unsigned)ceil_div(y_hi - y_lo, 4), (unsigned)ceil_div(nz, zpb));
    if (dtype == MCP_DTYPE_F32)
        step3d_kernel<float, 4><<<grid, block, 0, s>>>((const float*)in, (float*)out, pitch, plane_pitch, planes_alloc,
                                                       rows_alloc, z0, y0, d, h, w, z_lo, z_hi, y_lo, y_hi,
                                                       (float)coeff, zpb);
    else if (dtype == MCP_DTYPE_F64)
        step3d_kernel<double, 2><<<grid, block, 0, s>>>((const double*)in, (double*)out, pitch, plane_pitch,
                                                        planes_alloc, rows_alloc, z0, y0, d, h, w, z_lo, z_hi, y_lo,
                                                        y_hi, coeff, zpb);
    else
        MCP_REQUIRE(false, "mcp_step3d: bad dtype %d", dtype);
    MCP_CUDA_TRY(cudaGetLastError());
    return MCP_OK;
}

}  // extern "C"
