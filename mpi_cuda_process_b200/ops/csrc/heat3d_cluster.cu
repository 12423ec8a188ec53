// 3-D 7-point heat step with thread-block clusters: two CTAs that own y-adjacent tiles form a cluster and share the
// tile-boundary rows through distributed shared memory (DSMEM) instead of both loading them.
//
// Extension (the reference is 2-D and uses no shared memory at all, MDF_kernel.cu:10-22); built because the north star
// of BASELINE.json names "thread-block clusters share tile halos through DSMEM", and measured against the fused kernel
// of heat3d_fused.cu, whose CTAs each TMA-load a (32+2)-row box per plane (profiles/r2/heat3d_cluster.md).
//
// One launch = one generation of local planes [z_lo, z_hi) x rows [y_lo, y_hi) (same contract as mcp_step3d).
//  * cluster (1,2,1): CTA rank 0 owns rows [y, y+32), rank 1 rows [y+32, y+64) of one (x-strip, z-chunk).
//  * per plane each CTA TMA-loads 33 rows: its 32 rows and the ONE halo row on its outer side; the halo row on the
//    inner side is the partner's first / last row, read from the partner's shared memory (mapa + ld.shared::cluster).
//    All tensor maps here cover this GPU's memory only; there is no multicast.
//  * one barrier.cluster per plane orders "both boxes have landed" before the remote reads and "both CTAs are done with
//    plane j-2" before its ring slot is refilled; it also replaces the CTA barrier.
#include <algorithm>

#include "../../../include/mcp_b200.h"
#include "common.cuh"
#include "ptx.cuh"
#include "stencil_ops.cuh"

namespace mcp {

constexpr int kcThreads = 256;  // 8 warps x 4 rows
constexpr int kcTileRows = 32;
constexpr int kcBoxRows = kcTileRows + 1;
constexpr int kcRowBytes = 512 + 32;  // 128 fp32 (64 fp64) cells + 16-byte halo on both sides
constexpr int kcSlots = 4;
constexpr int kcBoxBytes = kcBoxRows * kcRowBytes;
constexpr int kcSlotBytes = (kcBoxBytes + 127) / 128 * 128;
constexpr int kcSmemBytes = kcSlots * kcSlotBytes + 64;

struct Cluster3dParams {
    long long pitch, plane_pitch;
    int planes_alloc, rows_alloc;
    int z0, y0, d, h, w;
    int z_lo, z_hi, y_lo, y_hi;
    int z_per_block;
    double coeff;
};

__device__ __forceinline__ unsigned cluster_ctarank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `local_smem_addr` in the shared memory of CTA `rank` of this cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_smem_addr, unsigned rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ uint4 ld_dsmem_v4(uint32_t cluster_addr) {
    uint4 v;
    asm volatile("ld.shared::cluster.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(cluster_addr)
                 : "memory");
    return v;
}

template <class T>
struct VC;
template <>
struct VC<float> {
    using type = float4;
    static constexpr int VEC = 4;
};
template <>
struct VC<double> {
    using type = double2;
    static constexpr int VEC = 2;
};

template <class T>
__global__ void __cluster_dims__(1, 2, 1) __launch_bounds__(kcThreads, 2)
step3d_cluster_kernel(const __grid_constant__ CUtensorMap tm_in, T* __restrict__ out,
                      const __grid_constant__ Cluster3dParams p) {
    using V = typename VC<T>::type;
    constexpr int VEC = VC<T>::VEC;
    constexpr int XS = 32 * VEC;
    constexpr int HALO = 16 / (int)sizeof(T);

    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kcSlots * kcSlotBytes);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned rank = cluster_ctarank();  // 0 = upper tile of the pair, 1 = lower tile

    if (tid == 0) {
        for (int s = 0; s < kcSlots; ++s) ptx::mbar_init(&bars[s], 1);
        ptx::mbar_fence_init();
        ptx::prefetch_tensormap(&tm_in);
    }
    cluster_sync_all();  // barriers initialised in both CTAs before anything is issued

    const int x0 = blockIdx.x * XS;
    const int ylo = p.y_lo + (int)blockIdx.y * kcTileRows;  // first own row (blockIdx.y = 2 * pair + rank)
    const int zb = p.z_lo + (int)blockIdx.z * p.z_per_block;
    const int ze = min(p.z_hi, zb + p.z_per_block);
    const int n_planes = (ze - zb) + 2;  // planes zb-1 .. ze (uniform over the cluster)
    // box row b of this CTA is local row ylo - 1 + b (rank 0: outer halo row first) or ylo + b (rank 1: it comes last)
    const int box_y = rank == 0 ? ylo - 1 : ylo;
    const int own0 = rank == 0 ? 1 : 0;  // box row of own row 0

    auto issue = [&](int j) {  // plane zb - 1 + j into slot j % kcSlots
        const uint32_t st = (uint32_t)j % kcSlots;
        ptx::mbar_arrive_expect_tx(&bars[st], kcBoxBytes);
        ptx::tma_load_3d(smem + st * kcSlotBytes, &tm_in, &bars[st], x0 - HALO, box_y, zb - 1 + j);
    };
    if (tid == 0)
        for (int j = 0; j < min(n_planes, 3); ++j) issue(j);

    const int r0 = warp * 4;  // own rows r0 .. r0+3
    const int cx = x0 + lane * VEC;
    const int colb = (HALO + lane * VEC) * (int)sizeof(T);
    unsigned xfix = 0, yfix = 0, ystore = 0;
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
        const int gx = cx + k;
        if (gx <= 0 || gx >= p.w - 1) xfix |= 1u << k;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ly = ylo + r0 + i, gy = p.y0 + ly;
        if (gy <= 0 || gy >= p.h - 1) yfix |= 1u << i;
        if (ly < p.y_hi && cx < (int)p.pitch) ystore |= 1u << i;
    }
    const uint32_t smem_base = ptx::smem_u32(smem);
    const uint32_t partner_base = mapa_shared(smem_base, rank ^ 1u);

    // rows r0-1 .. r0+4 of plane j: own box, except the one row that belongs to the partner's tile
    auto load6 = [&](int j, V (&dst)[6]) {
        const uint32_t slot_off = ((uint32_t)j % kcSlots) * kcSlotBytes + colb;
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const int row = r0 - 1 + i;  // own-row index, -1 .. 32
            uint4 raw;
            if (rank == 0 && row == kcTileRows)      // below our tile: the partner's own row 0 (its box row 0)
                raw = ld_dsmem_v4(partner_base + slot_off + 0 * kcRowBytes);
            else if (rank == 1 && row == -1)         // above our tile: the partner's own row 31 (its box row 32)
                raw = ld_dsmem_v4(partner_base + slot_off + kcTileRows * kcRowBytes);
            else
                raw = ptx::ld_shared_v4(smem + slot_off + (row + own0) * kcRowBytes);
            memcpy(&dst[i], &raw, 16);
        }
    };

    const T coeff = (T)p.coeff;
    V U[4], Cc[6], Dn[6];
    // planes 0 (-> U) and 1 (-> C)
    ptx::mbar_wait(&bars[0], 0);
    ptx::mbar_wait(&bars[1 % kcSlots], 0);
    cluster_sync_all();  // both CTAs' planes 0 and 1 have landed
    load6(0, Cc);
#pragma unroll
    for (int i = 0; i < 4; ++i) U[i] = Cc[i + 1];
    load6(1, Cc);

    for (int s = 0; s < ze - zb; ++s) {
        const int pz = zb + s;   // plane being computed
        const int jD = s + 2;    // ring index of plane pz + 1
        ptx::mbar_wait(&bars[jD % kcSlots], (jD / kcSlots) & 1u);
        // Both CTAs hold plane jD, and every thread of the cluster has finished iteration s - 1 (its reads of planes
        // <= jD - 1 from shared memory): the slot of plane jD - 3 may be refilled.
        cluster_sync_all();
        if (tid == 0 && jD + 1 < n_planes) issue(jD + 1);
        load6(jD, Dn);
        const int gz = p.z0 + pz;
        const bool zfix = (gz <= 0 || gz >= p.d - 1);
        const uint8_t* rowC = smem + ((s + 1) % kcSlots) * kcSlotBytes;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const T* cw = reinterpret_cast<const T*>(rowC + (r0 + i + own0) * kcRowBytes + colb);
            const T wl = cw[-1], er = cw[VEC];
            const T* pc = reinterpret_cast<const T*>(&Cc[i + 1]);
            const T* pn = reinterpret_cast<const T*>(&Cc[i]);
            const T* ps = reinterpret_cast<const T*>(&Cc[i + 2]);
            const T* pu = reinterpret_cast<const T*>(&U[i]);
            const T* pd = reinterpret_cast<const T*>(&Dn[i + 1]);
            V o;
            T* po = reinterpret_cast<T*>(&o);
            const bool rfix = zfix || ((yfix >> i) & 1u);
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                const T W = k == 0 ? wl : pc[k - 1];
                const T E = k == VEC - 1 ? er : pc[k + 1];
                const T v = heat3d_cell(pc[k], pn[k], ps[k], W, E, pu[k], pd[k], coeff);
                po[k] = (rfix || ((xfix >> k) & 1u)) ? pc[k] : v;
            }
            if ((ystore >> i) & 1u) {
                uint4 bits;
                memcpy(&bits, &o, 16);
                ptx::st_global_v4(out + (long long)pz * p.plane_pitch + (long long)(ylo + r0 + i) * p.pitch + cx, bits);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) U[i] = Cc[i + 1];
#pragma unroll
        for (int i = 0; i < 6; ++i) Cc[i] = Dn[i];
    }
    cluster_sync_all();  // nobody leaves while its partner may still read its shared memory
}

}  // namespace mcp

using namespace mcp;

extern "C" int mcp_step3d_cluster(int dtype, const void* in, void* out, int64_t pitch, int64_t plane_pitch,
                                  int planes_alloc, int rows_alloc, int z0, int y0, int d, int h, int w, int z_lo,
                                  int z_hi, int y_lo, int y_hi, double coeff, void* stream) {
    MCP_REQUIRE(in && out && in != out, "mcp_step3d_cluster: null or aliased buffers");
    MCP_REQUIRE(dtype == MCP_DTYPE_F32 || dtype == MCP_DTYPE_F64, "mcp_step3d_cluster: fp32 or fp64");
    MCP_REQUIRE(pitch % 32 == 0 && pitch >= w && plane_pitch == pitch * rows_alloc,
                "mcp_step3d_cluster: pitch must be a multiple of 32 and planes dense (plane_pitch = pitch * rows_alloc)");
    MCP_REQUIRE(z_lo >= 1 && z_hi <= planes_alloc - 1 && y_lo >= 1 && y_hi <= rows_alloc - 1,
                "mcp_step3d_cluster: the updated box needs one layer of allocated cells around it");
    MCP_REQUIRE(((uintptr_t)in % 128 == 0), "mcp_step3d_cluster: input must be 128-byte aligned");
    if (z_lo >= z_hi || y_lo >= y_hi) return MCP_OK;
    const int esz = dtype == MCP_DTYPE_F64 ? 8 : 4;
    CUtensorMap tm;
    uint64_t dims[3] = {(uint64_t)pitch, (uint64_t)rows_alloc, (uint64_t)planes_alloc};
    uint64_t strides[2] = {(uint64_t)pitch * esz, (uint64_t)plane_pitch * esz};
    uint32_t box[3] = {(uint32_t)(kcRowBytes / esz), (uint32_t)kcBoxRows, 1u};
    int rc = encode_tensor_map(&tm, dtype, 3, const_cast<void*>(in), dims, strides, box);
    if (rc) return rc;
    Cluster3dParams p;
    p.pitch = pitch; p.plane_pitch = plane_pitch;
    p.planes_alloc = planes_alloc; p.rows_alloc = rows_alloc;
    p.z0 = z0; p.y0 = y0; p.d = d; p.h = h; p.w = w;
    p.z_lo = z_lo; p.z_hi = z_hi; p.y_lo = y_lo; p.y_hi = y_hi;
    p.coeff = coeff;
    const int vec = dtype == MCP_DTYPE_F64 ? 2 : 4;
    const int nsx = (int)ceil_div(w, 32 * vec);
    const int nby = (int)ceil_div(y_hi - y_lo, kcTileRows);
    const int nz = z_hi - z_lo;
    p.z_per_block = std::max(16, (int)ceil_div(nz, 8));
    dim3 grid((unsigned)nsx, (unsigned)((nby + 1) / 2 * 2), (unsigned)ceil_div(nz, p.z_per_block));
    const void* fn = dtype == MCP_DTYPE_F64 ? (const void*)step3d_cluster_kernel<double>
                                            : (const void*)step3d_cluster_kernel<float>;
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kcSmemBytes));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kcThreads);
    cfg.dynamicSmemBytes = kcSmemBytes;
    cfg.stream = (cudaStream_t)stream;
    void* outp = out;
    void* args[3] = {(void*)&tm, (void*)&outp, (void*)&p};
    MCP_CUDA_TRY(cudaLaunchKernelExC(&cfg, fn, args));  // cluster dims come from __cluster_dims__
    return MCP_OK;
}
