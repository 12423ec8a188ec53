// Fused halo-exchange + 3-D 7-point heat kernel (pencil decomposition) for sm_100a.
//
// Extension named by BASELINE.json config 3 (3-D 7-point heat equation, 1024^3, 2-D pencil decomposition over
// 8 GPUs); the reference itself is strictly 2-D (MDF_kernel.cu:13-18).  Same communication idea as the 2-D
// kernel, which replaces the reference's MPI_Send/MPI_Recv + staging copies (MDF_kernel.cu:161-183): the CTA that
// computes a cell on a pencil face stores it twice -- locally and straight into the neighbour GPU's ghost layer
// through an NVLink-mapped pointer (z-faces: whole rows of a plane; y-faces: one strided row per plane, stored
// directly by the computing CTA, no pack kernel) -- then bumps a system-scope release counter; tiles that read a
// ghost layer acquire-spin on their local counter first.
//
// Compute scheme (2.5-D blocking): a CTA owns a tile of 32 rows x 128 columns (fp64: 64) and marches along z over
// a chunk of planes.  TMA (cp.async.bulk.tensor.3d) streams (32+2) x (128+8) boxes of successive planes into a
// 4-slot shared-memory ring with mbarrier completion; each thread keeps its 4 rows x 16 bytes of the planes
// z-1, z, z+1 in registers, so a plane is read from shared memory once.  One generation per pass (no temporal
// blocking: the 3-D working set does not fit the register file), device-side grid barrier between generations.
// Arithmetic: oracle.heat3d_update_cells, every operation individually rounded.
#include <algorithm>
#include <new>
#include <type_traits>
#include <vector>

#include "../../../include/mcp_b200.h"
#include "common.cuh"
#include "ptx.cuh"
#include "stencil_ops.cuh"

namespace mcp {

constexpr int k3Threads = 256;                 // 8 warps x 4 rows
constexpr int k3TileRows = 32;
constexpr int k3BoxRows = k3TileRows + 2;
constexpr int k3RowBytes = 512 + 32;           // 128 fp32 (64 fp64) cells + 16-byte halo on both sides
constexpr int k3Slots = 4;
constexpr int k3BoxBytes = k3BoxRows * k3RowBytes;             // 18496 bytes delivered by one TMA box
constexpr int k3SlotBytes = (k3BoxBytes + 127) / 128 * 128;   // ring slots stay 128-byte aligned for TMA
constexpr int k3SmemBytes = k3Slots * k3SlotBytes + 64;

struct Fused3dParams {
    void* buf[2];
    void* nbr_buf[4][2];  // z-, z+, y-, y+
    const unsigned long long* my_flags;
    unsigned long long* nbr_flag[4];
    unsigned long long flag_base[4];
    unsigned long long* grid_bar;
    unsigned long long* tile_ctr;
    unsigned long long bar_base, ctr_base, timeout_ns;
    int* status;
    const int* tiles;  // packed (zc << 20 | by << 10 | sx), boundary tiles first
    long long pitch, plane_pitch;
    int d, h, w;
    int z0, y0;         // global coordinates of local index 0 (i.e. of the low ghost layer)
    int nz, ny;         // owned extents; owned local indices are 1..nz, 1..ny
    int py_alloc, pz_alloc;
    int nbr_n[4];
    int expect[4];
    int n_tiles, zchunk;
    int generations, first_parity;
    double coeff;
};

__device__ __forceinline__ bool aborted3(const Fused3dParams& p) { return *((volatile int*)p.status) != 0; }

__device__ __noinline__ bool wait_ctr3(const Fused3dParams& p, const unsigned long long* ctr, unsigned long long need,
                                       bool sys) {
    auto ld = [&]() { return sys ? ptx::ld_acquire_sys(ctr) : ptx::ld_acquire_gpu(ctr); };
    if (ld() >= need) return true;
    const unsigned long long t0 = ptx::globaltimer_ns();
    unsigned spins = 0;
    while (ld() < need) {
        if ((++spins & 255u) == 0) {
            if (aborted3(p)) return false;
            if (ptx::globaltimer_ns() - t0 > p.timeout_ns) {
                atomicCAS(p.status, 0, MCP_ERR_TIMEOUT);
                return false;
            }
        }
    }
    return true;
}

template <class T>
struct V3;
template <>
struct V3<float> {
    using type = float4;
    static constexpr int VEC = 4;
};
template <>
struct V3<double> {
    using type = double2;
    static constexpr int VEC = 2;
};

template <class T>
__global__ void __launch_bounds__(k3Threads, 2)
heat3d_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ Fused3dParams p) {
    using V = typename V3<T>::type;
    constexpr int VEC = V3<T>::VEC;
    constexpr int XS = 32 * VEC;  // output columns per tile
    constexpr int HALO = 16 / (int)sizeof(T);

    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + k3Slots * k3SlotBytes);
    __shared__ int s_ticket, s_ok;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < k3Slots; ++s) ptx::mbar_init(&bars[s], 1);
        ptx::mbar_fence_init();
        ptx::prefetch_tensormap(&tmA);
        ptx::prefetch_tensormap(&tmB);
    }
    __syncthreads();

    uint32_t use = 0;  // number of slot uses so far (slot = use % k3Slots, parity = (use / k3Slots) & 1)
    unsigned long long epoch = 0;
    const T coeff = (T)p.coeff;

    for (int gen = 0; gen < p.generations; ++gen) {
        const int par = (p.first_parity + gen) & 1;
        const CUtensorMap* tm_in = par ? &tmB : &tmA;
        T* out = reinterpret_cast<T*>(p.buf[par ^ 1]);
        const unsigned long long ticket0 =
            p.ctr_base + (unsigned long long)gen * (unsigned long long)(p.n_tiles + gridDim.x);

        for (;;) {
            __syncthreads();  // everybody is done with s_ticket and with the ring of the previous tile
            if (tid == 0) {
                const unsigned long long v = atomicAdd(p.tile_ctr, 1ull) - ticket0;
                s_ticket = (v < (unsigned long long)p.n_tiles && !aborted3(p)) ? (int)v : -1;
            }
            __syncthreads();
            const int q = s_ticket;  // uniform across the CTA
            if (q < 0) break;
            const int packed = __ldg(&p.tiles[q]);
            const int zc = packed >> 20, by = (packed >> 10) & 1023, sx = packed & 1023;
            const int zlo = 1 + zc * p.zchunk, zhi = min(p.nz + 1, zlo + p.zchunk);  // owned planes [zlo, zhi)
            const int ylo = 1 + by * k3TileRows;                                      // first output row
            const int x0 = sx * XS;
            const bool f_zm = (zlo == 1), f_zp = (zhi == p.nz + 1);
            const bool f_ym = (ylo == 1), f_yp = (ylo + k3TileRows > p.ny);

            // ghost layers this tile reads were written by the neighbours during their previous generation
            if (tid == 0) {
                bool ok = true;
                const bool touch[4] = {f_zm, f_zp, f_ym, f_yp};
                for (int f = 0; f < 4 && ok; ++f)
                    if (touch[f] && p.nbr_buf[f][0] != nullptr)
                        ok = wait_ctr3(p, &p.my_flags[f],
                                       p.flag_base[f] + (unsigned long long)gen * (unsigned long long)p.expect[f], true);
                ptx::fence_proxy_async();
                s_ok = ok && !aborted3(p);
            }
            __syncthreads();
            if (!s_ok) continue;  // uniform: thread 0 decided

            auto issue = [&](uint32_t u, int pz) {
                const uint32_t st = u % k3Slots;
                ptx::mbar_arrive_expect_tx(&bars[st], k3BoxBytes);
                ptx::tma_load_3d(smem + st * k3SlotBytes, tm_in, &bars[st], x0 - HALO, ylo - 1, pz);
            };
            const int n_planes = (zhi - zlo) + 2;  // planes zlo-1 .. zhi
            if (tid == 0) {
                const int npro = n_planes < k3Slots ? n_planes : k3Slots;
                for (int i = 0; i < npro; ++i) issue(use + i, zlo - 1 + i);
            }

            // this thread: rows r0..r0+3 of the tile (box rows r0+1..r0+4), columns x0 + lane*VEC ..
            const int r0 = warp * 4;
            const int cx = x0 + lane * VEC;
            const int colb = (HALO + lane * VEC) * (int)sizeof(T);  // byte offset of the vector inside a box row
            unsigned xfix = 0;
#pragma unroll
            for (int k = 0; k < VEC; ++k) {
                const int gx = cx + k;
                if (gx <= 0 || gx >= p.w - 1) xfix |= 1u << k;
            }
            unsigned yfix = 0, ystore = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int ly = ylo + r0 + i;        // local row
                const int gy = p.y0 + ly;
                if (gy <= 0 || gy >= p.h - 1) yfix |= 1u << i;
                if (ly <= p.ny && cx < (int)p.pitch) ystore |= 1u << i;
            }

            V U[4], Cc[6], Dn[6];
            auto load6 = [&](uint32_t u, V (&dst)[6]) {
                const uint8_t* base = smem + (u % k3Slots) * k3SlotBytes + colb;
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const uint4 raw = ptx::ld_shared_v4(base + (r0 + i) * k3RowBytes);
                    memcpy(&dst[i], &raw, 16);
                }
            };
            // planes zlo-1 (-> U) and zlo (-> C)
            ptx::mbar_wait(&bars[use % k3Slots], (use / k3Slots) & 1u);
            load6(use, Cc);
#pragma unroll
            for (int i = 0; i < 4; ++i) U[i] = Cc[i + 1];
            ptx::mbar_wait(&bars[(use + 1) % k3Slots], ((use + 1) / k3Slots) & 1u);
            load6(use + 1, Cc);

            // GEN = the tile touches a pencil face (halo pushes) or holds fixed cells (masks); interior tiles
            // run the lean instantiation: no masks, no push tests, unconditional stores.
            auto zloop = [&](auto gen_tag) {
            constexpr bool GEN = decltype(gen_tag)::value;
            for (int s = 0; s < zhi - zlo; ++s) {
                const int pz = zlo + s;                // plane being computed
                const uint32_t uC = use + 1 + s, uD = use + 2 + s;
                ptx::mbar_wait(&bars[uD % k3Slots], (uD / k3Slots) & 1u);
                load6(uD, Dn);
                const int gz = p.z0 + pz;
                const bool zfix = (gz <= 0 || gz >= p.d - 1);
                const uint8_t* rowC = smem + (uC % k3Slots) * k3SlotBytes;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const T* cw = reinterpret_cast<const T*>(rowC + (r0 + i + 1) * k3RowBytes + colb);
                    const T wl = cw[-1], er = cw[VEC];
                    const T* pc = reinterpret_cast<const T*>(&Cc[i + 1]);
                    const T* pn = reinterpret_cast<const T*>(&Cc[i]);
                    const T* ps = reinterpret_cast<const T*>(&Cc[i + 2]);
                    const T* pu = reinterpret_cast<const T*>(&U[i]);
                    const T* pd = reinterpret_cast<const T*>(&Dn[i + 1]);
                    V o;
                    T* po = reinterpret_cast<T*>(&o);
                    const bool rfix = GEN && (zfix || ((yfix >> i) & 1u));
#pragma unroll
                    for (int k = 0; k < VEC; ++k) {
                        const T W = k == 0 ? wl : pc[k - 1];
                        const T E = k == VEC - 1 ? er : pc[k + 1];
                        const T v = heat3d_cell(pc[k], pn[k], ps[k], W, E, pu[k], pd[k], coeff);
                        po[k] = (GEN && (rfix || ((xfix >> k) & 1u))) ? pc[k] : v;
                    }
                    if (!GEN || ((ystore >> i) & 1u)) {
                        const int ly = ylo + r0 + i;
                        uint4 bits;
                        memcpy(&bits, &o, 16);
                        const long long off = (long long)ly * p.pitch + cx;
                        ptx::st_global_v4(out + (long long)pz * p.plane_pitch + off, bits);
                        // halo pushes: z-faces are whole rows of the first/last owned plane, y-faces one row per plane
                        if (GEN && pz == 1 && p.nbr_buf[0][0] != nullptr)
                            ptx::st_global_v4(reinterpret_cast<T*>(p.nbr_buf[0][par ^ 1]) +
                                                  (long long)(p.nbr_n[0] + 1) * p.plane_pitch + off, bits);
                        if (GEN && pz == p.nz && p.nbr_buf[1][0] != nullptr)
                            ptx::st_global_v4(reinterpret_cast<T*>(p.nbr_buf[1][par ^ 1]) + off, bits);
                        if (GEN && ly == 1 && p.nbr_buf[2][0] != nullptr)
                            ptx::st_global_v4(reinterpret_cast<T*>(p.nbr_buf[2][par ^ 1]) +
                                                  (long long)pz * p.plane_pitch +
                                                  (long long)(p.nbr_n[2] + 1) * p.pitch + cx, bits);
                        if (GEN && ly == p.ny && p.nbr_buf[3][0] != nullptr)
                            ptx::st_global_v4(reinterpret_cast<T*>(p.nbr_buf[3][par ^ 1]) +
                                                  (long long)pz * p.plane_pitch + cx, bits);
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) U[i] = Cc[i + 1];
#pragma unroll
                for (int i = 0; i < 6; ++i) Cc[i] = Dn[i];
                __syncthreads();  // slot of plane pz-1 (use + s) is free: nobody reads it any more
                if (tid == 0 && s + k3Slots < n_planes) issue(use + s + k3Slots, zlo - 1 + s + k3Slots);
            }
            };
            const bool lean = !f_zm && !f_zp && !f_ym && !f_yp && (p.z0 + zlo >= 1) && (p.z0 + zhi - 1 <= p.d - 2) &&
                              (p.y0 + ylo >= 1) && (p.y0 + ylo + k3TileRows - 1 <= p.h - 2) && (x0 >= 1) &&
                              (x0 + XS - 1 <= p.w - 2);
            if (lean) zloop(std::false_type{}); else zloop(std::true_type{});
            use += n_planes;

            // publish the pushed halo layers: all threads' peer stores -> one release per face
            __syncthreads();
            if (tid == 0) {
                const bool touch[4] = {f_zm, f_zp, f_ym, f_yp};
                bool fenced = false;  // one system-scope fence (a round trip to the peers), then relaxed signals
                for (int f = 0; f < 4; ++f)
                    if (touch[f] && p.nbr_buf[f][0] != nullptr) {
                        if (!fenced) ptx::fence_acq_rel_sys();
                        fenced = true;
                        ptx::red_add_relaxed_sys(p.nbr_flag[f], 1ull);
                    }
            }
        }

        if (gen + 1 < p.generations) {
            // grid barrier between generations
            __syncthreads();
            if (tid == 0) {
                ptx::fence_proxy_async();
                __threadfence();
                const unsigned long long target = p.bar_base + (epoch + 1ull) * gridDim.x;
                ptx::red_add_release_gpu(p.grid_bar, 1ull);
                wait_ctr3(p, p.grid_bar, target, false);
                ptx::fence_proxy_async();
            }
            __syncthreads();
            ++epoch;
        }
        if (aborted3(p)) break;
    }
}

}  // namespace mcp

using namespace mcp;

struct mcp_fused3d_plan {
    mcp_fused3d_desc d;
    CUtensorMap tm[2];
    unsigned long long* grid_bar;
    int* status;
    int* tiles_dev;
    int n_tiles, zchunk, grid, parity;
    unsigned long long flag_base[4], bar_base, ctr_base;
    int64_t launches;
};

namespace {

int vec3(int dtype) { return dtype == MCP_DTYPE_F64 ? 2 : 4; }

// Number of tiles of a block of (nz, ny) owned planes/rows that touch each face, for a given tiling.
void face_tile_counts(int nz, int ny, int w, int dtype, int zchunk, int out[4]) {
    const int nsx = (int)ceil_div(w, 32 * vec3(dtype));
    const int nby = (int)ceil_div(ny, k3TileRows);
    const int nzc = (int)ceil_div(nz, zchunk);
    out[0] = out[1] = nby * nsx;
    out[2] = out[3] = nzc * nsx;
}

}  // namespace

extern "C" {

int mcp_fused3d_create(const mcp_fused3d_desc* desc, mcp_fused3d_plan** plan_out) {
    MCP_REQUIRE(desc && plan_out, "mcp_fused3d_create: null argument");
    const mcp_fused3d_desc& d = *desc;
    MCP_REQUIRE(d.dtype == MCP_DTYPE_F32 || d.dtype == MCP_DTYPE_F64, "heat3d takes fp32 or fp64");
    MCP_REQUIRE(d.nz >= 1 && d.ny >= 1 && d.w >= 1, "empty block");
    MCP_REQUIRE(d.pitch >= d.w && d.pitch % 32 == 0, "pitch must be a multiple of 32 and >= w");
    MCP_REQUIRE(d.py_alloc >= d.ny + 2 && d.pz_alloc >= d.nz + 2, "allocation smaller than block + ghost layers");
    MCP_REQUIRE(d.buf[0] && d.buf[1], "null buffer");
    mcp_fused3d_plan* pl = new (std::nothrow) mcp_fused3d_plan();
    MCP_REQUIRE(pl, "out of host memory");
    pl->d = d;
    pl->parity = 0;
    pl->launches = 0;
    pl->bar_base = pl->ctr_base = 0;
    for (int f = 0; f < 4; ++f) pl->flag_base[f] = 0;
    pl->grid_bar = nullptr;
    pl->status = nullptr;
    pl->tiles_dev = nullptr;
    int sm = 0;
    int rc = device_sm_count(&sm);
    if (rc) { delete pl; return rc; }
    pl->grid = 2 * sm;
    pl->zchunk = d.zchunk > 0 ? d.zchunk : 64;

    const int esz = d.dtype == MCP_DTYPE_F64 ? 8 : 4;
    for (int b = 0; b < 2; ++b) {
        uint64_t dims[3] = {(uint64_t)d.pitch, (uint64_t)d.py_alloc, (uint64_t)d.pz_alloc};
        uint64_t strides[2] = {(uint64_t)d.pitch * esz, (uint64_t)d.pitch * d.py_alloc * esz};
        uint32_t box[3] = {(uint32_t)(k3RowBytes / esz), (uint32_t)k3BoxRows, 1u};
        rc = encode_tensor_map(&pl->tm[b], d.dtype, 3, d.buf[b], dims, strides, box);
        if (rc) { delete pl; return rc; }
    }
    // tile list, pencil-face tiles first so their pushes overlap the interior
    const int nsx = (int)ceil_div(d.w, 32 * vec3(d.dtype));
    const int nby = (int)ceil_div(d.ny, k3TileRows);
    const int nzc = (int)ceil_div(d.nz, pl->zchunk);
    MCP_REQUIRE(nsx < 1024 && nby < 1024 && nzc < 2048, "block too large for the tile encoding");
    std::vector<int> face, inner;
    for (int zc = 0; zc < nzc; ++zc)
        for (int by = 0; by < nby; ++by)
            for (int sx = 0; sx < nsx; ++sx) {
                const bool touches = zc == 0 || zc == nzc - 1 || by == 0 || by == nby - 1;
                (touches ? face : inner).push_back((zc << 20) | (by << 10) | sx);
            }
    face.insert(face.end(), inner.begin(), inner.end());
    pl->n_tiles = (int)face.size();
    cudaError_t e;
    if ((e = cudaMalloc(&pl->grid_bar, 512)) != cudaSuccess || (e = cudaMemset(pl->grid_bar, 0, 512)) != cudaSuccess ||
        (e = cudaMalloc(&pl->status, 256)) != cudaSuccess || (e = cudaMemset(pl->status, 0, 256)) != cudaSuccess ||
        (e = cudaMalloc(&pl->tiles_dev, face.size() * sizeof(int))) != cudaSuccess ||
        (e = cudaMemcpy(pl->tiles_dev, face.data(), face.size() * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess) {
        set_error("mcp_fused3d_create: scratch allocation failed: %s", cudaGetErrorString(e));
        mcp_fused3d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    *plan_out = pl;
    return MCP_OK;
}

int mcp_fused3d_face_tiles(int nz, int ny, int w, int dtype, int zchunk, int* counts4) {
    MCP_REQUIRE(counts4, "null argument");
    face_tile_counts(nz, ny, w, dtype, zchunk > 0 ? zchunk : 64, counts4);
    return MCP_OK;
}

int mcp_fused3d_run(mcp_fused3d_plan* pl, int generations, void* stream_) {
    MCP_REQUIRE(pl && generations >= 0, "mcp_fused3d_run: bad arguments");
    if (generations == 0) return MCP_OK;
    const mcp_fused3d_desc& d = pl->d;
    Fused3dParams p;
    memset(&p, 0, sizeof(p));
    for (int b = 0; b < 2; ++b) {
        p.buf[b] = d.buf[b];
        for (int f = 0; f < 4; ++f) p.nbr_buf[f][b] = d.nbr_buf[f][b];
    }
    p.my_flags = (const unsigned long long*)d.flags;
    for (int f = 0; f < 4; ++f) {
        p.nbr_flag[f] = (unsigned long long*)d.nbr_flag[f];
        p.flag_base[f] = pl->flag_base[f];
        p.nbr_n[f] = d.nbr_n[f];
        p.expect[f] = d.expect[f];
    }
    p.grid_bar = pl->grid_bar;
    p.tile_ctr = pl->grid_bar + 16;
    p.bar_base = pl->bar_base;
    p.ctr_base = pl->ctr_base;
    p.timeout_ns = 20ull * 1000000000ull;
    p.status = pl->status;
    p.tiles = pl->tiles_dev;
    p.pitch = d.pitch;
    p.plane_pitch = d.pitch * d.py_alloc;
    p.d = d.d; p.h = d.h; p.w = d.w;
    p.z0 = d.z_start - 1;
    p.y0 = d.y_start - 1;
    p.nz = d.nz; p.ny = d.ny;
    p.py_alloc = d.py_alloc; p.pz_alloc = d.pz_alloc;
    p.n_tiles = pl->n_tiles;
    p.zchunk = pl->zchunk;
    p.generations = generations;
    p.first_parity = pl->parity;
    p.coeff = d.coeff;

    void* args[3] = {(void*)&pl->tm[0], (void*)&pl->tm[1], (void*)&p};
    const void* fn = d.dtype == MCP_DTYPE_F64 ? (const void*)heat3d_fused_kernel<double>
                                              : (const void*)heat3d_fused_kernel<float>;
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, k3SmemBytes));
    int per_sm = 0;
    MCP_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, k3Threads, k3SmemBytes));
    MCP_REQUIRE(per_sm >= 1, "heat3d kernel does not fit on an SM");
    const int grid = std::min(pl->grid, per_sm * (pl->grid / 2));
    MCP_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(k3Threads), args, k3SmemBytes, (cudaStream_t)stream_));
    pl->launches++;
    pl->bar_base += (unsigned long long)(generations - 1) * grid;
    pl->ctr_base += (unsigned long long)generations * (unsigned long long)(pl->n_tiles + grid);
    for (int f = 0; f < 4; ++f) pl->flag_base[f] += (unsigned long long)generations * d.expect[f];
    pl->parity = (pl->parity + generations) & 1;
    return MCP_OK;
}

int mcp_fused3d_parity(const mcp_fused3d_plan* pl) { return pl ? pl->parity : 0; }

int mcp_fused3d_set_parity(mcp_fused3d_plan* pl, int parity) {
    MCP_REQUIRE(pl, "null plan");
    pl->parity = parity & 1;
    return MCP_OK;
}

int mcp_fused3d_status(mcp_fused3d_plan* pl) {
    MCP_REQUIRE(pl, "null plan");
    int st = 0;
    MCP_CUDA_TRY(cudaMemcpy(&st, pl->status, sizeof(int), cudaMemcpyDeviceToHost));
    if (st != 0) set_error("fused3d kernel recorded status %d", st);
    return st;
}

int64_t mcp_fused3d_launches(const mcp_fused3d_plan* pl) { return pl ? pl->launches : 0; }

int mcp_fused3d_destroy(mcp_fused3d_plan* pl) {
    if (!pl) return MCP_OK;
    if (pl->grid_bar) cudaFree(pl->grid_bar);
    if (pl->status) cudaFree(pl->status);
    if (pl->tiles_dev) cudaFree(pl->tiles_dev);
    delete pl;
    return MCP_OK;
}

}  // extern "C"
