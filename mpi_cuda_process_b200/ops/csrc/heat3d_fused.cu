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
// z-1, z, z+1 in registers, so a plane is read from shared memory once (plus the two neighbour cells of a lane's
// vector).  One generation per pass (no temporal
// blocking: the 3-D working set does not fit the register file), device-side grid barrier between generations.
// Optional fused residual: per-tile (sum d^2, max |d|) -> fixed-order reduction by CTA 0 behind the grid barrier ->
// all-reduce over the peers' inboxes (NVLink), history + stop decision on the device (run-to-tolerance in one launch).
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
#ifndef MCP_H3_SLOTS
#define MCP_H3_SLOTS 4
#endif
#ifndef MCP_H3_PERIOD
#define MCP_H3_PERIOD 1
#endif
// Plane ring and planes per __syncthreads (the ring is refilled k3Period slots at a time).  Measured on B200
// (profiles/r2/heat3d_ring.md): 4 slots / 1 plane per barrier is the fastest; 6 slots halve or third the barriers but
// the larger shared-memory footprint costs more than the barriers saved.
constexpr int k3Slots = MCP_H3_SLOTS;
constexpr int k3Period = MCP_H3_PERIOD;
constexpr int k3BoxBytes = k3BoxRows * k3RowBytes;             // 18496 bytes delivered by one TMA box
constexpr int k3SlotBytes = (k3BoxBytes + 127) / 128 * 128;   // ring slots stay 128-byte aligned for TMA
constexpr int k3SmemBytes = k3Slots * k3SlotBytes + 64;
constexpr int k3MaxRanks = MCP_MAX_RANKS;
constexpr int k3ResRing = MCP_RES_RING;
constexpr int k3MaxResSlots = 16384;

// One 32-byte block per plan that a host reads back with ONE copy after a run (same layout as the 2-D kernel's).
struct Result3d {
    int status;       // 0 or the MCP_ERR_* code a kernel recorded
    int stop_word;    // run-to-tolerance: 0 = keep going, else the number of generations this launch executes
    int gens_done;    // generations actually executed by the last launch
    int pad;
    double last_sumsq, last_max;
};

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
    // fused residual (generations >= res_first_gen accumulate it)
    int res_first_gen, res_slot0, stop_mode;
    double stop_thresh;
    double* partials;  // [2][n_tiles][2]
    double* res_out;   // [k3MaxResSlots][2]
    Result3d* result;
    int world, rank;
    double* peer_inbox[k3MaxRanks];
    unsigned long long* peer_res_flag[k3MaxRanks];
    const unsigned long long* res_flag;
    unsigned long long res_seq_base;
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

// v = *p (shared memory) if pred, else v unchanged; one predicated LDS, no branch.
__device__ __forceinline__ void ld_shared_if(float& v, const float* p, bool pred) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q ld.shared.f32 %0, [%1];\n\t}"
                 : "+f"(v) : "r"(ptx::smem_u32(p)), "r"((int)pred) : "memory");
}
__device__ __forceinline__ void ld_shared_if(double& v, const double* p, bool pred) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %2, 0;\n\t@q ld.shared.f64 %0, [%1];\n\t}"
                 : "+d"(v) : "r"(ptx::smem_u32(p)), "r"((int)pred) : "memory");
}

__device__ __forceinline__ void accum3(const float4& a, const float4& b, double& ss, double& mx) {
    // differences and squares in fp32 (the field's precision), one conversion per vector, fp64 across vectors
    const float d0 = __fsub_rn(a.x, b.x), d1 = __fsub_rn(a.y, b.y), d2 = __fsub_rn(a.z, b.z), d3 = __fsub_rn(a.w, b.w);
    float q = __fmul_rn(d0, d0);
    q = __fmaf_rn(d1, d1, q);
    q = __fmaf_rn(d2, d2, q);
    q = __fmaf_rn(d3, d3, q);
    ss += (double)q;
    mx = fmax(mx, (double)fmaxf(fmaxf(fabsf(d0), fabsf(d1)), fmaxf(fabsf(d2), fabsf(d3))));
}
__device__ __forceinline__ void accum3(const double2& a, const double2& b, double& ss, double& mx) {
    const double d0 = a.x - b.x, d1 = a.y - b.y;
    ss = fma(d0, d0, ss);
    ss = fma(d1, d1, ss);
    mx = fmax(mx, fmax(fabs(d0), fabs(d1)));
}

// Fixed-order reduction of one generation's per-tile partials by ONE warp, the all-reduce over the peers' inboxes,
// the residual history, the result block and the stop decision (same protocol as the 2-D kernel).
__device__ __noinline__ void reduce_and_exchange3(const Fused3dParams& p, const double* partials, const int gen,
                                                  const int lane) {
    double a = 0.0, b = 0.0;
    for (int i = lane; i < p.n_tiles; i += 32) {
        a += __ldcg(&partials[2 * i]);
        b = fmax(b, __ldcg(&partials[2 * i + 1]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
    }
    if (lane == 0) {
        double sa = a, sb = b;
        if (p.world > 1 && p.peer_inbox[0] != nullptr) {
            const unsigned long long seq = p.res_seq_base + (unsigned long long)(gen - p.res_first_gen);
            const size_t off = ((size_t)(seq % k3ResRing) * k3MaxRanks + (size_t)p.rank) * 2;
            for (int r = 0; r < p.world; ++r) {
                p.peer_inbox[r][off] = sa;
                p.peer_inbox[r][off + 1] = sb;
            }
            ptx::fence_acq_rel_sys();
            for (int r = 0; r < p.world; ++r) ptx::red_add_relaxed_sys(p.peer_res_flag[r], 1ull);
            if (wait_ctr3(p, p.res_flag, (seq + 1ull) * (unsigned long long)p.world, true)) {
                const volatile double* in = p.peer_inbox[p.rank] + (size_t)(seq % k3ResRing) * k3MaxRanks * 2;
                sa = 0.0;
                sb = 0.0;
                for (int r = 0; r < p.world; ++r) {  // rank order: the same bits on every rank
                    sa += in[2 * r];
                    sb = fmax(sb, in[2 * r + 1]);
                }
            }
        }
        const int slot = p.res_slot0 + (gen - p.res_first_gen);
        if (slot < k3MaxResSlots) {
            p.res_out[2 * slot] = sa;
            p.res_out[2 * slot + 1] = sb;
        }
        p.result->last_sumsq = sa;
        p.result->last_max = sb;
        if (p.stop_mode != 0 && *((volatile int*)&p.result->stop_word) == 0) {
            const bool conv = p.stop_mode == 1 ? (sb <= p.stop_thresh) : (sa <= p.stop_thresh);
            if (conv) {
                *((volatile int*)&p.result->stop_word) = gen + 2;  // one more generation runs while this is published
                __threadfence();
            }
        }
    }
    __syncwarp();
}

// RESK: the instantiation that can accumulate the residual (launched only when a residual is asked for, so that the
// plain kernel keeps its registers for the planes).
template <class T, bool RESK>
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
    __shared__ double s_red[2][k3Threads / 32];
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

    int executed = 0;
    for (int gen = 0; gen < p.generations; ++gen) {
        const bool res_now = RESK && gen >= p.res_first_gen;
        if (RESK && p.stop_mode != 0 && gen > 0) {
            // decided from the global residual of generation gen-2 and published before barrier gen-1
            const int sgen = *((volatile int*)&p.result->stop_word);
            if (sgen != 0 && gen >= sgen) break;
        }
        double* partials = p.partials + (size_t)(gen & 1) * p.n_tiles * 2;
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
            int issued = n_planes < k3Slots ? n_planes : k3Slots;  // planes put in flight so far (tid 0 acts on it)
            if (tid == 0)
                for (int i = 0; i < issued; ++i) issue(use + i, zlo - 1 + i);

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
            double rs = 0.0, rm = 0.0;
            auto zloop = [&](auto gen_tag, auto res_tag) {
            constexpr bool GEN = decltype(gen_tag)::value;
            constexpr bool RES = decltype(res_tag)::value;
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
                    const T* pc = reinterpret_cast<const T*>(&Cc[i + 1]);
                    // West / east neighbours of the lane's vector.  Measured on B200 (profiles/r2/heat3d_neighbours.md):
                    // reading them from the box in shared memory (a 4-way bank conflict, 26 M conflicts per launch)
                    // is 5 % FASTER than taking them from the adjacent lanes with shuffles + predicated halo loads
                    // for lanes 0 / 31 (442 K conflicts): the kernel is bound by barriers and latency, not by
                    // shared-memory bandwidth, and the shuffles cost issue slots.  The shuffle form stays selectable.
#ifndef MCP_H3_SHUFFLE_NEIGHBOURS
                    const T wl = cw[-1], er = cw[VEC];
#else
                    T wl = __shfl_up_sync(0xffffffffu, pc[VEC - 1], 1);
                    T er = __shfl_down_sync(0xffffffffu, pc[0], 1);
                    ld_shared_if(wl, cw - 1, lane == 0);    // predicated loads: no divergent branch in the plane loop
                    ld_shared_if(er, cw + VEC, lane == 31);
#endif
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
                        if (RES) accum3(o, Cc[i + 1], rs, rm);
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
                // Every k3Period planes: all warps are past plane pz in shared memory (it lives on in registers), so
                // the ring may hold planes pz+1 .. pz+k3Slots; put the missing ones in flight.
                if ((s % k3Period) == k3Period - 1 && s + 1 < zhi - zlo) {
                    __syncthreads();
                    const int upto = min(n_planes, s + 2 + k3Slots);
                    if (tid == 0)
                        for (int i = issued; i < upto; ++i) issue(use + i, zlo - 1 + i);
                    issued = max(issued, upto);
                }
            }
            };
            const bool lean = !f_zm && !f_zp && !f_ym && !f_yp && (p.z0 + zlo >= 1) && (p.z0 + zhi - 1 <= p.d - 2) &&
                              (p.y0 + ylo >= 1) && (p.y0 + ylo + k3TileRows - 1 <= p.h - 2) && (x0 >= 1) &&
                              (x0 + XS - 1 <= p.w - 2);
            bool did_res = false;
            if constexpr (RESK) {
                if (res_now) {
                    did_res = true;
                    if (lean) zloop(std::false_type{}, std::true_type{}); else zloop(std::true_type{}, std::true_type{});
                }
            }
            if (!did_res) {
                if (lean) zloop(std::false_type{}, std::false_type{}); else zloop(std::true_type{}, std::false_type{});
            }
            use += n_planes;
            if (res_now) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    rs += __shfl_xor_sync(0xffffffffu, rs, o);
                    rm = fmax(rm, __shfl_xor_sync(0xffffffffu, rm, o));
                }
                if (lane == 0) {
                    s_red[0][warp] = rs;
                    s_red[1][warp] = rm;
                }
            }

            // publish the pushed halo layers: all threads' peer stores -> one release per face
            __syncthreads();
            if (tid == 0) {
                if (res_now) {  // warp order: the same bits whichever CTA ran the tile
                    double a = 0.0, b = 0.0;
                    for (int k = 0; k < k3Threads / 32; ++k) {
                        a += s_red[0][k];
                        b = fmax(b, s_red[1][k]);
                    }
                    partials[2 * q] = a;
                    partials[2 * q + 1] = b;
                }
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

        if (res_now || gen + 1 < p.generations) {
            // grid barrier between generations (and in front of the reduction of a residual generation)
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
        if (res_now && blockIdx.x == 0) {
            if (warp == 0) reduce_and_exchange3(p, partials, gen, lane);
            __syncthreads();
        }
        executed = gen + 1;
    }
    if (blockIdx.x == 0 && tid == 0) p.result->gens_done = executed;
}

}  // namespace mcp

using namespace mcp;

struct mcp_fused3d_plan {
    mcp_fused3d_desc d;
    CUtensorMap tm[2];
    unsigned long long* grid_bar;
    Result3d* result;
    double* partials;
    double* res_out;
    int* tiles_dev;
    int n_tiles, zchunk, grid, parity;
    unsigned long long flag_base[4], bar_base, ctr_base, res_seq, timeout_ns;
    int res_slots_used, last_generations, last_stop_reason;
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
    for (int f = 0; f < 4; ++f) {
        MCP_REQUIRE((d.nbr_buf[f][0] == nullptr) == (d.nbr_buf[f][1] == nullptr), "nbr_buf[%d] must be both set or both null", f);
        if (d.nbr_buf[f][0] != nullptr)
            MCP_REQUIRE(d.flags != nullptr && d.nbr_flag[f] != nullptr && d.expect[f] >= 1 && d.nbr_n[f] >= 1,
                        "neighbour %d needs flags, nbr_flag, expect >= 1 and nbr_n >= 1", f);
    }
    // tile counts first: nothing is allocated yet if the block does not fit the tile encoding
    const int zchunk = d.zchunk > 0 ? d.zchunk : 64;
    const int nsx = (int)ceil_div(d.w, 32 * vec3(d.dtype));
    const int nby = (int)ceil_div(d.ny, k3TileRows);
    const int nzc = (int)ceil_div(d.nz, zchunk);
    MCP_REQUIRE(nsx < 1024 && nby < 1024 && nzc < 2048, "block too large for the tile encoding");
    mcp_fused3d_plan* pl = new (std::nothrow) mcp_fused3d_plan();
    MCP_REQUIRE(pl, "out of host memory");
    pl->d = d;
    pl->parity = 0;
    pl->launches = 0;
    pl->bar_base = pl->ctr_base = 0;
    for (int f = 0; f < 4; ++f) pl->flag_base[f] = 0;
    pl->grid_bar = nullptr;
    pl->result = nullptr;
    pl->partials = nullptr;
    pl->res_out = nullptr;
    pl->tiles_dev = nullptr;
    pl->res_seq = 0;
    pl->timeout_ns = 20ull * 1000000000ull;
    pl->res_slots_used = pl->last_generations = 0;
    pl->last_stop_reason = MCP_STOP_NONE;
    int sm = 0;
    int rc = device_sm_count(&sm);
    if (rc) { delete pl; return rc; }
    pl->grid = 2 * sm;
    pl->zchunk = zchunk;

    const int esz = d.dtype == MCP_DTYPE_F64 ? 8 : 4;
    for (int b = 0; b < 2; ++b) {
        uint64_t dims[3] = {(uint64_t)d.pitch, (uint64_t)d.py_alloc, (uint64_t)d.pz_alloc};
        uint64_t strides[2] = {(uint64_t)d.pitch * esz, (uint64_t)d.pitch * d.py_alloc * esz};
        uint32_t box[3] = {(uint32_t)(k3RowBytes / esz), (uint32_t)k3BoxRows, 1u};
        rc = encode_tensor_map(&pl->tm[b], d.dtype, 3, d.buf[b], dims, strides, box);
        if (rc) { delete pl; return rc; }
    }
    // tile list, pencil-face tiles first so their pushes overlap the interior
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
        (e = cudaMalloc(&pl->result, 256)) != cudaSuccess || (e = cudaMemset(pl->result, 0, 256)) != cudaSuccess ||
        (e = cudaMalloc(&pl->partials, face.size() * 4 * sizeof(double))) != cudaSuccess ||
        (e = cudaMalloc(&pl->res_out, (size_t)k3MaxResSlots * 2 * sizeof(double))) != cudaSuccess ||
        (e = cudaMemset(pl->res_out, 0, (size_t)k3MaxResSlots * 2 * sizeof(double))) != cudaSuccess ||
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

static int run3d_impl(mcp_fused3d_plan* pl, int generations, int residual_mode, int stop_mode, double stop_thresh,
                      cudaStream_t stream) {
    const mcp_fused3d_desc& d = pl->d;
    pl->last_generations = 0;
    pl->last_stop_reason = MCP_STOP_NONE;
    if (residual_mode != MCP_RES_NONE) pl->res_slots_used = 0;
    if (generations == 0) return MCP_OK;
    const int requested = generations;
    if (stop_mode != 0) generations = std::min(generations, k3MaxResSlots);
    const bool exchange = d.world > 1 && d.res_inbox != nullptr;
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
    p.timeout_ns = pl->timeout_ns;
    p.result = pl->result;
    p.status = &pl->result->status;
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
    // MCP_RES_LAST: every generation carries the residual; MCP_RES_FINAL: only the last one
    p.res_first_gen = residual_mode == MCP_RES_LAST ? 0 : (residual_mode == MCP_RES_FINAL ? generations - 1 : generations);
    p.res_slot0 = pl->res_slots_used;
    p.stop_mode = stop_mode;
    p.stop_thresh = stop_thresh;
    p.partials = pl->partials;
    p.res_out = pl->res_out;
    p.world = d.world > 0 ? d.world : 1;
    p.rank = d.rank;
    p.res_flag = (const unsigned long long*)d.res_flag;
    p.res_seq_base = pl->res_seq;
    if (exchange)
        for (int r = 0; r < p.world; ++r) {
            p.peer_inbox[r] = (double*)d.peer_res_inbox[r];
            p.peer_res_flag[r] = (unsigned long long*)d.peer_res_flag[r];
        }
    if (stop_mode != 0) MCP_CUDA_TRY(cudaMemsetAsync(&pl->result->stop_word, 0, 2 * sizeof(int), stream));

    void* args[3] = {(void*)&pl->tm[0], (void*)&pl->tm[1], (void*)&p};
    const bool resk = residual_mode != MCP_RES_NONE;
    const void* fn = d.dtype == MCP_DTYPE_F64
                         ? (resk ? (const void*)heat3d_fused_kernel<double, true> : (const void*)heat3d_fused_kernel<double, false>)
                         : (resk ? (const void*)heat3d_fused_kernel<float, true> : (const void*)heat3d_fused_kernel<float, false>);
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, k3SmemBytes));
    int per_sm = 0;
    MCP_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, k3Threads, k3SmemBytes));
    MCP_REQUIRE(per_sm >= 1, "heat3d kernel does not fit on an SM");
    const int grid = std::min(pl->grid, per_sm * (pl->grid / 2));
    MCP_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(k3Threads), args, k3SmemBytes, stream));
    pl->launches++;

    int executed = generations;
    if (stop_mode != 0) {
        // the number of generations that ran is decided on the device: wait for it before touching the counters
        MCP_CUDA_TRY(cudaStreamSynchronize(stream));
        Result3d rb;
        MCP_CUDA_TRY(cudaMemcpy(&rb, pl->result, sizeof(rb), cudaMemcpyDeviceToHost));
        executed = rb.gens_done;
        MCP_REQUIRE(executed >= 0 && executed <= generations, "device reported %d executed generations of %d", executed,
                    generations);
        pl->last_stop_reason = rb.stop_word != 0 ? MCP_STOP_CONVERGED
                                                 : (generations < requested ? MCP_STOP_SLOT_LIMIT : MCP_STOP_MAX_GENERATIONS);
    }
    const int res_gens = std::max(0, executed - p.res_first_gen);
    // a barrier after every generation but the last, and after the last one too when it carries the residual
    unsigned long long barriers = (unsigned long long)std::max(0, executed - 1);
    if (executed > 0 && executed - 1 >= p.res_first_gen) barriers += 1;
    pl->bar_base += barriers * grid;
    pl->ctr_base += (unsigned long long)executed * (unsigned long long)(pl->n_tiles + grid);
    for (int f = 0; f < 4; ++f) pl->flag_base[f] += (unsigned long long)executed * d.expect[f];
    pl->parity = (pl->parity + executed) & 1;
    pl->res_slots_used = std::min(k3MaxResSlots, pl->res_slots_used + res_gens);
    if (exchange) pl->res_seq += (unsigned long long)res_gens;
    pl->last_generations = executed;
    return MCP_OK;
}

int mcp_fused3d_run(mcp_fused3d_plan* pl, int generations, void* stream_) {
    MCP_REQUIRE(pl && generations >= 0, "mcp_fused3d_run: bad arguments");
    return run3d_impl(pl, generations, MCP_RES_NONE, 0, 0.0, (cudaStream_t)stream_);
}

int mcp_fused3d_run_residual(mcp_fused3d_plan* pl, int generations, int residual_mode, void* stream_) {
    MCP_REQUIRE(pl && generations >= 0, "mcp_fused3d_run_residual: bad arguments");
    MCP_REQUIRE(residual_mode == MCP_RES_NONE || residual_mode == MCP_RES_LAST || residual_mode == MCP_RES_FINAL,
                "residual_mode must be MCP_RES_NONE, MCP_RES_LAST or MCP_RES_FINAL");
    MCP_REQUIRE(residual_mode != MCP_RES_LAST || generations <= k3MaxResSlots,
                "at most %d generations per call with a residual history", k3MaxResSlots);
    return run3d_impl(pl, generations, residual_mode, 0, 0.0, (cudaStream_t)stream_);
}

int mcp_fused3d_run_to_tolerance(mcp_fused3d_plan* pl, int max_generations, int norm, double tol, int* generations_done,
                                 void* stream_) {
    MCP_REQUIRE(pl && max_generations >= 1, "mcp_fused3d_run_to_tolerance: bad arguments");
    MCP_REQUIRE(norm == MCP_NORM_LINF || norm == MCP_NORM_L2, "norm must be MCP_NORM_LINF or MCP_NORM_L2");
    MCP_REQUIRE(tol >= 0.0, "negative tolerance");
    const mcp_fused3d_desc& d = pl->d;
    bool neighbours = false;
    for (int f = 0; f < 4; ++f) neighbours = neighbours || d.nbr_buf[f][0] != nullptr;
    MCP_REQUIRE(!neighbours || d.res_inbox != nullptr,
                "run_to_tolerance across ranks needs the fused residual exchange (res_inbox): a local residual would "
                "stop the ranks in different generations");
    int rc = run3d_impl(pl, max_generations, MCP_RES_LAST, norm == MCP_NORM_LINF ? 1 : 2,
                        norm == MCP_NORM_L2 ? tol * tol : tol, (cudaStream_t)stream_);
    if (generations_done) *generations_done = pl->last_generations;
    return rc;
}

int mcp_fused3d_stop_reason(const mcp_fused3d_plan* pl) { return pl ? pl->last_stop_reason : MCP_STOP_NONE; }

int mcp_fused3d_residuals(mcp_fused3d_plan* pl, double* out, int max_entries) {
    MCP_REQUIRE(pl && out, "null argument");
    const int n = std::min(pl->res_slots_used, max_entries);
    if (n > 0) MCP_CUDA_TRY(cudaMemcpy(out, pl->res_out, (size_t)n * 2 * sizeof(double), cudaMemcpyDeviceToHost));
    return n;
}

int mcp_fused3d_result_async(mcp_fused3d_plan* pl, void* host32, void* stream_) {
    MCP_REQUIRE(pl && host32, "null argument");
    static_assert(sizeof(Result3d) == 32, "result block layout");
    MCP_CUDA_TRY(cudaMemcpyAsync(host32, pl->result, sizeof(Result3d), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    return MCP_OK;
}

int mcp_fused3d_set_timeout_ms(mcp_fused3d_plan* pl, long long ms) {
    MCP_REQUIRE(pl && ms >= 1, "bad timeout");
    pl->timeout_ns = (unsigned long long)ms * 1000000ull;
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
    MCP_CUDA_TRY(cudaMemcpy(&st, &pl->result->status, sizeof(int), cudaMemcpyDeviceToHost));
    if (st != 0) set_error("fused3d kernel recorded status %d", st);
    return st;
}

int64_t mcp_fused3d_launches(const mcp_fused3d_plan* pl) { return pl ? pl->launches : 0; }

int mcp_fused3d_destroy(mcp_fused3d_plan* pl) {
    if (!pl) return MCP_OK;
    if (pl->grid_bar) cudaFree(pl->grid_bar);
    if (pl->result) cudaFree(pl->result);
    if (pl->partials) cudaFree(pl->partials);
    if (pl->res_out) cudaFree(pl->res_out);
    if (pl->tiles_dev) cudaFree(pl->tiles_dev);
    delete pl;
    return MCP_OK;
}

}  // extern "C"
