// Fused halo-exchange + 2-D stencil (+ residual) persistent kernel for sm_100a.
//
// What it replaces in the reference: the per-generation sequence of MDF_kernel.cu:155-222 --
// full-grid H2D copy, middle_kernel, w-2 one-element MPI_Recv, second H2D copy, border_kernel,
// cudaDeviceSynchronize, full-grid D2H copy, w-2 one-element MPI_Send -- and the kernels
// run_mdf/middle_kernel/border_kernel (MDF_kernel.cu:10-70), game_of_life (kernel.cu:10-113).
//
// Design (B200-first, not a translation):
//  * One cooperative launch runs many generations.  148 CTAs x 16 warps stay resident.
//  * The slab is cut into tiles = (column strip of 32 lanes x 16 B) x (row chunk).  A tile belongs to ONE
//    warp; warps never synchronise with each other inside a sweep.
//  * Each warp owns a 4-stage shared-memory ring fed by TMA (cp.async.bulk.tensor.2d, 6 rows x 512 B
//    per stage, mbarrier completion).  TMA's out-of-bounds fill supplies the cells outside the allocation.
//  * Temporal blocking in registers: a warp streams down the rows and carries TB generations as a
//    wavefront (3-row window per level, compile-time rotated), so HBM sees one read + one write per
//    TB generations.  Left/right neighbours come from warp shuffles.  Each update is the reference's
//    exact rounding sequence, so results are bit-identical to a generation-at-a-time Jacobi.
//  * Tiles on the slab cut store their top/bottom `ghost` output rows twice: locally and straight into
//    the neighbour GPU's ghost rows through an NVLink-mapped peer pointer, then bump a system-scope
//    release counter in the neighbour's flag pad.  Tiles that read ghost rows acquire-spin on their local
//    counter first.  Boundary tiles are scheduled first so the transfer overlaps the interior tiles.
//  * Sweeps are separated by a device-side grid barrier (no host round trip, no NCCL, no MPI).
//  * Optional fused residual: per-tile (sum d^2, max |d|) in fp64 -> fixed-order reduction (deterministic).
#include <cooperative_groups.h>

#include <stdlib.h>

#include <algorithm>
#include <type_traits>
#include <cmath>
#include <new>
#include <vector>

#include "../../../include/mcp_b200.h"
#include "common.cuh"
#include "ptx.cuh"
#include "stencil_ops.cuh"

namespace mcp {

constexpr int kWarpsPerCta = 16;
constexpr int kThreads = kWarpsPerCta * 32;
constexpr int kRowsPerStage = 6;  // multiple of 3: the 3-row register window rotates at compile time
constexpr int kStages = 4;
constexpr int kRowBytes = 512;  // 32 lanes x 16 B
constexpr int kStageBytes = kRowsPerStage * kRowBytes;
constexpr int smem_bytes_for(int warps) { return warps * kStages * kStageBytes + warps * kStages * 8; }
constexpr int kSmemBytes = smem_bytes_for(kWarpsPerCta);
// TB = 8 keeps 96 registers of row windows per thread: run 12 warps x 168 registers instead of 16 x 128.
constexpr int kWarpsTb8 = 8;
__host__ __device__ constexpr int warps_for_tb(int tb) { return tb >= 8 ? kWarpsTb8 : kWarpsPerCta; }
// Register budget per operator: kind 0 = scalar float cells (12 window registers per level), 1 = Life int32 (cells
// + horizontal sums, 24 per level), 2 = bit-packed Life (cells + two sum planes, 36 per level).  When the windows of
// TB levels no longer fit 128 registers the CTA runs 8 warps x up to 255 registers instead of 16 x 128.
__host__ __device__ constexpr int warps_for(int kind, int tb) {
    if (kind == 2 && tb == 2) return 12;  // 148 registers: 12 warps x 32 x 148 fits the register file
    return (kind == 2 && tb >= 2) || (kind == 1 && tb >= 4) ? kWarpsTb8 : warps_for_tb(tb);
}
constexpr int kMaxResSlots = 16384;
constexpr int kMaxRanks = MCP_MAX_RANKS;
constexpr int kResRing = MCP_RES_RING;
constexpr int kMaxStripFlags = MCP_MAX_STRIPS;

struct Fused2dParams {
    void* buf[2];
    void* up_buf[2];
    void* dn_buf[2];
    const unsigned long long* my_flags;  // [0] arrivals from up, [1] arrivals from down
    // per-strip halo flags (optional): [dir][strip] = number of sweeps whose boundary tile of that strip the neighbour
    // has finished and pushed; a tile waits for strips s-1, s, s+1 only instead of for the neighbour's whole edge
    const unsigned long long* strip_flags;
    unsigned long long* up_strip_flag;  // up neighbour's [1][*] row (we are its down side)
    unsigned long long* dn_strip_flag;  // down neighbour's [0][*] row
    unsigned long long seq_base;        // sweeps completed by every rank before this launch
    int prev_n_strips;                  // strip count of the previous launch (0 = none): sweep 0 waits for all of them
    unsigned long long* up_flag;
    unsigned long long* dn_flag;
    unsigned long long* grid_bar;
    unsigned long long* tile_ctr;
    unsigned* tile_done;  // [n_tiles] last sweep (1-based, within this launch) each tile has completed
    unsigned* tile_ns;    // [n_tiles] trace (debug_flags & 4): duration of the tile in the last sweep it ran, in ns
    int* status;
    double* partials;  // [2][n_tiles][2]
    double* res_out;   // [kMaxResSlots][2]
    const int* chunk_y;  // [n_chunks + 1] local row boundaries of the row chunks (host-built schedule)
    // cross-rank residual all-reduce over NVLink: every rank stores its (sumsq, max) into slot [seq % ring][rank] of
    // every rank's inbox, bumps that rank's arrival counter, then sums its own inbox in rank order
    double* peer_inbox[kMaxRanks];
    // NVLS variant: `nvls_local` is this rank's copy of a multicast-mapped pad, `nvls_mc` its multicast address
    // (layout: u64 arrival counter @0, ring of (sumsq f64, max-bits u64) pairs @128)
    unsigned char* nvls_local;
    unsigned char* nvls_mc;
    unsigned long long* peer_res_flag[kMaxRanks];
    const unsigned long long* res_flag;
    unsigned long long res_seq_base;
    int* stop_word;    // 0 = keep going, else total number of sweeps this launch executes
    int* sweeps_done;  // sweeps actually executed by this launch
    double stop_thresh;
    int stop_mode;  // 0 none, 1: stop when max|d| <= thresh, 2: stop when sum d^2 <= thresh
    int world, rank;
    unsigned long long flag_base_up, flag_base_dn;
    unsigned long long bar_base;
    unsigned long long ctr_base;
    unsigned long long timeout_ns;
    long long pitch;
    int rows_alloc, h, w, row0;
    int own_lo, own_hi;
    int ylo, yhi, xlo, xhi;
    int G, up_ghost_row;
    int n_strips, n_chunks;
    int sweeps, first_parity, res_slot0;
    int debug_flags;
    int sync_mode;  // 1 = grid barrier between sweeps, 2 = per-tile dataflow flags (no barrier)
};

// ------------------------------------------------------------------------------------------ operators

// Column masks of the scalar-cell operators: one bit per vector component.
template <int VEC_>
struct ComponentMask {
    using Mask = unsigned;
    static constexpr int CPE = 1;  // cells per storage element
    static __device__ __forceinline__ Mask full_mask() { return 0xfu; }
    static __device__ __forceinline__ bool any(Mask m) { return m != 0u; }
    // component v of the lane's vector is cell cx + v
    static __device__ __forceinline__ Mask col_mask(int cx, int xlo, int xhi) {
        Mask m = 0;
#pragma unroll
        for (int v = 0; v < VEC_; ++v) {
            const int gx = cx + v;
            if (gx < xlo || gx > xhi) m |= 1u << v;
        }
        return m;
    }
};

struct MdfF32 : ComponentMask<4> {
    using T = float;
    using V = float4;
    static constexpr int VEC = 4;
    struct Row {
        V v;
    };
    static __device__ __forceinline__ Row make_row(const V& raw) { return Row{raw}; }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        const float wl = __shfl_up_sync(0xffffffffu, c.v.w, 1);
        const float er = __shfl_down_sync(0xffffffffu, c.v.x, 1);
        V o;
        o.x = mdf_cell(c.v.x, n.v.x, s.v.x, wl, c.v.y);
        o.y = mdf_cell(c.v.y, n.v.y, s.v.y, c.v.x, c.v.z);
        o.z = mdf_cell(c.v.z, n.v.z, s.v.z, c.v.y, c.v.w);
        o.w = mdf_cell(c.v.w, n.v.w, s.v.w, c.v.z, er);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        o.z = (m & 4u) ? keep.z : upd.z;
        o.w = (m & 8u) ? keep.w : upd.w;
        return o;
    }
    // residual contribution of 4 cells: differences and their squares in fp32 (the field's own precision),
    // one conversion per vector, accumulation across rows in fp64
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const float d0 = __fsub_rn(a.x, b.x), d1 = __fsub_rn(a.y, b.y);
        const float d2 = __fsub_rn(a.z, b.z), d3 = __fsub_rn(a.w, b.w);
        float q = __fmul_rn(d0, d0);
        q = __fmaf_rn(d1, d1, q);
        q = __fmaf_rn(d2, d2, q);
        q = __fmaf_rn(d3, d3, q);
        ss += (double)q;
        const float m = fmaxf(fmaxf(fabsf(d0), fabsf(d1)), fmaxf(fabsf(d2), fabsf(d3)));
        mx = fmax(mx, (double)m);
    }
};

struct MdfF64 : ComponentMask<2> {
    using T = double;
    using V = double2;
    static constexpr int VEC = 2;
    struct Row {
        V v;
    };
    static __device__ __forceinline__ Row make_row(const V& raw) { return Row{raw}; }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        const double wl = __shfl_up_sync(0xffffffffu, c.v.y, 1);
        const double er = __shfl_down_sync(0xffffffffu, c.v.x, 1);
        V o;
        o.x = mdf_cell(c.v.x, n.v.x, s.v.x, wl, c.v.y);
        o.y = mdf_cell(c.v.y, n.v.y, s.v.y, c.v.x, er);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        return o;
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const double d0 = a.x - b.x, d1 = a.y - b.y;
        ss = fma(d0, d0, ss); ss = fma(d1, d1, ss);
        mx = fmax(mx, fmax(fabs(d0), fabs(d1)));
    }
};

struct LifeI32 : ComponentMask<4> {
    using T = int;
    using V = int4;
    static constexpr int VEC = 4;
    struct Row {
        V v;   // cells
        V hs;  // horizontal 3-sums (left + self + right)
    };
    static __device__ __forceinline__ Row make_row(const V& raw) {
        const int wl = __shfl_up_sync(0xffffffffu, raw.w, 1);
        const int er = __shfl_down_sync(0xffffffffu, raw.x, 1);
        Row r;
        r.v = raw;
        const int xy = raw.x + raw.y, zw = raw.z + raw.w;
        r.hs.x = wl + xy;
        r.hs.y = xy + raw.z;
        r.hs.z = raw.y + zw;
        r.hs.w = zw + er;
        return r;
    }
    static __device__ __forceinline__ int rule(int n, int c) { return life_rule(n, c); }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        V o;
        o.x = rule(n.hs.x + s.hs.x + c.hs.x - c.v.x, c.v.x);
        o.y = rule(n.hs.y + s.hs.y + c.hs.y - c.v.y, c.v.y);
        o.z = rule(n.hs.z + s.hs.z + c.hs.z - c.v.z, c.v.z);
        o.w = rule(n.hs.w + s.hs.w + c.hs.w - c.v.w, c.v.w);
        return o;
    }
    static __device__ __forceinline__ V select(unsigned m, const V& keep, const V& upd) {
        V o;
        o.x = (m & 1u) ? keep.x : upd.x;
        o.y = (m & 2u) ? keep.y : upd.y;
        o.z = (m & 4u) ? keep.z : upd.z;
        o.w = (m & 8u) ? keep.w : upd.w;
        return o;
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const int d0 = a.x - b.x, d1 = a.y - b.y, d2 = a.z - b.z, d3 = a.w - b.w;
        const int q = d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
        ss += (double)q;
        mx = fmax(mx, q ? 1.0 : 0.0);
    }
};

// Bit-packed Game of Life: one cell per bit, 32 cells per 32-bit word (cell x = bit x%32 of word x/32), so a lane's
// 16-byte vector holds 128 cells and a warp row 4096.  Same rule as LifeI32 (kernel.cu:61,66), evaluated for 32 cells
// at a time with bit-sliced adders.  The optional fast path of SURVEY section 7.1; the reference stores one int per cell.
struct LifeB1 {
    using T = unsigned;
    using V = uint4;
    static constexpr int VEC = 4;   // storage elements per lane
    static constexpr int CPE = 32;  // cells per storage element
    using Mask = uint4;             // one bit per cell: 1 = keep the old value
    struct Row {
        V v;       // cells
        V s0, s1;  // per cell: (s1:s0) = west + self + east (0..3)
    };
    static __device__ __forceinline__ Mask full_mask() {
        return make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    }
    static __device__ __forceinline__ bool any(const Mask& m) { return (m.x | m.y | m.z | m.w) != 0u; }
    // bits of the word holding cells [c0, c0+32) that lie outside [xlo, xhi]
    static __device__ __forceinline__ unsigned word_mask(int c0, int xlo, int xhi) {
        unsigned keep = 0u;  // bits inside [xlo, xhi] -> evolve
        const int lo = max(xlo, c0), hi = min(xhi, c0 + 31);
        if (lo <= hi) {
            const int n = hi - lo + 1;
            const unsigned run = n >= 32 ? 0xffffffffu : ((1u << n) - 1u);
            keep = run << (lo - c0);
        }
        return ~keep;
    }
    static __device__ __forceinline__ Mask col_mask(int cx, int xlo, int xhi) {
        // cx = index of the lane's first storage element; element e holds cells 32e .. 32e+31
        return make_uint4(word_mask((cx + 0) * 32, xlo, xhi), word_mask((cx + 1) * 32, xlo, xhi),
                          word_mask((cx + 2) * 32, xlo, xhi), word_mask((cx + 3) * 32, xlo, xhi));
    }
    static __device__ __forceinline__ unsigned maj(unsigned a, unsigned b, unsigned c) { return (a & b) | (c & (a ^ b)); }
    static __device__ __forceinline__ Row make_row(const V& raw) {
        const unsigned prevw = __shfl_up_sync(0xffffffffu, raw.w, 1);    // word to the west of .x
        const unsigned nextx = __shfl_down_sync(0xffffffffu, raw.x, 1);  // word to the east of .w
        Row r;
        r.v = raw;
        const unsigned c[4] = {raw.x, raw.y, raw.z, raw.w};
        const unsigned wprev[4] = {prevw, raw.x, raw.y, raw.z};
        const unsigned enext[4] = {raw.y, raw.z, raw.w, nextx};
        unsigned s0[4], s1[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned W = (c[k] << 1) | (wprev[k] >> 31);  // west neighbour of every cell of the word
            const unsigned E = (c[k] >> 1) | (enext[k] << 31);  // east neighbour
            s0[k] = W ^ c[k] ^ E;
            s1[k] = maj(W, c[k], E);
        }
        r.s0 = make_uint4(s0[0], s0[1], s0[2], s0[3]);
        r.s1 = make_uint4(s1[0], s1[1], s1[2], s1[3]);
        return r;
    }
    static __device__ __forceinline__ unsigned word(unsigned n0, unsigned n1, unsigned c0, unsigned c1, unsigned s0,
                                                    unsigned s1, unsigned centre) {
        // total (0..9, centre included) = l + 2*t1 + 4*t2 + 8*t3
        const unsigned l = n0 ^ c0 ^ s0, cl = maj(n0, c0, s0);
        const unsigned h = n1 ^ c1 ^ s1, ch = maj(n1, c1, s1);
        const unsigned t1 = h ^ cl, c2 = h & cl;
        const unsigned t2 = ch ^ c2, t3 = ch & c2;
        // alive next iff total == 3, or total == 4 and alive now  (= 3 neighbours, or 2 neighbours and alive)
        return ~t3 & ((l & t1 & ~t2) | (~l & ~t1 & t2 & centre));
    }
    static __device__ __forceinline__ V update(const Row& n, const Row& c, const Row& s) {
        return make_uint4(word(n.s0.x, n.s1.x, c.s0.x, c.s1.x, s.s0.x, s.s1.x, c.v.x),
                          word(n.s0.y, n.s1.y, c.s0.y, c.s1.y, s.s0.y, s.s1.y, c.v.y),
                          word(n.s0.z, n.s1.z, c.s0.z, c.s1.z, s.s0.z, s.s1.z, c.v.z),
                          word(n.s0.w, n.s1.w, c.s0.w, c.s1.w, s.s0.w, s.s1.w, c.v.w));
    }
    static __device__ __forceinline__ V select(const Mask& m, const V& keep, const V& upd) {
        return make_uint4((keep.x & m.x) | (upd.x & ~m.x), (keep.y & m.y) | (upd.y & ~m.y),
                          (keep.z & m.z) | (upd.z & ~m.z), (keep.w & m.w) | (upd.w & ~m.w));
    }
    static __device__ __forceinline__ void accum(const V& a, const V& b, double& ss, double& mx) {
        const int q = __popc(a.x ^ b.x) + __popc(a.y ^ b.y) + __popc(a.z ^ b.z) + __popc(a.w ^ b.w);
        ss += (double)q;  // changed cells: (a-b)^2 is 0 or 1
        mx = fmax(mx, q ? 1.0 : 0.0);
    }
};

template <class V>
__device__ __forceinline__ V bitcast_v(const uint4& u) {
    static_assert(sizeof(V) == 16, "vector must be 16 bytes");
    V v;
    memcpy(&v, &u, 16);
    return v;
}
template <class V>
__device__ __forceinline__ uint4 bitcast_u(const V& v) {
    uint4 u;
    memcpy(&u, &v, 16);
    return u;
}

__host__ __device__ constexpr int mod3(int x) { return ((x % 3) + 3) % 3; }

// ------------------------------------------------------------------------------------------ waits

__device__ __forceinline__ bool aborted(const Fused2dParams& p) { return *((volatile int*)p.status) != 0; }

// Lane-0 style bounded acquire spin on a monotonic counter.  Returns false on abort/timeout.
__device__ __noinline__ bool wait_counter_sys(const Fused2dParams& p, const unsigned long long* ctr,
                                              unsigned long long need) {
    if (ptx::ld_acquire_sys(ctr) >= need) return true;
    const unsigned long long t0 = ptx::globaltimer_ns();
    unsigned spins = 0;
    while (ptx::ld_acquire_sys(ctr) < need) {
        if ((++spins & 255u) == 0) {
            if (aborted(p)) return false;
            if (ptx::globaltimer_ns() - t0 > p.timeout_ns) {
                atomicCAS(p.status, 0, MCP_ERR_TIMEOUT);
                return false;
            }
            __nanosleep(64);
        }
    }
    return true;
}

__device__ __noinline__ void grid_barrier(const Fused2dParams& p, unsigned long long& epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_proxy_async();  // our generic-proxy stores are read by other CTAs' TMA loads next sweep
        __threadfence();
        const unsigned long long target = p.bar_base + (epoch + 1ull) * gridDim.x;
        ptx::red_add_release_gpu(p.grid_bar, 1ull);
        const unsigned long long t0 = ptx::globaltimer_ns();
        unsigned spins = 0;
        while (ptx::ld_acquire_gpu(p.grid_bar) < target) {
            if ((++spins & 1023u) == 0) {
                if (aborted(p)) break;
                if (ptx::globaltimer_ns() - t0 > p.timeout_ns) {
                    atomicCAS(p.status, 0, MCP_ERR_TIMEOUT);
                    break;
                }
            }
        }
        ptx::fence_proxy_async();
    }
    __syncthreads();
    ++epoch;
}

// ------------------------------------------------------------------------------------------ tile

// 16-byte store with a predicate instead of a branch (keeps the row loop free of control flow).
__device__ __forceinline__ void st_global_v4_pred(void* p, const uint4& v, bool pred) {
    asm volatile(
        "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %5, 0;\n\t"
        "@q st.global.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
        ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"((int)pred)
        : "memory");
}
__device__ __forceinline__ uint4 ld_global_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}

// One shared-memory stage = kRowsPerStage input rows pushed through the TB-level register wavefront.
// MODE 0: the lean path used by almost every iteration -- no masks, no range checks, no branches.
// MODE 1: edge strips away from the top/bottom of the tile: per-lane column masks only (fixed ring / pad columns).
// MODE 2: general: row and column masks, output rows may fall outside [y0, y1), useless pre-roll levels skipped.
template <class Op, int TB, int MODE, bool RES>
__device__ __forceinline__ void stage_rows(typename Op::Row (&L)[TB][3], const uint8_t* sp,
                                           typename Op::T*& optr, const long long pitch, const int cbase,
                                           const int n_in, const int gy_base, const int ylo, const int yhi,
                                           const typename Op::Mask colfix, const bool lane_valid,
                                           const bool lane_store,
                                           double& rs, double& rm) {
    using V = typename Op::V;
    constexpr bool GENERAL = (MODE == 2);
#pragma unroll
    for (int j = 0; j < kRowsPerStage; ++j) {
        L[0][mod3(j)] = Op::make_row(bitcast_v<V>(ptx::ld_shared_v4(sp + j * kRowBytes)));
#pragma unroll
        for (int t = 1; t <= TB; ++t) {
            const int sc = mod3(j - t), sn = mod3(j - t - 1), ss = mod3(j - t + 1);
            // Pre-roll: level t's first meaningful row appears at input counter 2t; anything earlier would be
            // computed from rows that do not exist and is never read by a meaningful cell -- skip it.
            if (GENERAL && (cbase + j < 2 * t)) continue;
            V res = Op::update(L[t - 1][sn], L[t - 1][sc], L[t - 1][ss]);
            if (GENERAL) {
                const int gy = gy_base + j - t;
                const typename Op::Mask m = (gy < ylo || gy > yhi) ? Op::full_mask() : colfix;
                res = Op::select(m, L[t - 1][sc].v, res);
            } else if (MODE == 1) {
                res = Op::select(colfix, L[t - 1][sc].v, res);
            }
            if (t < TB) {
                L[t < TB ? t : 0][sc] = Op::make_row(res);
            } else {
                bool row_ok = true;
                if (GENERAL) {
                    const int c = cbase + j;
                    row_ok = (c >= 2 * TB) && (c < n_in);
                }
                if (RES) {
                    if (row_ok && lane_valid) Op::accum(res, L[TB - 1][sc].v, rs, rm);
                }
                st_global_v4_pred(optr, bitcast_u<V>(res), lane_store && row_ok);
            }
        }
        optr += pitch;  // one output row per input row, whether or not it was stored
    }
}

template <class Op, int TB, bool RES>
__device__ __forceinline__ void process_tile(const Fused2dParams& p, const CUtensorMap* tm_in,
                                             typename Op::T* __restrict__ out, typename Op::T* up_out,
                                             typename Op::T* dn_out, const int x_ld, const int y0, const int y1,
                                             uint8_t* my_smem, uint64_t* bars, uint32_t& it, const int lane,
                                             double& rs, double& rm) {
    using T = typename Op::T;
    using V = typename Op::V;
    using Row = typename Op::Row;
    constexpr int VEC = Op::VEC;  // storage elements per lane (x_ld, cx, pitch are in elements)
    constexpr int CPE = Op::CPE;  // cells per element (32 for the bit-packed operator)
    constexpr int M = (TB + VEC * CPE - 1) / (VEC * CPE);  // margin lanes whose cells go stale within TB levels

    const int ys = y0 - TB;
    const int n_in = (y1 - y0) + 2 * TB;
    const int n_it = (n_in + kRowsPerStage - 1) / kRowsPerStage;
    const int cx = x_ld + lane * VEC;

    const typename Op::Mask colfix = Op::col_mask(cx, p.xlo, p.xhi);
    const bool col_masked = (x_ld * CPE < p.xlo) || ((x_ld + 32 * VEC) * CPE - 1 > p.xhi);
    const bool lane_valid = (lane >= M) && (lane < 32 - M);
    const bool lane_store = lane_valid && (cx < (int)p.pitch);

    Row L[TB][3];
#pragma unroll
    for (int t = 0; t < TB; ++t)
#pragma unroll
        for (int k = 0; k < 3; ++k) L[t][k] = Op::make_row(bitcast_v<V>(make_uint4(0, 0, 0, 0)));

    auto issue = [&](uint32_t k, int row) {
        const uint32_t st = k % kStages;
        ptx::mbar_arrive_expect_tx(&bars[st], kStageBytes);
        ptx::tma_load_2d(my_smem + st * kStageBytes, tm_in, &bars[st], x_ld, row);
    };

    if (lane == 0) {
        const int npro = n_it < kStages ? n_it : kStages;
        for (int i = 0; i < npro; ++i) issue(it + i, ys + i * kRowsPerStage);
    }

    // output pointer of the row produced by input-row counter c: local row ys + c - TB
    T* optr = out + (long long)(ys - TB) * p.pitch + cx;

    for (int i = 0; i < n_it; ++i) {
        const uint32_t st = it % kStages, par = (it / kStages) & 1u;
        ptx::mbar_wait(&bars[st], par);
        const uint8_t* sp = my_smem + st * kStageBytes + lane * 16;
        const int cbase = i * kRowsPerStage;
        const int gy_base = p.row0 + ys + cbase;  // global row of this stage's first input row
        const bool row_special = (gy_base - TB < p.ylo) || (gy_base + kRowsPerStage > p.yhi) || (cbase < 2 * TB) ||
                                 (cbase + kRowsPerStage > n_in);
        if (row_special)
            stage_rows<Op, TB, 2, RES>(L, sp, optr, p.pitch, cbase, n_in, gy_base, p.ylo, p.yhi, colfix, lane_valid,
                                       lane_store, rs, rm);
        else if (col_masked)
            stage_rows<Op, TB, 1, RES>(L, sp, optr, p.pitch, cbase, n_in, gy_base, p.ylo, p.yhi, colfix, lane_valid,
                                       lane_store, rs, rm);
        else
            stage_rows<Op, TB, 0, RES>(L, sp, optr, p.pitch, cbase, n_in, gy_base, p.ylo, p.yhi, colfix, lane_valid,
                                       lane_store, rs, rm);
        __syncwarp();
        if (lane == 0 && i + kStages < n_it) issue(it + kStages, ys + (i + kStages) * kRowsPerStage);
        ++it;
    }

    // Halo push: re-read the boundary rows this lane just stored (same thread, same addresses) and store them
    // into the neighbour GPU's ghost rows over NVLink.  Kept out of the row loop so interior tiles pay nothing.
    if (lane_store) {
        if (up_out != nullptr) {
            const int r1 = min(y1, p.own_lo + p.G);
            for (int r = y0; r < r1; ++r) {
                const uint4 v = ld_global_v4(out + (long long)r * p.pitch + cx);
                ptx::st_global_v4(up_out + (long long)(p.up_ghost_row + (r - p.own_lo)) * p.pitch + cx, v);
            }
        }
        if (dn_out != nullptr) {
            const int r0 = max(y0, p.own_hi - p.G);
            for (int r = r0; r < y1; ++r) {
                const uint4 v = ld_global_v4(out + (long long)r * p.pitch + cx);
                ptx::st_global_v4(dn_out + (long long)(r - (p.own_hi - p.G)) * p.pitch + cx, v);
            }
        }
    }
}

// Tile order within a sweep: chunks from the outside in (0, nc-1, 1, nc-2, ...), all strips of a chunk together.
// The slab-cut chunks come first so their halo pushes overlap the rest of the sweep, and a chunk's neighbours in
// the previous sweep were finished equally early, so the barrier-free path never has an early ticket waiting for
// a late one.
__host__ __device__ __forceinline__ void decode_tile(int q, int ns, int nc, int& chunk, int& strip) {
    const int ci = q / ns;
    strip = q - ci * ns;
    chunk = (ci & 1) ? nc - 1 - (ci >> 1) : (ci >> 1);
}

// Inverse of decode_tile.
__host__ __device__ __forceinline__ int encode_tile(int chunk, int strip, int ns, int nc) {
    const int ci = (chunk <= nc - 1 - chunk) ? 2 * chunk : 2 * (nc - 1 - chunk) + 1;
    return ci * ns + strip;
}

// Bounded acquire spin at GPU scope on a per-tile completion word.
__device__ __noinline__ bool wait_tile_done(const Fused2dParams& p, const unsigned* word, unsigned need) {
    auto ld = [&]() {
        unsigned v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(word) : "memory");
        return v;
    };
    if (ld() >= need) return true;
    const unsigned long long t0 = ptx::globaltimer_ns();
    unsigned spins = 0;
    while (ld() < need) {
        if ((++spins & 255u) == 0) {
            if (aborted(p)) return false;
            if (ptx::globaltimer_ns() - t0 > p.timeout_ns) {
                atomicCAS(p.status, 0, MCP_ERR_TIMEOUT);
                return false;
            }
        }
    }
    return true;
}

template <class Op, int TB, bool RES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
fused2d_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ Fused2dParams p) {
    using T = typename Op::T;
    constexpr int VEC = Op::VEC;
    constexpr int M = (TB + VEC * Op::CPE - 1) / (VEC * Op::CPE);
    constexpr int OUTW = (32 - 2 * M) * VEC;  // storage elements written per strip

    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* my_smem = smem + warp * (kStages * kStageBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + WARPS * kStages * kStageBytes) + warp * kStages;
    __shared__ double red_ss[WARPS], red_mx[WARPS];

    if (lane == 0) {
        for (int s = 0; s < kStages; ++s) ptx::mbar_init(&bars[s], 1);
        ptx::mbar_fence_init();
        ptx::prefetch_tensormap(&tmA);
        ptx::prefetch_tensormap(&tmB);
    }
    __syncthreads();

    const int NW = gridDim.x * WARPS;
    const int n_tiles = p.n_strips * p.n_chunks;
    uint32_t it = 0;

    const bool dataflow = !RES && p.sync_mode == 2;
    // One tile of one sweep: wait for what it reads, run it, publish what it wrote.
    auto run_tile = [&](const int sw, const int q, double* partials) {
        const int par = (p.first_parity + sw) & 1;
        const CUtensorMap* tm_in = par ? &tmB : &tmA;
        T* out = reinterpret_cast<T*>(p.buf[par ^ 1]);
        int chunk, strip;
        decode_tile(q, p.n_strips, p.n_chunks, chunk, strip);
        const int y0 = __ldg(&p.chunk_y[chunk]);
        const int y1 = __ldg(&p.chunk_y[chunk + 1]);
        const int x_ld = strip * OUTW - M * VEC;
        const bool top = (chunk == 0), bot = (chunk == p.n_chunks - 1);
        T* up_out = top ? reinterpret_cast<T*>(p.up_buf[par ^ 1]) : nullptr;
        T* dn_out = bot ? reinterpret_cast<T*>(p.dn_buf[par ^ 1]) : nullptr;

        if (dataflow && sw > 0 && lane < 9) {
            // Dataflow instead of a grid barrier: this tile reads what its (up to) 9 neighbours wrote in the
            // previous sweep and overwrites what they read then -- both are settled once they have finished it.
            // One lane per neighbour, so the nine acquire loads are in flight together.
            const int c2 = chunk + lane / 3 - 1, s2 = strip + lane % 3 - 1;
            if (c2 >= 0 && c2 < p.n_chunks && s2 >= 0 && s2 < p.n_strips)
                wait_tile_done(p, &p.tile_done[encode_tile(c2, s2, p.n_strips, p.n_chunks)], (unsigned)sw);
        }
        // Ghost rows this tile reads were written by the neighbour GPU during its previous sweep.
        if (p.strip_flags != nullptr) {
            const unsigned long long need = p.seq_base + (unsigned long long)sw;  // sweeps the neighbour must have done
            if (need > 0) {
                const bool regroup = (sw == 0 && p.prev_n_strips != p.n_strips);  // strips were cut differently
#pragma unroll
                for (int dir = 0; dir < 2; ++dir) {
                    if ((dir == 0 ? up_out : dn_out) == nullptr) continue;
                    const unsigned long long* row = p.strip_flags + dir * kMaxStripFlags;
                    if (regroup) {
                        for (int s2 = lane; s2 < p.prev_n_strips; s2 += 32) wait_counter_sys(p, &row[s2], need);
                    } else if (lane < 3) {
                        const int s2 = strip + lane - 1;
                        if (s2 >= 0 && s2 < p.n_strips) wait_counter_sys(p, &row[s2], need);
                    }
                }
            }
        } else if (lane == 0) {
            bool ok = true;
            if (up_out != nullptr)
                ok = wait_counter_sys(p, &p.my_flags[0], p.flag_base_up + (unsigned long long)sw * p.n_strips);
            if (ok && dn_out != nullptr)
                ok = wait_counter_sys(p, &p.my_flags[1], p.flag_base_dn + (unsigned long long)sw * p.n_strips);
        }
        __syncwarp();
        if (lane == 0) ptx::fence_proxy_async();  // generic-proxy stores (other SMs, peers) -> our TMA loads
        __syncwarp();
        if (aborted(p)) return;

        double rs = 0.0, rm = 0.0;
        const bool trace = (p.debug_flags & 4) != 0;
        const unsigned long long t_begin = trace ? ptx::globaltimer_ns() : 0ull;
        process_tile<Op, TB, RES>(p, tm_in, out, up_out, dn_out, x_ld, y0, y1, my_smem, bars, it, lane, rs, rm);
        if (trace && lane == 0) p.tile_ns[q] = (unsigned)(ptx::globaltimer_ns() - t_begin);

        if (RES) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                rs += __shfl_xor_sync(0xffffffffu, rs, o);
                rm = fmax(rm, __shfl_xor_sync(0xffffffffu, rm, o));
            }
            if (lane == 0) {
                partials[2 * q] = rs;
                partials[2 * q + 1] = rm;
            }
        }
        // Publish: all lanes' stores (local rows, peer ghost rows) -> one release per consumer.
        __syncwarp();
        if (lane == 0) {
            if (dataflow) {
                ptx::fence_proxy_async();
                __threadfence();
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&p.tile_done[q]), "r"((unsigned)(sw + 1))
                             : "memory");
            }
            const unsigned long long done_seq = p.seq_base + (unsigned long long)sw + 1ull;
            if (up_out != nullptr) {
                ptx::fence_acq_rel_sys();
                if (p.up_strip_flag != nullptr) ptx::st_release_sys(&p.up_strip_flag[strip], done_seq);
                ptx::red_add_release_sys(p.up_flag, 1ull);
            }
            if (dn_out != nullptr) {
                ptx::fence_acq_rel_sys();
                if (p.dn_strip_flag != nullptr) ptx::st_release_sys(&p.dn_strip_flag[strip], done_seq);
                ptx::red_add_release_sys(p.dn_flag, 1ull);
            }
        }
    };

    if (dataflow) {
        // ---- barrier-free path: one ticket stream over all sweeps; ticket t = (sweep t / n_tiles, tile t % n_tiles).
        // Tickets are handed out in order and a tile only ever waits for lower tickets, so the lowest unfinished
        // ticket can always run: no deadlock as long as every CTA is resident (cooperative launch).
        const unsigned long long total = (unsigned long long)p.sweeps * (unsigned long long)n_tiles;
        auto grab = [&]() -> unsigned long long {
            unsigned long long v = 0;
            if (lane == 0) v = atomicAdd(p.tile_ctr, 1ull) - p.ctr_base;
            return __shfl_sync(0xffffffffu, v, 0);
        };
        unsigned long long t = grab();
        while (t < total) {
            const unsigned long long t_next = grab();  // latency hidden behind this tile
            if (!aborted(p)) {
                const int sw = (int)(t / (unsigned long long)n_tiles);
                const int q = (int)(t - (unsigned long long)sw * n_tiles);
                run_tile(sw, q, nullptr);
            }
            t = t_next;
        }
        return;
    }

    // ---- barrier path: sweeps separated by a device-side grid barrier (always used with the fused residual, so
    // that CTA 0 can reduce the per-tile partials of a finished sweep)
    unsigned long long epoch = 0;
    int executed = 0;
    for (int sw = 0; sw < p.sweeps; ++sw) {
        if (RES && p.stop_mode != 0 && sw > 0) {
            // decided by CTA 0 from the global residual of sweep sw-2 and published before barrier sw-1
            const int s = *((volatile int*)p.stop_word);
            if (s != 0 && sw >= s) break;
        }
        double* partials = p.partials + (size_t)(sw & 1) * n_tiles * 2;

        // Dynamic tile queue: one monotonic counter; every warp grabs until it draws a ticket past the end, so
        // a sweep consumes exactly n_tiles + NW tickets (the host advances ctr_base by that per sweep).
        const unsigned long long ticket0 = p.ctr_base + (unsigned long long)sw * (unsigned long long)(n_tiles + NW);
        auto grab = [&]() -> int {
            unsigned long long v = 0;
            if (lane == 0) v = atomicAdd(p.tile_ctr, 1ull) - ticket0;
            v = __shfl_sync(0xffffffffu, v, 0);
            return v < (unsigned long long)n_tiles ? (int)v : n_tiles;
        };
        int q = grab();
        while (q < n_tiles) {
            const int q_next = grab();
            if (!aborted(p)) run_tile(sw, q, partials);
            q = q_next;
        }

        if (RES || sw + 1 < p.sweeps) grid_barrier(p, epoch);
        if (aborted(p)) break;

        if (RES && blockIdx.x == 0) {
            // fixed-order reduction of the per-tile partials -> one (sumsq, max) per sweep
            double a = 0.0, b = 0.0;
            for (int i = threadIdx.x; i < n_tiles; i += WARPS * 32) {
                a += __ldcg(&partials[2 * i]);
                b = fmax(b, __ldcg(&partials[2 * i + 1]));
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a += __shfl_xor_sync(0xffffffffu, a, o);
                b = fmax(b, __shfl_xor_sync(0xffffffffu, b, o));
            }
            if (lane == 0) {
                red_ss[warp] = a;
                red_mx[warp] = b;
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                double sa = 0.0, sb = 0.0;
                for (int k = 0; k < WARPS; ++k) {
                    sa += red_ss[k];
                    sb = fmax(sb, red_mx[k]);
                }
                if (p.world > 1 && p.nvls_mc != nullptr) {
                    // NVLS all-reduce: every rank stores its pair into its OWN copy of the slot, one multimem.red
                    // bumps the arrival counter of every GPU through the switch, and multimem.ld_reduce pulls the
                    // sum / max of all copies through the switch in one load each.
                    const unsigned long long seq = p.res_seq_base + (unsigned long long)sw;
                    const size_t off = 128 + (size_t)(seq % kResRing) * 16;
                    *reinterpret_cast<volatile double*>(p.nvls_local + off) = sa;
                    *reinterpret_cast<volatile unsigned long long*>(p.nvls_local + off + 8) =
                        (unsigned long long)__double_as_longlong(sb);  // sb >= 0: bit pattern order == value order
                    ptx::fence_acq_rel_sys();
                    ptx::multimem_red_add_release_u64(reinterpret_cast<unsigned long long*>(p.nvls_mc), 1ull);
                    if (wait_counter_sys(p, reinterpret_cast<const unsigned long long*>(p.nvls_local),
                                         (seq + 1ull) * (unsigned long long)p.world)) {
                        sa = ptx::multimem_ld_reduce_add_f64(reinterpret_cast<const double*>(p.nvls_mc + off));
                        sb = __longlong_as_double((long long)ptx::multimem_ld_reduce_max_u64(
                            reinterpret_cast<const unsigned long long*>(p.nvls_mc + off + 8)));
                    }
                } else if (p.world > 1 && p.peer_inbox[0] != nullptr) {
                    // all-reduce over peer memory: push, signal, wait for everybody, sum in rank order
                    const unsigned long long seq = p.res_seq_base + (unsigned long long)sw;
                    const size_t off = ((size_t)(seq % kResRing) * kMaxRanks + (size_t)p.rank) * 2;
                    for (int r = 0; r < p.world; ++r) {
                        p.peer_inbox[r][off] = sa;
                        p.peer_inbox[r][off + 1] = sb;
                    }
                    ptx::fence_acq_rel_sys();
                    for (int r = 0; r < p.world; ++r) ptx::red_add_release_sys(p.peer_res_flag[r], 1ull);
                    if (wait_counter_sys(p, p.res_flag, (seq + 1ull) * (unsigned long long)p.world)) {
                        const volatile double* in = p.peer_inbox[p.rank] + (size_t)(seq % kResRing) * kMaxRanks * 2;
                        sa = 0.0;
                        sb = 0.0;
                        for (int r = 0; r < p.world; ++r) {
                            sa += in[2 * r];
                            sb = fmax(sb, in[2 * r + 1]);
                        }
                    }
                }
                const int slot = p.res_slot0 + sw;
                if (slot < kMaxResSlots) {
                    p.res_out[2 * slot] = sa;
                    p.res_out[2 * slot + 1] = sb;
                }
                if (p.stop_mode != 0 && *((volatile int*)p.stop_word) == 0) {
                    const bool conv = p.stop_mode == 1 ? (sb <= p.stop_thresh) : (sa <= p.stop_thresh);
                    if (conv) {
                        *((volatile int*)p.stop_word) = sw + 2;  // one more sweep runs while this is published
                        __threadfence();
                    }
                }
            }
            __syncthreads();
        }
        executed = sw + 1;
    }
    if (RES && blockIdx.x == 0 && threadIdx.x == 0) *p.sweeps_done = executed;
}

// ------------------------------------------------------------------------------------------ host plan

}  // namespace mcp

using namespace mcp;

struct mcp_fused2d_plan {
    mcp_fused2d_desc d;
    int rows_alloc;
    int sm_count;
    int grid;
    int parity;
    CUtensorMap tm[2];
    unsigned long long* grid_bar;
    int* status;
    double* partials;
    size_t partials_cap_tiles;
    double* res_out;
    int* chunk_tab;                 // device: kNumTb tables of kMaxChunks + 1 row boundaries
    int* stop_words;                // device: [0] stop word, [1] sweeps done
    unsigned* tile_done;            // device: [partials_cap_tiles]
    unsigned* tile_ns;              // device: [partials_cap_tiles] per-tile trace
    int last_tiles;                 // tile count of the last launch
    unsigned long long res_seq;     // sweeps whose residual has been exchanged so far
    unsigned long long sweep_seq;   // sweeps completed so far (identical on every rank)
    int last_ns;                    // strip count of the previous launch
    int last_generations;           // generations executed by the last run call
    std::vector<int> chunk_host[4];  // per time block 1,2,4,8
    unsigned long long flag_base_up, flag_base_dn, bar_base, ctr_base;
    int res_slots_used;
    int64_t launches;
    unsigned long long timeout_ns;
    int debug_flags;
    int sync_mode;  // 0 = choose per launch, 1 = barrier, 2 = dataflow
};

namespace {

int vec_of(int dtype) { return dtype == MCP_DTYPE_F64 ? 2 : 4; }

int op_kind(const mcp_fused2d_desc& d) { return d.dtype == MCP_DTYPE_B1 ? 2 : (d.op == MCP_OP_LIFE ? 1 : 0); }

int cpe_of(int dtype) { return dtype == MCP_DTYPE_B1 ? 32 : 1; }  // cells per storage element

// output columns (cells) per strip
int outw_rt(int dtype, int tb) {
    const int vec = vec_of(dtype), cpe = cpe_of(dtype);
    return (32 - 2 * ((tb + vec * cpe - 1) / (vec * cpe))) * vec * cpe;
}

int max_tb(int op, int dtype) {
    if (op == MCP_OP_LIFE) return 4;
    return 8;
}

constexpr int kMaxChunks = 4096;

int tb_index(int tb) { return tb == 1 ? 0 : tb == 2 ? 1 : tb == 4 ? 2 : 3; }

// Row-chunk schedule for the dynamic tile queue.  Tiles are drawn in chunk order, so chunk heights follow a
// guided self-scheduling curve: tall chunks first (the 2*TB redundant pre-roll rows per chunk stay cheap),
// short chunks last (the tail of a sweep is one *short* tile).  Used for experiments through MCP_GSS_*; the
// default schedule (build_chunks) is uniform because the measured gain of the taper was within noise.
void gss_chunks(int rows, int own_lo, int ns, double NW, int tb, int G, double k, int rmin_mult, int rmax_mult,
                std::vector<int>* out) {
    std::vector<int>& y = *out;
    y.clear();
    y.push_back(own_lo);
    const int rmin = std::max(G, rmin_mult * tb);
    const int rmax = std::max(rmin, rmax_mult * tb);
    int done = 0;
    while (done < rows) {
        const int rem = rows - done;
        int r = (int)((double)rem * ns / (k * NW));  // remaining work ~ k tiles per warp
        r = std::max(rmin, std::min(rmax, r));
        // a tile reads r + 2*TB rows in stages of kRowsPerStage: make that a whole number of stages
        r -= (r + 2 * tb) % kRowsPerStage;
        if (r < rmin) r += kRowsPerStage;
        if (rem - r < rmin) r = rem;  // no sliver at the end (every chunk >= ghost depth)
        if ((int)y.size() >= kMaxChunks) r = rem;
        done += r;
        y.push_back(own_lo + done);
    }
}

void build_chunks(const mcp_fused2d_plan* pl, int tb, std::vector<int>* out, int* n_strips) {
    const mcp_fused2d_desc& d = pl->d;
    const int outw = outw_rt(d.dtype, tb);
    const int ns = (int)ceil_div(d.w, outw);
    const int G = std::max(d.ghost, 1);
    const int own_lo = d.ghost;
    const int rows = d.rows_own;
    *n_strips = ns;
    std::vector<int>& y = *out;
    if (d.chunk_rows > 0) {
        y.clear();
        y.push_back(own_lo);
        const int nc = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(ceil_div(rows, d.chunk_rows),
                                                                                   std::max(1, rows / G)), kMaxChunks));
        for (int c = 1; c <= nc; ++c) y.push_back(own_lo + (int)((long long)rows * c / nc));
        return;
    }
    const int NW = pl->grid * warps_for(op_kind(d), tb);
    const char* e_k = getenv("MCP_GSS_K");
    const char* e_rmin = getenv("MCP_GSS_RMIN");
    const char* e_rmax = getenv("MCP_GSS_RMAX");
    if (e_k || e_rmin || e_rmax) {  // manual override for experiments
        gss_chunks(rows, own_lo, ns, NW, tb, G, e_k ? atof(e_k) : 3.0, e_rmin ? atoi(e_rmin) : 8,
                   e_rmax ? atoi(e_rmax) : 48, &y);
        return;
    }
    // Two overheads pull in opposite directions: every chunk re-reads and partly recomputes 2*TB pre-roll rows and
    // refills its TMA pipeline (~(TB + 6) row-times per tile), and the sweep ends with about half a tile of idle
    // time per warp.  With `share` rows of work per warp, overhead(r) ~ (TB + 6)/r + r/(2*share) is minimal at
    // r* = sqrt(2*share*(TB + 6)); the measured optimum on B200 follows this within noise (profiles/).
    const double share = (double)rows * ns / NW;
    const int rmin = std::max(G, 4 * tb);
    int r = (int)std::lround(std::sqrt(2.0 * share * (tb + 6)));
    r = std::max(rmin, std::min(64 * tb, r));
    r -= (r + 2 * tb) % kRowsPerStage;  // whole number of TMA stages per tile
    if (r < rmin) r += kRowsPerStage;
    y.clear();
    y.push_back(own_lo);
    int done = 0;
    while (done < rows) {
        int take = std::min(r, rows - done);
        if (rows - done - take < rmin) take = rows - done;  // fold a sliver into the last chunk
        if ((int)y.size() >= kMaxChunks) take = rows - done;
        done += take;
        y.push_back(own_lo + done);
    }
}

void choose_geometry(const mcp_fused2d_plan* pl, int tb, int* n_strips, int* n_chunks) {
    const std::vector<int>& y = pl->chunk_host[tb_index(tb)];
    *n_strips = (int)ceil_div(pl->d.w, outw_rt(pl->d.dtype, tb));
    *n_chunks = (int)y.size() - 1;
}

template <class Op, int TB>
int launch_variant(mcp_fused2d_plan* pl, const Fused2dParams& p, bool res, cudaStream_t stream) {
    constexpr int W = warps_for(std::is_same<Op, LifeB1>::value ? 2 : (std::is_same<Op, LifeI32>::value ? 1 : 0), TB);
    constexpr int smem = smem_bytes_for(W);
    void* args[3] = {(void*)&pl->tm[0], (void*)&pl->tm[1], (void*)&p};
    const void* fn = res ? (const void*)fused2d_kernel<Op, TB, true, W> : (const void*)fused2d_kernel<Op, TB, false, W>;
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCP_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(pl->grid), dim3(W * 32), args, smem, stream));
    pl->launches++;
    return MCP_OK;
}

template <class Op>
int launch_tb(mcp_fused2d_plan* pl, int tb, const Fused2dParams& p, bool res, cudaStream_t stream) {
    switch (tb) {
        case 1: return launch_variant<Op, 1>(pl, p, res, stream);
        case 2: return launch_variant<Op, 2>(pl, p, res, stream);
        case 4: return launch_variant<Op, 4>(pl, p, res, stream);
        default: break;
    }
    set_error("unsupported time block %d", tb);
    return MCP_ERR_UNSUPPORTED;
}

int launch_any(mcp_fused2d_plan* pl, int tb, const Fused2dParams& p, bool res, cudaStream_t stream) {
    const mcp_fused2d_desc& d = pl->d;
    if (d.op == MCP_OP_MDF && d.dtype == MCP_DTYPE_F32) {
        if (tb == 8) return launch_variant<MdfF32, 8>(pl, p, res, stream);
        return launch_tb<MdfF32>(pl, tb, p, res, stream);
    }
    if (d.op == MCP_OP_MDF && d.dtype == MCP_DTYPE_F64) {
        if (tb == 8) return launch_variant<MdfF64, 8>(pl, p, res, stream);
        return launch_tb<MdfF64>(pl, tb, p, res, stream);
    }
    if (d.op == MCP_OP_LIFE && d.dtype == MCP_DTYPE_I32) return launch_tb<LifeI32>(pl, tb, p, res, stream);
    if (d.op == MCP_OP_LIFE && d.dtype == MCP_DTYPE_B1) return launch_tb<LifeB1>(pl, tb, p, res, stream);
    set_error("unsupported op/dtype combination %d/%d", d.op, d.dtype);
    return MCP_ERR_UNSUPPORTED;
}

}  // namespace

extern "C" {

int mcp_fused2d_create(const mcp_fused2d_desc* desc, mcp_fused2d_plan** plan_out) {
    MCP_REQUIRE(desc && plan_out, "mcp_fused2d_create: null argument");
    const mcp_fused2d_desc& d = *desc;
    MCP_REQUIRE(d.op == MCP_OP_MDF || d.op == MCP_OP_LIFE, "bad op %d", d.op);
    MCP_REQUIRE((d.op == MCP_OP_LIFE) == (d.dtype == MCP_DTYPE_I32 || d.dtype == MCP_DTYPE_B1),
                "op %d does not take dtype %d", d.op, d.dtype);
    MCP_REQUIRE(d.h >= 1 && d.w >= 1 && d.rows_own >= 1, "empty grid");
    MCP_REQUIRE(d.ghost >= 1 && d.ghost <= 8, "ghost depth %d outside [1,8]", d.ghost);
    MCP_REQUIRE(d.pitch * cpe_of(d.dtype) >= d.w && d.pitch % 32 == 0,
                "pitch %lld (elements) must be a multiple of 32 and cover w", (long long)d.pitch);
    MCP_REQUIRE(d.buf[0] && d.buf[1], "null field buffer");
    MCP_REQUIRE(((uintptr_t)d.buf[0] % 128 == 0) && ((uintptr_t)d.buf[1] % 128 == 0), "buffers must be 128-byte aligned");
    MCP_REQUIRE((d.up_buf[0] == nullptr) == (d.up_buf[1] == nullptr), "up_buf must be both set or both null");
    MCP_REQUIRE((d.dn_buf[0] == nullptr) == (d.dn_buf[1] == nullptr), "dn_buf must be both set or both null");
    if (d.up_buf[0] || d.dn_buf[0]) {
        MCP_REQUIRE(d.flags != nullptr, "flags required when neighbours exist");
        MCP_REQUIRE(d.rows_own >= d.ghost, "a slab of %d rows is thinner than the ghost depth %d", d.rows_own, d.ghost);
    }
    MCP_REQUIRE(!d.up_buf[0] || d.up_flag, "up_flag required with up_buf");
    MCP_REQUIRE(!d.dn_buf[0] || d.dn_flag, "dn_flag required with dn_buf");
    if (d.strip_flags != nullptr) {
        MCP_REQUIRE(!d.up_buf[0] || d.up_strip_flag, "up_strip_flag required with strip_flags and up_buf");
        MCP_REQUIRE(!d.dn_buf[0] || d.dn_strip_flag, "dn_strip_flag required with strip_flags and dn_buf");
        MCP_REQUIRE(ceil_div(d.w, outw_rt(d.dtype, 1)) <= kMaxStripFlags, "grid too wide for per-strip halo flags");
    }

    mcp_fused2d_plan* pl = new (std::nothrow) mcp_fused2d_plan();
    MCP_REQUIRE(pl, "out of host memory");
    pl->d = d;
    pl->rows_alloc = d.rows_own + 2 * d.ghost;
    pl->parity = 0;
    pl->flag_base_up = pl->flag_base_dn = pl->bar_base = pl->ctr_base = 0;
    pl->res_slots_used = 0;
    pl->launches = 0;
    pl->timeout_ns = 20ull * 1000000000ull;
    pl->debug_flags = getenv("MCP_DEBUG_FLAGS") ? atoi(getenv("MCP_DEBUG_FLAGS")) : 0;
    pl->sync_mode = getenv("MCP_SYNC_MODE") ? atoi(getenv("MCP_SYNC_MODE")) : 0;
    pl->grid_bar = nullptr;
    pl->status = nullptr;
    pl->partials = nullptr;
    pl->res_out = nullptr;
    pl->chunk_tab = nullptr;
    pl->stop_words = nullptr;
    pl->tile_done = nullptr;
    pl->tile_ns = nullptr;
    pl->last_tiles = 0;
    pl->res_seq = 0;
    pl->sweep_seq = 0;
    pl->last_ns = 0;
    pl->last_generations = 0;

    int rc = device_sm_count(&pl->sm_count);
    if (rc) { delete pl; return rc; }
    pl->grid = pl->sm_count;
    if (d.max_ctas > 0) pl->grid = std::min(pl->grid, d.max_ctas);

    const int esz = d.dtype == MCP_DTYPE_F64 ? 8 : 4;
    for (int b = 0; b < 2; ++b) {
        uint64_t dims[2] = {(uint64_t)d.pitch, (uint64_t)pl->rows_alloc};
        uint64_t strides[1] = {(uint64_t)d.pitch * esz};
        uint32_t box[2] = {(uint32_t)(kRowBytes / esz), (uint32_t)kRowsPerStage};
        rc = encode_tensor_map(&pl->tm[b], d.dtype == MCP_DTYPE_B1 ? (int)MCP_DTYPE_I32 : d.dtype, 2, d.buf[b], dims,
                               strides, box);
        if (rc) { delete pl; return rc; }
    }
    // scratch: grid barrier counter, status word, residual partials + history
    const int min_outw = outw_rt(d.dtype, max_tb(d.op, d.dtype));
    const size_t max_strips = (size_t)ceil_div(d.w, min_outw);
    const size_t max_chunks = (size_t)std::max(1, d.rows_own / std::max(d.ghost, 1));
    pl->partials_cap_tiles = max_strips * std::min<size_t>(max_chunks, 4096);
    cudaError_t e;
    if ((e = cudaMalloc(&pl->grid_bar, 256)) != cudaSuccess || (e = cudaMemset(pl->grid_bar, 0, 256)) != cudaSuccess ||
        (e = cudaMalloc(&pl->status, 256)) != cudaSuccess || (e = cudaMemset(pl->status, 0, 256)) != cudaSuccess ||
        (e = cudaMalloc(&pl->partials, pl->partials_cap_tiles * 4 * sizeof(double))) != cudaSuccess ||
        (e = cudaMalloc(&pl->res_out, kMaxResSlots * 2 * sizeof(double))) != cudaSuccess ||
        (e = cudaMemset(pl->res_out, 0, kMaxResSlots * 2 * sizeof(double))) != cudaSuccess) {
        set_error("mcp_fused2d_create: scratch allocation failed: %s", cudaGetErrorString(e));
        mcp_fused2d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    if ((e = cudaMalloc(&pl->stop_words, 256)) != cudaSuccess || (e = cudaMemset(pl->stop_words, 0, 256)) != cudaSuccess) {
        set_error("mcp_fused2d_create: stop word allocation failed: %s", cudaGetErrorString(e));
        mcp_fused2d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    if ((e = cudaMalloc(&pl->tile_ns, pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess ||
        (e = cudaMemset(pl->tile_ns, 0, pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess ||
        (e = cudaMalloc(&pl->tile_done, pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess) {
        set_error("mcp_fused2d_create: tile flag allocation failed: %s", cudaGetErrorString(e));
        mcp_fused2d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    if ((e = cudaMalloc(&pl->chunk_tab, 4 * (kMaxChunks + 1) * sizeof(int))) != cudaSuccess) {
        set_error("mcp_fused2d_create: chunk table allocation failed: %s", cudaGetErrorString(e));
        mcp_fused2d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    for (int tb = 1; tb <= 8; tb *= 2) {
        int ns = 0;
        std::vector<int>& y = pl->chunk_host[tb_index(tb)];
        build_chunks(pl, tb, &y, &ns);
        e = cudaMemcpy(pl->chunk_tab + tb_index(tb) * (kMaxChunks + 1), y.data(), y.size() * sizeof(int),
                       cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            set_error("mcp_fused2d_create: chunk table upload failed: %s", cudaGetErrorString(e));
            mcp_fused2d_destroy(pl);
            return MCP_ERR_CUDA;
        }
    }
    *plan_out = pl;
    return MCP_OK;
}

int mcp_fused2d_geometry(const mcp_fused2d_plan* pl, int time_block, int* n_strips, int* n_chunks, int* grid,
                         int* smem_bytes) {
    MCP_REQUIRE(pl, "null plan");
    int ns, nc;
    choose_geometry(pl, time_block, &ns, &nc);
    if (n_strips) *n_strips = ns;
    if (n_chunks) *n_chunks = nc;
    if (grid) *grid = pl->grid;
    if (smem_bytes) *smem_bytes = smem_bytes_for(warps_for(op_kind(pl->d), time_block));
    return MCP_OK;
}

int mcp_fused2d_chunks(const mcp_fused2d_plan* pl, int time_block, int* rows_out, int max_entries) {
    MCP_REQUIRE(pl && rows_out, "null argument");
    const std::vector<int>& y = pl->chunk_host[tb_index(time_block)];
    const int n = std::min((int)y.size() - 1, max_entries);
    for (int i = 0; i < n; ++i) rows_out[i] = y[i + 1] - y[i];
    return n;
}

static int run_impl(mcp_fused2d_plan* pl, int generations, int time_block, bool res, int stop_mode,
                    double stop_thresh, cudaStream_t stream) {
    const mcp_fused2d_desc& d = pl->d;
    const int tbmax = std::min(max_tb(d.op, d.dtype), d.ghost);
    int tb = 1;
    while (tb * 2 <= time_block && tb * 2 <= tbmax) tb *= 2;
    if (res) pl->res_slots_used = 0;
    pl->last_generations = 0;
    const bool exchange = d.world > 1 && (d.res_inbox != nullptr || d.nvls_mc != nullptr);

    int remaining = generations;
    while (remaining > 0) {
        while (tb > remaining) tb >>= 1;
        const int sweeps = remaining / tb;
        int ns, nc;
        choose_geometry(pl, tb, &ns, &nc);
        MCP_REQUIRE((size_t)ns * nc <= pl->partials_cap_tiles, "tile count %d x %d exceeds scratch capacity", ns, nc);

        Fused2dParams p;
        memset(&p, 0, sizeof(p));
        for (int b = 0; b < 2; ++b) {
            p.buf[b] = d.buf[b];
            p.up_buf[b] = d.up_buf[b];
            p.dn_buf[b] = d.dn_buf[b];
        }
        p.my_flags = (const unsigned long long*)d.flags;
        p.up_flag = (unsigned long long*)d.up_flag;
        p.dn_flag = (unsigned long long*)d.dn_flag;
        p.grid_bar = pl->grid_bar;
        p.tile_ctr = pl->grid_bar + 16;  // separate 128-byte line of the same zeroed allocation
        p.ctr_base = pl->ctr_base;
        p.tile_done = pl->tile_done;
        p.tile_ns = pl->tile_ns;
        pl->last_tiles = ns * nc;
        p.status = pl->status;
        p.partials = pl->partials;
        p.res_out = pl->res_out;
        p.chunk_y = pl->chunk_tab + tb_index(tb) * (kMaxChunks + 1);
        p.flag_base_up = pl->flag_base_up;
        p.flag_base_dn = pl->flag_base_dn;
        p.bar_base = pl->bar_base;
        p.timeout_ns = pl->timeout_ns;
        p.pitch = d.pitch;
        p.rows_alloc = pl->rows_alloc;
        p.h = d.h;
        p.w = d.w;
        p.row0 = d.row_start - d.ghost;
        p.own_lo = d.ghost;
        p.own_hi = d.ghost + d.rows_own;
        p.ylo = d.ylo; p.yhi = d.yhi; p.xlo = d.xlo; p.xhi = d.xhi;
        p.G = d.ghost;
        p.up_ghost_row = d.ghost + d.up_rows_own;
        p.n_strips = ns;
        p.n_chunks = nc;
        p.sweeps = sweeps;
        p.first_parity = pl->parity;
        p.res_slot0 = pl->res_slots_used;
        p.debug_flags = pl->debug_flags;
        p.world = d.world > 0 ? d.world : 1;
        p.rank = d.rank;
        p.res_flag = (const unsigned long long*)d.res_flag;
        p.res_seq_base = pl->res_seq;
        if (exchange)
            for (int r = 0; r < p.world; ++r) {
                p.peer_inbox[r] = (double*)d.peer_res_inbox[r];
                p.peer_res_flag[r] = (unsigned long long*)d.peer_res_flag[r];
            }
        p.strip_flags = (const unsigned long long*)d.strip_flags;
        p.up_strip_flag = (unsigned long long*)d.up_strip_flag;
        p.dn_strip_flag = (unsigned long long*)d.dn_strip_flag;
        p.seq_base = pl->sweep_seq;
        p.prev_n_strips = pl->last_ns;
        p.nvls_local = (unsigned char*)d.nvls_local;
        p.nvls_mc = (unsigned char*)d.nvls_mc;
        p.stop_word = pl->stop_words;
        p.sweeps_done = pl->stop_words + 1;
        p.stop_mode = stop_mode;
        p.stop_thresh = stop_thresh;
        if (stop_mode != 0) MCP_CUDA_TRY(cudaMemsetAsync(pl->stop_words, 0, 2 * sizeof(int), stream));
        // Sweep synchronisation: per-tile dataflow flags remove the barrier and the idle tail of every sweep, which
        // pays when a warp has few tiles per sweep; with many tiles per warp the plain barrier is cheaper.
        int sync_mode = pl->sync_mode;
        if (sync_mode == 0) {
            const double tiles_per_warp = (double)ns * nc / ((double)pl->grid * warps_for(op_kind(d), tb));
            sync_mode = (tb < 8 || tiles_per_warp < 5.0) ? 2 : 1;
        }
        if (res) sync_mode = 1;
        p.sync_mode = sync_mode;
        if (sync_mode == 2) MCP_CUDA_TRY(cudaMemsetAsync(pl->tile_done, 0, (size_t)ns * nc * sizeof(unsigned), stream));

        int rc = launch_any(pl, tb, p, res, stream);
        if (rc) return rc;

        int executed = sweeps;
        if (stop_mode != 0) {
            // the number of sweeps that ran is decided on the device: wait for it before touching the counters
            MCP_CUDA_TRY(cudaStreamSynchronize(stream));
            int words[2] = {0, 0};
            MCP_CUDA_TRY(cudaMemcpy(words, pl->stop_words, sizeof(words), cudaMemcpyDeviceToHost));
            executed = words[1];
            MCP_REQUIRE(executed >= 0 && executed <= sweeps, "device reported %d executed sweeps of %d", executed, sweeps);
        }
        const unsigned long long nw = (unsigned long long)pl->grid * warps_for(op_kind(d), tb);
        if (sync_mode == 1) {
            const unsigned long long barriers = res ? (unsigned long long)executed : (unsigned long long)(executed - 1);
            pl->bar_base += barriers * (unsigned long long)pl->grid;
            pl->ctr_base += (unsigned long long)executed * ((unsigned long long)ns * nc + nw);
        } else {
            pl->ctr_base += (unsigned long long)executed * ((unsigned long long)ns * nc) + nw;
        }
        pl->flag_base_up += (unsigned long long)executed * ns;
        pl->flag_base_dn += (unsigned long long)executed * ns;
        pl->parity = (pl->parity + executed) & 1;
        pl->sweep_seq += (unsigned long long)executed;
        pl->last_ns = ns;
        if (res) {
            pl->res_slots_used = std::min(kMaxResSlots, pl->res_slots_used + executed);
            if (exchange) pl->res_seq += (unsigned long long)executed;
        }
        pl->last_generations += executed * tb;
        if (stop_mode != 0) break;  // tolerance runs use one block size; the caller decides how to continue
        remaining -= sweeps * tb;
    }
    return MCP_OK;
}

int mcp_fused2d_run(mcp_fused2d_plan* pl, int generations, int time_block, int residual_mode, void* stream_) {
    MCP_REQUIRE(pl, "null plan");
    MCP_REQUIRE(generations >= 0, "negative generation count");
    MCP_REQUIRE(time_block >= 1, "time_block must be >= 1");
    return run_impl(pl, generations, time_block, residual_mode != MCP_RES_NONE, 0, 0.0, (cudaStream_t)stream_);
}

int mcp_fused2d_run_to_tolerance(mcp_fused2d_plan* pl, int max_generations, int time_block, int norm, double tol,
                                 int* generations_done, void* stream_) {
    MCP_REQUIRE(pl, "null plan");
    MCP_REQUIRE(max_generations >= 1 && time_block >= 1, "bad generation count or time block");
    MCP_REQUIRE(norm == MCP_NORM_LINF || norm == MCP_NORM_L2, "norm must be MCP_NORM_LINF or MCP_NORM_L2");
    MCP_REQUIRE(tol >= 0.0, "negative tolerance");
    const double thresh = norm == MCP_NORM_L2 ? tol * tol : tol;  // the kernel compares the sum of squares
    int rc = run_impl(pl, max_generations, time_block, true, norm == MCP_NORM_LINF ? 1 : 2, thresh,
                      (cudaStream_t)stream_);
    if (generations_done) *generations_done = pl->last_generations;
    return rc;
}

// Queue position <-> (chunk, strip), the same functions the kernel uses (host build of decode_tile/encode_tile).
int mcp_fused2d_tile_at(int n_strips, int n_chunks, int q, int* chunk, int* strip) {
    MCP_REQUIRE(chunk && strip && n_strips >= 1 && n_chunks >= 1 && q >= 0 && q < n_strips * n_chunks, "bad arguments");
    decode_tile(q, n_strips, n_chunks, *chunk, *strip);
    return encode_tile(*chunk, *strip, n_strips, n_chunks);
}

// Device-free planning: the tile schedule a plan WOULD use (chunk heights in queue order, strip count) for a slab of
// `rows_own` rows and `w` cells on `grid` CTAs.  Pure host arithmetic -- usable on a box without a GPU (tests).
int mcp_fused2d_plan_schedule(int op, int dtype, int w, int rows_own, int ghost, int time_block, int grid,
                              int chunk_rows, int* rows_out, int max_entries, int* n_strips) {
    MCP_REQUIRE(rows_out && n_strips, "null argument");
    MCP_REQUIRE(w >= 1 && rows_own >= 1 && ghost >= 1 && grid >= 1, "bad geometry");
    MCP_REQUIRE(time_block == 1 || time_block == 2 || time_block == 4 || time_block == 8, "time_block must be 1, 2, 4 or 8");
    mcp_fused2d_plan tmp;
    memset(&tmp.d, 0, sizeof(tmp.d));
    tmp.d.op = op;
    tmp.d.dtype = dtype;
    tmp.d.w = w;
    tmp.d.rows_own = rows_own;
    tmp.d.ghost = ghost;
    tmp.d.chunk_rows = chunk_rows;
    tmp.grid = grid;
    std::vector<int> y;
    int ns = 0;
    build_chunks(&tmp, time_block, &y, &ns);
    *n_strips = ns;
    const int n = std::min((int)y.size() - 1, max_entries);
    for (int i = 0; i < n; ++i) rows_out[i] = y[i + 1] - y[i];
    return (int)y.size() - 1;
}

// Trace (MCP_DEBUG_FLAGS & 4): duration in ns of every tile in the last sweep it ran, in queue order, with its
// (chunk, strip) so that a host tool can see which tiles are the critical path.
int mcp_fused2d_tile_trace(mcp_fused2d_plan* pl, int time_block, uint32_t* ns_out, int* chunk_out, int* strip_out,
                           int max_entries) {
    MCP_REQUIRE(pl && ns_out, "null argument");
    int nstr, nch;
    choose_geometry(pl, time_block, &nstr, &nch);
    const int n = std::min(std::min(nstr * nch, pl->last_tiles), max_entries);
    if (n > 0) MCP_CUDA_TRY(cudaMemcpy(ns_out, pl->tile_ns, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    for (int q = 0; q < n; ++q) {
        const int ci = q / nstr;
        if (strip_out) strip_out[q] = q - ci * nstr;
        if (chunk_out) chunk_out[q] = (ci & 1) ? nch - 1 - (ci >> 1) : (ci >> 1);
    }
    return n;
}

int mcp_fused2d_expected_flags(const mcp_fused2d_plan* pl, uint64_t* up_dn) {
    MCP_REQUIRE(pl && up_dn, "null argument");
    up_dn[0] = pl->flag_base_up;
    up_dn[1] = pl->flag_base_dn;
    return MCP_OK;
}

int mcp_fused2d_parity(const mcp_fused2d_plan* pl) { return pl ? pl->parity : 0; }

int mcp_fused2d_set_parity(mcp_fused2d_plan* pl, int parity) {
    MCP_REQUIRE(pl, "null plan");
    pl->parity = parity & 1;
    return MCP_OK;
}

int mcp_fused2d_residuals(mcp_fused2d_plan* pl, double* out, int max_entries) {
    MCP_REQUIRE(pl && out, "null argument");
    const int n = std::min(pl->res_slots_used, max_entries);
    if (n > 0) {
        cudaError_t e = cudaMemcpy(out, pl->res_out, (size_t)n * 2 * sizeof(double), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            set_error("mcp_fused2d_residuals: %s", cudaGetErrorString(e));
            return MCP_ERR_CUDA;
        }
    }
    return n;
}

int mcp_fused2d_status(mcp_fused2d_plan* pl) {
    MCP_REQUIRE(pl, "null plan");
    int st = 0;
    MCP_CUDA_TRY(cudaMemcpy(&st, pl->status, sizeof(int), cudaMemcpyDeviceToHost));
    if (st != 0) set_error("fused2d kernel recorded status %d (%s)", st, st == MCP_ERR_TIMEOUT ? "flag/barrier wait timed out" : "error");
    return st;
}

int64_t mcp_fused2d_launches(const mcp_fused2d_plan* pl) { return pl ? pl->launches : 0; }

int mcp_fused2d_destroy(mcp_fused2d_plan* pl) {
    if (!pl) return MCP_OK;
    if (pl->grid_bar) cudaFree(pl->grid_bar);
    if (pl->status) cudaFree(pl->status);
    if (pl->partials) cudaFree(pl->partials);
    if (pl->res_out) cudaFree(pl->res_out);
    if (pl->chunk_tab) cudaFree(pl->chunk_tab);
    if (pl->stop_words) cudaFree(pl->stop_words);
    if (pl->tile_done) cudaFree(pl->tile_done);
    if (pl->tile_ns) cudaFree(pl->tile_ns);
    delete pl;
    return MCP_OK;
}

}  // extern "C"
