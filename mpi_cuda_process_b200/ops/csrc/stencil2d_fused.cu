// Fused halo-exchange + 2-D stencil (+ residual) persistent kernel for sm_100a.
//
// What it replaces in the reference: the per-generation sequence of MDF_kernel.cu:155-222 --
// full-grid H2D copy, middle_kernel, w-2 one-element MPI_Recv, second H2D copy, border_kernel,
// cudaDeviceSynchronize, full-grid D2H copy, w-2 one-element MPI_Send -- and the kernels
// run_mdf/middle_kernel/border_kernel (MDF_kernel.cu:10-70), game_of_life (kernel.cu:10-113).
//
// Design (B200-first, not a translation):
//  * One cooperative launch runs many generations.  148 CTAs x 8..16 warps stay resident.
//  * The slab is cut into tiles = (column strip of 32 lanes x 16 B) x (row chunk).  A tile belongs to ONE
//    warp; warps never synchronise with each other inside a sweep.
//  * Each warp owns a 4-stage shared-memory ring fed by TMA (cp.async.bulk.tensor.2d, 6 rows x 512 B
//    per stage, mbarrier completion).  TMA's out-of-bounds fill supplies the cells outside the allocation.
//  * Temporal blocking in registers: a warp streams down the rows and carries TB generations as a
//    wavefront (3-row window per level, compile-time rotated), so HBM sees one read + one write per
//    TB generations.  Left/right neighbours come from warp shuffles.  Each update is the reference's
//    exact rounding sequence (fp32: on the packed fp32 pipe, FADD2 / FFMA2), so results are bit-identical to
//    a generation-at-a-time Jacobi.
//  * Tiles on the slab cut store their top/bottom `ghost` output rows twice: locally and straight into
//    the neighbour GPU's ghost rows through an NVLink-mapped peer pointer, then ONE system-scope fence and
//    relaxed signal writes into the neighbour's flag pad (per-strip sequence flags + an arrival counter).
//    Tiles that read ghost rows acquire-spin (bounded) on their local flags first.  Boundary tiles are
//    scheduled first so the transfer overlaps the interior tiles.
//  * Between sweeps: per-tile dataflow flags (a tile of sweep k+1 waits for its nine neighbours of sweep k;
//    one ticket stream over all sweeps; the next tile's first TMA stages are put in flight from the current
//    tile's tail), or a device-side grid barrier where a finished sweep must be reduced.  No host round
//    trip, no NCCL, no MPI.  Before a launch ends it drains the halo pushes addressed to this rank.
//  * Optional fused residual: per-tile (sum d^2, max |d|) in fp64 -> fixed-order reduction by one warp ->
//    all-reduce over the peers' inboxes (or NVLS multimem) -> history, stop decision, 32-byte result block.
#include <algorithm>
#include <type_traits>
#include <cmath>
#include <new>
#include <vector>

#include "../../../include/mcp_b200.h"
#include "common.cuh"
#include "ptx.cuh"
#include "stencil2d_ops.cuh"

namespace mcp {

constexpr int kWarpsPerCta = 16;
#ifndef MCP_ROWS_PER_STAGE
#define MCP_ROWS_PER_STAGE 6
#endif
#ifndef MCP_STAGES
#define MCP_STAGES 4
#endif
constexpr int kRowsPerStage = MCP_ROWS_PER_STAGE;  // multiple of 3: the 3-row register window rotates at compile time
constexpr int kStages = MCP_STAGES;
constexpr int kRowBytes = 512;  // 32 lanes x 16 B
constexpr int kStageBytes = kRowsPerStage * kRowBytes;
constexpr int smem_bytes_for(int warps) { return warps * kStages * kStageBytes + warps * kStages * 8; }
// TB = 8 keeps 96 registers of row windows per thread: 8 warps x up to 255 registers instead of 16 x 128 (12 warps x
// 168 registers were measured too: no gain, profiles/README.md).
constexpr int kWarpsTb8 = 8;
#ifndef MCP_WARPS_F32_TB8
#define MCP_WARPS_F32_TB8 8
#endif
constexpr int kWarpsF32Tb8 = MCP_WARPS_F32_TB8;  // fp32 MDF at TB = 8 (experiment knob)
__host__ __device__ constexpr int warps_for_tb(int tb) { return tb >= 8 ? kWarpsTb8 : kWarpsPerCta; }
// Register budget per operator: kind 0 = scalar float cells (12 window registers per level), 1 = Life int32 (cells
// + horizontal sums, 24 per level), 2 = bit-packed Life (cells + two sum planes, 36 per level).  When the windows of
// TB levels no longer fit 128 registers the CTA runs 8 warps x up to 255 registers instead of 16 x 128.
__host__ __device__ constexpr int warps_for(int kind, int tb) {
    if (kind == 0 && tb >= 8) return kWarpsF32Tb8;
    if (kind == 3) return warps_for_tb(tb);  // fp64 MDF
    if (kind == 2 && tb == 2) return 12;  // 148 registers: 12 warps x 32 x 148 fits the register file
    return (kind == 2 && tb >= 2) || (kind == 1 && tb >= 4) ? kWarpsTb8 : warps_for_tb(tb);
}
constexpr int kMaxResSlots = 16384;
constexpr int kMaxRanks = MCP_MAX_RANKS;
constexpr int kResRing = MCP_RES_RING;
constexpr int kMaxStripFlags = MCP_MAX_STRIPS;

// One 32-byte block per plan that a host reads back with ONE copy after a run.
struct ResultBlock {
    int status;       // 0 or the MCP_ERR_* code a kernel recorded
    int stop_word;    // run-to-tolerance: 0 = keep going, else the number of sweeps this launch executes
    int sweeps_done;  // sweeps actually executed by the last launch
    int pad;
    double last_sumsq, last_max;  // global residual of the last residual sweep
};

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
    unsigned* tile_ns;    // trace (debug_flags & 4), ns summed over the launch's sweeps: [0][q] row loop,
                          // [1][q] waiting for dependencies, [2][q] halo push + publish; planes of trace_cap words
    int trace_cap;
    int* status;
    double* partials;  // [2][n_tiles][2]
    double* res_out;   // [kMaxResSlots][2]
    ResultBlock* result;  // status / stop word / executed sweeps / last global residual, read back in one copy
    unsigned* res_count;  // tiles of the residual sweep that have delivered their partial (barrier-free path)
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
    int n_strips, n_chunks, chunk_r;  // a tile = one column strip x one row chunk of chunk_r rows (chunk_row)
    int sweeps, first_parity, res_slot0;
    int res_first_sweep;  // sweeps >= this accumulate the residual (0 = every sweep, sweeps - 1 = the last one only)
    int debug_flags;
    int prefetch;   // barrier-free path: put the next tile's first TMA stages in flight from the current tile's tail
    int sync_mode;  // 1 = grid barrier between sweeps, 2 = per-tile dataflow flags (no barrier)
};

// operators: stencil2d_ops.cuh


// Schedule arithmetic shared by the host plan and the kernel (no tables: a tile start costs no dependent loads).
// n_chunks - 1 chunks of chunk_r rows and a last one that takes the remainder (between half and 1.5 x chunk_r).
__host__ __device__ __forceinline__ int chunk_row_of(int own_lo, int rows, int n_chunks, int chunk_r, int c) {
    return c < n_chunks ? own_lo + c * chunk_r : own_lo + rows;
}

// ------------------------------------------------------------------------------------------ waits

__device__ __forceinline__ bool aborted(const Fused2dParams& p) { return *((volatile int*)p.status) != 0; }

__device__ __forceinline__ int chunk_row(const Fused2dParams& p, int c) {
    return chunk_row_of(p.own_lo, p.own_hi - p.own_lo, p.n_chunks, p.chunk_r, c);
}

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

// Bounded acquire spin at GPU scope on a per-(chunk, strip) completion word.
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

// TMA completion wait.  The fast path is one try_wait; a wait that keeps failing (a faulted or never-issued copy)
// gives up with MCP_ERR_TIMEOUT instead of spinning until the watchdog.
__device__ __forceinline__ void mbar_wait_bounded(const Fused2dParams& p, uint64_t* bar, uint32_t parity) {
    if (ptx::mbar_try_wait(bar, parity)) return;
    unsigned spins = 0;
    unsigned long long t0 = 0;
    while (!ptx::mbar_try_wait(bar, parity)) {
        if ((++spins & 0x3fffu) == 0) {
            const unsigned long long now = ptx::globaltimer_ns();
            if (t0 == 0) t0 = now;
            if (now - t0 > p.timeout_ns) {
                atomicCAS(p.status, 0, MCP_ERR_TIMEOUT);
                return;
            }
        }
    }
}

// Test hook (debug_flags & 8): pseudo-random stalls of 5..200 us in front of halo pushes and publishes, so that the
// flag protocol is exercised with ranks and tiles far out of step.
__device__ __noinline__ void injected_delay(const Fused2dParams& p, unsigned salt) {
    unsigned h = (unsigned)p.rank * 0x9e3779b9u ^ salt * 0x85ebca6bu;
    h ^= h >> 15; h *= 0x2c1b3c6du; h ^= h >> 12; h *= 0x297a2d39u; h ^= h >> 15;
    if ((h & 7u) != 0u) return;  // one event in eight is delayed
    const unsigned long long dur = 5000ull + (unsigned long long)((h >> 3) % 195000u);
    const unsigned long long t0 = ptx::globaltimer_ns();
    while (ptx::globaltimer_ns() - t0 < dur) __nanosleep(1000);
}

__device__ __noinline__ void grid_barrier(const Fused2dParams& p, unsigned long long& epoch, int ncta) {
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::fence_proxy_async();  // our generic-proxy stores are read by other CTAs' TMA loads next sweep
        __threadfence();
        const unsigned long long target = p.bar_base + (epoch + 1ull) * (unsigned long long)ncta;
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
            typename Op::Out res = Op::update(L[t - 1][sn], L[t - 1][sc], L[t - 1][ss]);
            if (GENERAL) {
                const int gy = gy_base + j - t;
                const typename Op::Mask m = (gy < ylo || gy > yhi) ? Op::full_mask() : colfix;
                res = Op::sel(m, L[t - 1][sc], res);
            } else if (MODE == 1) {
                res = Op::sel(colfix, L[t - 1][sc], res);
            }
            if (t < TB) {
                L[t < TB ? t : 0][sc] = Op::to_row(res);
            } else {
                bool row_ok = true;
                if (GENERAL) {
                    const int c = cbase + j;
                    row_ok = (c >= 2 * TB) && (c < n_in);
                }
                if (RES) {
                    if (row_ok && lane_valid) Op::acc(res, L[TB - 1][sc], rs, rm);
                }
                st_global_v4_pred(optr, bitcast_u<V>(Op::to_vec(res)), lane_store && row_ok);
            }
        }
        optr += pitch;  // one output row per input row, whether or not it was stored
    }
}

// Tile order within a sweep: row chunks from the outside in (0, nc-1, 1, nc-2, ...), all strips of a chunk
// together.  The chunks on the slab cuts come first so that their halo rows leave early and the transfer overlaps
// the rest of the sweep; a chunk's neighbours in the previous sweep were drawn equally early, so on the barrier-free
// path an early ticket never waits for a late one.
__host__ __device__ __forceinline__ void decode_tile(int q, int ns, int ng, int& group, int& strip) {
    const int gi = q / ns;
    strip = q - gi * ns;
    group = (gi & 1) ? ng - 1 - (gi >> 1) : (gi >> 1);
}

// Inverse of decode_tile.
__host__ __device__ __forceinline__ int encode_tile(int group, int strip, int ns, int ng) {
    const int gi = (group <= ng - 1 - group) ? 2 * group : 2 * (ng - 1 - group) + 1;
    return gi * ns + strip;
}

// Ghost rows this tile reads were written by the neighbour GPU during its previous sweep: wait for its flags.
__device__ __forceinline__ void wait_ghost(const Fused2dParams& p, const int dir, const int sw, const int strip,
                                           const int lane) {
    const unsigned long long need = p.seq_base + (unsigned long long)sw;  // sweeps the neighbour must have done
    if (need == 0) return;
    if (p.strip_flags != nullptr) {
        const unsigned long long* row = p.strip_flags + dir * kMaxStripFlags;
        if (sw == 0 && p.prev_n_strips != p.n_strips) {  // the previous launch cut its strips differently
            for (int s2 = lane; s2 < p.prev_n_strips; s2 += 32) wait_counter_sys(p, &row[s2], need);
        } else if (lane < 3) {
            const int s2 = strip + lane - 1;
            if (s2 >= 0 && s2 < p.n_strips) wait_counter_sys(p, &row[s2], need);
        }
    } else if (lane == 0) {
        wait_counter_sys(p, &p.my_flags[dir],
                         (dir == 0 ? p.flag_base_up : p.flag_base_dn) + (unsigned long long)sw * p.n_strips);
    }
}

// The tile this warp will run next (barrier-free path).  While the current tile consumes its last kStages stages, the
// ring slots they free are refilled with the FIRST stages of the next tile -- if a non-blocking look at its nine
// dependency flags finds them all set.  The next tile then starts with its data in flight (no TMA refill bubble), and
// the publish fence of the current tile overlaps that flight.  Nothing here ever blocks, so the ticket-order argument
// for freedom from deadlock is untouched; a failed probe just falls back to the ordinary start of a tile.
struct NextTile {
    const CUtensorMap* tm_in;
    int valid;   // 0 = no prefetch (last ticket, ghost-row dependencies, grid-barrier path)
    int sw, chunk, strip, x_ld, ys, n_it;
};

template <class Op, int TB, bool RES>
__device__ __forceinline__ void process_tile(const Fused2dParams& p, const CUtensorMap* tm_in,
                                             typename Op::T* __restrict__ out, typename Op::T* up_out,
                                             typename Op::T* dn_out, const int x_ld, const int y0, const int y1,
                                             uint8_t* my_smem, uint64_t* bars, uint32_t& it, const int lane,
                                             double& rs, double& rm, const NextTile& nx, int& pre) {
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
        for (int i = pre; i < npro; ++i) issue(it + i, ys + i * kRowsPerStage);  // `pre` stages are in flight already
    }
    pre = 0;
    // probe of the next tile's dependency flags: loads issued one stage before their values are looked at
    unsigned probe = 0xffffffffu;
    const int i_probe = n_it - kStages - 1;
    const unsigned* probe_ptr = nullptr;
    if (nx.valid && nx.sw > 0 && lane < 9) {
        const int c2 = nx.chunk + lane / 3 - 1, s2 = nx.strip + lane % 3 - 1;
        if (c2 >= 0 && c2 < p.n_chunks && s2 >= 0 && s2 < p.n_strips) probe_ptr = &p.tile_done[c2 * p.n_strips + s2];
    }
    bool next_ok = false;

    // output pointer of the row produced by input-row counter c: local row ys + c - TB
    T* optr = out + (long long)(ys - TB) * p.pitch + cx;

    for (int i = 0; i < n_it; ++i) {
        const uint32_t st = it % kStages, par = (it / kStages) & 1u;
        mbar_wait_bounded(p, &bars[st], par);
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
        if (i + kStages < n_it) {
            if (lane == 0) issue(it + kStages, ys + (i + kStages) * kRowsPerStage);
            if (i == i_probe && probe_ptr != nullptr)
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(probe) : "l"(probe_ptr) : "memory");
        } else if (nx.valid) {
            const int k = i + kStages - n_it;  // stage of the NEXT tile that fits the slot just freed
            if (k == 0) {
                if (i_probe < 0 && probe_ptr != nullptr)  // tile too short for the early probe
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(probe) : "l"(probe_ptr) : "memory");
                next_ok = __all_sync(0xffffffffu, probe >= (unsigned)nx.sw);
                if (next_ok && lane == 0) ptx::fence_proxy_async();
                __syncwarp();
            }
            if (next_ok && k < nx.n_it) {
                if (lane == 0) {
                    const uint32_t st2 = (it + kStages) % kStages;
                    ptx::mbar_arrive_expect_tx(&bars[st2], kStageBytes);
                    ptx::tma_load_2d(my_smem + st2 * kStageBytes, nx.tm_in, &bars[st2], nx.x_ld,
                                     nx.ys + k * kRowsPerStage);
                }
                pre = k + 1;
            }
        }
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

// Fixed-order reduction of the per-tile (sum d^2, max |d|) partials of one sweep by ONE warp, then the cross-rank
// all-reduce over NVLink (P2P inbox or NVLS multimem), the stop decision and the result block.  The order never
// depends on which warp runs it or on timing, so every rank and every run obtains the same bits.
__device__ __noinline__ void reduce_and_exchange(const Fused2dParams& p, const double* partials, const int n_tiles,
                                                 const int sw, const int lane) {
    double a = 0.0, b = 0.0;
    for (int i = lane; i < n_tiles; i += 32) {
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
        if (p.world > 1 && p.nvls_mc != nullptr) {
            // NVLS all-reduce: every rank stores its pair into its OWN copy of the slot, one multimem.red
            // bumps the arrival counter of every GPU through the switch, and multimem.ld_reduce pulls the
            // sum / max of all copies through the switch in one load each.
            const unsigned long long seq = p.res_seq_base + (unsigned long long)(sw - p.res_first_sweep);
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
            const unsigned long long seq = p.res_seq_base + (unsigned long long)(sw - p.res_first_sweep);
            const size_t off = ((size_t)(seq % kResRing) * kMaxRanks + (size_t)p.rank) * 2;
            for (int r = 0; r < p.world; ++r) {
                p.peer_inbox[r][off] = sa;
                p.peer_inbox[r][off + 1] = sb;
            }
            ptx::fence_acq_rel_sys();  // one fence, then relaxed signals (a release pattern for every peer)
            for (int r = 0; r < p.world; ++r) ptx::red_add_relaxed_sys(p.peer_res_flag[r], 1ull);
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
        const int slot = p.res_slot0 + (sw - p.res_first_sweep);
        if (slot < kMaxResSlots) {
            p.res_out[2 * slot] = sa;
            p.res_out[2 * slot + 1] = sb;
        }
        p.result->last_sumsq = sa;
        p.result->last_max = sb;
        if (p.stop_mode != 0 && *((volatile int*)p.stop_word) == 0) {
            const bool conv = p.stop_mode == 1 ? (sb <= p.stop_thresh) : (sa <= p.stop_thresh);
            if (conv) {
                *((volatile int*)p.stop_word) = sw + 2;  // one more sweep runs while this is published
                __threadfence();
            }
        }
    }
    __syncwarp();
}

// Before the kernel ends: every halo row the neighbours push towards this rank during this launch has landed (their
// boundary chunks of the LAST sweep raise the flags this waits for).  After that the host may overwrite the slab
// (next input) or launch again without any host-side barrier between the ranks.
__device__ __noinline__ void drain_incoming(const Fused2dParams& p, const int lane, const int sweeps_run) {
    const unsigned long long need = p.seq_base + (unsigned long long)sweeps_run;
#pragma unroll
    for (int dir = 0; dir < 2; ++dir) {
        if ((dir == 0 ? p.up_buf[0] : p.dn_buf[0]) == nullptr) continue;
        if (p.strip_flags != nullptr) {
            const unsigned long long* row = p.strip_flags + dir * kMaxStripFlags;
            for (int s2 = lane; s2 < p.n_strips; s2 += 32) wait_counter_sys(p, &row[s2], need);
        } else if (lane == 0) {
            wait_counter_sys(p, &p.my_flags[dir],
                             (dir == 0 ? p.flag_base_up : p.flag_base_dn) + (unsigned long long)sweeps_run * p.n_strips);
        }
    }
    __syncwarp();
}

// The kernel body; `cta` / `ncta` are this CTA's index and the CTA count of ITS rank (the whole grid in a normal
// launch; a slice of it when several ranks are emulated inside one launch, see fused2d_group_kernel).
template <class Op, int TB, bool RES, int WARPS>
__device__ __forceinline__ void fused2d_body(const CUtensorMap* tmA, const CUtensorMap* tmB, const Fused2dParams& p,
                                             const int cta, const int ncta) {
    using T = typename Op::T;
    constexpr int VEC = Op::VEC;
    constexpr int M = (TB + VEC * Op::CPE - 1) / (VEC * Op::CPE);
    constexpr int OUTW = (32 - 2 * M) * VEC;  // storage elements written per strip

    extern __shared__ __align__(1024) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* my_smem = smem + warp * (kStages * kStageBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + WARPS * kStages * kStageBytes) + warp * kStages;

    if (lane == 0) {
        for (int s = 0; s < kStages; ++s) ptx::mbar_init(&bars[s], 1);
        ptx::mbar_fence_init();
        ptx::prefetch_tensormap(tmA);
        ptx::prefetch_tensormap(tmB);
    }
    __syncthreads();

    const int NW = ncta * WARPS;
    const int n_tiles = p.n_strips * p.n_chunks;
    uint32_t it = 0;

    const bool dataflow = p.sync_mode == 2;
    // One tile of one sweep.  RES_NOW: this sweep accumulates the residual (a compile-time copy of the row loop).
    int pre = 0;  // stages of the tile about to start that the previous tile has already put in flight
    auto run_tile = [&](const int sw, const int q, double* partials, auto res_now, const NextTile& nx) {
        constexpr bool RES_NOW = decltype(res_now)::value;
        const int par = (p.first_parity + sw) & 1;
        const CUtensorMap* tm_in = par ? tmB : tmA;
        T* out = reinterpret_cast<T*>(p.buf[par ^ 1]);
        int chunk, strip;
        decode_tile(q, p.n_strips, p.n_chunks, chunk, strip);
        const int y0 = chunk_row(p, chunk);
        const int y1 = chunk_row(p, chunk + 1);
        const int x_ld = strip * OUTW - M * VEC;
        const bool top = (chunk == 0), bot = (chunk == p.n_chunks - 1);
        T* up_out = top ? reinterpret_cast<T*>(p.up_buf[par ^ 1]) : nullptr;
        T* dn_out = bot ? reinterpret_cast<T*>(p.dn_buf[par ^ 1]) : nullptr;

        const bool trace = (p.debug_flags & 4) != 0;
        const unsigned long long t_enter = trace ? ptx::globaltimer_ns() : 0ull;
        if (pre > 0) {
            // dependencies were found settled (and fenced) when the previous tile put our first stages in flight
        } else if (dataflow && sw > 0 && lane < 9) {
            // Dataflow instead of a grid barrier: this tile reads what its (up to) 9 neighbours wrote in the
            // previous sweep and overwrites what they read then -- both are settled once they have finished it.
            // One lane per neighbour, so the nine acquire loads are in flight together.
            const int c2 = chunk + lane / 3 - 1, s2 = strip + lane % 3 - 1;
            if (c2 >= 0 && c2 < p.n_chunks && s2 >= 0 && s2 < p.n_strips)
                wait_tile_done(p, &p.tile_done[c2 * p.n_strips + s2], (unsigned)sw);
        }
        if (pre == 0) {
            if (up_out != nullptr) wait_ghost(p, 0, sw, strip, lane);
            if (dn_out != nullptr) wait_ghost(p, 1, sw, strip, lane);
            __syncwarp();
            if (lane == 0) ptx::fence_proxy_async();  // generic-proxy stores (other SMs, peers) -> our TMA loads
            __syncwarp();
            if (aborted(p)) return;
        }

        double rs = 0.0, rm = 0.0;
        const unsigned long long t_begin = trace ? ptx::globaltimer_ns() : 0ull;
        if ((p.debug_flags & 8) && (top || bot)) injected_delay(p, (unsigned)(sw * 131071 + strip * 2 + 0));
        process_tile<Op, TB, RES_NOW>(p, tm_in, out, up_out, dn_out, x_ld, y0, y1, my_smem, bars, it, lane, rs, rm, nx,
                                      pre);
        const unsigned long long t_end = trace ? ptx::globaltimer_ns() : 0ull;

        if (RES_NOW) {
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
        if ((p.debug_flags & 8) && (top || bot)) injected_delay(p, (unsigned)(sw * 131071 + strip * 2 + 1));
        if (lane == 0) {
            const bool cut = up_out != nullptr || dn_out != nullptr;
            if (dataflow) ptx::fence_proxy_async();
            if (cut) {
                // ONE system-scope fence covers the local rows, the rows pushed to both neighbours and every flag
                // below: with NVLink stores outstanding each fence is a round trip to the peer (measured ~3.5 us;
                // four of them per boundary tile made its publish 29 us, profiles/r2/)
                ptx::fence_acq_rel_sys();
                if (dataflow)
                    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(&p.tile_done[chunk * p.n_strips + strip]),
                                 "r"((unsigned)(sw + 1))
                                 : "memory");
                const unsigned long long done_seq = p.seq_base + (unsigned long long)sw + 1ull;
                if (up_out != nullptr) {
                    if (p.up_strip_flag != nullptr) ptx::st_relaxed_sys(&p.up_strip_flag[strip], done_seq);
                    ptx::red_add_relaxed_sys(p.up_flag, 1ull);
                }
                if (dn_out != nullptr) {
                    if (p.dn_strip_flag != nullptr) ptx::st_relaxed_sys(&p.dn_strip_flag[strip], done_seq);
                    ptx::red_add_relaxed_sys(p.dn_flag, 1ull);
                }
            } else if (dataflow) {
                asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&p.tile_done[chunk * p.n_strips + strip]),
                             "r"((unsigned)(sw + 1))
                             : "memory");
            }
            if (trace) {
                p.tile_ns[q] += (unsigned)(t_end - t_begin);
                p.tile_ns[p.trace_cap + q] += (unsigned)(t_begin - t_enter);
                p.tile_ns[2 * p.trace_cap + q] += (unsigned)(ptx::globaltimer_ns() - t_end);
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
                NextTile nx;
                nx.valid = 0;
                if (t_next < total && p.prefetch) {
                    nx.sw = (int)(t_next / (unsigned long long)n_tiles);
                    decode_tile((int)(t_next - (unsigned long long)nx.sw * n_tiles), p.n_strips, p.n_chunks, nx.chunk,
                                nx.strip);
                    // tiles that read ghost rows wait for NVLink flags at their own start: no prefetch for those
                    const bool ghost = (nx.chunk == 0 && p.up_buf[0] != nullptr) ||
                                       (nx.chunk == p.n_chunks - 1 && p.dn_buf[0] != nullptr);
                    if (!ghost) {
                        const int ny0 = chunk_row(p, nx.chunk), ny1 = chunk_row(p, nx.chunk + 1);
                        nx.valid = 1;
                        nx.tm_in = ((p.first_parity + nx.sw) & 1) ? tmB : tmA;
                        nx.x_ld = nx.strip * OUTW - M * VEC;
                        nx.ys = ny0 - TB;
                        nx.n_it = (ny1 - ny0 + 2 * TB + kRowsPerStage - 1) / kRowsPerStage;
                    }
                }
                bool res_now = false;
                if constexpr (RES) res_now = sw >= p.res_first_sweep;
                if constexpr (RES) if (res_now) {
                    // only the LAST sweep may carry the residual on this path (host-enforced): the last tile to
                    // finish it reduces the partials and runs the cross-rank exchange
                    run_tile(sw, q, p.partials, std::true_type{}, nx);
                    unsigned last = 0;
                    __syncwarp();
                    if (lane == 0) {
                        __threadfence();
                        last = (atomicAdd(p.res_count, 1u) + 1u == (unsigned)n_tiles) ? 1u : 0u;
                    }
                    last = __shfl_sync(0xffffffffu, last, 0);
                    if (last) {
                        __threadfence();
                        reduce_and_exchange(p, p.partials, n_tiles, sw, lane);
                        if (lane == 0) {
                            *p.res_count = 0u;
                            p.result->sweeps_done = p.sweeps;
                        }
                    }
                }
                if (!res_now) run_tile(sw, q, nullptr, std::false_type{}, nx);
            }
            t = t_next;
        }
        if (cta == 0 && warp == 0 && !aborted(p)) drain_incoming(p, lane, p.sweeps);
        return;
    }

    // ---- barrier path: sweeps separated by a device-side grid barrier (used when every sweep carries the residual:
    // history and run-to-tolerance, where the stop decision of sweep k must reach every CTA before sweep k+2)
    unsigned long long epoch = 0;
    int executed = 0;
    NextTile no_next;
    no_next.valid = 0;
    for (int sw = 0; sw < p.sweeps; ++sw) {
        const bool res_now = RES && sw >= p.res_first_sweep;
        if (RES && p.stop_mode != 0 && sw > 0) {
            // decided from the global residual of sweep sw-2 and published before barrier sw-1
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
            if (!aborted(p)) {
                if constexpr (RES) {
                    if (res_now)
                        run_tile(sw, q, partials, std::true_type{}, no_next);
                    else
                        run_tile(sw, q, partials, std::false_type{}, no_next);
                } else {
                    run_tile(sw, q, partials, std::false_type{}, no_next);
                }
            }
            q = q_next;
        }

        if (res_now || sw + 1 < p.sweeps) grid_barrier(p, epoch, ncta);
        if (aborted(p)) break;

        if (RES && res_now && cta == 0) {
            if (warp == 0) reduce_and_exchange(p, partials, n_tiles, sw, lane);
            __syncthreads();
        }
        executed = sw + 1;
    }
    if (cta == 0 && warp == 0) {
        if (lane == 0) p.result->sweeps_done = executed;
        if (!aborted(p)) drain_incoming(p, lane, executed);
    }
}

template <class Op, int TB, bool RES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
fused2d_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ Fused2dParams p) {
    fused2d_body<Op, TB, RES, WARPS>(&tmA, &tmB, p, (int)blockIdx.x, (int)gridDim.x);
}

// Several ranks inside ONE cooperative launch on ONE GPU: CTA b works for rank b % n_ranks with the parameters and
// tensor maps of that rank read from global memory.  The ranks' kernels wait on each other's flags exactly as they do
// across GPUs (only the pointer values differ), which separate launches sharing one GPU must never do.  Used by the
// tests to exercise the whole cross-rank protocol (strip flags, residual exchange, drain, timeouts, injected delays)
// on a single-GPU box.
struct GroupRankArgs {
    alignas(64) CUtensorMap tm[2];
    Fused2dParams p;
};

template <class Op, int TB, bool RES, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) fused2d_group_kernel(const GroupRankArgs* args, const int n_ranks) {
    const int r = (int)blockIdx.x % n_ranks;
    fused2d_body<Op, TB, RES, WARPS>(&args[r].tm[0], &args[r].tm[1], args[r].p, (int)blockIdx.x / n_ranks,
                                     (int)gridDim.x / n_ranks);
}

// ------------------------------------------------------------------------------------------ host plan

}  // namespace mcp

using namespace mcp;

namespace {

constexpr int kMaxChunks = 4096;

// Tile schedule of one (time block, sweep-synchronisation mode) pair.
struct Geometry {
    std::vector<int> chunk_y;  // n_chunks + 1 local row boundaries
    int n_strips = 0, chunk_r = 1;
    int n_chunks() const { return (int)chunk_y.size() - 1; }
};

}  // namespace

struct mcp_fused2d_plan {
    mcp_fused2d_desc d;
    int rows_alloc;
    int sm_count;
    int grid;
    int parity;
    CUtensorMap tm[2];
    unsigned long long* grid_bar;   // device: [0] grid barrier, [16] tile ticket counter, [32] residual tile counter
    ResultBlock* result;            // device
    double* partials;
    size_t partials_cap_tiles;
    double* res_out;
    unsigned* tile_done;            // device: [partials_cap_tiles]
    unsigned* tile_ns;              // device: [partials_cap_tiles] per-tile trace
    GroupRankArgs* group_args;      // device: parameters of the ranks of an emulated multi-rank launch
    int last_tiles;                 // tile count of the last launch
    unsigned long long res_seq;     // sweeps whose residual has been exchanged so far
    unsigned long long sweep_seq;   // sweeps completed so far (identical on every rank)
    int last_ns;                    // strip count of the previous launch
    int last_generations;           // generations executed by the last run call
    int last_stop_reason;           // run-to-tolerance: MCP_STOP_*
    Geometry geo[4][2];             // [time block 1,2,4,8][0 = grid barrier, 1 = dataflow]
    unsigned long long flag_base_up, flag_base_dn, bar_base, ctr_base;
    int res_slots_used;
    int64_t launches;
    unsigned long long timeout_ns;
    int debug_flags;
    int sync_mode;  // 0 = choose per launch, 1 = barrier, 2 = dataflow
    int prefetch;
};

namespace {

int vec_of(int dtype) { return dtype == MCP_DTYPE_F64 ? 2 : 4; }

int op_kind(const mcp_fused2d_desc& d) {
    if (d.dtype == MCP_DTYPE_B1) return 2;
    if (d.op == MCP_OP_LIFE) return 1;
    return d.dtype == MCP_DTYPE_F64 ? 3 : 0;
}

int cpe_of(int dtype) { return dtype == MCP_DTYPE_B1 ? 32 : 1; }  // cells per storage element

// output columns (cells) per strip
int outw_rt(int dtype, int tb) {
    const int vec = vec_of(dtype), cpe = cpe_of(dtype);
    return (32 - 2 * ((tb + vec * cpe - 1) / (vec * cpe))) * vec * cpe;
}

int max_tb(int op, int dtype) {
    (void)dtype;
    if (op == MCP_OP_LIFE) return 4;
    return 8;
}

int tb_index(int tb) { return tb == 1 ? 0 : tb == 2 ? 1 : tb == 4 ? 2 : 3; }

// Chunks of `r` rows; the last one takes the remainder when that is at least max(G, r/2) rows, else the remainder is
// folded into the chunk before it (so the last chunk is never a long straggler); every chunk holds >= G rows.
// r + 2*TB is made a whole number of TMA stages, so a tile neither loads nor pushes rows it does not need.
void fill_geometry(int rows, int own_lo, int G, int tb, int r, Geometry* g) {
    r = std::max(std::max(r, G), 1);
    if (r < rows) {
        r -= (r + 2 * tb) % kRowsPerStage;
        while (r < G) r += kRowsPerStage;
    }
    int nc = std::max(1, (rows + r - 1) / r);
    if (nc > 1 && rows - (nc - 1) * r < std::max(G, r / 2)) nc -= 1;
    if (nc > kMaxChunks) {
        nc = kMaxChunks;
        r = rows / nc;
    }
    g->chunk_r = r;
    g->chunk_y.clear();
    for (int c = 0; c <= nc; ++c) g->chunk_y.push_back(chunk_row_of(own_lo, rows, nc, r, c));
}

// Tile height.  Two overheads pull in opposite directions: every tile re-reads and partly recomputes 2*TB pre-roll rows
// and refills its TMA pipeline (~(TB + 6) row-times per tile), and a sweep ends with about half a tile of idle time
// per warp.  With `share` rows of work per warp, overhead(r) ~ (TB + 6)/r + r/(2*share) is minimal at
// r* = sqrt(2*share*(TB + 6)).
int formula_rows(const mcp_fused2d_desc& d, int grid, int tb, int ns) {
    const int NW = grid * warps_for(op_kind(d), tb);
    const double share = (double)d.rows_own * ns / NW;
    return (int)std::lround(std::sqrt(2.0 * share * (tb + 6)));
}

// Schedule for the grid-barrier path: one chunk = one tile.
void build_barrier_geometry(const mcp_fused2d_desc& d, int grid, int tb, Geometry* g) {
    const int ns = (int)ceil_div(d.w, outw_rt(d.dtype, tb));
    const int G = std::max(d.ghost, 1);
    g->n_strips = ns;
    int r = d.chunk_rows;
    if (r <= 0) r = std::max(std::max(G, 4 * tb), std::min(64 * tb, formula_rows(d, grid, tb, ns)));
    fill_geometry(d.rows_own, d.ghost, G, tb, r, g);
}

// Schedule for the barrier-free path: a tile waits for its nine neighbours of the previous sweep and publishes itself.
void build_dataflow_geometry(const mcp_fused2d_desc& d, int grid, int tb, Geometry* g) {
    const int ns = (int)ceil_div(d.w, outw_rt(d.dtype, tb));
    const int G = std::max(d.ghost, 1), rows = d.rows_own;
    g->n_strips = ns;
    int r = d.chunk_rows;
    // measured on B200 (profiles/): never below 16 time blocks -- a tile pays ~2 us of TMA refill and a publish on top
    // of the pre-roll, which short tiles cannot amortise
    if (r <= 0) {
        r = std::max(16 * tb, std::min(64 * tb, formula_rows(d, grid, tb, ns)));
        // at least two tiles per warp and sweep: a warp that draws a tile of the NEXT sweep then finds the tiles it
        // depends on (drawn a tile-time earlier) finished instead of idling for them (measured cliff: 1.6 tiles per
        // warp run at 0.65 of the rate of 2.0, profiles/)
        const int NW = grid * warps_for(op_kind(d), tb);
        const int cap = (int)((long long)rows * ns / (2LL * NW));
        r = std::max(std::max(G, 4 * tb), std::min(r, cap));
    }
    fill_geometry(rows, d.ghost, G, tb, r, g);
}

void build_geometry(const mcp_fused2d_desc& d, int grid, int tb, int mode, Geometry* g) {
    if (mode == 0)
        build_barrier_geometry(d, grid, tb, g);
    else
        build_dataflow_geometry(d, grid, tb, g);
}

template <class Op, int TB>
const void* kernel_ptr(bool res, bool group) {
    constexpr int W = warps_for(std::is_same<Op, LifeB1>::value ? 2 : std::is_same<Op, LifeI32>::value ? 1 : std::is_same<Op, MdfF64>::value ? 3 : 0, TB);
    if (group) {
        if constexpr (std::is_same<Op, MdfF32>::value && (TB == 4 || TB == 8)) {
            return res ? (const void*)fused2d_group_kernel<Op, TB, true, W> : (const void*)fused2d_group_kernel<Op, TB, false, W>;
        } else {
            return nullptr;
        }
    }
    return res ? (const void*)fused2d_kernel<Op, TB, true, W> : (const void*)fused2d_kernel<Op, TB, false, W>;
}

template <class Op, int TB>
constexpr int kernel_warps() {
    return warps_for(std::is_same<Op, LifeB1>::value ? 2 : std::is_same<Op, LifeI32>::value ? 1 : std::is_same<Op, MdfF64>::value ? 3 : 0, TB);
}

template <class Op>
int pick_tb(int tb, bool res, bool group, const void** fn, int* warps) {
    switch (tb) {
        case 1: *fn = kernel_ptr<Op, 1>(res, group); *warps = kernel_warps<Op, 1>(); return MCP_OK;
        case 2: *fn = kernel_ptr<Op, 2>(res, group); *warps = kernel_warps<Op, 2>(); return MCP_OK;
        case 4: *fn = kernel_ptr<Op, 4>(res, group); *warps = kernel_warps<Op, 4>(); return MCP_OK;
        default: break;
    }
    set_error("unsupported time block %d", tb);
    return MCP_ERR_UNSUPPORTED;
}

int pick_kernel(const mcp_fused2d_desc& d, int tb, bool res, bool group, const void** fn, int* warps,
                bool scalar = false) {
    if (scalar && !group && d.op == MCP_OP_MDF && d.dtype == MCP_DTYPE_F32 && tb == 8) {
        *fn = kernel_ptr<MdfF32Scalar, 8>(res, false);
        *warps = kernel_warps<MdfF32Scalar, 8>();
        return MCP_OK;
    }
    if (d.op == MCP_OP_MDF && d.dtype == MCP_DTYPE_F32) {
        if (tb == 8) { *fn = kernel_ptr<MdfF32, 8>(res, group); *warps = kernel_warps<MdfF32, 8>(); return MCP_OK; }
        return pick_tb<MdfF32>(tb, res, group, fn, warps);
    }
    if (d.op == MCP_OP_MDF && d.dtype == MCP_DTYPE_F64) {
        if (tb == 8) { *fn = kernel_ptr<MdfF64, 8>(res, group); *warps = kernel_warps<MdfF64, 8>(); return MCP_OK; }
        return pick_tb<MdfF64>(tb, res, group, fn, warps);
    }
    if (d.op == MCP_OP_LIFE && d.dtype == MCP_DTYPE_I32) return pick_tb<LifeI32>(tb, res, group, fn, warps);
    if (d.op == MCP_OP_LIFE && d.dtype == MCP_DTYPE_B1) return pick_tb<LifeB1>(tb, res, group, fn, warps);
    set_error("unsupported op/dtype combination %d/%d", d.op, d.dtype);
    return MCP_ERR_UNSUPPORTED;
}

// What one launch does, decided on the host before the kernel parameters are filled in.
struct LaunchSpec {
    int tb, sweeps, mode;  // mode: 0 = grid barrier, 1 = dataflow
    bool res;              // residual instantiation of the kernel
    bool scalar = false;   // fp32 MDF, TB = 8: scalar operator (see prefers_scalar_barrier)
    int res_first_sweep;
    int stop_mode;
    double stop_thresh;
};

int fill_params(mcp_fused2d_plan* pl, const LaunchSpec& ls, Fused2dParams* out) {
    const mcp_fused2d_desc& d = pl->d;
    const Geometry& g = pl->geo[tb_index(ls.tb)][ls.mode];
    const int ns = g.n_strips;
    MCP_REQUIRE((size_t)ns * g.n_chunks() <= pl->partials_cap_tiles, "tile count %d x %d exceeds scratch capacity", ns,
                g.n_chunks());
    const bool exchange = d.world > 1 && (d.res_inbox != nullptr || d.nvls_mc != nullptr);
    Fused2dParams& p = *out;
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
    p.tile_ctr = pl->grid_bar + 16;  // separate 128-byte lines of the same zeroed allocation
    p.res_count = reinterpret_cast<unsigned*>(pl->grid_bar + 32);
    p.ctr_base = pl->ctr_base;
    p.tile_done = pl->tile_done;
    p.tile_ns = pl->tile_ns;
    p.trace_cap = (int)pl->partials_cap_tiles;
    pl->last_tiles = ns * g.n_chunks();
    p.result = pl->result;
    p.status = &pl->result->status;
    p.stop_word = &pl->result->stop_word;
    p.partials = pl->partials;
    p.res_out = pl->res_out;
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
    p.n_chunks = g.n_chunks();
    p.chunk_r = g.chunk_r;
    p.sweeps = ls.sweeps;
    p.first_parity = pl->parity;
    p.res_slot0 = pl->res_slots_used;
    p.res_first_sweep = ls.res ? ls.res_first_sweep : ls.sweeps;
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
    p.stop_mode = ls.stop_mode;
    p.stop_thresh = ls.stop_thresh;
    p.sync_mode = ls.mode == 0 ? 1 : 2;
    p.prefetch = pl->prefetch;
    return MCP_OK;
}

// Host bookkeeping after a launch that executed `executed` sweeps (all counters are monotonic across launches).
void account(mcp_fused2d_plan* pl, const LaunchSpec& ls, int executed, int warps) {
    const mcp_fused2d_desc& d = pl->d;
    const Geometry& g = pl->geo[tb_index(ls.tb)][ls.mode];
    const unsigned long long tiles = (unsigned long long)g.n_strips * g.n_chunks();
    const unsigned long long nw = (unsigned long long)pl->grid * warps;
    const int res_sweeps = ls.res ? std::max(0, executed - ls.res_first_sweep) : 0;
    if (ls.mode == 0) {
        // a barrier after every sweep but the last, and after the last one too when it carries the residual
        unsigned long long barriers = (unsigned long long)std::max(0, executed - 1);
        if (executed > 0 && ls.res && executed - 1 >= ls.res_first_sweep) barriers += 1;
        pl->bar_base += barriers * (unsigned long long)pl->grid;
        pl->ctr_base += (unsigned long long)executed * (tiles + nw);
    } else {
        pl->ctr_base += (unsigned long long)executed * tiles + nw;
    }
    pl->flag_base_up += (unsigned long long)executed * g.n_strips;
    pl->flag_base_dn += (unsigned long long)executed * g.n_strips;
    pl->parity = (pl->parity + executed) & 1;
    pl->sweep_seq += (unsigned long long)executed;
    pl->last_ns = g.n_strips;
    if (ls.res) {
        pl->res_slots_used = std::min(kMaxResSlots, pl->res_slots_used + res_sweeps);
        const bool exchange = d.world > 1 && (d.res_inbox != nullptr || d.nvls_mc != nullptr);
        if (exchange) pl->res_seq += (unsigned long long)res_sweeps;
    }
    pl->last_generations += executed * ls.tb;
}

// fp32 MDF at TB = 8 on a slab with many tiles per warp: the scalar operator behind the grid barrier is the fastest
// combination measured (16384^2 on one GPU, same box: scalar + barrier 3282, packed + barrier 3105, packed + dataflow
// 3081, scalar + dataflow 3034 GLUPS); on small slabs packed + dataflow wins by far (2048 rows: 2740 vs 1880 / 2130).
bool prefers_scalar_barrier(const mcp_fused2d_plan* pl, int tb) {
    const mcp_fused2d_desc& d = pl->d;
    if (pl->sync_mode != 0 || d.op != MCP_OP_MDF || d.dtype != MCP_DTYPE_F32 || tb != 8) return false;
    if (d.up_buf[0] != nullptr || d.dn_buf[0] != nullptr) return false;  // measured on single-GPU slabs only
    const Geometry& g = pl->geo[tb_index(tb)][0];
    const double tiles_per_warp = (double)g.n_strips * g.n_chunks() / ((double)pl->grid * warps_for(0, tb));
    return tiles_per_warp >= 5.0;
}

int choose_mode(const mcp_fused2d_plan* pl, int tb, bool every_sweep_residual) {
    if (every_sweep_residual) return 0;  // CTA 0 reduces a finished sweep behind the grid barrier
    if (pl->sync_mode == 1) return 0;
    if (prefers_scalar_barrier(pl, tb)) return 0;
    // One arrival COUNTER per neighbour cannot tell sweeps apart once the neighbour runs barrier-free: its early
    // strips of sweep k+1 would be counted for a late strip of sweep k (found by the injected-delay stress test).
    // Without per-strip flags the ranks therefore keep the grid barrier, which orders a rank's sweeps.
    const bool neighbours = pl->d.up_buf[0] != nullptr || pl->d.dn_buf[0] != nullptr;
    if (neighbours && pl->d.strip_flags == nullptr) return 0;
    return 1;
}

// The time block a run with the requested value executes first: the largest power of two within the request, the
// operator's limit and the ghost depth (the rule of run_impl), so that the introspection calls describe that schedule.
int effective_tb(const mcp_fused2d_plan* pl, int time_block) {
    const int tbmax = std::min(max_tb(pl->d.op, pl->d.dtype), pl->d.ghost);
    int tb = 1;
    while (tb * 2 <= time_block && tb * 2 <= tbmax) tb *= 2;
    return tb;
}

}  // namespace

extern "C" {

int mcp_fused2d_destroy(mcp_fused2d_plan* pl);

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
        // strips are narrowest (most numerous) at the largest time block this plan can run
        MCP_REQUIRE(ceil_div(d.w, outw_rt(d.dtype, std::min(max_tb(d.op, d.dtype), d.ghost))) <= kMaxStripFlags,
                    "grid too wide for per-strip halo flags");
    }
    MCP_REQUIRE(d.chunk_rows >= 0, "negative chunk_rows");

    mcp_fused2d_plan* pl = new (std::nothrow) mcp_fused2d_plan();
    MCP_REQUIRE(pl, "out of host memory");
    pl->d = d;
    pl->rows_alloc = d.rows_own + 2 * d.ghost;
    pl->parity = 0;
    pl->flag_base_up = pl->flag_base_dn = pl->bar_base = pl->ctr_base = 0;
    pl->res_slots_used = 0;
    pl->launches = 0;
    pl->timeout_ns = 20ull * 1000000000ull;
    pl->debug_flags = 0;
    pl->sync_mode = 0;
    pl->prefetch = 1;
    pl->grid_bar = nullptr;
    pl->result = nullptr;
    pl->partials = nullptr;
    pl->res_out = nullptr;
    pl->tile_done = nullptr;
    pl->tile_ns = nullptr;
    pl->group_args = nullptr;
    pl->last_tiles = 0;
    pl->res_seq = 0;
    pl->sweep_seq = 0;
    pl->last_ns = 0;
    pl->last_generations = 0;
    pl->last_stop_reason = MCP_STOP_NONE;

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
    // scratch: counters, result block, residual partials + history, per-chunk flags, schedule tables
    const int min_outw = outw_rt(d.dtype, max_tb(d.op, d.dtype));
    const size_t max_strips = (size_t)ceil_div(d.w, min_outw);
    const size_t max_chunks = (size_t)std::max(1, d.rows_own / std::max(d.ghost, 1));
    pl->partials_cap_tiles = max_strips * std::min<size_t>(max_chunks, kMaxChunks);
    cudaError_t e;
    if ((e = cudaMalloc(&pl->grid_bar, 512)) != cudaSuccess || (e = cudaMemset(pl->grid_bar, 0, 512)) != cudaSuccess ||
        (e = cudaMalloc(&pl->result, 256)) != cudaSuccess || (e = cudaMemset(pl->result, 0, 256)) != cudaSuccess ||
        (e = cudaMalloc(&pl->partials, pl->partials_cap_tiles * 4 * sizeof(double))) != cudaSuccess ||
        (e = cudaMalloc(&pl->res_out, kMaxResSlots * 2 * sizeof(double))) != cudaSuccess ||
        (e = cudaMemset(pl->res_out, 0, kMaxResSlots * 2 * sizeof(double))) != cudaSuccess ||
        (e = cudaMalloc(&pl->tile_ns, 3 * pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess ||
        (e = cudaMemset(pl->tile_ns, 0, 3 * pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess ||
        (e = cudaMalloc(&pl->tile_done, pl->partials_cap_tiles * sizeof(unsigned))) != cudaSuccess) {
        set_error("mcp_fused2d_create: scratch allocation failed: %s", cudaGetErrorString(e));
        mcp_fused2d_destroy(pl);
        return MCP_ERR_CUDA;
    }
    for (int tb = 1; tb <= 8; tb *= 2)
        for (int mode = 0; mode < 2; ++mode) build_geometry(d, pl->grid, tb, mode, &pl->geo[tb_index(tb)][mode]);
    *plan_out = pl;
    return MCP_OK;
}

int mcp_fused2d_set_option(mcp_fused2d_plan* pl, int option, long long value) {
    MCP_REQUIRE(pl, "null plan");
    switch (option) {
        case MCP_OPT_TIMEOUT_MS:
            MCP_REQUIRE(value >= 1, "timeout must be >= 1 ms");
            pl->timeout_ns = (unsigned long long)value * 1000000ull;
            return MCP_OK;
        case MCP_OPT_DEBUG_FLAGS:
            pl->debug_flags = (int)value;
            return MCP_OK;
        case MCP_OPT_SYNC_MODE:
            MCP_REQUIRE(value >= 0 && value <= 2, "sync mode must be 0 (choose), 1 (grid barrier) or 2 (dataflow)");
            pl->sync_mode = (int)value;
            return MCP_OK;
        case MCP_OPT_PREFETCH:
            pl->prefetch = value != 0;
            return MCP_OK;
        default: break;
    }
    set_error("unknown option %d", option);
    return MCP_ERR_ARG;
}

int mcp_fused2d_geometry(const mcp_fused2d_plan* pl, int time_block, int* n_strips, int* n_chunks, int* grid,
                         int* smem_bytes) {
    MCP_REQUIRE(pl, "null plan");
    time_block = effective_tb(pl, time_block);
    const Geometry& g = pl->geo[tb_index(time_block)][choose_mode(pl, time_block, false)];
    if (n_strips) *n_strips = g.n_strips;
    if (n_chunks) *n_chunks = g.n_chunks();
    if (grid) *grid = pl->grid;
    if (smem_bytes) *smem_bytes = smem_bytes_for(warps_for(op_kind(pl->d), time_block));
    return MCP_OK;
}

// Heights of the row chunks (= tiles) of the default schedule, in slab order.
int mcp_fused2d_chunks(const mcp_fused2d_plan* pl, int time_block, int* rows_out, int max_entries) {
    MCP_REQUIRE(pl && rows_out, "null argument");
    time_block = effective_tb(pl, time_block);
    const Geometry& g = pl->geo[tb_index(time_block)][choose_mode(pl, time_block, false)];
    const int n = std::min(g.n_chunks(), max_entries);
    for (int i = 0; i < n; ++i) rows_out[i] = g.chunk_y[i + 1] - g.chunk_y[i];
    return n;
}

static int launch(mcp_fused2d_plan* pl, const LaunchSpec& ls, cudaStream_t stream, int* warps_out) {
    Fused2dParams p;
    int rc = fill_params(pl, ls, &p);
    if (rc) return rc;
    const void* fn = nullptr;
    int warps = 0;
    rc = pick_kernel(pl->d, ls.tb, ls.res, false, &fn, &warps, ls.scalar);
    if (rc) return rc;
    if (ls.stop_mode != 0) MCP_CUDA_TRY(cudaMemsetAsync(&pl->result->stop_word, 0, 2 * sizeof(int), stream));
    if (ls.mode == 1)
        MCP_CUDA_TRY(cudaMemsetAsync(pl->tile_done, 0, (size_t)p.n_strips * p.n_chunks * sizeof(unsigned), stream));
    if (pl->debug_flags & 4)
        MCP_CUDA_TRY(cudaMemsetAsync(pl->tile_ns, 0, 3 * pl->partials_cap_tiles * sizeof(unsigned), stream));
    const int smem = smem_bytes_for(warps);
    void* args[3] = {(void*)&pl->tm[0], (void*)&pl->tm[1], (void*)&p};
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCP_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(pl->grid), dim3(warps * 32), args, smem, stream));
    pl->launches++;
    *warps_out = warps;
    return MCP_OK;
}

static int run_impl(mcp_fused2d_plan* pl, int generations, int time_block, int residual_mode, int stop_mode,
                    double stop_thresh, cudaStream_t stream) {
    int tb = effective_tb(pl, time_block);
    if (residual_mode != MCP_RES_NONE) pl->res_slots_used = 0;
    pl->last_generations = 0;
    pl->last_stop_reason = MCP_STOP_NONE;

    int remaining = generations;
    while (remaining > 0) {
        while (tb > remaining) tb >>= 1;
        LaunchSpec ls;
        ls.tb = tb;
        ls.sweeps = remaining / tb;
        const bool last_launch = remaining - ls.sweeps * tb == 0;
        // MCP_RES_LAST: every sweep carries the residual; MCP_RES_FINAL: only the very last sweep of the run
        ls.res = residual_mode == MCP_RES_LAST || (residual_mode == MCP_RES_FINAL && last_launch);
        ls.res_first_sweep = residual_mode == MCP_RES_FINAL ? ls.sweeps - 1 : 0;
        if (stop_mode != 0) ls.sweeps = std::min(ls.sweeps, kMaxResSlots);
        ls.stop_mode = stop_mode;
        ls.stop_thresh = stop_thresh;
        ls.mode = choose_mode(pl, tb, ls.res && ls.res_first_sweep < ls.sweeps - 1);
        if (stop_mode != 0) ls.mode = 0;
        ls.scalar = ls.mode == 0 && prefers_scalar_barrier(pl, tb);

        int warps = 0;
        int rc = launch(pl, ls, stream, &warps);
        if (rc) return rc;

        int executed = ls.sweeps;
        if (stop_mode != 0) {
            // the number of sweeps that ran is decided on the device: wait for it before touching the counters
            MCP_CUDA_TRY(cudaStreamSynchronize(stream));
            ResultBlock rb;
            MCP_CUDA_TRY(cudaMemcpy(&rb, pl->result, sizeof(rb), cudaMemcpyDeviceToHost));
            executed = rb.sweeps_done;
            MCP_REQUIRE(executed >= 0 && executed <= ls.sweeps, "device reported %d executed sweeps of %d", executed,
                        ls.sweeps);
            if (rb.stop_word != 0)
                pl->last_stop_reason = MCP_STOP_CONVERGED;
            else
                pl->last_stop_reason = ls.sweeps < remaining / tb ? MCP_STOP_SLOT_LIMIT : MCP_STOP_MAX_GENERATIONS;
        }
        account(pl, ls, executed, warps);
        if (stop_mode != 0) break;  // tolerance runs use one block size; the caller decides how to continue
        remaining -= ls.sweeps * tb;
    }
    return MCP_OK;
}

int mcp_fused2d_run(mcp_fused2d_plan* pl, int generations, int time_block, int residual_mode, void* stream_) {
    MCP_REQUIRE(pl, "null plan");
    MCP_REQUIRE(generations >= 0, "negative generation count");
    MCP_REQUIRE(time_block >= 1, "time_block must be >= 1");
    MCP_REQUIRE(residual_mode == MCP_RES_NONE || residual_mode == MCP_RES_LAST || residual_mode == MCP_RES_FINAL,
                "residual_mode must be MCP_RES_NONE, MCP_RES_LAST or MCP_RES_FINAL");
    return run_impl(pl, generations, time_block, residual_mode, 0, 0.0, (cudaStream_t)stream_);
}

int mcp_fused2d_run_to_tolerance(mcp_fused2d_plan* pl, int max_generations, int time_block, int norm, double tol,
                                 int* generations_done, void* stream_) {
    MCP_REQUIRE(pl, "null plan");
    MCP_REQUIRE(max_generations >= 1 && time_block >= 1, "bad generation count or time block");
    MCP_REQUIRE(norm == MCP_NORM_LINF || norm == MCP_NORM_L2, "norm must be MCP_NORM_LINF or MCP_NORM_L2");
    MCP_REQUIRE(tol >= 0.0, "negative tolerance");
    const mcp_fused2d_desc& d = pl->d;
    // Every rank must take the same stop decision in the same sweep, or the ranks that go on spin on halo flags
    // that never come.  That needs the in-kernel all-reduce, and a reduction whose bits agree on every rank.
    const bool neighbours = d.up_buf[0] != nullptr || d.dn_buf[0] != nullptr;
    MCP_REQUIRE(!neighbours || d.res_inbox != nullptr || d.nvls_mc != nullptr,
                "run_to_tolerance across ranks needs the fused residual exchange (res_inbox or nvls_mc): a local "
                "residual would stop the ranks in different sweeps");
    MCP_REQUIRE(!(neighbours && d.nvls_mc != nullptr && norm == MCP_NORM_L2),
                "the NVLS sum is associated by the switch and may round differently per rank: use MCP_NORM_LINF with "
                "nvls_mc, or the peer-inbox exchange for MCP_NORM_L2");
    const double thresh = norm == MCP_NORM_L2 ? tol * tol : tol;  // the kernel compares the sum of squares
    int rc = run_impl(pl, max_generations, time_block, MCP_RES_LAST, norm == MCP_NORM_LINF ? 1 : 2, thresh,
                      (cudaStream_t)stream_);
    if (generations_done) *generations_done = pl->last_generations;
    return rc;
}

int mcp_fused2d_stop_reason(const mcp_fused2d_plan* pl) { return pl ? pl->last_stop_reason : MCP_STOP_NONE; }

// Emulated multi-rank launch: the plans of n ranks that live on THIS GPU advance together inside one cooperative
// kernel (see fused2d_group_kernel).  One launch, fixed generation count, generations % time_block == 0.
int mcp_fused2d_run_group(mcp_fused2d_plan** plans, int n, int generations, int time_block, int residual_mode,
                          void* stream_) {
    MCP_REQUIRE(plans && n >= 1 && n <= kMaxRanks, "bad plan list");
    cudaStream_t stream = (cudaStream_t)stream_;
    mcp_fused2d_plan* p0 = plans[0];
    MCP_REQUIRE(p0, "null plan");
    const int tb = effective_tb(p0, time_block);
    MCP_REQUIRE(generations >= tb && generations % tb == 0, "generations must be a positive multiple of the time block");
    MCP_REQUIRE(residual_mode == MCP_RES_NONE || residual_mode == MCP_RES_LAST || residual_mode == MCP_RES_FINAL,
                "bad residual mode");
    const int ncta = p0->sm_count / n;
    MCP_REQUIRE(ncta >= 1, "more ranks than SMs");
    LaunchSpec ls;
    ls.tb = tb;
    ls.sweeps = generations / tb;
    ls.res = residual_mode != MCP_RES_NONE;
    ls.res_first_sweep = residual_mode == MCP_RES_FINAL ? ls.sweeps - 1 : 0;
    ls.stop_mode = 0;
    ls.stop_thresh = 0.0;
    // same rules as choose_mode, minus the scalar-operator preference (the group kernel is built for the packed one)
    const bool every = ls.res && ls.res_first_sweep < ls.sweeps - 1;
    const bool coarse = p0->d.strip_flags == nullptr && (p0->d.up_buf[0] != nullptr || p0->d.dn_buf[0] != nullptr);
    ls.mode = (every || coarse || p0->sync_mode == 1) ? 0 : 1;
    ls.scalar = false;
    const void* fn = nullptr;
    int warps = 0;
    int rc = pick_kernel(p0->d, tb, ls.res, true, &fn, &warps);
    if (rc) return rc;
    MCP_REQUIRE(fn != nullptr, "the emulated multi-rank launch is built for fp32 MDF with time block 4 or 8 only");
    std::vector<GroupRankArgs> host((size_t)n);
    for (int r = 0; r < n; ++r) {
        mcp_fused2d_plan* pl = plans[r];
        MCP_REQUIRE(pl && pl->d.op == p0->d.op && pl->d.dtype == p0->d.dtype && pl->d.ghost == p0->d.ghost,
                    "plans of one group must share operator, dtype and ghost depth");
        pl->grid = ncta;  // tile schedules and ticket accounting of this launch use the rank's share of the GPU
        if (residual_mode != MCP_RES_NONE) pl->res_slots_used = 0;
        pl->last_generations = 0;
        rc = fill_params(pl, ls, &host[r].p);
        if (rc) return rc;
        host[r].tm[0] = pl->tm[0];
        host[r].tm[1] = pl->tm[1];
        if (ls.mode == 1)
            MCP_CUDA_TRY(cudaMemsetAsync(pl->tile_done, 0,
                                         (size_t)host[r].p.n_strips * host[r].p.n_chunks * sizeof(unsigned), stream));
    }
    if (p0->group_args == nullptr) MCP_CUDA_TRY(cudaMalloc(&p0->group_args, sizeof(GroupRankArgs) * kMaxRanks));
    MCP_CUDA_TRY(cudaMemcpyAsync(p0->group_args, host.data(), sizeof(GroupRankArgs) * n, cudaMemcpyHostToDevice, stream));
    MCP_CUDA_TRY(cudaStreamSynchronize(stream));  // `host` goes out of scope; the copy source must stay valid
    const int smem = smem_bytes_for(warps);
    const GroupRankArgs* dargs = p0->group_args;
    int n_arg = n;
    void* args[2] = {(void*)&dargs, (void*)&n_arg};
    MCP_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    MCP_CUDA_TRY(cudaLaunchCooperativeKernel(fn, dim3(ncta * n), dim3(warps * 32), args, smem, stream));
    for (int r = 0; r < n; ++r) {
        plans[r]->launches++;
        account(plans[r], ls, ls.sweeps, warps);
    }
    return MCP_OK;
}

// Queue position <-> (group, strip), the same functions the kernel uses (host build of decode_tile/encode_tile).
int mcp_fused2d_tile_at(int n_strips, int n_chunks, int q, int* chunk, int* strip) {
    MCP_REQUIRE(chunk && strip && n_strips >= 1 && n_chunks >= 1 && q >= 0 && q < n_strips * n_chunks, "bad arguments");
    decode_tile(q, n_strips, n_chunks, *chunk, *strip);
    return encode_tile(*chunk, *strip, n_strips, n_chunks);
}

// Device-free planning: the schedule a plan WOULD use for a slab of `rows_own` rows and `w` cells on `grid` CTAs --
// chunk heights in slab order and the strip count.  mode 0 = grid-barrier schedule, 1 = barrier-free (dataflow)
// schedule.  Pure host arithmetic (tests run it on a box without a GPU).  Returns the chunk count.
int mcp_fused2d_plan_schedule2(int op, int dtype, int w, int rows_own, int ghost, int time_block, int grid,
                               int chunk_rows, int mode, int* rows_out, int max_entries, int* n_strips) {
    MCP_REQUIRE(rows_out && n_strips, "null argument");
    MCP_REQUIRE(w >= 1 && rows_own >= 1 && ghost >= 1 && grid >= 1, "bad geometry");
    MCP_REQUIRE(time_block == 1 || time_block == 2 || time_block == 4 || time_block == 8, "time_block must be 1, 2, 4 or 8");
    MCP_REQUIRE(mode == 0 || mode == 1, "mode must be 0 or 1");
    mcp_fused2d_desc d;
    memset(&d, 0, sizeof(d));
    d.op = op;
    d.dtype = dtype;
    d.w = w;
    d.rows_own = rows_own;
    d.ghost = ghost;
    d.chunk_rows = chunk_rows;
    Geometry g;
    build_geometry(d, grid, time_block, mode, &g);
    *n_strips = g.n_strips;
    const int n = std::min(g.n_chunks(), max_entries);
    for (int i = 0; i < n; ++i) rows_out[i] = g.chunk_y[i + 1] - g.chunk_y[i];
    return g.n_chunks();
}

int mcp_fused2d_plan_schedule(int op, int dtype, int w, int rows_own, int ghost, int time_block, int grid,
                              int chunk_rows, int* rows_out, int max_entries, int* n_strips) {
    return mcp_fused2d_plan_schedule2(op, dtype, w, rows_own, ghost, time_block, grid, chunk_rows, 0, rows_out,
                                      max_entries, n_strips);
}

// Trace (MCP_OPT_DEBUG_FLAGS & 4): per tile, in queue order, the ns spent over all sweeps of the last launch in
// the row loop (which = 0), waiting for dependencies (1) and in halo push + publish (2), with its (chunk, strip), so
// that a host tool can see which tiles are the critical path and where ranks wait for each other.
int mcp_fused2d_tile_trace2(mcp_fused2d_plan* pl, int time_block, int which, uint32_t* ns_out, int max_entries) {
    MCP_REQUIRE(pl && ns_out && which >= 0 && which <= 2, "bad argument");
    time_block = effective_tb(pl, time_block);
    const Geometry& g = pl->geo[tb_index(time_block)][choose_mode(pl, time_block, false)];
    const int n = std::min(std::min(g.n_strips * g.n_chunks(), pl->last_tiles), max_entries);
    if (n > 0)
        MCP_CUDA_TRY(cudaMemcpy(ns_out, pl->tile_ns + (size_t)which * pl->partials_cap_tiles, (size_t)n * sizeof(uint32_t),
                                cudaMemcpyDeviceToHost));
    return n;
}

int mcp_fused2d_tile_trace(mcp_fused2d_plan* pl, int time_block, uint32_t* ns_out, int* chunk_out, int* strip_out,
                           int max_entries) {
    MCP_REQUIRE(pl && ns_out, "null argument");
    time_block = effective_tb(pl, time_block);
    const Geometry& g = pl->geo[tb_index(time_block)][choose_mode(pl, time_block, false)];
    const int nstr = g.n_strips, ng = g.n_chunks();
    const int n = std::min(std::min(nstr * ng, pl->last_tiles), max_entries);
    if (n > 0) MCP_CUDA_TRY(cudaMemcpy(ns_out, pl->tile_ns, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    for (int q = 0; q < n; ++q) {
        int group, strip;
        decode_tile(q, nstr, ng, group, strip);
        if (strip_out) strip_out[q] = strip;
        if (chunk_out) chunk_out[q] = group;
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

// One asynchronous 32-byte copy of the plan's result block (status, stop word, executed sweeps, last global
// residual) into pinned host memory, ordered after the runs already enqueued on `stream`.
int mcp_fused2d_result_async(mcp_fused2d_plan* pl, void* host32, void* stream_) {
    MCP_REQUIRE(pl && host32, "null argument");
    static_assert(sizeof(ResultBlock) == 32, "result block layout");
    MCP_CUDA_TRY(cudaMemcpyAsync(host32, pl->result, sizeof(ResultBlock), cudaMemcpyDeviceToHost, (cudaStream_t)stream_));
    return MCP_OK;
}

int mcp_fused2d_status(mcp_fused2d_plan* pl) {
    MCP_REQUIRE(pl, "null plan");
    int st = 0;
    MCP_CUDA_TRY(cudaMemcpy(&st, &pl->result->status, sizeof(int), cudaMemcpyDeviceToHost));
    if (st != 0) set_error("fused2d kernel recorded status %d (%s)", st, st == MCP_ERR_TIMEOUT ? "flag/barrier wait timed out" : "error");
    return st;
}

// Clear a recorded kernel error (after a timeout the plan's counters no longer match the peers': destroy it, or
// reset every rank's plan and flag pads together).
int mcp_fused2d_clear_status(mcp_fused2d_plan* pl) {
    MCP_REQUIRE(pl, "null plan");
    MCP_CUDA_TRY(cudaMemset(&pl->result->status, 0, sizeof(int)));
    return MCP_OK;
}

int64_t mcp_fused2d_launches(const mcp_fused2d_plan* pl) { return pl ? pl->launches : 0; }

int mcp_fused2d_destroy(mcp_fused2d_plan* pl) {
    if (!pl) return MCP_OK;
    if (pl->grid_bar) cudaFree(pl->grid_bar);
    if (pl->result) cudaFree(pl->result);
    if (pl->partials) cudaFree(pl->partials);
    if (pl->res_out) cudaFree(pl->res_out);
    if (pl->tile_done) cudaFree(pl->tile_done);
    if (pl->tile_ns) cudaFree(pl->tile_ns);
    if (pl->group_args) cudaFree(pl->group_args);
    delete pl;
    return MCP_OK;
}

}  // extern "C"
