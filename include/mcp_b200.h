/* mcp_b200.h -- C ABI of libmcp_b200.so, the sm_100a native layer of mpi_cuda_process_b200.
 *
 * Every function returns 0 on success and a negative MCP_ERR_* code on failure; the message is
 * available from mcp_last_error() (thread-local).  Pointers named *_dev are device pointers of the
 * current CUDA device; `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *
 * "Reference" citations name the file:line of Rodrigovicente/MPI-CUDA-Process whose function the
 * entry point provides (M: = MDF_kernel.cu, K: = kernel.cu).
 */
#ifndef MCP_B200_H
#define MCP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* return codes (also what mcp_fused2d_status / mcp_fused3d_status and the result blocks carry) */
#define MCP_OK 0
#define MCP_ERR_CUDA (-1)        /* a CUDA runtime / driver call failed; mcp_last_error() has its message */
#define MCP_ERR_ARG (-2)         /* an argument or descriptor was refused before anything was launched */
#define MCP_ERR_UNSUPPORTED (-3) /* no kernel instantiation for this operator / dtype / time block */
#define MCP_ERR_TIMEOUT (-4)     /* an in-kernel flag, barrier or TMA wait ran into MCP_OPT_TIMEOUT_MS */
#define MCP_ERR_STATE (-5)       /* the object is not in a state that allows the call */

/* element kinds */
#define MCP_DTYPE_F32 0
#define MCP_DTYPE_F64 1
#define MCP_DTYPE_I32 2
#define MCP_DTYPE_B1 3  /* bit-packed cells: 32 per 32-bit word, cell x = bit x%32 of word x/32 (life only) */

/* stencil operators */
#define MCP_OP_MDF 0  /* 2-D 5-point heat/Jacobi update, M:10-22 */
#define MCP_OP_LIFE 1 /* Game of Life B3/S23, K:10-68 */

#define MCP_MAX_RANKS 16
#define MCP_RES_RING 64
#define MCP_MAX_STRIPS 2048
#define MCP_NORM_LINF 1
#define MCP_NORM_L2 2

/* residual modes of the fused kernel */
#define MCP_RES_NONE 0
#define MCP_RES_LAST 1  /* residual of the last generation of every time block (a device-side grid barrier per block) */
#define MCP_RES_FINAL 2 /* residual of the very last generation of the run only: every other time block runs the
                         * residual-free row loop on the barrier-free path; still ONE launch */

/* mcp_fused2d_set_option */
#define MCP_OPT_TIMEOUT_MS 1  /* bound of every in-kernel flag / barrier / TMA wait (default 20000) */
#define MCP_OPT_DEBUG_FLAGS 2 /* bit 2 (4): per-tile timing trace; bit 3 (8): pseudo-random delays in front of halo
                               * pushes and publishes (protocol stress tests) */
#define MCP_OPT_SYNC_MODE 3   /* 0 = choose, 1 = device-side grid barrier between sweeps, 2 = per-chunk dataflow flags */
#define MCP_OPT_PREFETCH 4    /* dataflow: 1 (default) = a tile's tail puts the next tile's first TMA stages in flight */

/* mcp_fused2d_stop_reason: why the last run-to-tolerance ended */
#define MCP_STOP_NONE 0
#define MCP_STOP_CONVERGED 1       /* the residual met the tolerance (the run ends one time block later) */
#define MCP_STOP_MAX_GENERATIONS 2 /* max_generations reached first */
#define MCP_STOP_SLOT_LIMIT 3      /* the residual history table (16384 time blocks per call) is full: call again */

/* ---- runtime (reference: bring-up/teardown in main(), M:114-144,223-234) ---------------------- */
const char* mcp_last_error(void);
int mcp_version(void);
int mcp_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem);

/* Symmetric arena: one allocation per rank, zero-filled, shareable with peer processes through
 * CUDA IPC so that kernels can store straight into a neighbour GPU's ghost rows over NVLink.
 * Replaces the replicated full-grid cudaMalloc of M:141-144 and the host staging of M:161,171,177. */
int mcp_arena_alloc(size_t bytes, void** ptr);
int mcp_arena_free(void* ptr);
int mcp_ipc_handle_size(void);
int mcp_ipc_get_handle(void* ptr, void* handle_out);
int mcp_ipc_open_handle(const void* handle, void** ptr_out);
int mcp_ipc_close_handle(void* ptr);
int mcp_enable_peer_access(int peer_device);
/* Conveniences for hosts that do not link the CUDA runtime themselves (see examples/solve_mdf.c). */
int mcp_set_device(int device);
int mcp_device_synchronize(void);
int mcp_copy_to_host(void* dst_host, const void* src_dev, size_t bytes);
int mcp_copy_to_device(void* dst_dev, const void* src_host, size_t bytes);

/* ---- initialisation and boundary conditions (reference: create_universe, M:88-99, K:131-146) --- */
/* Fill local rows [0, rows_alloc) of a pitched buffer from the global init rule: global row of local
 * row r is row0 + r; cells outside the h x w domain (ghost rows of edge ranks, pad columns) are 0. */
int mcp_init_ring2d(int dtype, void* u_dev, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                    double interior, void* stream);
/* Synthetic MDF fields (exactly reproducible on the host): kind 1 = interior uniform in [interior, ring) from
 * hash32(i*w+j, seed); kind 2 = hot box (value ring) over the central third.  The Dirichlet ring stays at `ring`. */
int mcp_init_shape2d(int dtype, int kind, void* u_dev, int64_t pitch, int rows_alloc, int row0, int h, int w,
                     double ring, double interior, uint32_t seed, void* stream);
/* Seeded synthetic life field: ring dead, interior alive iff hash32(i*w+j, seed) < prob * 2^32. */
int mcp_init_life_hash(int32_t* u_dev, int64_t pitch, int rows_alloc, int row0, int h, int w, double prob,
                       uint32_t seed, void* stream);
/* Same field, bit-packed (MCP_DTYPE_B1): pitch in 32-bit words. */
int mcp_init_life_hash_bits(uint32_t* u_dev, int64_t pitch_words, int rows_alloc, int row0, int h, int w, double prob,
                            uint32_t seed, void* stream);
/* Re-impose the fixed ring on owned rows (the "else -> 100 / 0" branches of M:35,43,59,67, K:86,102,110). */
int mcp_apply_ring2d(int dtype, void* u_dev, int64_t pitch, int rows_alloc, int row0, int h, int w, double ring,
                     void* stream);
/* 3-D: boundary shell = ring, interior = interior; local planes/rows offset by (z0, y0). */
int mcp_init_shell3d(int dtype, void* u_dev, int64_t pitch, int64_t plane_pitch, int planes_alloc, int rows_alloc,
                     int z0, int y0, int d, int h, int w, double ring, double interior, void* stream);

/* ---- plain one-generation kernels (baseline path; reference: middle_kernel/border_kernel M:24-70, K:70-113) */
/* Updates local rows [row_lo, row_hi) of `in` into `out`; reads rows row_lo-1 .. row_hi.  Cells with global
 * row/col outside [ylo,yhi] x [xlo,xhi] keep their value. */
int mcp_step2d(int op, int dtype, const void* in_dev, void* out_dev, int64_t pitch, int rows_alloc, int row0, int h,
               int w, int row_lo, int row_hi, int ylo, int yhi, int xlo, int xhi, void* stream);
/* 3-D 7-point heat step on local planes [z_lo,z_hi) x rows [y_lo,y_hi). */
int mcp_step3d(int dtype, const void* in_dev, void* out_dev, int64_t pitch, int64_t plane_pitch, int planes_alloc,
               int rows_alloc, int z0, int y0, int d, int h, int w, int z_lo, int z_hi, int y_lo, int y_hi,
               double coeff, void* stream);
/* Same contract, thread-block-cluster version: pairs of CTAs that own y-adjacent 32-row tiles form a (1,2,1) cluster,
 * each TMA-loads its own rows plus its OUTER halo row and reads the inner halo row from its partner's shared memory
 * (mapa + ld.shared::cluster), one barrier.cluster per plane.  Needs dense planes (plane_pitch = pitch * rows_alloc),
 * pitch % 32 == 0, a 128-byte aligned input and one allocated layer around the updated box.  Measured against the
 * fused kernel in profiles/r2/heat3d_cluster.md.  No counterpart in the reference (no shared memory there at all). */
int mcp_step3d_cluster(int dtype, const void* in_dev, void* out_dev, int64_t pitch, int64_t plane_pitch,
                       int planes_alloc, int rows_alloc, int z0, int y0, int d, int h, int w, int z_lo, int z_hi,
                       int y_lo, int y_hi, double coeff, void* stream);

/* ---- fused halo + stencil (+ residual) persistent kernel -------------------------------------------
 * One cooperative launch advances `generations` generations: TMA-fed shared-memory row ring per warp,
 * `time_block` generations per HBM pass held in registers, boundary rows stored straight into the
 * neighbour GPUs' ghost rows (peer pointers) and signalled with system-scope release counters.
 * Reference: the whole generation loop M:155-222 (kernels + MPI halo + staging copies). */
typedef struct mcp_fused2d_desc {
    int op;          /* MCP_OP_* */
    int dtype;       /* MCP_DTYPE_* */
    int h, w;        /* global grid */
    int row_start;   /* first owned global row */
    int rows_own;    /* owned rows */
    int ghost;       /* ghost rows above and below the owned rows in every buffer (>= time_block) */
    int64_t pitch;   /* storage elements per row, multiple of 32 (MCP_DTYPE_B1: 32-bit words of 32 cells) */
    void* buf[2];    /* local double buffer, each (ghost + rows_own + ghost) x pitch */
    void* up_buf[2]; /* rank-1's buffers mapped into this process, or NULL at the top of the chain */
    void* dn_buf[2]; /* rank+1's buffers, or NULL at the bottom */
    int up_rows_own; /* rows owned by rank-1 (locates its bottom ghost rows) */
    uint64_t* flags;   /* local: [0] arrivals from rank-1, [1] arrivals from rank+1 (8-byte counters, zeroed) */
    uint64_t* up_flag; /* address of rank-1's flags[1] mapped here, or NULL */
    uint64_t* dn_flag; /* address of rank+1's flags[0] mapped here, or NULL */
    int ylo, yhi, xlo, xhi; /* cells outside this inclusive global box keep their value */
    int chunk_rows;  /* rows per tile (0 = choose) */
    int max_ctas;    /* 0 = one CTA per SM */
    /* Fused global residual all-reduce over NVLink (optional; world <= 1 or res_inbox == NULL = local only).
     * Every rank owns an inbox ring of MCP_RES_RING x MCP_MAX_RANKS (sumsq, max) pairs and an arrival counter,
     * both zeroed, and passes the addresses of ALL ranks' inboxes/counters (its own included) mapped into this
     * process.  Results are summed in rank order, so every rank obtains the same bits. */
    int world, rank;
    double* res_inbox;                        /* local ring */
    uint64_t* res_flag;                       /* local arrival counter */
    double* peer_res_inbox[MCP_MAX_RANKS];    /* [r] = rank r's ring as seen from here */
    uint64_t* peer_res_flag[MCP_MAX_RANKS];   /* [r] = rank r's counter as seen from here */
    /* NVLS variant of the same all-reduce (takes precedence when set): a zeroed pad of 128 + 16*MCP_RES_RING bytes
     * mapped on every rank AND bound to a multicast object; nvls_local = this rank's unicast address of its copy,
     * nvls_mc = the multicast address.  The arrival counter is bumped on all GPUs by one multimem.red, the sum and
     * max come back from one multimem.ld_reduce each (reduced inside NVSwitch).  The max is order independent; the
     * sum's association order is the switch's, so use the P2P variant when bit-identical sums across ranks matter. */
    void* nvls_local;
    void* nvls_mc;
    /* Per-strip halo flags (needed for the barrier-free schedule between ranks; without them a plan with neighbours
     * keeps the device-side grid barrier between sweeps, because one arrival counter per neighbour cannot tell the
     * sweeps of a neighbour apart once that neighbour overlaps them): a zeroed array of 2 x MCP_MAX_STRIPS
     * 64-bit words per rank ([0][s] written by rank-1, [1][s] by rank+1), and the addresses of the rows this rank
     * writes: rank-1's [1][*] and rank+1's [0][*].  A boundary tile then waits only for the neighbour's boundary
     * tiles of strips s-1, s, s+1 of the previous sweep. */
    uint64_t* strip_flags;
    uint64_t* up_strip_flag;
    uint64_t* dn_strip_flag;
} mcp_fused2d_desc;

typedef struct mcp_fused2d_plan mcp_fused2d_plan;

int mcp_fused2d_create(const mcp_fused2d_desc* desc, mcp_fused2d_plan** plan_out);
/* Advance `generations` generations; the state ends in buffer mcp_fused2d_parity(plan). */
int mcp_fused2d_run(mcp_fused2d_plan* plan, int generations, int time_block, int residual_mode, void* stream);
/* Run-to-tolerance in ONE launch: advance until the global residual of a time block's last generation satisfies
 * max|u'-u| <= tol (MCP_NORM_LINF) or ||u'-u||_2 <= tol (MCP_NORM_L2), at most max_generations.  The decision is
 * taken on the device from the fused all-reduce; the run ends one time block after the block that converged.
 * Synchronises the stream (the executed count is a device result).  Extension: the reference has a fixed
 * generation count only (M:153,157,186). */
int mcp_fused2d_run_to_tolerance(mcp_fused2d_plan* plan, int max_generations, int time_block, int norm, double tol,
                                 int* generations_done, void* stream);
/* Why the last mcp_fused2d_run_to_tolerance ended: MCP_STOP_*.  A caller that sees MCP_STOP_SLOT_LIMIT has not
 * converged yet and may simply call again. */
int mcp_fused2d_stop_reason(const mcp_fused2d_plan* plan);
/* Plan options (MCP_OPT_*); take effect from the next run. */
int mcp_fused2d_set_option(mcp_fused2d_plan* plan, int option, long long value);
/* Several ranks on ONE GPU in one cooperative launch (tests, single-GPU boxes): plans[r] is rank r's plan, created
 * with its neighbours' buffers and flag words as "peer" pointers exactly as across GPUs.  CTA b works for rank
 * b % n, so the ranks run concurrently and wait on each other's flags like kernels on different GPUs do -- which
 * separate launches sharing one GPU must never do.  One launch; generations must be a multiple of the time block.
 * Built for fp32 MDF with time block 4 or 8.  Plans used here must always be advanced through this call.
 * No counterpart in the reference (it cannot run two ranks on one GPU: no cudaSetDevice, M:114-121). */
int mcp_fused2d_run_group(mcp_fused2d_plan** plans, int n, int generations, int time_block, int residual_mode,
                          void* stream);
/* Queue position q -> (group, strip) of the sweep's tile order; returns the position computed back from them. */
int mcp_fused2d_tile_at(int n_strips, int n_chunks, int q, int* chunk, int* strip);
/* Device-free planning (pure host arithmetic, works without a GPU): chunk heights in queue order and strip count
 * that a plan would use for a slab of rows_own x w cells on `grid` CTAs.  Returns the chunk count. */
int mcp_fused2d_plan_schedule(int op, int dtype, int w, int rows_own, int ghost, int time_block, int grid,
                              int chunk_rows, int* rows_out, int max_entries, int* n_strips);
/* Same with the schedule kind: mode 0 = grid-barrier schedule, 1 = barrier-free (dataflow) schedule. */
int mcp_fused2d_plan_schedule2(int op, int dtype, int w, int rows_own, int ghost, int time_block, int grid,
                               int chunk_rows, int mode, int* rows_out, int max_entries, int* n_strips);
/* Tile trace (mcp_fused2d_set_option(plan, MCP_OPT_DEBUG_FLAGS, 4) before the run): ns that every tile spent in its
 * row loop, summed over the sweeps of the last launch, in queue order, with its chunk and strip index.  Returns the
 * number of entries.  mcp_fused2d_tile_trace2 selects the quantity: 0 = row loop, 1 = waiting for dependencies
 * (previous-sweep tiles, neighbour GPUs' halo flags), 2 = halo push + publish. */
int mcp_fused2d_tile_trace(mcp_fused2d_plan* plan, int time_block, uint32_t* ns_out, int* chunk_out, int* strip_out,
                           int max_entries);
int mcp_fused2d_tile_trace2(mcp_fused2d_plan* plan, int time_block, int which, uint32_t* ns_out, int max_entries);
/* Protocol check: the number of halo arrivals this rank should have received so far from rank-1 and rank+1.
 * After a synchronised run the local counters flags[0], flags[1] must equal these (when that neighbour exists). */
int mcp_fused2d_expected_flags(const mcp_fused2d_plan* plan, uint64_t* up_dn);
int mcp_fused2d_parity(const mcp_fused2d_plan* plan);
int mcp_fused2d_set_parity(mcp_fused2d_plan* plan, int parity);
/* Residual history of the last run with MCP_RES_LAST: per time block (sum of squares, max abs) of u' - u --
 * global over all ranks when the descriptor enabled the fused all-reduce, else over this rank's owned cells.  Returns the number of entries written (<= max_entries). */
int mcp_fused2d_residuals(mcp_fused2d_plan* plan, double* sumsq_max_pairs, int max_entries);
/* 0 = healthy; otherwise the MCP_ERR_* code a kernel recorded (e.g. halo-flag wait timeout). */
int mcp_fused2d_status(mcp_fused2d_plan* plan);
int mcp_fused2d_clear_status(mcp_fused2d_plan* plan);
/* Enqueue ONE 32-byte copy of the plan's result block into (pinned) host memory behind the runs already on
 * `stream`: { int32 status; int32 stop_word; int32 sweeps_done; int32 pad; double last_sumsq; double last_max } --
 * the kernel status and the global residual of the last residual sweep without a blocking call. */
int mcp_fused2d_result_async(mcp_fused2d_plan* plan, void* host32, void* stream);
/* Number of kernel launches issued by the plan so far. */
int64_t mcp_fused2d_launches(const mcp_fused2d_plan* plan);
int mcp_fused2d_geometry(const mcp_fused2d_plan* plan, int time_block, int* n_strips, int* n_chunks, int* grid,
                         int* smem_bytes);
/* Heights of the row chunks of the tile schedule for `time_block` (queue order = chunk order). Returns the count. */
int mcp_fused2d_chunks(const mcp_fused2d_plan* plan, int time_block, int* rows_out, int max_entries);
int mcp_fused2d_destroy(mcp_fused2d_plan* plan);

/* ---- fused halo + 3-D 7-point heat kernel (pencil decomposition) -------------------------------------
 * Extension (BASELINE.json config 3); the reference is strictly 2-D (M:13-18).  One cooperative launch advances
 * `generations` generations: TMA-fed plane ring, z-march with the three planes of a cell in registers, pencil-face
 * cells stored straight into the four neighbour GPUs' ghost layers (z-faces: rows of a plane, y-faces: one strided
 * row per plane) with system-scope release counters, device-side grid barrier between generations.
 * Every rank allocates buffers of the SAME pz_alloc x py_alloc x pitch shape (sized for the largest block) so that
 * strides agree across ranks; owned local indices are 1..nz (planes) and 1..ny (rows), ghost depth 1. */
typedef struct mcp_fused3d_desc {
    int dtype;            /* MCP_DTYPE_F32 or MCP_DTYPE_F64 */
    int d, h, w;          /* global grid (planes, rows, columns) */
    int z_start, nz;      /* first owned global plane, owned planes */
    int y_start, ny;      /* first owned global row, owned rows */
    int64_t pitch;        /* elements per row, multiple of 32 */
    int py_alloc, pz_alloc;
    void* buf[2];
    void* nbr_buf[4][2];  /* z-, z+, y-, y+ neighbours' buffers mapped here, or NULL */
    int nbr_n[4];         /* nz of the z-/z+ neighbours, ny of the y-/y+ neighbours */
    uint64_t* flags;      /* local: 4 arrival counters (from z-, z+, y-, y+), zeroed */
    uint64_t* nbr_flag[4];/* the counter of each neighbour that THIS rank bumps (its opposite-face counter) */
    int expect[4];        /* arrivals per generation expected from each neighbour (mcp_fused3d_face_tiles) */
    double coeff;         /* diffusion number, 1/6 = 3-D stability limit */
    int zchunk;           /* planes per tile (0 = 64) */
    /* Fused global residual all-reduce over NVLink, as in mcp_fused2d_desc (world <= 1 or res_inbox == NULL = local). */
    int world, rank;
    double* res_inbox;
    uint64_t* res_flag;
    double* peer_res_inbox[MCP_MAX_RANKS];
    uint64_t* peer_res_flag[MCP_MAX_RANKS];
} mcp_fused3d_desc;
typedef struct mcp_fused3d_plan mcp_fused3d_plan;
int mcp_fused3d_create(const mcp_fused3d_desc* desc, mcp_fused3d_plan** plan_out);
/* Tiles of a block that touch its z-, z+, y-, y+ faces: what a neighbour must expect per generation. */
int mcp_fused3d_face_tiles(int nz, int ny, int w, int dtype, int zchunk, int* counts4);
int mcp_fused3d_run(mcp_fused3d_plan* plan, int generations, void* stream);
/* Same with the fused residual: MCP_RES_LAST = (sum of squares, max abs) of u' - u of EVERY generation (history read
 * with mcp_fused3d_residuals), MCP_RES_FINAL = of the last generation only; reduced per tile, in fixed order by one
 * warp behind the device-side grid barrier, then all-reduced over the peers' inboxes.  One launch. */
int mcp_fused3d_run_residual(mcp_fused3d_plan* plan, int generations, int residual_mode, void* stream);
/* Run-to-tolerance in ONE launch (stop decision on the device, identical on every rank; the run ends one generation
 * after the one that converged).  Synchronises the stream.  Extension: the reference is 2-D with a fixed count. */
int mcp_fused3d_run_to_tolerance(mcp_fused3d_plan* plan, int max_generations, int norm, double tol, int* generations_done,
                                 void* stream);
int mcp_fused3d_stop_reason(const mcp_fused3d_plan* plan);
int mcp_fused3d_residuals(mcp_fused3d_plan* plan, double* sumsq_max_pairs, int max_entries);
/* One asynchronous 32-byte copy of { int32 status, stop_word, gens_done, pad; double last_sumsq, last_max }. */
int mcp_fused3d_result_async(mcp_fused3d_plan* plan, void* host32, void* stream);
int mcp_fused3d_set_timeout_ms(mcp_fused3d_plan* plan, long long ms);
int mcp_fused3d_parity(const mcp_fused3d_plan* plan);
int mcp_fused3d_set_parity(mcp_fused3d_plan* plan, int parity);
int mcp_fused3d_status(mcp_fused3d_plan* plan);
int64_t mcp_fused3d_launches(const mcp_fused3d_plan* plan);
int mcp_fused3d_destroy(mcp_fused3d_plan* plan);

/* ---- residual of two fields (unfused; two-pass, fixed order, deterministic) ------------------------
 * out_host[0] = sum over the row range of (a-b)^2 in fp64, out_host[1] = max |a-b|.
 * `scratch_dev` holds at least mcp_residual2d_scratch_bytes() bytes. */
size_t mcp_residual2d_scratch_bytes(void);
int mcp_residual2d(int dtype, const void* a_dev, const void* b_dev, int64_t pitch, int w, int row_lo, int row_hi,
                   void* scratch_dev, double* out_dev, void* stream);

/* Same result as mcp_residual2d, for callers that cannot allocate device memory at all (a host whose GPU memory
 * is entirely owned by another allocator, a fixed-budget service, a plug-in forbidden to allocate after
 * start-up, a loop that cannot tolerate the synchronisation cudaMalloc may imply): the scratch is a fixed array
 * built into the library image, so the caller passes only inputs and the 2-double output.
 * Limits and rules:
 *   - at most MCP_NOALLOC_MAX_ROWS rows (row_hi - row_lo) per call; more is refused with MCP_ERR_ARG and a
 *     message -- split the range or use mcp_residual2d;
 *   - the array (1 MiB) is held by every GPU that loads the library, until the process exits;
 *   - one array per GPU is shared by all calls and keeps the previous call's contents (nothing needs zeroing:
 *     each call rewrites every entry it reads); calls that overlap on one GPU race -- serialise them;
 *   - the first call on each device looks up the array's device address (with that device current); make
 *     that first call outside any stream capture.
 * Makes no cudaMalloc/cudaFree call and uses no allocating library routine.
 * No counterpart in the reference (which has no residual and no allocation-free path). */
#define MCP_NOALLOC_MAX_ROWS 65536
int mcp_residual2d_noalloc_max_rows(void);
int mcp_residual2d_noalloc(int dtype, const void* a_dev, const void* b_dev, int64_t pitch, int w, int row_lo,
                           int row_hi, double* out_dev, void* stream);

/* ---- gather + pack (reference: print_array, K:115-129; no gather exists in the reference) -------- */
/* Cast/pack owned rows [row_lo,row_hi) x [0,w) of a pitched buffer into a dense destination (which may be a
 * peer pointer on rank 0): mode 0 = raw copy, 1 = fp64 -> fp32, 2 = int32 cells -> print_array bytes
 * ('0' if cell == 1 else ' ', '\n' after each row; dst row stride w+1), 3 = bit-packed life words (dtype
 * MCP_DTYPE_B1, pitch in words) -> one int32 per cell.  16-byte loads and stores when w % 4 == 0. */
int mcp_pack_rows(int dtype, int mode, const void* src_dev, int64_t pitch, int w, int row_lo, int row_hi,
                  void* dst_dev, int64_t dst_row0, void* stream);

/* Same for a 3-D block (pencils): owned planes [z_lo, z_lo+nz) x rows [y_lo, y_lo+ny) x [0,w) of a
 * (planes, rows, pitch) buffer into a dense (D, dst_h, w) destination at plane dst_z0, row dst_y0.
 * mode 0 = raw copy, 1 = fp64 -> fp32.  A 2-D slab is the case nz = 1. */
int mcp_pack_block(int dtype, int mode, const void* src_dev, int64_t pitch, int64_t plane_pitch, int z_lo, int nz,
                   int y_lo, int ny, int w, void* dst_dev, int64_t dst_h, int64_t dst_z0, int64_t dst_y0, void* stream);

/* Position-dependent 64-bit checksum of rows [row_lo,row_hi) x [0,w) (global rows grow0 ...): the wrapping sum over
 * cells of mix64(bits ^ golden * global index), written to the device word out_dev.  Order independent, so the
 * per-rank checksums of a decomposed field add up (mod 2^64) to the checksum of the whole field: fields of benchmark
 * size can be compared bit for bit across rank counts and backends without a gather.  No counterpart in the reference. */
int mcp_checksum2d(int dtype, const void* u_dev, int64_t pitch, int w, int row_lo, int row_hi, int64_t grow0,
                   void* out_dev, void* stream);

/* ---- CPU reference (host code; works without a GPU) ------------------------------------------------ */
/* glibc rand() stream exactly as the reference consumes it (K:141-142): fills an h x w int32 grid. */
int mcp_host_life_init_reference(int32_t* u, int h, int w, float prob);
/* One generation on the host with the literal mixed-precision arithmetic of M:20 (fp32 adds, fp64 final fma). */
int mcp_host_mdf_step_f32(const float* in, float* out, int h, int w, int strict_fp64);
int mcp_host_mdf_step_f64(const double* in, double* out, int h, int w);
int mcp_host_life_step(const int32_t* in, int32_t* out, int h, int w, int fixed_ring);
/* print_array (K:115-129) into a byte buffer of size 2 + h*(w+1). Returns bytes written. */
int64_t mcp_host_print_array(const int32_t* u, int w, int64_t size, char* out, int64_t out_cap);

#ifdef __cplusplus
}
#endif
#endif /* MCP_B200_H */
