/* The reference's program shape -- an MPI program, one rank per GPU, `generations height width` on the stdin of every
 * rank after the three prompts (MDF_kernel.cu:103-131) -- on the native layer, without Python: row slabs, the
 * symmetric arena exported and mapped with CUDA IPC, halo rows pushed GPU to GPU from inside ONE kernel launch per rank.
 * MPI is used exactly as far as the reference uses it (Init, rank, size, blocking Send/Recv, Finalize) and only to
 * exchange the 64-byte IPC handles and the final checksums; baseline/mpi_shim provides those calls on this box.
 *
 *   gcc -O2 -I include -I baseline/mpi_shim examples/solve_mdf_mpi.c baseline/mpi_shim/mpi_shim.c \
 *       -L mpi_cuda_process_b200/ops/_lib -lmcp_b200 -Wl,-rpath,$PWD/mpi_cuda_process_b200/ops/_lib -o /tmp/solve_mdf_mpi
 *   echo "1000 16384 16384" | python baseline/mpi_shim/mcp_mpirun.py -np 8 /tmp/solve_mdf_mpi
 *
 * Rank 0 prints one JSON line with the 64-bit checksum of the global field (the ranks' checksums add up), which equals
 * `Solver.checksum()` of the same run and the CPU oracle's (tests/test_zzz_native_mpi_example.py).
 * Arena layout per rank (the one solver.py uses): halo counters @0, per-strip halo flags @32768, buffers @65536. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mpi.h>

#include "mcp_b200.h"

#define CHECK(call)                                                                         \
    do {                                                                                    \
        if ((call) < 0) {                                                                   \
            fprintf(stderr, "rank %d: %s failed: %s\n", rank, #call, mcp_last_error());     \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)

enum { PAD_BYTES = 65536, STRIP_FLAGS_AT = 32768, HANDLE_BYTES = 64, TB = 8 };

static int rows_of(int h, int size, int r) { return h / size + (r < h % size ? 1 : 0); }

static int start_of(int h, int size, int r) {
    int s = 0;
    for (int i = 0; i < r; ++i) s += rows_of(h, size, i);
    return s;
}

/* Blocking pairwise exchange of one IPC handle with a neighbour; the lower rank of a pair sends first. */
static void swap_handles(int rank, int peer, const char* mine, char* theirs) {
    if (rank < peer) {
        MPI_Send(mine, HANDLE_BYTES, MPI_CHAR, peer, 7, MPI_COMM_WORLD);
        MPI_Recv(theirs, HANDLE_BYTES, MPI_CHAR, peer, 7, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
    } else {
        MPI_Recv(theirs, HANDLE_BYTES, MPI_CHAR, peer, 7, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
        MPI_Send(mine, HANDLE_BYTES, MPI_CHAR, peer, 7, MPI_COMM_WORLD);
    }
}

int main(int argc, char** argv) {
    int g, h, w, rank = 0, size = 1;
    printf("Enter desired number of generations:\n");
    if (scanf("%d", &g) != 1) return 2;
    printf("Enter desired height of universe:\n");
    if (scanf("%d", &h) != 1) return 2;
    printf("Enter desired width of universe:\n");
    if (scanf("%d", &w) != 1) return 2;

    MPI_Init(&argc, &argv);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    if (size > MCP_MAX_RANKS || h / size < TB || g % TB != 0) {
        if (rank == 0) fprintf(stderr, "need <= %d ranks, >= %d rows per rank and a multiple of %d generations\n",
                               MCP_MAX_RANKS, TB, TB);
        MPI_Finalize();
        return 2;
    }
    CHECK(mcp_set_device(rank));                               /* one rank per GPU (the reference: all on GPU 0) */
    if (mcp_ipc_handle_size() != HANDLE_BYTES) return 3;

    const int ghost = TB, rows_own = rows_of(h, size, rank), row_start = start_of(h, size, rank);
    const int max_rows = rows_of(h, size, 0);                  /* every rank uses the layout of the largest slab */
    const int64_t pitch = (w + 31) / 32 * 32;
    const size_t buf_bytes = ((size_t)(max_rows + 2 * ghost) * (size_t)pitch * sizeof(float) + 4095) / 4096 * 4096;
    const int rows_alloc = rows_own + 2 * ghost;

    char* arena = NULL;
    CHECK(mcp_arena_alloc(PAD_BYTES + 2 * buf_bytes, (void**)&arena));                 /* zero-filled */
    float* buf[2] = {(float*)(arena + PAD_BYTES), (float*)(arena + PAD_BYTES + buf_bytes)};
    for (int b = 0; b < 2; ++b)
        CHECK(mcp_init_ring2d(MCP_DTYPE_F32, buf[b], pitch, rows_alloc, row_start - ghost, h, w, 100.0, 0.0, NULL));
    CHECK(mcp_device_synchronize());

    /* export this rank's arena, map the neighbours' (rank-1 = up, rank+1 = down) */
    char mine[HANDLE_BYTES], theirs[HANDLE_BYTES];
    char *up = NULL, *dn = NULL;
    CHECK(mcp_ipc_get_handle(arena, mine));
    for (int phase = 0; phase < 2; ++phase) {                  /* pairs (even, even+1), then pairs (odd, odd+1) */
        const int lower = (rank % 2 == phase) ? rank : rank - 1;
        const int peer = lower == rank ? rank + 1 : rank - 1;
        if (lower < 0 || lower + 1 >= size) continue;
        swap_handles(rank, peer, mine, theirs);
        CHECK(mcp_ipc_open_handle(theirs, (void**)(peer < rank ? &up : &dn)));
    }

    mcp_fused2d_desc d;
    memset(&d, 0, sizeof(d));
    d.op = MCP_OP_MDF;
    d.dtype = MCP_DTYPE_F32;
    d.h = h; d.w = w;
    d.row_start = row_start; d.rows_own = rows_own; d.ghost = ghost; d.pitch = pitch;
    d.ylo = 1; d.yhi = h - 2; d.xlo = 1; d.xhi = w - 2;        /* the Dirichlet ring keeps its value */
    d.world = size; d.rank = rank;
    d.flags = (uint64_t*)arena;                                /* [0] arrivals from up, [1] arrivals from down */
    d.strip_flags = (uint64_t*)(arena + STRIP_FLAGS_AT);       /* [0][s] written by up, [1][s] by down */
    for (int b = 0; b < 2; ++b) {
        d.buf[b] = buf[b];
        d.up_buf[b] = up ? up + PAD_BYTES + (size_t)b * buf_bytes : NULL;
        d.dn_buf[b] = dn ? dn + PAD_BYTES + (size_t)b * buf_bytes : NULL;
    }
    if (up) {
        d.up_rows_own = rows_of(h, size, rank - 1);
        d.up_flag = (uint64_t*)(up + 8);                                                /* up's flags[1] */
        d.up_strip_flag = (uint64_t*)(up + STRIP_FLAGS_AT + MCP_MAX_STRIPS * 8);        /* up's [1][*] row */
    }
    if (dn) {
        d.dn_flag = (uint64_t*)dn;                                                      /* down's flags[0] */
        d.dn_strip_flag = (uint64_t*)(dn + STRIP_FLAGS_AT);                             /* down's [0][*] row */
    }
    mcp_fused2d_plan* plan = NULL;
    CHECK(mcp_fused2d_create(&d, &plan));
    MPI_Barrier(MPI_COMM_WORLD);                               /* every field is initialised and every arena mapped */

    CHECK(mcp_fused2d_run(plan, g, TB, MCP_RES_NONE, NULL));   /* ONE cooperative launch; halo exchange inside */
    CHECK(mcp_device_synchronize());
    CHECK(mcp_fused2d_status(plan));

    /* checksum of the owned rows on the device; rank 0 adds the ranks' values */
    uint64_t* sum_dev = (uint64_t*)(arena + 1024);             /* a free word of the pad */
    uint64_t sum = 0, total = 0;
    const float* cur = buf[mcp_fused2d_parity(plan)];
    CHECK(mcp_checksum2d(MCP_DTYPE_F32, cur, pitch, w, ghost, ghost + rows_own, row_start, sum_dev, NULL));
    CHECK(mcp_device_synchronize());
    CHECK(mcp_copy_to_host(&sum, sum_dev, sizeof(sum)));
    if (rank == 0) {
        total = sum;
        for (int r = 1; r < size; ++r) {
            uint64_t other = 0;
            MPI_Recv(&other, 8, MPI_CHAR, r, 9, MPI_COMM_WORLD, MPI_STATUS_IGNORE);
            total += other;
        }
        printf("{\"ranks\": %d, \"grid\": [%d, %d], \"generations\": %d, \"launches_per_rank\": %lld, "
               "\"checksum\": \"0x%016llx\"}\n", size, h, w, g, (long long)mcp_fused2d_launches(plan),
               (unsigned long long)total);
    } else {
        MPI_Send(&sum, 8, MPI_CHAR, 0, 9, MPI_COMM_WORLD);
    }

    MPI_Barrier(MPI_COMM_WORLD);                               /* nobody unmaps an arena a peer may still use */
    mcp_fused2d_destroy(plan);
    if (up) mcp_ipc_close_handle(up);
    if (dn) mcp_ipc_close_handle(dn);
    mcp_arena_free(arena);
    MPI_Finalize();
    return 0;
}
