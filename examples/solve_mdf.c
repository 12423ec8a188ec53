/* Pure-C use of the native layer (no CUDA headers, no Python): the reference's whole program -- read g, h, w from
 * stdin after its three prompts (MDF_kernel.cu:103-112), initialise the ring, iterate, look at the result -- on one
 * GPU through libmcp_b200.so.
 *
 *   gcc -O2 -I include examples/solve_mdf.c -L mpi_cuda_process_b200/ops/_lib -lmcp_b200 \
 *       -Wl,-rpath,$PWD/mpi_cuda_process_b200/ops/_lib -o /tmp/solve_mdf
 *   echo "100 256 256" | /tmp/solve_mdf
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "mcp_b200.h"

#define CHECK(call)                                                        \
    do {                                                                   \
        if ((call) < 0) {                                                  \
            fprintf(stderr, "%s failed: %s\n", #call, mcp_last_error());   \
            return 1;                                                      \
        }                                                                  \
    } while (0)

int main(void) {
    int g, h, w;
    printf("Enter desired number of generations:\n");
    if (scanf("%d", &g) != 1) return 2;
    printf("Enter desired height of universe:\n");
    if (scanf("%d", &h) != 1) return 2;
    printf("Enter desired width of universe:\n");
    if (scanf("%d", &w) != 1) return 2;

    const int ghost = 8;                                   /* = time block: 8 generations per pass over the field */
    const int64_t pitch = (w + 31) / 32 * 32;
    const int rows_alloc = h + 2 * ghost;
    const size_t buf_bytes = (size_t)rows_alloc * pitch * sizeof(float);

    CHECK(mcp_set_device(0));
    void* arena = NULL;
    CHECK(mcp_arena_alloc(2 * buf_bytes, &arena));         /* zero-filled */
    float* buf0 = (float*)arena;
    float* buf1 = (float*)((char*)arena + buf_bytes);
    CHECK(mcp_init_ring2d(MCP_DTYPE_F32, buf0, pitch, rows_alloc, -ghost, h, w, 100.0, 0.0, NULL));
    CHECK(mcp_init_ring2d(MCP_DTYPE_F32, buf1, pitch, rows_alloc, -ghost, h, w, 100.0, 0.0, NULL));

    mcp_fused2d_desc d;
    memset(&d, 0, sizeof(d));
    d.op = MCP_OP_MDF;
    d.dtype = MCP_DTYPE_F32;
    d.h = h; d.w = w;
    d.row_start = 0; d.rows_own = h; d.ghost = ghost; d.pitch = pitch;
    d.buf[0] = buf0; d.buf[1] = buf1;
    d.ylo = 1; d.yhi = h - 2; d.xlo = 1; d.xhi = w - 2;    /* the Dirichlet ring keeps its value */
    mcp_fused2d_plan* plan = NULL;
    CHECK(mcp_fused2d_create(&d, &plan));
    CHECK(mcp_fused2d_run(plan, g, 8, MCP_RES_LAST, NULL));
    CHECK(mcp_device_synchronize());
    CHECK(mcp_fused2d_status(plan));

    double res[2 * 16384];
    const int nres = mcp_fused2d_residuals(plan, res, 16384);
    const float* cur = mcp_fused2d_parity(plan) ? buf1 : buf0;
    float* row = (float*)malloc((size_t)pitch * sizeof(float));
    CHECK(mcp_copy_to_host(row, cur + (size_t)(ghost + h / 2) * pitch, (size_t)pitch * sizeof(float)));
    printf("after %d generations: u[%d][1] = %.9g, u[%d][%d] = %.9g, last max|du| = %.9g (%lld launches)\n", g, h / 2,
           row[1], h / 2, w / 2, row[w / 2], nres > 0 ? res[2 * nres - 1] : 0.0, (long long)mcp_fused2d_launches(plan));
    free(row);
    mcp_fused2d_destroy(plan);
    mcp_arena_free(arena);
    return 0;
}
