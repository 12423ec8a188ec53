/* Implementation of baseline/mpi_shim/mpi.h over named pipes.
 * The launcher exports MCP_MPI_RANK, MCP_MPI_SIZE and MCP_MPI_DIR; pipe "<dir>/p_<src>_<dst>" carries the messages
 * from src to dst as (tag, nbytes, payload) records.  A send blocks once the pipe buffer is full (like an MPI
 * rendezvous), a receive blocks until the next record arrives; records of one pair are delivered in order. */
#ifndef _GNU_SOURCE
#define _GNU_SOURCE /* F_SETPIPE_SZ */
#endif
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#define MAXR 64
static int g_rank = 0, g_size = 1, g_init = 0;
static int g_out[MAXR], g_in[MAXR];
static char g_dir[512];

static int dt_size(MPI_Datatype dt) {
    switch (dt) {
        case MPI_CHAR: return 1;
        case MPI_INT: return 4;
        case MPI_FLOAT: return 4;
        case MPI_DOUBLE: return 8;
        default: return 1;
    }
}

static int full_write(int fd, const void* p, size_t n) {
    const char* c = (const char*)p;
    while (n > 0) {
        ssize_t k = write(fd, c, n);
        if (k < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        c += k;
        n -= (size_t)k;
    }
    return 0;
}

static int full_read(int fd, void* p, size_t n) {
    char* c = (char*)p;
    while (n > 0) {
        ssize_t k = read(fd, c, n);
        if (k < 0) {
            if (errno == EINTR) continue;
            return -1;
        }
        if (k == 0) return -1; /* peer closed */
        c += k;
        n -= (size_t)k;
    }
    return 0;
}

static int open_pipe(int src, int dst, int flags) {
    char path[640];
    snprintf(path, sizeof(path), "%s/p_%d_%d", g_dir, src, dst);
    return open(path, flags);
}

int MPI_Init(int* argc, char*** argv) {
    (void)argc;
    (void)argv;
    const char* r = getenv("MCP_MPI_RANK");
    const char* s = getenv("MCP_MPI_SIZE");
    const char* d = getenv("MCP_MPI_DIR");
    g_rank = r ? atoi(r) : 0;
    g_size = s ? atoi(s) : 1;
    if (g_size > MAXR) return MPI_ERR_OTHER;
    snprintf(g_dir, sizeof(g_dir), "%s", d ? d : "/tmp");
    for (int i = 0; i < MAXR; ++i) g_out[i] = g_in[i] = -1;
    /* O_RDWR never blocks at open time on Linux FIFOs and keeps the pipe alive until both sides are done */
    for (int p = 0; p < g_size; ++p) {
        if (g_size == 1) break;
        g_out[p] = open_pipe(g_rank, p, O_RDWR);
        g_in[p] = open_pipe(p, g_rank, O_RDWR);
        if (g_out[p] < 0 || g_in[p] < 0) return MPI_ERR_OTHER;
#ifdef F_SETPIPE_SZ
        /* "eager buffer": a border row of one-element messages (12 B each) must fit without a matching receive,
         * as it would in a real MPI's eager protocol (the reference sends at the end of a generation and receives
         * at the start of the next one, MDF_kernel.cu:166-183).  Best effort; the default is 64 KiB. */
        (void)fcntl(g_out[p], F_SETPIPE_SZ, 1 << 20);
#endif
    }
    g_init = 1;
    return MPI_SUCCESS;
}

int MPI_Finalize(void) {
    for (int p = 0; p < MAXR; ++p) {
        if (g_out[p] >= 0) close(g_out[p]);
        if (g_in[p] >= 0) close(g_in[p]);
    }
    g_init = 0;
    return MPI_SUCCESS;
}

int MPI_Comm_size(MPI_Comm comm, int* size) {
    (void)comm;
    *size = g_size;
    return MPI_SUCCESS;
}

int MPI_Comm_rank(MPI_Comm comm, int* rank) {
    (void)comm;
    *rank = g_rank;
    return MPI_SUCCESS;
}

int MPI_Get_processor_name(char* name, int* resultlen) {
    if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) strcpy(name, "localhost");
    name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
    *resultlen = (int)strlen(name);
    return MPI_SUCCESS;
}

int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm) {
    (void)comm;
    if (!g_init || dest < 0 || dest >= g_size || g_out[dest] < 0) return MPI_ERR_OTHER;
    int hdr[2] = {tag, count * dt_size(dt)};
    if (full_write(g_out[dest], hdr, sizeof(hdr))) return MPI_ERR_OTHER;
    if (full_write(g_out[dest], buf, (size_t)hdr[1])) return MPI_ERR_OTHER;
    return MPI_SUCCESS;
}

int MPI_Recv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status* status) {
    (void)comm;
    if (!g_init || source < 0 || source >= g_size || g_in[source] < 0) return MPI_ERR_OTHER;
    int hdr[2];
    if (full_read(g_in[source], hdr, sizeof(hdr))) return MPI_ERR_OTHER;
    const int cap = count * dt_size(dt);
    if (hdr[0] != tag || hdr[1] > cap) return MPI_ERR_OTHER; /* in-order, single-tag traffic only */
    if (full_read(g_in[source], buf, (size_t)hdr[1])) return MPI_ERR_OTHER;
    if (status) {
        status->MPI_SOURCE = source;
        status->MPI_TAG = hdr[0];
        status->MPI_ERROR = MPI_SUCCESS;
        status->count_bytes = hdr[1];
    }
    return MPI_SUCCESS;
}

int MPI_Barrier(MPI_Comm comm) {
    /* dissemination over the pipes: everybody reports to rank 0, rank 0 releases everybody */
    char c = 0;
    if (g_size == 1) return MPI_SUCCESS;
    if (g_rank == 0) {
        for (int p = 1; p < g_size; ++p)
            if (MPI_Recv(&c, 1, MPI_CHAR, p, -77, comm, MPI_STATUS_IGNORE)) return MPI_ERR_OTHER;
        for (int p = 1; p < g_size; ++p)
            if (MPI_Send(&c, 1, MPI_CHAR, p, -77, comm)) return MPI_ERR_OTHER;
    } else {
        if (MPI_Send(&c, 1, MPI_CHAR, 0, -77, comm)) return MPI_ERR_OTHER;
        if (MPI_Recv(&c, 1, MPI_CHAR, 0, -77, comm, MPI_STATUS_IGNORE)) return MPI_ERR_OTHER;
    }
    return MPI_SUCCESS;
}
