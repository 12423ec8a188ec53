/* Minimal MPI stand-in for this box (no MPI is installed, no network to install one).
 * Exactly the seven calls the reference makes (MDF_kernel.cu:124-131,168,182,234 / kernel.cu:171-178,215,229,281)
 * over named pipes between processes started by baseline/mpi_shim/mcp_mpirun.py.  Blocking, in-order, tag-checked.
 * Purpose: prove that the reference's sources compile unmodified against an mpi.h, and give the host-staged
 * baseline a process-to-process transport with MPI semantics.  NOT a general MPI. */
#ifndef MCP_MPI_SHIM_H
#define MCP_MPI_SHIM_H
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct MPI_Status {
    int MPI_SOURCE, MPI_TAG, MPI_ERROR;
    int count_bytes;
} MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_MAX_PROCESSOR_NAME 256
#define MPI_CHAR 1
#define MPI_INT 4
#define MPI_FLOAT 14
#define MPI_DOUBLE 18
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Get_processor_name(char* name, int* resultlen);
int MPI_Send(const void* buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status* status);
int MPI_Barrier(MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif
