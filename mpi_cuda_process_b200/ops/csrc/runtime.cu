// Runtime bring-up pieces of the C-ABI library: error string, device query, TMA descriptor
// encoding, symmetric arena (cudaMalloc + CUDA IPC) and peer mapping.
// Reference counterpart: stream/MPI/alloc bring-up and teardown in main()
// (MDF_kernel.cu:114-144, 223-234), which never selects a device (D2) and never checks an error (D15).
#include <stdarg.h>

#include "../../../include/mcp_b200.h"
#include "common.cuh"

namespace mcp {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
        set_error("cuTensorMapEncodeTiled entry point not available: %s", cudaGetErrorString(e));
        return nullptr;
    }
    fn = reinterpret_cast<PFN_encodeTiled>(p);
    return fn;
}

int encode_tensor_map(CUtensorMap* out, int dtype, int rank, void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box) {
    PFN_encodeTiled fn = get_encode_fn();
    if (!fn) return MCP_ERR_CUDA;
    CUtensorMapDataType dt = dtype == MCP_F64   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64
                             : dtype == MCP_I32 ? CU_TENSOR_MAP_DATA_TYPE_INT32
                                                : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    cuuint64_t gdim[5], gstr[5];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        es[i] = 1;
        if (i + 1 < rank) gstr[i] = strides_bytes[i];
    }
    CUresult r = fn(out, dt, (cuuint32_t)rank, base, gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u)", (int)r,
                  rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 1), box[0],
                  rank > 1 ? box[1] : 1);
        return MCP_ERR_CUDA;
    }
    return MCP_OK;
}

int device_sm_count(int* out) {
    int dev = 0;
    MCP_CUDA_TRY(cudaGetDevice(&dev));
    MCP_CUDA_TRY(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
    return MCP_OK;
}

}  // namespace mcp

using namespace mcp;

extern "C" {

const char* mcp_last_error(void) { return g_err; }

int mcp_version(void) { return 100; }

int mcp_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem) {
    int dev = 0;
    MCP_CUDA_TRY(cudaGetDevice(&dev));
    cudaDeviceProp p;
    MCP_CUDA_TRY(cudaGetDeviceProperties(&p, dev));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc_major) *cc_major = p.major;
    if (cc_minor) *cc_minor = p.minor;
    if (total_mem) *total_mem = p.totalGlobalMem;
    return MCP_OK;
}

// ---- symmetric arena: one cudaMalloc per rank, shared with peers through CUDA IPC -----------------

int mcp_arena_alloc(size_t bytes, void** ptr) {
    MCP_REQUIRE(ptr && bytes > 0, "mcp_arena_alloc: bad arguments");
    MCP_CUDA_TRY(cudaMalloc(ptr, bytes));
    cudaError_t e = cudaMemset(*ptr, 0, bytes);
    if (e != cudaSuccess) {
        cudaFree(*ptr);
        *ptr = nullptr;
        MCP_CUDA_TRY(e);
    }
    return MCP_OK;
}

int mcp_arena_free(void* ptr) {
    if (ptr) MCP_CUDA_TRY(cudaFree(ptr));
    return MCP_OK;
}

int mcp_ipc_handle_size(void) { return (int)sizeof(cudaIpcMemHandle_t); }

int mcp_ipc_get_handle(void* ptr, void* handle_out) {
    MCP_REQUIRE(ptr && handle_out, "mcp_ipc_get_handle: null argument");
    cudaIpcMemHandle_t h;
    MCP_CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle_out, &h, sizeof(h));
    return MCP_OK;
}

int mcp_ipc_open_handle(const void* handle, void** ptr_out) {
    MCP_REQUIRE(handle && ptr_out, "mcp_ipc_open_handle: null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    MCP_CUDA_TRY(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return MCP_OK;
}

int mcp_ipc_close_handle(void* ptr) {
    if (ptr) MCP_CUDA_TRY(cudaIpcCloseMemHandle(ptr));
    return MCP_OK;
}

// Small conveniences for hosts that do not link the CUDA runtime themselves (pure C, ctypes, ...).
int mcp_copy_to_host(void* dst_host, const void* src_dev, size_t bytes) {
    MCP_REQUIRE(dst_host && src_dev, "mcp_copy_to_host: null argument");
    MCP_CUDA_TRY(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
    return MCP_OK;
}

int mcp_copy_to_device(void* dst_dev, const void* src_host, size_t bytes) {
    MCP_REQUIRE(dst_dev && src_host, "mcp_copy_to_device: null argument");
    MCP_CUDA_TRY(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
    return MCP_OK;
}

int mcp_device_synchronize(void) {
    MCP_CUDA_TRY(cudaDeviceSynchronize());
    return MCP_OK;
}

int mcp_set_device(int device) {
    MCP_CUDA_TRY(cudaSetDevice(device));
    return MCP_OK;
}

// Same-process peer mapping (used by single-process multi-GPU tests and tools).
int mcp_enable_peer_access(int peer_device) {
    int dev = 0, can = 0;
    MCP_CUDA_TRY(cudaGetDevice(&dev));
    if (dev == peer_device) return MCP_OK;
    MCP_CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, peer_device));
    MCP_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
    cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        return MCP_OK;
    }
    MCP_CUDA_TRY(e);
    return MCP_OK;
}

}  // extern "C"
