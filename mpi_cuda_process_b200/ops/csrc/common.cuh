// Shared host-side helpers for the C-ABI library: checked CUDA calls and a last-error string.
// The reference checks no return code anywhere; every entry point here returns
// 0 on success or a negative code, with the message available from mcp_last_error().
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../../include/mcp_b200.h"  // MCP_OK, MCP_ERR_* (the public return codes)

namespace mcp {

void set_error(const char* fmt, ...);

#define MCP_CUDA_TRY(expr)                                                                            \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            ::mcp::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return MCP_ERR_CUDA;                                                                      \
        }                                                                                             \
    } while (0)

#define MCP_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            ::mcp::set_error(__VA_ARGS__); \
            return MCP_ERR_ARG;           \
        }                                 \
    } while (0)

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
inline int64_t ceil_div(int64_t x, int64_t m) { return (x + m - 1) / m; }

// element kinds used across the API
enum : int { MCP_F32 = 0, MCP_F64 = 1, MCP_I32 = 2 };
inline int dtype_size(int dt) { return dt == MCP_F64 ? 8 : 4; }

// Encode a tiled tensor map through the driver entry point (no link-time libcuda dependency).
// dims/strides innermost first; strides in bytes for dims 1..rank-1.
int encode_tensor_map(CUtensorMap* out, int dtype, int rank, void* base, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box);

int device_sm_count(int* out);

}  // namespace mcp
