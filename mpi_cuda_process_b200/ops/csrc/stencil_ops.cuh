// Cell-level update rules shared by the fused and the plain kernels.  Each is the reference's arithmetic
// with the rounding sequence pinned by explicit intrinsics (no contraction, no re-association), so a GPU
// result can be compared bit for bit with oracle.py.
#pragma once
#include <cuda_runtime.h>

namespace mcp {

// MDF_kernel.cu:20: new = 0.25*(E + W + N + S - 4*C) + C, operand order right, left, up, down.
// Reference SASS: FADD, FADD, FADD, FFMA(C,-4,s), then DFMA(s,0.25,C) in fp64 rounded to fp32; the
// all-fp32 fma below is bit-identical because 0.25*s is exact (tests/test_oracle.py).
__device__ __forceinline__ float mdf_cell(float C, float N, float S, float W, float E) {
    float s = __fadd_rn(E, W);
    s = __fadd_rn(s, N);
    s = __fadd_rn(s, S);
    s = __fmaf_rn(C, -4.0f, s);
    return __fmaf_rn(s, 0.25f, C);
}
// literal mixed-precision form (the --strict-fp64-accum switch)
__device__ __forceinline__ float mdf_cell_strict(float C, float N, float S, float W, float E) {
    float s = __fadd_rn(E, W);
    s = __fadd_rn(s, N);
    s = __fadd_rn(s, S);
    s = __fmaf_rn(C, -4.0f, s);
    return (float)__fma_rn((double)s, 0.25, (double)C);
}
__device__ __forceinline__ double mdf_cell(double C, double N, double S, double W, double E) {
    double s = __dadd_rn(E, W);
    s = __dadd_rn(s, N);
    s = __dadd_rn(s, S);
    s = __fma_rn(C, -4.0, s);
    return __fma_rn(s, 0.25, C);
}

// kernel.cu:61,66: B3/S23 over the 8-neighbourhood.
__device__ __forceinline__ int life_rule(int n, int c) { return (n == 3 || (n == 2 && c != 0)) ? 1 : 0; }

// 3-D 7-point FTCS step, every operation individually rounded (oracle.heat3d_update_cells).
__device__ __forceinline__ float heat3d_cell(float C, float N, float S, float W, float E, float U, float D,
                                             float coeff) {
    float s = __fadd_rn(E, W);
    s = __fadd_rn(s, N);
    s = __fadd_rn(s, S);
    s = __fadd_rn(s, U);
    s = __fadd_rn(s, D);
    const float d = __fsub_rn(s, __fmul_rn(6.0f, C));
    return __fadd_rn(C, __fmul_rn(coeff, d));
}
__device__ __forceinline__ double heat3d_cell(double C, double N, double S, double W, double E, double U, double D,
                                              double coeff) {
    double s = __dadd_rn(E, W);
    s = __dadd_rn(s, N);
    s = __dadd_rn(s, S);
    s = __dadd_rn(s, U);
    s = __dadd_rn(s, D);
    const double d = __dsub_rn(s, __dmul_rn(6.0, C));
    return __dadd_rn(C, __dmul_rn(coeff, d));
}

}  // namespace mcp
