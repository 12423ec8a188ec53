// Host-side reference code: CPU oracle in C++ and the glibc rand() initialiser.
// Counterparts: create_universe (kernel.cu:131-146), the update expressions (MDF_kernel.cu:20,
// kernel.cu:61,66) under the intended semantics of SURVEY.md 2.8, print_array (kernel.cu:115-129).
// Built twice: into libmcp_b200.so (GPU box) and into libmcp_host.so (no CUDA needed).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

extern "C" {

int mcp_host_life_init_reference(int32_t* u, int h, int w, float prob) {
    // Same consumption order as kernel.cu:134-145: one rand() per interior cell, row-major; the stream is the
    // process-wide unseeded glibc generator, so re-seed to the default (1) to make the call repeatable.
    srand(1);
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            const int64_t k = (int64_t)i * w + j;
            if (i == 0 || j == 0 || i == h - 1 || j == w - 1) {
                u[k] = 0;
            } else {
                const float rp = (float)rand() / RAND_MAX;
                u[k] = rp > prob ? 0 : 1;
            }
        }
    return 0;
}

int mcp_host_mdf_step_f32(const float* in, float* out, int h, int w, int strict_fp64) {
    memcpy(out, in, (size_t)h * w * sizeof(float));
#pragma omp parallel for schedule(static)
    for (int i = 1; i < h - 1; ++i)
        for (int j = 1; j < w - 1; ++j) {
            const int64_t k = (int64_t)i * w + j;
            const float C = in[k];
            float s = in[k + 1] + in[k - 1];
            s = s + in[k - w];
            s = s + in[k + w];
            s = __builtin_fmaf(C, -4.0f, s);
            out[k] = strict_fp64 ? (float)__builtin_fma((double)s, 0.25, (double)C) : __builtin_fmaf(s, 0.25f, C);
        }
    return 0;
}

int mcp_host_mdf_step_f64(const double* in, double* out, int h, int w) {
    memcpy(out, in, (size_t)h * w * sizeof(double));
#pragma omp parallel for schedule(static)
    for (int i = 1; i < h - 1; ++i)
        for (int j = 1; j < w - 1; ++j) {
            const int64_t k = (int64_t)i * w + j;
            const double C = in[k];
            double s = in[k + 1] + in[k - 1];
            s = s + in[k - w];
            s = s + in[k + w];
            s = __builtin_fma(C, -4.0, s);
            out[k] = __builtin_fma(s, 0.25, C);
        }
    return 0;
}

int mcp_host_life_step(const int32_t* in, int32_t* out, int h, int w, int fixed_ring) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            const int64_t k = (int64_t)i * w + j;
            if (fixed_ring && (i == 0 || j == 0 || i == h - 1 || j == w - 1)) {
                out[k] = in[k];
                continue;
            }
            int n = 0;
            for (int di = -1; di <= 1; ++di)
                for (int dj = -1; dj <= 1; ++dj) {
                    if (!di && !dj) continue;
                    const int y = i + di, x = j + dj;
                    if (y >= 0 && y < h && x >= 0 && x < w) n += in[(int64_t)y * w + x];
                }
            out[k] = (n == 3 || (n == 2 && in[k])) ? 1 : 0;
        }
    return 0;
}

int64_t mcp_host_print_array(const int32_t* u, int w, int64_t size, char* out, int64_t cap) {
    int64_t n = 0;
    auto put = [&](char c) {
        if (n < cap) out[n] = c;
        ++n;
    };
    put('\n');
    for (int64_t i = 0; i < size; ++i) {
        put(u[i] == 1 ? '0' : ' ');
        if ((i + 1) % w == 0) put('\n');
    }
    put('\n');
    return n;
}

}  // extern "C"
