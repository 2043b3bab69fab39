// harness_common.h -- shared bits of the test_spmv / test_ilu0 harnesses (host code only).
#pragma once
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "reaspasol_gpu.h"
#include "rsg_host.h"

struct HarnessArgs {
    const char *matrix = nullptr;
    char precision = 'd'; // d | s | m   (the reference switches with a compile-time #define FLOAT)
    int repeat = 50;      // REPEAT in GPU/spmv.cu:174
    int warmup = 0;       // the reference has none
    int algo = 0;
    int gpus = 1;         // --gpus=N: row-block product over N devices of this process (rsg_spmv_multi_*)
    bool stats = false;   // extra lines after the reference's own output
    bool expand_symmetric = false;
};

inline bool parse_args(int argc, char **argv, HarnessArgs *a)
{
    if (argc < 2) {
        fprintf(stderr,
                "-- Usage examples --\n"
                "  %s inline_1.mtx type: run with inline_1 matrix in matrix market format\n",
                argv[0]);
        return false;
    }
    a->matrix = argv[1];
#ifdef FLOAT
    a->precision = 's';
#endif
    for (int i = 2; i < argc; ++i) {
        std::string s(argv[i]);
        auto val = [&](const char *key) -> const char * {
            size_t n = strlen(key);
            return s.compare(0, n, key) == 0 ? s.c_str() + n : nullptr;
        };
        if (const char *v = val("--precision=")) a->precision = v[0];
        else if (const char *v = val("--repeat=")) a->repeat = atoi(v);
        else if (const char *v = val("--warmup=")) a->warmup = atoi(v);
        else if (const char *v = val("--gpus=")) a->gpus = atoi(v);
        else if (s == "--gpus" && i + 1 < argc) a->gpus = atoi(argv[++i]);
        else if (const char *v = val("--kernel=")) {
            a->algo = !strcmp(v, "rowblock") ? RSG_SPMV_ROWBLOCK : !strcmp(v, "merge") ? RSG_SPMV_MERGE
                    : !strcmp(v, "vector") ? RSG_SPMV_VECTOR : !strcmp(v, "scalar") ? RSG_SPMV_SCALAR
                    : !strcmp(v, "tile") ? RSG_SPMV_TILE : !strcmp(v, "window") ? RSG_SPMV_WINDOW : RSG_SPMV_AUTO;
        } else if (s == "--stats") a->stats = true;
        else if (s == "--expand-symmetric") a->expand_symmetric = true;
    }
    if (a->precision != 'd' && a->precision != 's' && a->precision != 'm') a->precision = 'd';
    if (a->repeat < 1) a->repeat = 1;
    if (a->gpus < 1) a->gpus = 1;
    return true;
}

#define HARNESS_CUDA(expr)                                                                         \
    do {                                                                                           \
        cudaError_t e_ = (expr);                                                                   \
        if (e_ != cudaSuccess) fprintf(stderr, "CUDA failure %s:%d: '%s'\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
    } while (0)
#define HARNESS_RSG(expr)                                                                          \
    do {                                                                                           \
        int s_ = (expr);                                                                           \
        if (s_ != RSG_OK) fprintf(stderr, "libreaspasol_gpu failure %s:%d: '%s'\n", __FILE__, __LINE__, rsg_status_string(s_)); \
    } while (0)

template <typename T> T *to_device(const T *src, size_t n)
{
    T *d = nullptr;
    HARNESS_CUDA(cudaMalloc(&d, sizeof(T) * (n ? n : 1)));
    if (n) HARNESS_CUDA(cudaMemcpy(d, src, sizeof(T) * n, cudaMemcpyHostToDevice));
    return d;
}
