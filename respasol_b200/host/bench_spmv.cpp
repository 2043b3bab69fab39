// bench_spmv -- the reference's registered-benchmark sweep (/root/reference/bench_spmv.cc:51-55,113-117 over the 36
// matrix names of bench_consts.h:8-46) on the GPU path: BMMatrixDouble/<i> and BMMatrixSingle/<i>, i = 0..35 (the
// reference registers DenseRange(0,36), one index past the table -- not reproduced), each a loop of SpMV products
// with x from dlarnv (bench_spmv.cc:28-29), timed per product in milliseconds.
//
// No google/benchmark here (the vendored tree is not part of the path): the console table and the
// --benchmark_format=json / --benchmark_out=<file> output follow google-benchmark's schema (context + benchmarks[]
// with name, run_name, run_type, repetitions, iterations, real_time, cpu_time, time_unit), which is what
// benchmark/tools/compare.py reads -- so a run of this binary can be compared against a run of the reference's.
//
// Matrices: <mat-dir>/<group>/<name>.mtx is loaded when it exists (MAT_DIR = "../matrices/", bench_consts.h:6);
// otherwise -- no network, the SuiteSparse download scripts cannot run -- a synthetic stand-in of the same shape
// class is generated: rows and entries per row from the SuiteSparse index (approximate), structure class banded /
// 5-point / power-law.  --scale=F shrinks the stand-ins (tests).
//
// Flags: --benchmark_filter=<regex> --benchmark_format=console|json --benchmark_out=<file> --benchmark_list_tests
//        --mat-dir=<dir> --scale=<0..1] --iterations=N --warmup=N --cpu (time the OpenMP host product instead)
//        --kernel=auto|tile|merge|vector|rowblock|stream|pipe|scalar (RSG_OPT_SPMV_ALGO) --reduce=N (RSG_OPT_SPMV_REDUCE)
#include <chrono>
#include <cmath>
#include <regex>

#include "harness_common.h"

namespace {

struct MatSpec {
    const char *name;  // as in bench_consts.h
    char cls;          // 'B' banded(per_row, band), '5' five-point grid, 'P' power-law
    int rows;          // SuiteSparse index, approximate
    long long entries; // stored entries of the full (expanded) matrix, approximate
    int band;          // half bandwidth of the stand-in ('B')
};

// sizes: SuiteSparse Matrix Collection index (rounded); full (symmetric-expanded) entry counts
const MatSpec kMats[36] = {
    {"moderate/2cubes_sphere", 'B', 101492, 1647264, 3000},    {"moderate/ASIC_320ks", 'P', 321671, 1316085, 0},
    {"moderate/Baumann", 'B', 112211, 748331, 2500},           {"moderate/cfd2", 'B', 123440, 3085406, 2000},
    {"moderate/crashbasis", 'B', 160000, 1750416, 800},        {"moderate/dc1", 'P', 116835, 766396, 0},
    {"moderate/ct20stif", 'B', 52329, 2600295, 1500},          {"moderate/Dubcova3", 'B', 146689, 3636643, 4000},
    {"moderate/ecology2", '5', 999999, 4995991, 0},            {"moderate/FEM_3D_thermal2", 'B', 147900, 3489300, 2500},
    {"moderate/G2_circuit", '5', 150102, 726674, 0},           {"moderate/Goodwin_095", 'B', 100037, 3226066, 3000},
    {"moderate/matrix-new_3", 'B', 125329, 893984, 2000},      {"moderate/offshore", 'B', 259789, 4242673, 6000},
    {"moderate/para-10", 'B', 155924, 2094873, 3000},          {"moderate/parabolic_fem", 'B', 525825, 3674625, 1500},
    {"moderate/ss1", 'B', 205282, 845089, 1000},               {"moderate/stomach", 'B', 213360, 3021648, 4000},
    {"moderate/thermomech_TK", 'B', 102158, 711558, 2000},     {"moderate/tmt_unsym", '5', 917825, 4584801, 0},
    {"moderate/xenon2", 'B', 157464, 3866688, 3000},           {"big/af_shell10", 'B', 1508065, 52259885, 4000},
    {"big/af_shell2", 'B', 504855, 17562051, 3000},            {"big/atmosmodd", 'B', 1270432, 8814880, 11000},
    {"big/atmosmodl", 'B', 1489752, 10319760, 12000},          {"big/cage13", 'B', 445315, 7479343, 30000},
    {"big/CurlCurl_2", 'B', 806529, 8921789, 8000},            {"big/dielFilterV2real", 'B', 1157456, 48538952, 20000},
    {"big/Geo_1438", 'B', 1437960, 63156690, 12000},           {"big/Hook_1498", 'B', 1498023, 60917445, 12000},
    {"big/ML_Laplace", 'B', 377002, 27689972, 2000},           {"big/nlpkkt80", 'B', 1062400, 28704672, 15000},
    {"big/Serena", 'B', 1391349, 64531701, 12000},             {"big/Si87H76", 'B', 240369, 10661631, 20000},
    {"big/StocF-1465", 'B', 1465137, 21005389, 10000},         {"big/Transport", 'B', 1602111, 23500731, 10000},
};

struct Options {
    std::string filter = ".*", format = "console", out, mat_dir = "../matrices/";
    double scale = 1.0;
    int iterations = 50, warmup = 3;
    bool list = false, cpu = false;
    int algo = 0; // rsg_spmv_algo (--kernel=auto|tile|merge|vector|rowblock|stream|pipe)
    int reduce = 0; // RSG_OPT_SPMV_REDUCE (--reduce=lanes per row, 0 = heuristic)
};

struct Result {
    std::string name, source;
    int iterations = 0, rows = 0;
    long long stored = 0;
    double ms = 0; // per product
    bool error = false;
    std::string message;
};

bool load_or_generate(const Options &o, const MatSpec &s, CSR *A, std::string *source)
{
    const std::string path = o.mat_dir + s.name + ".mtx";
    if (FILE *f = fopen(path.c_str(), "rb")) {
        fclose(f);
        *source = path;
        return rsg_load_matrix_market(path.c_str(), A, 0, 0, RSG_LOADER_BUGCOMPAT) == 1;
    }
    const int rows = std::max(64, (int)std::lround(s.rows * o.scale));
    const int per_row = std::max(1, (int)std::lround((double)s.entries / s.rows));
    *source = "synthetic";
    if (s.cls == '5') {
        const int N = std::max(8, (int)std::lround(std::sqrt((double)rows)));
        return rsg_gen_laplace2d(N, A) == 1;
    }
    if (s.cls == 'P') // circuit matrices: short power-law rows plus a few dense rows / columns
        return rsg_gen_powerlaw(rows, 2.2, 64, 4, std::max(8, rows / 3), 4, A) == 1;
    return rsg_gen_banded(rows, per_row, std::max(per_row, (int)std::lround(s.band * std::sqrt(o.scale))), 7, A) == 1;
}

template <typename T> Result run_one(const Options &o, const std::string &bench, int index)
{
    Result r;
    r.name = bench + "/" + std::to_string(index);
    CSR A;
    if (!load_or_generate(o, kMats[index], &A, &r.source)) { r.error = true; r.message = "failed to load or generate the matrix"; return r; }
    const int m = A.m, n = A.n, stored = rsg_csr_stored_nnz(&A);
    r.rows = m; r.stored = stored;
    std::vector<double> xd((size_t)n);
    rsg_dlarnv_u01(n, xd.data()); // LAPACKE_dlarnv(1, {0,0,0,1}, ...), bench_spmv.cc:28-29
    std::vector<T> vals((size_t)stored), x((size_t)n), y((size_t)m);
    for (int i = 0; i < stored; ++i) vals[(size_t)i] = (T)A.values[i];
    for (int i = 0; i < n; ++i) x[(size_t)i] = (T)xd[(size_t)i];
    r.iterations = o.iterations;
    if (o.cpu) {
        auto product = [&]() {
#pragma omp parallel for schedule(static)
            for (int i = 0; i < m; ++i) {
                T s = 0;
                for (int k = A.rowptr[i]; k < A.rowptr[i + 1]; ++k) s += vals[(size_t)k] * x[(size_t)A.colidx[k]];
                y[(size_t)i] = s;
            }
        };
        for (int i = 0; i < o.warmup; ++i) product();
        auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < o.iterations; ++i) product();
        r.ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / o.iterations;
    } else {
        int *dRow = to_device(A.rowptr, (size_t)m + 1), *dCol = to_device(A.colidx, (size_t)stored);
        T *dVal = to_device(vals.data(), (size_t)stored), *dX = to_device(x.data(), (size_t)n), *dY = to_device(y.data(), (size_t)m);
        rsg_handle_t h = nullptr;
        int st = rsg_create(&h, 0);
        if (st == RSG_OK) st = rsg_set_option(h, RSG_OPT_SPMV_ALGO, o.algo);
        if (st == RSG_OK) st = rsg_set_option(h, RSG_OPT_SPMV_REDUCE, o.reduce);
        if (st == RSG_OK) st = rsg_spmv_analysis(h, m, n, stored, 0, dRow, dCol);
        auto product = [&]() -> int {
            if constexpr (sizeof(T) == 8) return spmv_csr_d(h, m, n, stored, 0, dRow, dCol, (const double *)dVal, 1.0, (const double *)dX, 0.0, (double *)dY);
            else return spmv_csr_s(h, m, n, stored, 0, dRow, dCol, (const float *)dVal, 1.0f, (const float *)dX, 0.0f, (float *)dY);
        };
        for (int i = 0; i < o.warmup && st == RSG_OK; ++i) st = product();
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        for (int i = 0; i < o.iterations && st == RSG_OK; ++i) st = product();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        r.ms = ms / o.iterations;
        if (st != RSG_OK || cudaGetLastError() != cudaSuccess) { r.error = true; r.message = rsg_status_string(st); }
        cudaEventDestroy(e0); cudaEventDestroy(e1);
        if (h) rsg_destroy(h);
        cudaFree(dRow); cudaFree(dCol); cudaFree(dVal); cudaFree(dX); cudaFree(dY);
    }
    rsg_free_csr(&A);
    return r;
}

void write_json(FILE *f, const Options &o, const std::vector<Result> &rs, const char *exe)
{
    int dev = 0;
    cudaDeviceProp prop{};
    const bool gpu = cudaGetDeviceProperties(&prop, dev) == cudaSuccess;
    fprintf(f, "{\n  \"context\": {\n    \"executable\": \"%s\",\n    \"library\": \"libreaspasol_gpu %d\",\n", exe, rsg_version());
    fprintf(f, "    \"device\": \"%s\",\n    \"path\": \"%s\",\n    \"scale\": %g,\n    \"library_build_type\": \"release\"\n  },\n",
            o.cpu ? "host (OpenMP)" : (gpu ? prop.name : "none"), o.cpu ? "cpu" : "gpu", o.scale);
    fprintf(f, "  \"benchmarks\": [\n");
    for (size_t i = 0; i < rs.size(); ++i) {
        const Result &r = rs[i];
        fprintf(f, "    {\n      \"name\": \"%s\",\n      \"run_name\": \"%s\",\n      \"run_type\": \"iteration\",\n", r.name.c_str(), r.name.c_str());
        fprintf(f, "      \"repetitions\": 1,\n      \"repetition_index\": 0,\n      \"threads\": 1,\n      \"iterations\": %d,\n", r.iterations);
        if (r.error) fprintf(f, "      \"error_occurred\": true,\n      \"error_message\": \"%s\",\n", r.message.c_str());
        fprintf(f, "      \"real_time\": %.9e,\n      \"cpu_time\": %.9e,\n      \"time_unit\": \"ms\",\n", r.ms, r.ms);
        fprintf(f, "      \"rows\": %d,\n      \"stored\": %lld,\n      \"matrix\": \"%s\",\n      \"source\": \"%s\"\n    }%s\n", r.rows, r.stored,
                kMats[atoi(r.name.c_str() + r.name.find('/') + 1)].name, r.source.c_str(), i + 1 < rs.size() ? "," : "");
    }
    fprintf(f, "  ]\n}\n");
}

} // namespace

int main(int argc, char **argv)
{
    Options o;
    for (int i = 1; i < argc; ++i) {
        std::string s(argv[i]);
        auto val = [&](const char *key) -> const char * {
            size_t n = strlen(key);
            return s.compare(0, n, key) == 0 ? s.c_str() + n : nullptr;
        };
        if (const char *v = val("--benchmark_filter=")) o.filter = v;
        else if (const char *v = val("--benchmark_format=")) o.format = v;
        else if (const char *v = val("--benchmark_out=")) o.out = v;
        else if (const char *v = val("--mat-dir=")) { o.mat_dir = v; if (!o.mat_dir.empty() && o.mat_dir.back() != '/') o.mat_dir += '/'; }
        else if (const char *v = val("--scale=")) o.scale = atof(v);
        else if (const char *v = val("--iterations=")) o.iterations = atoi(v);
        else if (const char *v = val("--warmup=")) o.warmup = atoi(v);
        else if (s == "--benchmark_list_tests" || s == "--benchmark_list_tests=true") o.list = true;
        else if (s == "--cpu") o.cpu = true;
        else if (const char *v = val("--reduce=")) o.reduce = atoi(v);
        else if (const char *v = val("--kernel=")) {
            o.algo = !strcmp(v, "tile") ? RSG_SPMV_TILE : !strcmp(v, "merge") ? RSG_SPMV_MERGE : !strcmp(v, "vector") ? RSG_SPMV_VECTOR
                   : !strcmp(v, "rowblock") ? RSG_SPMV_ROWBLOCK : !strcmp(v, "stream") ? RSG_SPMV_STREAM : !strcmp(v, "pipe") ? RSG_SPMV_PIPE
                   : !strcmp(v, "scalar") ? RSG_SPMV_SCALAR : RSG_SPMV_AUTO;
        }
        else if (val("--benchmark_")) { /* other google-benchmark flags are accepted and ignored */ }
        else { fprintf(stderr, "%s: unknown flag %s\n", argv[0], argv[i]); return 1; }
    }
    if (!(o.scale > 0.0) || o.scale > 1.0) o.scale = 1.0;
    if (o.iterations < 1) o.iterations = 1;
    std::regex re;
    try { re = std::regex(o.filter); } catch (const std::regex_error &) { fprintf(stderr, "bad --benchmark_filter\n"); return 1; }
    std::vector<Result> results;
    const bool console = o.format != "json";
    if (console && !o.list) {
        printf("%-28s %14s %14s %12s\n", "Benchmark", "Time", "CPU", "Iterations");
        printf("------------------------------------------------------------------------\n");
    }
    for (int pass = 0; pass < 2; ++pass) {
        const std::string bench = pass == 0 ? "BMMatrixDouble" : "BMMatrixSingle";
        for (int i = 0; i < 36; ++i) {
            const std::string name = bench + "/" + std::to_string(i);
            if (!std::regex_search(name, re)) continue;
            if (o.list) { printf("%s\n", name.c_str()); continue; }
            Result r = pass == 0 ? run_one<double>(o, bench, i) : run_one<float>(o, bench, i);
            if (console) {
                if (r.error) printf("%-28s ERROR OCCURRED: '%s'\n", r.name.c_str(), r.message.c_str());
                else printf("%-28s %11.4f ms %11.4f ms %12d  %s rows=%d stored=%lld (%s)\n", r.name.c_str(), r.ms, r.ms, r.iterations,
                            kMats[i].name, r.rows, r.stored, r.source.c_str());
                fflush(stdout);
            }
            results.push_back(r);
        }
    }
    if (o.list) return 0;
    if (!console) write_json(stdout, o, results, argv[0]);
    if (!o.out.empty()) {
        FILE *f = fopen(o.out.c_str(), "w");
        if (!f) { fprintf(stderr, "cannot write %s\n", o.out.c_str()); return 1; }
        write_json(f, o, results, argv[0]);
        fclose(f);
    }
    return 0;
}
