// kmer_searcher -- drop-in replacement for the reference executable
// (kmer_searcher/kmer_searcher.cpp:232-375):
//     kmer_searcher <kmer_lib> <input_fastx> <output_dir> <k> <threads>
// Either path may be /dev/stdin.  Writes <output_dir>/output.bin and
// <output_dir>/kmer_frequency.bin in the reference's formats (rows in input
// order, indices ascending).  All lookups run on the GPU through the C ABI of
// libfedrann_b200.so; `threads` is validated as the reference does and
// otherwise unused.  There is no CPU fallback: without an sm_100 GPU the tool
// fails with exit code 1.
#include <cuda_runtime.h>
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/fedrann_b200.h"

static bool slurp(const char *path, std::string &out) {
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    char buf[1 << 16];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, n);
    fclose(f);
    return true;
}

static int parse_int(const char *s, bool &ok) {
    char *end = nullptr;
    long v = strtol(s, &end, 10);
    ok = end != s;
    return (int)v;
}

#define CK(call)                                                                       \
    do {                                                                               \
        int _rc = (call);                                                              \
        if (_rc != FR_OK) {                                                            \
            fprintf(stderr, "kmer_searcher: %s failed (%d): %s\n", #call, _rc, fr_last_error(ctx)); \
            return 1;                                                                  \
        }                                                                              \
    } while (0)
#define CU(call)                                                                       \
    do {                                                                               \
        cudaError_t _e = (call);                                                       \
        if (_e != cudaSuccess) {                                                       \
            fprintf(stderr, "kmer_searcher: %s: %s\n", #call, cudaGetErrorString(_e)); \
            return 1;                                                                  \
        }                                                                              \
    } while (0)

int main(int argc, char *argv[]) {
    if (argc != 6) {
        fprintf(stderr, "Usage: %s <kmer_lib> <input_fastq> <output_dir> <k> <threads>\n", argv[0]);
        return 1;
    }
    bool ok = false;
    const int K = parse_int(argv[4], ok);
    if (!ok || K <= 0 || K > 31) { fprintf(stderr, "Invalid k value: %s\n", argv[4]); return 1; }
    const int threads = parse_int(argv[5], ok);
    if (!ok || threads <= 0) { fprintf(stderr, "Invalid number of threads: %s\n", argv[5]); return 1; }
    const std::string out_dir = argv[3];
    struct stat sb;
    if (stat(out_dir.c_str(), &sb) != 0 && mkdir(out_dir.c_str(), 0777) != 0) {
        fprintf(stderr, "Cannot create output directory: %s\n", out_dir.c_str());
        return 1;
    }
    std::string lib_text, reads_text;
    slurp(argv[1], lib_text);   // a missing library file is an empty library (kmer_searcher.cpp:262)
    if (!slurp(argv[2], reads_text)) {
        fprintf(stderr, "Error reading input: Cannot open input file: %s\n", argv[2]);
        return 1;
    }
    // host-side parsing
    int64_t n_codes = 0;
    std::vector<uint64_t> codes(lib_text.size() / 2 + 1);
    if (fr_library_parse(lib_text.data(), (int64_t)lib_text.size(), K, codes.data(), (int64_t)codes.size(), &n_codes))
        return 1;
    int64_t R = 0, B = 0, NL = 0;
    if (fr_fastx_parse(reads_text.data(), (int64_t)reads_text.size(), nullptr, nullptr, nullptr, &R, &B, &NL)) return 1;
    std::vector<uint8_t> bases((size_t)B + 16);
    std::vector<int64_t> offsets((size_t)R + 1);
    std::vector<char> names_nl((size_t)NL + 1);
    if (fr_fastx_parse(reads_text.data(), (int64_t)reads_text.size(), bases.data(), offsets.data(), names_nl.data(), &R,
                       &B, &NL))
        return 1;
    std::string names;
    std::vector<int64_t> name_off((size_t)R + 1, 0);
    {
        int64_t r = 0;
        for (int64_t i = 0; i < NL; ++i) {
            if (names_nl[(size_t)i] == '\n') name_off[(size_t)++r] = (int64_t)names.size();
            else names.push_back(names_nl[(size_t)i]);
        }
    }
    // device
    fr_ctx *ctx = nullptr;
    if (fr_create(0, &ctx) != FR_OK) {
        fprintf(stderr, "kmer_searcher: no sm_100 GPU available (this build has no CPU path)\n");
        return 1;
    }
    uint64_t *d_codes = nullptr;
    uint8_t *d_bases = nullptr;
    int64_t *d_offsets = nullptr, *d_indptr = nullptr;
    uint32_t *d_docfreq = nullptr;
    int32_t *d_indices = nullptr;
    CU(cudaMalloc(&d_codes, (size_t)(n_codes > 0 ? n_codes : 1) * 8));
    CU(cudaMalloc(&d_bases, (size_t)B + 16));
    CU(cudaMalloc(&d_offsets, (size_t)(R + 1) * 8));
    CU(cudaMalloc(&d_indptr, (size_t)(R + 1) * 8));
    CU(cudaMalloc(&d_docfreq, (size_t)(n_codes > 0 ? n_codes : 1) * 4));
    CU(cudaMemcpy(d_codes, codes.data(), (size_t)n_codes * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_bases, bases.data(), (size_t)B, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_offsets, offsets.data(), (size_t)(R + 1) * 8, cudaMemcpyHostToDevice));
    CU(cudaMemset(d_docfreq, 0, (size_t)(n_codes > 0 ? n_codes : 1) * 4));
    CK(fr_table_build(ctx, d_codes, n_codes, K, nullptr));
    int64_t nnz = 0;
    CK(fr_extract_count(ctx, d_bases, B, d_offsets, R, d_indptr, d_docfreq, &nnz, nullptr));
    CU(cudaMalloc(&d_indices, (size_t)(nnz > 0 ? nnz : 1) * 4));
    CK(fr_extract_fill(ctx, d_indices, nullptr));
    CU(cudaDeviceSynchronize());
    std::vector<int64_t> indptr((size_t)R + 1);
    std::vector<int32_t> indices((size_t)(nnz > 0 ? nnz : 1));
    std::vector<uint32_t> docfreq((size_t)(n_codes > 0 ? n_codes : 1));
    CU(cudaMemcpy(indptr.data(), d_indptr, (size_t)(R + 1) * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(indices.data(), d_indices, (size_t)nnz * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(docfreq.data(), d_docfreq, (size_t)n_codes * 4, cudaMemcpyDeviceToHost));
    if (fr_write_frequency_bin((out_dir + "/kmer_frequency.bin").c_str(), docfreq.data(), n_codes) != FR_OK) return 1;
    const int rc = fr_write_output_bin((out_dir + "/output.bin").c_str(), names.empty() ? "" : names.data(),
                                       name_off.data(), R, indptr.data(), indices.data());
    if (rc == FR_ERR_BAD_ID) { fprintf(stderr, "terminate: ID contains non-ASCII characters\n"); return 134; }
    if (rc == FR_ERR_IO) fprintf(stderr, "Failed to open output file: %s/output.bin\n", out_dir.c_str());
    fr_destroy(ctx);
    return rc == FR_OK ? 0 : 1;
}
