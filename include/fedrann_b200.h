/*
 * fedrann_b200.h -- C ABI of the B200-native FEDRANN hot path
 * (k-mer feature extraction -> sparse random projection -> exact cosine kNN).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch / C++ types.
 * One fr_ctx per GPU (per rank).  Device pointers are caller-owned (e.g.
 * torch.Tensor.data_ptr()); `stream` is a cudaStream_t passed as void*.
 * Every function returns FR_OK (0) or a negative FR_ERR_* code; the message of
 * the last error on a context is available through fr_last_error().  No C++
 * exception crosses this boundary.  Unless stated otherwise calls are
 * asynchronous on `stream`.
 *
 * Reference interfaces replaced (paths relative to the FEDRANN repository):
 *   kmer_searcher/kmer_searcher.cpp:232-375   the kmer_searcher executable
 *   fedrann/feature_extraction.py:167-292     get_feature_matrix (A_chunk.dot(P))
 *   fedrann/nearest_neighbors.py:22-55        NNDescent_ava.get_neighbors
 */
#ifndef FEDRANN_B200_H
#define FEDRANN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FR_OK 0
#define FR_ERR_CUDA (-1)      /* a CUDA runtime call failed                   */
#define FR_ERR_ARG (-2)       /* invalid argument                              */
#define FR_ERR_STATE (-3)     /* call order violated (e.g. fill before count)  */
#define FR_ERR_NOMEM (-4)     /* device allocation failed                      */
#define FR_ERR_NO_DEVICE (-5) /* no usable sm_100 device                       */
#define FR_ERR_IO (-6)        /* a file could not be opened / written           */
#define FR_ERR_BAD_ID (-7)    /* a read id holds a non-printable byte (the reference throws, kmer_searcher.cpp:112-116) */
#define FR_ERR_UNVERIFIED (-8) /* fr_table_set_sampler: the library is not such a sample (nothing changes) */

typedef struct fr_ctx fr_ctx;

int fr_version(void);
int fr_create(int device, fr_ctx **out);
void fr_destroy(fr_ctx *ctx);
const char *fr_last_error(const fr_ctx *ctx);
/* number of kernels this context has launched so far (bench.py: gpu_launches) */
int64_t fr_launch_count(const fr_ctx *ctx);

/* Per-kernel device timing (bench.py: roofline.achieved).  While enabled every
 * kernel launch of this context is bracketed by a CUDA event pair on its
 * stream.  fr_profile_read waits for the recorded events, writes one line
 * "<kernel> <launches> <total_ms>\n" per kernel into buf (NUL-terminated,
 * truncated to cap; *len_out = full length) and clears the records. */
int fr_profile_enable(fr_ctx *ctx, int on);
int fr_profile_read(fr_ctx *ctx, char *buf, int64_t cap, int64_t *len_out);

/* ---- host-side parsers (native equivalents of the reference's C++ I/O) ---- */

/* kmer_searcher.cpp:259-273.  Whitespace-separated tokens; wrong-length,
 * non-ACGT and duplicate tokens are skipped and consume no index.  Writes
 * index -> 2-bit code (A0 C1 G2 T3, first base most significant) to
 * codes_out (capacity cap); *n_out = number of unique k-mers (may exceed cap:
 * call again with a larger buffer). */
int fr_library_parse(const char *text, int64_t text_len, int k, uint64_t *codes_out,
                     int64_t cap, int64_t *n_out);

/* kmer_searcher.cpp:153-201.  FASTA/FASTQ text -> concatenated bases +
 * offsets + names.  Two-call protocol: with bases_out == NULL only the counts
 * are returned.  names_out receives the names separated by '\n'.  Large
 * inputs are parsed by several host threads (FASTA: ranges cut at header
 * lines; FASTQ: ranges at guessed record starts, used only when consistent
 * with the sequential reader); the records are always the sequential ones. */
int fr_fastx_parse(const char *text, int64_t text_len, uint8_t *bases_out, int64_t *offsets_out,
                   char *names_out, int64_t *n_reads, int64_t *n_bases, int64_t *names_len);

/* kmer_searcher.cpp:98-130 and 203-229: output.bin / kmer_frequency.bin
 * writers (rows in input order, indices ascending).  FR_ERR_IO when the file cannot be created,
 * FR_ERR_BAD_ID for a non-printable id byte; ids longer than 65535 bytes are truncated to
 * (length mod 65536) bytes exactly as the reference's uint16 cast does (:118-120). */
int fr_write_output_bin(const char *path, const char *names, const int64_t *name_offsets,
                        int64_t n_reads, const int64_t *indptr, const int32_t *indices);
int fr_write_frequency_bin(const char *path, const uint32_t *docfreq, int64_t n_features);

/* feature_extraction.py:108-132 (parse_kmer_searcher_output) and :295-302
 * (get_metadata): output.bin, held in memory, -> names + CSR.  Two-call
 * protocol: with indptr_out == NULL only n_records / nnz / names_len are
 * returned.  names_out receives the id bytes back to back, name_offsets_out
 * [n_records + 1] their bounds; indices keep the file's 64-bit values.
 * status_out: 0 ok, 1 header incomplete, 2 wrong magic, 3 unsupported version
 * (the version is returned in n_records), 4 truncated record -- the cases
 * the reference raises on. */
int fr_read_output_bin(const unsigned char *blob, int64_t len, char *names_out,
                       int64_t *name_offsets_out, int64_t *indptr_out, int64_t *indices_out,
                       int64_t *n_records, int64_t *nnz, int64_t *names_len, int *status_out);

/* ---- K2: k-mer library -> device lookup table (kmer_searcher.cpp:259-273) ---- */
/* d_codes: device u64[n], feature index -> k-mer code, all distinct. */
int fr_table_build(fr_ctx *ctx, const uint64_t *d_codes, int64_t n, int k, void *stream);

/* Optional, after fr_table_build: declare that the library is a hash-threshold sample -- every
 * key's canonical k-mer c (min of the k-mer and its reverse complement) satisfies
 * the sampling rule of csrc/common.cuh (`fr_sample_keep`: a 32-bit murmur-style mix of the
 * k-mer and the seed, compared with the top 32 bits of threshold; everything passes at
 * threshold = 2^64 - 1), which is how N1 (fr_kmer_count_*) samples.  The claim is
 * VERIFIED on the device for every key; FR_ERR_UNVERIFIED (and no change of behaviour) when it does not
 * hold.  When it holds, extraction rejects positions whose canonical k-mer fails the same test
 * in registers; results are identical either way. */
int fr_table_set_sampler(fr_ctx *ctx, uint64_t seed, uint64_t threshold, void *stream);

/* ---- K1+K3: reads -> CSR feature matrix (kmer_searcher.cpp:295-354) ---- */
/* d_bases: device u8 ASCII, 16-byte aligned, all reads concatenated;
 * d_offsets: device i64[n_reads+1].  Fills d_indptr (i64[n_reads+1]) and, when
 * non-NULL, ADDS per-feature document frequencies into d_docfreq (u32[F]).
 * Synchronises `stream` and returns the number of non-zeros in *nnz_out. */
int fr_extract_count(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                     int64_t n_reads, int64_t *d_indptr, uint32_t *d_docfreq, int64_t *nnz_out,
                     void *stream);
/* Writes the column indices (i32[nnz], ascending inside each row, rows in
 * input order) of the preceding fr_extract_count. */
int fr_extract_fill(fr_ctx *ctx, int32_t *d_indices, void *stream);
/* The same in ONE call that never synchronises (for callers that pipeline chunks of reads, e.g.
 * while later reads still cross PCIe): d_indices has room for `cap` entries; d_status is a device
 * i64[2] the caller zeroed before the first call.  On completion status[1] = number of non-zeros
 * and status[0] |= 1/2 (internal scratch too small) | 4 (more than cap non-zeros); when a bit is
 * set nothing was written and the caller repeats the chunk with fr_extract_count/fill. */
int fr_extract_async(fr_ctx *ctx, const uint8_t *d_bases, int64_t n_bases, const int64_t *d_offsets,
                     int64_t n_reads, int64_t *d_indptr, int32_t *d_indices, int64_t cap,
                     int64_t *d_status, void *stream);

/* ---- K4: E = A.P (feature_extraction.py:204-213, 136-140) ---- */
/* P in CSR (F = 2*nk rows, D columns): d_p_indptr i64[F+1], d_p_indices
 * i32[nnzP], d_p_data f32[nnzP].  d_E: f32[2*n_reads, D]; row 2r = forward,
 * row 2r+1 = index-swapped reverse orientation.  Accumulation is sequential
 * fp32 in ascending feature order (bit-exact with scipy's csr_matmat). */
int fr_project(fr_ctx *ctx, const int64_t *d_indptr, const int32_t *d_indices, int64_t n_reads,
               const int64_t *d_p_indptr, const int32_t *d_p_indices, const float *d_p_data,
               int64_t nk, int D, float *d_E, void *stream);

/* ---- K5: row normalisation for cosine kNN ---- */
/* d_En: f32[N, D] unit rows (zero rows stay zero); d_Eh: f16[N, Dp] copy
 * padded with zeros to Dp (multiple of 64, >= D); d_norms: f32[N]. */
int fr_normalize(fr_ctx *ctx, const float *d_E, int64_t N, int D, int Dp, float *d_En, void *d_Eh,
                 float *d_norms, void *stream);

/* ---- K6+K7: exact cosine kNN (nearest_neighbors.py:22-55) ---- */
/* Queries = database rows [q0, q0+nq); database = all n_db rows.  Output
 * d_idx i32[nq,k], d_dist f32[nq,k], ascending distance, ties by lower index.
 * mode 0 = tensor-core candidate pass + fp32 re-rank with certification
 * (uncertified rows fall back to the exact scan); mode 1 = exact fp32 scan. */
int fr_knn(fr_ctx *ctx, const float *d_En, const void *d_Eh, const float *d_norms, int64_t n_db,
           int D, int Dp, int64_t q0, int64_t nq, int k, int mode, int32_t *d_idx, float *d_dist,
           void *stream);
/* Optional, before fr_knn: the fp32 rows d_En are still being written by other work (e.g. an
 * all-gather on another stream) and are complete once `event` (a cudaEvent_t) has fired.  The
 * tensor-core pass reads only the fp16 copy and starts at once; the first consumer of d_En
 * waits for the event on fr_knn's stream.  One-shot: applies to the next fr_knn call only. */
int fr_knn_rows_ready_after(fr_ctx *ctx, void *event);
/* Incremental all-vs-all search (same result as fr_knn mode 0 with q0 = 0, nq = n_db): for callers whose
 * database rows become available in ranges -- e.g. run_fedrann_pipeline's reads still crossing PCIe
 * (fedrann/__main__.py:321-365 computes the whole matrix before get_neighbors, nearest_neighbors.py:22).
 * begin(): d_En / d_Eh / d_norms are the FULL-SIZE arrays ([n_db, D] fp32, [n_db, Dp] fp16, [n_db]); their rows
 * may still be unwritten.  bounds is a host array of n_phases + 1 cumulative row counts (bounds[0] = 0,
 * bounds[n_phases] = n_db).  phase(j), j = 0, 1, ... in order: rows [0, bounds[j+1]) of d_Eh and d_norms must
 * be complete on `stream`; it scores (new rows x all rows present) and (earlier rows x new rows).
 * finish(): needs every row of d_En; re-ranks, certifies and writes idx / dist [n_db, k].
 * FR_ERR_ARG from begin() when the shape is outside the tensor-core path (D > 512, k > 128, n_db < 2k):
 * use fr_knn then. */
int fr_knn_inc_begin(fr_ctx *ctx, const float *d_En, const void *d_Eh, const float *d_norms, int64_t n_db,
                     int D, int Dp, int k, int n_phases, const int64_t *bounds, void *stream);
int fr_knn_inc_phase(fr_ctx *ctx, int j, void *stream);
int fr_knn_inc_finish(fr_ctx *ctx, int32_t *d_idx, float *d_dist, void *stream);
/* statistics of the last fr_knn call: rows that needed the exact fallback */
int64_t fr_knn_last_fallback_rows(fr_ctx *ctx);
/* which path the last fr_knn / fr_knn_inc_begin took: 0 = tensor-core pass + fp32 re-rank, 1 = fp32 scan on the
 * CUDA cores (mode 1, or a shape outside the tensor-core path: D > 512, k > 128, n_db < 2k); -1 before any call */
int fr_knn_last_path(const fr_ctx *ctx);
/* Exact fp32 cosine kNN that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same result as fr_knn mode 1 (d_Eh is not needed): d_idx i32[nq,k], d_dist f32[nq,k],
 * ascending distance, ties by lower index; with fewer than k database rows the tail is idx -1 / dist +inf.
 * Its scratch -- one 64-bit key per (query, database row) -- is a file-scope static __device__ array of
 * 2 097 152 keys (16 MiB) that every GPU which loads the library holds until exit.  Limits, FR_ERR_ARG with a
 * message beyond them: nq * n_db <= 2 097 152, nq <= 4096, D <= 1024, 1 <= k <= 128.  The call makes no
 * cudaMalloc / cudaFree and uses no allocating library routine; the first call on a device resolves the array's
 * address (cudaGetSymbolAddress), so make one call before starting a graph capture.  All calls share the array:
 * nothing in it must start at zero (a call reads only the keys it wrote itself), but calls that overlap on one
 * GPU -- two streams, two threads -- race on it; serialise them.  Asynchronous on `stream`. */
int fr_knn_exact_noalloc(fr_ctx *ctx, const float *d_En, const float *d_norms, int64_t n_db, int D, int64_t q0,
                         int64_t nq, int k, int32_t *d_idx, float *d_dist, void *stream);

/* ---- multi-GPU: the embedding database exchanged through peer memory (NVLink), one rank per GPU ----
 * reads shard over the GPUs, each GPU searches its query rows against ALL rows.  The reference
 * is a single process (fedrann/feature_extraction.py:235-290 is its only parallel site), so there is no
 * interface to replace; a host program (C, or the Python host of this repo) drives it like this, once per rank:
 *   fr_exchange_create   allocates this rank's slab holding the WHOLE database (f32 rows [N, D], f16 rows
 *                        [N, Dp], norms [N]; N = row_bounds[world], rank r owns rows [row_bounds[r],
 *                        row_bounds[r+1])) and returns its CUDA IPC handle (FR_IPC_HANDLE_BYTES bytes);
 *   (the caller gives every rank every handle: any transport -- MPI, sockets, torch.distributed)
 *   fr_exchange_connect  maps the peers' slabs (handles of all `world` ranks, rank-major);
 *   per step:  write the local rows into the slab (fr_normalize with the fr_exchange_rows_* pointers offset to
 *              the local row range), then
 *     fr_exchange_pull     crosses a flag barrier over peer memory (every rank's rows are written) and queues
 *                          the peer-to-peer copies of every remote shard on a copy stream (copy engines: no SM
 *                          is taken from the search), fp16 rows first in ring order, each followed by a
 *                          "shard landed" flag, then norms + f32 rows;
 *     fr_knn               the NEXT fr_knn of the context on the slab's arrays consumes them as they land: the
 *                          tensor-core pass starts on the local shard and waits for a shard's flag only when
 *                          it reaches that shard's first tile; the re-rank waits for the f32 rows;
 *     fr_exchange_release  second barrier: no rank overwrites rows a peer may still be pulling.
 *   fr_exchange_destroy  after a barrier of the caller's own (peers must have released).
 * All calls but create / connect / destroy are asynchronous on `stream`. */
#define FR_MAX_RANKS 16
#define FR_IPC_HANDLE_BYTES 64
typedef struct fr_exchange fr_exchange;
int fr_exchange_create(fr_ctx *ctx, int rank, int world, const int64_t *row_bounds, int D, int Dp,
                       fr_exchange **out, unsigned char *handle_out);
int fr_exchange_connect(fr_exchange *x, const unsigned char *handles);
void *fr_exchange_rows_f32(fr_exchange *x);
void *fr_exchange_rows_f16(fr_exchange *x);
void *fr_exchange_norms(fr_exchange *x);
int fr_exchange_pull(fr_exchange *x, void *stream);
int fr_exchange_release(fr_exchange *x, void *stream);
void fr_exchange_destroy(fr_exchange *x);

/* ---- N1: canonical k-mer counting + seeded sampling ----
 * Replaces the pre-step of run_kmer_searcher (fedrann/count_kmers.py:76-123:
 * `jellyfish count -C`, `jellyfish dump -L <min-mult>`, the awk sampler).
 * A canonical k-mer c (min of the k-mer and its reverse complement, codes A0 C1
 * G2 T3) is sampled iff it passes `fr_sample_keep` (csrc/common.cuh: a 32-bit
 * murmur-style mix of the k-mer and the seed, compared with the top 32 bits of
 * threshold = fraction * 2^64; everything passes at 2^64 - 1); windows with a non-ACGT base or crossing a read boundary
 * are skipped.  begin() sizes and clears the count table (capacity = slots, a
 * power of two is chosen >= capacity), add() may be called once per batch of
 * reads and returns the number of distinct sampled k-mers so far (FR_ERR_NOMEM
 * when the table is full: begin again with a larger capacity), export() writes
 * the (code, count) pairs with count >= min_multiplicity in unspecified order;
 * *n_out may exceed cap (nothing is written beyond cap). */
int fr_kmer_count_begin(fr_ctx *ctx, int k, int64_t capacity, void *stream);
int fr_kmer_count_add(fr_ctx *ctx, const uint8_t *d_bases, const int64_t *d_offsets, int64_t n_reads,
                      uint64_t threshold, uint64_t seed, int64_t *distinct_out, void *stream);
int fr_kmer_count_export(fr_ctx *ctx, int min_multiplicity, uint64_t *d_codes_out, uint32_t *d_counts_out,
                         int64_t cap, int64_t *n_out, void *stream);

/* ---- N2: overlaps.tsv (fedrann/__main__.py:261-300 + DataFrame.to_csv :385) ----
 * One line per (row, neighbour != row) in row order then rank order:
 * query_name, query_orientation (+/- for even/odd row), target_name,
 * target_orientation, neighbor_rank (position in the unfiltered list), distance
 * (shortest decimal that round-trips the float32, as numpy/pandas print it).
 * names: read names of rows 0,2,4,... concatenated, name_offsets i64[n_rows/2+1].
 * Negative neighbour ids wrap like Python's read_names[-1].  append != 0 skips the
 * header and appends (rank-ordered multi-GPU output). */
int fr_write_overlaps_tsv(const char *path, int append, const char *names, const int64_t *name_offsets,
                          int64_t n_rows_total, int64_t row0, int64_t n_rows, int k, const int32_t *idx,
                          const float *dist, int64_t *lines_out);

#ifdef __cplusplus
}
#endif
#endif
