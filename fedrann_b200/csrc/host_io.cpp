// Native host-side I/O of the path: the k-mer library parser, the FASTA/FASTQ
// reader and the output.bin / kmer_frequency.bin writers.  These restate the
// host parts of kmer_searcher.cpp (library load :259-273, read_sequences
// :153-201, write_to_file :98-130, writeKmerStatsToBinary :203-229) so that the
// replacement executable and the Python host share one implementation.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_set>
#include <thread>
#include <vector>

#include "../../include/fedrann_b200.h"

namespace {
inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }

inline int base_code(char c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return -1;
    }
}
}  // namespace

extern "C" int fr_library_parse(const char *text, int64_t text_len, int k, uint64_t *codes_out, int64_t cap,
                                int64_t *n_out) {
    if (!text || text_len < 0 || k < 1 || k > 31 || !n_out) return FR_ERR_ARG;
    std::unordered_set<uint64_t> seen;
    seen.reserve((size_t)(text_len / (k + 1)) + 16);
    int64_t n = 0, i = 0;
    while (i < text_len) {
        while (i < text_len && is_space(text[i])) ++i;
        if (i >= text_len) break;
        const int64_t s = i;
        while (i < text_len && !is_space(text[i])) ++i;
        if (i - s != k) continue;                      // :266 wrong length
        uint64_t code = 0;
        bool ok = true;
        for (int64_t j = s; j < i; ++j) {
            int c = base_code(text[j]);
            if (c < 0) { ok = false; break; }          // :147 non-ACGT
            code = (code << 2) | (uint64_t)c;
        }
        if (!ok) continue;
        if (!seen.insert(code).second) continue;       // :268 duplicate
        if (codes_out && n < cap) codes_out[n] = code;
        ++n;
    }
    *n_out = n;
    return FR_OK;
}

namespace {

struct FastxCount { int64_t R = 0, B = 0, NL = 0; };

// The record state machine of read_sequences (kmer_searcher.cpp:153-201) over text[beg, end).
// For FASTA a range that starts at a header line is independent of what precedes it, so large
// inputs are parsed by several threads (ranges cut at "\n>"); with emit set the outputs are
// written at the given record / base / name offsets.
// A FASTQ record that starts inside the range takes its three following lines even where they lie past
// `end` (lines are read up to text_end); *stop = where the state machine stands, between records, when it
// leaves the range -- the next range is independent of this one exactly when it starts there.
FastxCount fastx_range(const char *text, int64_t beg, int64_t end, int64_t text_end, bool is_fastq, bool emit,
                       uint8_t *bases_out, int64_t *offsets_out, char *names_out, FastxCount at, int64_t *stop) {
    int64_t R = at.R, B = at.B, NL = at.NL;
    bool have_rec = false;  // FASTA: a record with a non-empty id is open
    int64_t pos = beg;
    auto next_line = [&](int64_t &ls, int64_t &le) -> bool {
        if (pos >= text_end) return false;
        ls = pos;
        const void *nl = memchr(text + pos, '\n', (size_t)(text_end - pos));
        le = nl ? (const char *)nl - text : text_end;
        pos = nl ? le + 1 : text_end;
        return true;
    };
    auto open_rec = [&](const char *id, int64_t idlen) {
        if (emit) {
            offsets_out[R] = B;
            memcpy(names_out + NL, id, (size_t)idlen);
            names_out[NL + idlen] = '\n';
        }
        NL += idlen + 1;
    };
    int64_t ls, le;
    while (pos < end && next_line(ls, le)) {
        if (le == ls) continue;                                                     // :168
        if (!is_fastq) {
            if (text[ls] == '>') {
                if (have_rec) { ++R; have_rec = false; }
                int64_t idend = ls + 1;
                while (idend < le && text[idend] != ' ' && text[idend] != '\t') ++idend;   // :176-181
                if (idend > ls + 1) { open_rec(text + ls + 1, idend - ls - 1); have_rec = true; }
            } else if (have_rec) {
                if (emit) memcpy(bases_out + B, text + ls, (size_t)(le - ls));      // :185
                B += le - ls;
            }
        } else if (text[ls] == '@') {                                               // :188-194
            open_rec(text + ls + 1, le - ls - 1);
            int64_t ss, se;
            if (next_line(ss, se)) {
                if (emit) memcpy(bases_out + B, text + ss, (size_t)(se - ss));
                B += se - ss;
            }
            int64_t a, b;
            next_line(a, b);
            next_line(a, b);
            ++R;
        }
    }
    if (!is_fastq && have_rec) ++R;
    if (stop) *stop = pos;
    FastxCount out;
    out.R = R - at.R; out.B = B - at.B; out.NL = NL - at.NL;
    return out;
}

}  // namespace

extern "C" int fr_fastx_parse(const char *text, int64_t text_len, uint8_t *bases_out, int64_t *offsets_out,
                              char *names_out, int64_t *n_reads, int64_t *n_bases, int64_t *names_len) {
    if (!text || text_len < 0 || !n_reads || !n_bases || !names_len) return FR_ERR_ARG;
    const bool emit = bases_out != nullptr;
    // the format is decided by the first line (:163-166)
    const bool is_fastq = text_len > 0 && text[0] == '@';
    // FASTA ranges: cut at header lines, one range per thread (FR_PARSE_THREADS / FR_PARSE_MIN_RANGE: tests)
    int64_t min_range = 8 << 20;
    int threads = (int)std::thread::hardware_concurrency();
    if (threads > 16) threads = 16;
    if (const char *e = getenv("FR_PARSE_THREADS")) threads = atoi(e);
    if (const char *e = getenv("FR_PARSE_MIN_RANGE")) min_range = atoll(e);
    if (min_range < 1) min_range = 1;
    if (threads < 1) threads = 1;
    if ((int64_t)threads > text_len / min_range) threads = (int)(text_len / min_range);
    if (threads < 1) threads = 1;
    // line [a, b) starting at p (b excludes the newline); false at the end of the text
    auto line_at = [&](int64_t p, int64_t &b, int64_t &next) -> bool {
        if (p >= text_len) return false;
        const void *nl = memchr(text + p, '\n', (size_t)(text_len - p));
        b = nl ? (const char *)nl - text : text_len;
        next = nl ? b + 1 : text_len;
        return true;
    };
    std::vector<int64_t> cut(1, 0);
    for (int t = 1; t < threads; ++t) {
        int64_t p = text_len / threads * t;
        if (p < cut.back()) p = cut.back();
        int64_t c = -1;
        if (!is_fastq) {
            while (p < text_len) {                       // next "\n>"
                const void *nl = memchr(text + p, '\n', (size_t)(text_len - p));
                if (!nl) break;
                p = (const char *)nl - text + 1;
                if (p < text_len && text[p] == '>') { c = p; break; }
            }
        } else {
            // FASTQ: a line start that looks like a record start -- '@' line, a non-empty line, a '+' line and
            // a line as long as the second (a quality line that begins with '@' fails the '+' test).  It is a
            // guess; the ranges are only used when every range ends where the next one starts (below).
            const void *nl = memchr(text + p, '\n', (size_t)(text_len - p));
            int64_t q = nl ? (const char *)nl - text + 1 : text_len;
            for (int tries = 0; tries < 64 && q < text_len && c < 0; ++tries) {
                int64_t b0, n0, b1, n1, b2, n2, b3, n3;
                if (!line_at(q, b0, n0)) break;
                if (text[q] == '@' && line_at(n0, b1, n1) && b1 > n0 && line_at(n1, b2, n2) && b2 > n1 &&
                    text[n1] == '+' && line_at(n2, b3, n3) && b3 - n2 == b1 - n0)
                    c = q;
                q = n0;
            }
        }
        if (c < 0) { if (is_fastq) continue; else break; }
        if (c > cut.back()) cut.push_back(c);
    }
    cut.push_back(text_len);
    int nr = (int)cut.size() - 1;
    std::vector<FastxCount> cnt(nr), at(nr + 1);
    std::vector<int64_t> stop(nr, 0);
    auto run = [&](bool do_emit) {
        std::vector<std::thread> th;
        for (int t = 1; t < nr; ++t)
            th.emplace_back([&, t] {
                cnt[t] = fastx_range(text, cut[t], cut[t + 1], text_len, is_fastq, do_emit, bases_out, offsets_out,
                                     names_out, at[t], &stop[t]);
            });
        cnt[0] = fastx_range(text, cut[0], cut[1], text_len, is_fastq, do_emit, bases_out, offsets_out, names_out,
                             at[0], &stop[0]);
        for (auto &x : th) x.join();
    };
    run(false);
    if (is_fastq) {
        // range t + 1 was parsed as if a record could start at cut[t + 1]; that is what the sequential reader
        // does only if range t stops exactly there.  Otherwise (odd input): one range, sequentially.
        bool ok = true;
        for (int t = 0; t + 1 < nr; ++t) ok = ok && stop[t] == cut[t + 1];
        if (!ok) {
            cut.assign({0, text_len});
            nr = 1;
            run(false);
        }
    }
    for (int t = 0; t < nr; ++t) {
        at[t + 1].R = at[t].R + cnt[t].R; at[t + 1].B = at[t].B + cnt[t].B; at[t + 1].NL = at[t].NL + cnt[t].NL;
    }
    if (emit) {
        run(true);
        offsets_out[0] = 0;
        offsets_out[at[nr].R] = at[nr].B;
    }
    *n_reads = at[nr].R; *n_bases = at[nr].B; *names_len = at[nr].NL;
    return FR_OK;
}

extern "C" int fr_write_output_bin(const char *path, const char *names, const int64_t *name_offsets, int64_t n_reads,
                                   const int64_t *indptr, const int32_t *indices) {
    if (!path || !name_offsets || !indptr || n_reads < 0) return FR_ERR_ARG;
    FILE *f = fopen(path, "wb");
    if (!f) return FR_ERR_IO;
    std::vector<char> buf;
    buf.reserve(1 << 22);
    auto put = [&](const void *p, size_t n) {
        const char *c = (const char *)p;
        buf.insert(buf.end(), c, c + n);
        if (buf.size() >= (1u << 22)) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    };
    const char magic[8] = {'K', 'M', 'E', 'R', 1, 0, 0, 0};
    put(magic, 8);
    uint64_t total = (uint64_t)n_reads;
    put(&total, 8);
    int rc = FR_OK;
    for (int64_t r = 0; r < n_reads && rc == FR_OK; ++r) {
        const char *id = names + name_offsets[r];
        const int64_t idlen = name_offsets[r + 1] - name_offsets[r];
        for (int64_t j = 0; j < idlen; ++j)
            if (id[j] < 32 || id[j] > 126) rc = FR_ERR_BAD_ID;    // :112-116 (the reference throws)
        // :118-120: the length is stored in 16 bits and exactly that many id bytes follow (an id longer
        // than 65535 bytes is truncated consistently, as the reference does)
        uint16_t l16 = (uint16_t)idlen;
        put(&l16, 2);
        put(id, (size_t)l16);
        uint32_t cnt = (uint32_t)(indptr[r + 1] - indptr[r]);
        put(&cnt, 4);
        // the row's indices widened to u64 in place at the end of the buffer (one resize per row, not one
        // insert per index: the reference's per-index ofstream::write, :125-128, is what made this file slow)
        const size_t at = buf.size(), nb = 8 * (size_t)cnt;
        buf.resize(at + nb);
        char *dst = buf.data() + at;
        for (int64_t t = indptr[r]; t < indptr[r + 1]; ++t, dst += 8) {
            const uint64_t v = (uint64_t)(uint32_t)indices[t];
            memcpy(dst, &v, 8);
        }
        if (buf.size() >= (1u << 22)) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    }
    if (!buf.empty()) fwrite(buf.data(), 1, buf.size(), f);
    fclose(f);
    return rc;
}

// feature_extraction.py:108-132, the reader of output.bin: one sequential pass (records are variable-length).
// Two-call protocol like fr_fastx_parse: with indptr_out == NULL only the sizes are returned.  status_out:
// 0 ok, 1 header shorter than 16 bytes, 2 wrong magic, 3 wrong version (returned in *n_records), 4 truncated record.
extern "C" int fr_read_output_bin(const unsigned char *blob, int64_t len, char *names_out, int64_t *name_offsets_out,
                                  int64_t *indptr_out, int64_t *indices_out, int64_t *n_records, int64_t *nnz,
                                  int64_t *names_len, int *status_out) {
    if (!blob || len < 0 || !n_records || !nnz || !names_len || !status_out) return FR_ERR_ARG;
    *status_out = 0; *n_records = 0; *nnz = 0; *names_len = 0;
    if (len < 16) { *status_out = 1; return FR_OK; }
    if (memcmp(blob, "KMER", 4) != 0) { *status_out = 2; return FR_OK; }
    if (blob[4] != 1) { *status_out = 3; *n_records = blob[4]; return FR_OK; }
    uint64_t total;
    memcpy(&total, blob + 8, 8);
    const bool emit = indptr_out != nullptr;
    int64_t pos = 16, nz = 0, nl = 0;
    if (emit) { indptr_out[0] = 0; name_offsets_out[0] = 0; }
    for (uint64_t r = 0; r < total; ++r) {
        if (pos + 2 > len) { *status_out = 4; return FR_OK; }
        uint16_t idlen;
        memcpy(&idlen, blob + pos, 2);
        pos += 2;
        if (pos + (int64_t)idlen + 4 > len) { *status_out = 4; return FR_OK; }
        if (emit) memcpy(names_out + nl, blob + pos, idlen);
        nl += idlen;
        pos += idlen;
        uint32_t cnt;
        memcpy(&cnt, blob + pos, 4);
        pos += 4;
        if (pos + 8 * (int64_t)cnt > len) { *status_out = 4; return FR_OK; }
        if (emit) {
            memcpy(indices_out + nz, blob + pos, 8 * (size_t)cnt);      // little-endian u64, as stored
            indptr_out[r + 1] = nz + cnt;
            name_offsets_out[r + 1] = nl;
        }
        nz += cnt;
        pos += 8 * (int64_t)cnt;
    }
    *n_records = (int64_t)total; *nnz = nz; *names_len = nl;
    return FR_OK;
}

extern "C" int fr_write_frequency_bin(const char *path, const uint32_t *docfreq, int64_t n_features) {
    FILE *f = fopen(path, "wb");
    if (!f) return FR_ERR_IO;
    for (int64_t i = 0; i < n_features; ++i) {
        if (!docfreq[i]) continue;                                   // :217
        uint64_t rec[2] = {(uint64_t)i, (uint64_t)docfreq[i]};
        fwrite(rec, 8, 2, f);
    }
    fclose(f);
    return FR_OK;
}

// ---- N2: overlaps.tsv ----
// get_output_dataframe (fedrann/__main__.py:261-300) + DataFrame.to_csv(sep="\t", index=False)
// (:385) restated as one streaming writer.  A float32 distance is printed the way numpy/pandas
// print it: the shortest decimal that round-trips, positional for 1e-4 <= |x| < 1e16 (with at
// least one digit after the point), scientific (d.ddde-XX) otherwise.
#include <charconv>

namespace {
inline int fmt_f32(float x, char *out) {
    if (x != x) { memcpy(out, "nan", 3); return 3; }     // pandas' na_rep is "" -- never produced by the path
    const float ax = x < 0 ? -x : x;
    char *p = out;
    if (ax != 0.0f && ((double)ax < 1e-4 || (double)ax >= 1e16)) {
        auto r = std::to_chars(p, p + 48, x, std::chars_format::scientific);
        return (int)(r.ptr - out);
    }
    auto r = std::to_chars(p, p + 48, x, std::chars_format::fixed);
    bool dot = false;
    for (char *q = p; q < r.ptr; ++q) dot |= (*q == '.');
    if (!dot) { *r.ptr++ = '.'; *r.ptr++ = '0'; }
    return (int)(r.ptr - out);
}
// csv.QUOTE_MINIMAL as pandas applies it: a field holding the delimiter, a quote or a line break
// is wrapped in double quotes, embedded quotes doubled
inline void put_field(std::vector<char> &buf, const char *s, size_t n) {
    bool q = false;
    for (size_t i = 0; i < n; ++i) q |= (s[i] == '\t' || s[i] == '"' || s[i] == '\n' || s[i] == '\r');
    if (!q) { buf.insert(buf.end(), s, s + n); return; }
    buf.push_back('"');
    for (size_t i = 0; i < n; ++i) { if (s[i] == '"') buf.push_back('"'); buf.push_back(s[i]); }
    buf.push_back('"');
}
}  // namespace

extern "C" int fr_write_overlaps_tsv(const char *path, int append, const char *names, const int64_t *name_offsets,
                                     int64_t n_rows_total, int64_t row0, int64_t n_rows, int k, const int32_t *idx,
                                     const float *dist, int64_t *lines_out) {
    if (!path || !names || !name_offsets || n_rows < 0 || k < 0 || (n_rows > 0 && (!idx || !dist))) return FR_ERR_ARG;
    FILE *f = fopen(path, append ? "ab" : "wb");
    if (!f) return FR_ERR_IO;
    std::vector<char> buf;
    buf.reserve((1u << 22) + 4096);
    if (!append) {
        static const char hdr[] = "query_name\tquery_orientation\ttarget_name\ttarget_orientation\tneighbor_rank\tdistance\n";
        buf.insert(buf.end(), hdr, hdr + sizeof(hdr) - 1);
    }
    int64_t lines = 0;
    char num[64];
    for (int64_t i = 0; i < n_rows; ++i) {
        const int64_t q = row0 + i;
        const char *qn = names + name_offsets[q >> 1];
        const size_t ql = (size_t)(name_offsets[(q >> 1) + 1] - name_offsets[q >> 1]);
        for (int j = 0; j < k; ++j) {
            int64_t t = idx[i * k + j];
            if (t == q) continue;                                 // :279-280
            if (t < 0) t += n_rows_total;                         // Python's read_names[-1]
            if (t < 0 || t >= n_rows_total) { fclose(f); return FR_ERR_ARG; }
            const char *tn = names + name_offsets[t >> 1];
            const size_t tl = (size_t)(name_offsets[(t >> 1) + 1] - name_offsets[t >> 1]);
            put_field(buf, qn, ql);
            buf.push_back('\t'); buf.push_back((q & 1) ? '-' : '+'); buf.push_back('\t');
            put_field(buf, tn, tl);
            buf.push_back('\t'); buf.push_back((t & 1) ? '-' : '+'); buf.push_back('\t');
            auto r = std::to_chars(num, num + 16, j);
            buf.insert(buf.end(), num, r.ptr);
            buf.push_back('\t');
            const int n = fmt_f32(dist[i * k + j], num);
            buf.insert(buf.end(), num, num + n);
            buf.push_back('\n');
            ++lines;
        }
        if (buf.size() >= (1u << 22)) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    }
    if (!buf.empty()) fwrite(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (lines_out) *lines_out = lines;
    return FR_OK;
}
