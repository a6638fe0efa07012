// Native host I/O of the `match_trie` stage: the reference's per-line Python readers and its
// five-writes-per-row writer are slower than the GPU pass by three orders of magnitude, so the file formats either side
// of the kernel are parsed / formatted / (de)compressed here with the host's threads.
//
//   sct_io_read_long_reads   <- read_long_reads   reference scTagger.py:602-628   (col 0 = name, col 3 = segment)
//   sct_io_read_fastqs       <- read_fastqs       reference scTagger.py:159-173   (line 0 = @name ..., line 1 = sequence)
//   sct_io_format_matches    <- the writer loop   reference scTagger.py:770-790   (name, e, n_ties, segment, barcodes)
//
// Host code only: nothing in this file touches the GPU, and nothing here computes a match.
#include "host_common.h"

#include <zlib.h>

using namespace scth;

struct sct_reads {
    std::vector<uint8_t> raw;           // whole (decompressed) file
    std::vector<uint8_t> name_bytes, seg_bytes;
    std::vector<int64_t> name_off, seg_off;
    int64_t n = 0;
};

struct sct_buffer {
    std::vector<std::vector<uint8_t>> blocks;   // per block of rows, in row order
    std::vector<uint8_t> flat;
    int64_t rows = 0;
};

namespace {

int read_whole_file(const char *path, bool gz, std::vector<uint8_t> &out) {
    if (!gz) {
        FILE *f = fopen(path, "rb");
        if (!f) return fail(SCT_E_IO, "cannot open %s: %s", path, strerror(errno));
        if (fseek(f, 0, SEEK_END) == 0) {
            const long sz = ftell(f);
            if (sz > 0) out.reserve((size_t)sz + out.size());
            fseek(f, 0, SEEK_SET);
        }
        const size_t step = (size_t)8 << 20;
        for (;;) {
            const size_t have = out.size();
            out.resize(have + step);
            const size_t got = fread(out.data() + have, 1, step, f);
            out.resize(have + got);
            if (got < step) break;
        }
        const bool bad = ferror(f);
        fclose(f);
        if (bad) return fail(SCT_E_IO, "read error on %s", path);
        return SCT_OK;
    }
    // gzip.open semantics: every member of a multi-member file is read (gzread does that); a file that is not gzip at
    // all is an error in Python (BadGzipFile), so transparent pass-through is refused
    gzFile g = gzopen(path, "rb");
    if (!g) return fail(SCT_E_IO, "cannot open %s: %s", path, strerror(errno));
    gzbuffer(g, 1u << 20);
    const size_t step = (size_t)16 << 20;
    bool first = true;
    for (;;) {
        const size_t have = out.size();
        out.resize(have + step);
        const int got = gzread(g, out.data() + have, (unsigned)step);
        if (got < 0) {
            int en = 0;
            const char *msg = gzerror(g, &en);
            std::string m = msg ? msg : "?";
            gzclose(g);
            return fail(SCT_E_IO, "gzip error on %s: %s", path, m.c_str());
        }
        out.resize(have + (size_t)got);
        if (first && got > 0 && gzdirect(g)) {
            gzclose(g);
            return fail(SCT_E_IO, "%s: not a gzipped file", path);
        }
        first = false;
        if ((size_t)got < step) break;
    }
    // A truncated stream makes gzread return the bytes inflated so far and only records Z_BUF_ERROR ("unexpected end of
    // file"); a corrupt tail shows as Z_DATA_ERROR here or as a CRC/length failure in gzclose.  Python's gzip.open
    // raises EOFError / BadGzipFile for these (the reference's readers, scTagger.py:605, 634), so they are SCT_E_IO.
    int en = Z_OK;
    const char *msg = gzerror(g, &en);
    if (en != Z_OK && en != Z_STREAM_END) {
        std::string m = (msg && *msg) ? msg : "unexpected end of file";
        gzclose(g);
        return fail(SCT_E_IO, "gzip error on %s: %s", path, m.c_str());
    }
    const int crc = gzclose(g);
    if (crc != Z_OK) return fail(SCT_E_IO, "gzip error on %s: %s", path, crc == Z_BUF_ERROR ? "unexpected end of file" : "corrupt stream");
    return SCT_OK;
}

// [a, b) pieces of `raw` that start and end at line boundaries, one per worker
std::vector<int64_t> line_aligned_cuts(const std::vector<uint8_t> &raw, int parts) {
    const int64_t sz = (int64_t)raw.size();
    std::vector<int64_t> cut(parts + 1, sz);
    cut[0] = 0;
    for (int p = 1; p < parts; ++p) {
        int64_t x = std::max(cut[p - 1], sz * p / parts);
        if (x < sz) {
            const void *nl = memchr(raw.data() + x, '\n', (size_t)(sz - x));
            x = nl ? (const uint8_t *)nl - raw.data() + 1 : sz;
        }
        cut[p] = x;
    }
    return cut;
}

inline int64_t count_lines(const uint8_t *p, const uint8_t *e) {
    int64_t n = 0;
    while (p < e) {
        const void *nl = memchr(p, '\n', (size_t)(e - p));
        ++n;
        if (!nl) break;
        p = (const uint8_t *)nl + 1;
    }
    return n;
}

struct PieceCount {
    int64_t lines = 0, seen = 0, name_bytes = 0, seg_bytes = 0, bad_line = -1;
};

}  // namespace

int read_input_file(const char *path, bool gz, int threads, std::vector<uint8_t> &out);   // defined below

// mode 0: read_long_reads (scTagger.py:602-628): >= 4 columns, column 0 = name, column 3 = segment.
// mode 1: the -i reader of extract_sr_bc_from_lr (scTagger.py:811-820): `_,_,p,seg = l.split('\t')` — exactly 4 columns
//         (anything else is the reference's ValueError), rows whose column 2 is `NA` are skipped, names are not kept.
// mode 2: the whitelist reader of extract_sr_bc_from_lr (scTagger.py:798): every line minus its last character
//         (`l[:-1]`: the newline — or, on a last line without one, a real character, as in the reference).
static int read_tsv(const char *path, int32_t threads, int mode, sct_reads **out) {
    if (!path || !out) return fail(SCT_E_ARG, "sct_io_read: null argument");
    *out = nullptr;
    sct_reads *r = new (std::nothrow) sct_reads;
    if (!r) return fail(SCT_E_NOMEM, "out of memory");
    try {
        const size_t pl = strlen(path);
        const bool gz = pl >= 3 && strcmp(path + pl - 3, ".gz") == 0;      // scTagger.py:605
        int rc = read_input_file(path, gz, threads, r->raw);
        if (rc != SCT_OK) {
            delete r;
            return rc;
        }
        const int nt = std::max(1, std::min<int>(threads, 64));
        HostPool pool(nt);
        const auto cut = line_aligned_cuts(r->raw, nt);
        std::vector<PieceCount> pc(nt);
        const uint8_t *base = r->raw.data();
        // pass 1: rows and bytes per piece.  A row is `line.rstrip('\n').split('\t')`; l[0] and l[3] are needed, fewer
        // than four columns is the reference's IndexError (scTagger.py:611-613)
        pool.run([&](int t) {
            const uint8_t *p = base + cut[t], *e = base + cut[t + 1];
            PieceCount c;
            while (p < e) {
                const uint8_t *nl = (const uint8_t *)memchr(p, '\n', (size_t)(e - p));
                const uint8_t *le = nl ? nl : e;
                if (mode == 2 && !nl) --le;                                // l[:-1] on a last line without '\n'
                while (le > p && le[-1] == '\r') --le;                     // text mode: "\r\n" reads as "\n"
                if (mode == 2) {
                    c.seg_bytes += le - p;
                    ++c.lines, ++c.seen;
                    if (!nl) break;
                    p = nl + 1;
                    continue;
                }
                const uint8_t *c0 = (const uint8_t *)memchr(p, '\t', (size_t)(le - p));
                const uint8_t *c1 = c0 ? (const uint8_t *)memchr(c0 + 1, '\t', (size_t)(le - c0 - 1)) : nullptr;
                const uint8_t *c2 = c1 ? (const uint8_t *)memchr(c1 + 1, '\t', (size_t)(le - c1 - 1)) : nullptr;
                const uint8_t *c3 = c2 ? (const uint8_t *)memchr(c2 + 1, '\t', (size_t)(le - c2 - 1)) : nullptr;
                ++c.seen;
                if (!c2 || (mode == 1 && c3)) {
                    if (c.bad_line < 0) c.bad_line = c.seen - 1;
                } else if (mode == 1 && c2 - c1 == 3 && c1[1] == 'N' && c1[2] == 'A') {
                    // p == 'NA': skipped (scTagger.py:821-822)
                } else {
                    if (mode == 0) c.name_bytes += c0 - p;
                    c.seg_bytes += (c3 ? c3 : le) - (c2 + 1);
                    ++c.lines;
                }
                if (!nl) break;
                p = nl + 1;
            }
            pc[t] = c;
        });
        int64_t n = 0, nb = 0, sb = 0, seen = 0;
        std::vector<int64_t> row0(nt + 1), nb0(nt + 1), sb0(nt + 1);
        for (int t = 0; t < nt; ++t) {
            if (pc[t].bad_line >= 0) {
                const long long line = (long long)(seen + pc[t].bad_line + 1);
                delete r;
                if (mode == 1)
                    return fail(SCT_E_FORMAT, "line %lld of %s does not have exactly 4 tab-separated columns (the "
                                "reference's ValueError on `_,_,p,seg = ...`)", line, path);
                return fail(SCT_E_FORMAT, "list index out of range (line %lld of %s has fewer than 4 tab-separated columns)",
                            line, path);
            }
            row0[t] = n, nb0[t] = nb, sb0[t] = sb;
            n += pc[t].lines, nb += pc[t].name_bytes, sb += pc[t].seg_bytes, seen += pc[t].seen;
        }
        row0[nt] = n, nb0[nt] = nb, sb0[nt] = sb;
        r->n = n;
        r->name_bytes.resize((size_t)nb + 1);
        r->seg_bytes.resize((size_t)sb + 1);
        r->name_off.resize((size_t)n + 1);
        r->seg_off.resize((size_t)n + 1);
        // pass 2: fill
        pool.run([&](int t) {
            const uint8_t *p = base + cut[t], *e = base + cut[t + 1];
            int64_t row = row0[t], no = nb0[t], so = sb0[t];
            while (p < e) {
                const uint8_t *nl = (const uint8_t *)memchr(p, '\n', (size_t)(e - p));
                const uint8_t *le = nl ? nl : e;
                if (mode == 2 && !nl) --le;
                while (le > p && le[-1] == '\r') --le;
                if (mode == 2) {
                    r->name_off[row] = 0;
                    r->seg_off[row] = so;
                    memcpy(r->seg_bytes.data() + so, p, (size_t)(le - p));
                    so += le - p;
                    ++row;
                    if (!nl) break;
                    p = nl + 1;
                    continue;
                }
                const uint8_t *c0 = (const uint8_t *)memchr(p, '\t', (size_t)(le - p));
                const uint8_t *c1 = (const uint8_t *)memchr(c0 + 1, '\t', (size_t)(le - c0 - 1));
                const uint8_t *c2 = (const uint8_t *)memchr(c1 + 1, '\t', (size_t)(le - c1 - 1));
                const uint8_t *c3 = (const uint8_t *)memchr(c2 + 1, '\t', (size_t)(le - c2 - 1));
                const uint8_t *se = c3 ? c3 : le;
                if (!(mode == 1 && c2 - c1 == 3 && c1[1] == 'N' && c1[2] == 'A')) {
                    r->name_off[row] = no;
                    r->seg_off[row] = so;
                    if (mode == 0) {
                        memcpy(r->name_bytes.data() + no, p, (size_t)(c0 - p));
                        no += c0 - p;
                    }
                    memcpy(r->seg_bytes.data() + so, c2 + 1, (size_t)(se - (c2 + 1)));
                    so += se - (c2 + 1);
                    ++row;
                }
                if (!nl) break;
                p = nl + 1;
            }
        });
        r->name_off[n] = nb;
        r->seg_off[n] = sb;
        std::vector<uint8_t>().swap(r->raw);
    } catch (const std::bad_alloc &) {
        delete r;
        return fail(SCT_E_NOMEM, "out of host memory reading %s", path);
    }
    *out = r;
    return SCT_OK;
}

extern "C" {

int sct_io_read_long_reads(const char *path, int32_t threads, sct_reads **out) { return read_tsv(path, threads, 0, out); }

int sct_io_read_lr_bc_segments(const char *path, int32_t threads, sct_reads **out) { return read_tsv(path, threads, 1, out); }

int sct_io_read_whitelist(const char *path, int32_t threads, sct_reads **out) { return read_tsv(path, threads, 2, out); }

// ---------------------------------------------------------------------------------------------------------------------
static inline bool py_space(uint8_t c) { return c == ' ' || (c >= 9 && c <= 13); }   // str.split() / str.rstrip() set

int sct_io_read_fastqs(const char *path, int32_t gzipped, int32_t threads, sct_reads **out) {
    if (!path || !out) return fail(SCT_E_ARG, "sct_io_read_fastqs: null argument");
    *out = nullptr;
    sct_reads *r = new (std::nothrow) sct_reads;
    if (!r) return fail(SCT_E_NOMEM, "out of memory");
    try {
        const size_t pl = strlen(path);
        const bool gz = gzipped || (pl >= 3 && strcmp(path + pl - 3, ".gz") == 0);      // scTagger.py:164
        int rc = read_input_file(path, gz, threads, r->raw);
        if (rc != SCT_OK) {
            delete r;
            return rc;
        }
        const int nt = std::max(1, std::min<int>(threads, 64));
        HostPool pool(nt);
        const auto cut = line_aligned_cuts(r->raw, nt);
        const uint8_t *base = r->raw.data();
        std::vector<int64_t> line0(nt + 1, 0);
        pool.run([&](int t) { line0[t + 1] = count_lines(base + cut[t], base + cut[t + 1]); });
        for (int t = 0; t < nt; ++t) line0[t + 1] += line0[t];
        const int64_t lines = line0[nt];
        const int64_t n_names = (lines + 3) / 4, n_seqs = (lines + 2) / 4;
        // the reference zips names with sequences (scTagger.py:319), so a trailing header without a sequence is dropped
        const int64_t n = std::min(n_names, n_seqs);
        std::vector<PieceCount> pc(nt);
        // walks the lines of piece t: fn(kind 0 = header / 1 = sequence, record, token begin, token end)
        auto walk = [&](int t, auto &&fn) {
            const uint8_t *p = base + cut[t], *e = base + cut[t + 1];
            int64_t idx = line0[t];
            while (p < e) {
                const uint8_t *nl = (const uint8_t *)memchr(p, '\n', (size_t)(e - p));
                const uint8_t *le = nl ? nl : e;
                const int k = (int)(idx & 3);
                if (k == 0) {
                    const uint8_t *a = p;
                    while (a < le && py_space(*a)) ++a;
                    const uint8_t *b = a;
                    while (b < le && !py_space(*b)) ++b;
                    fn(0, idx >> 2, a, b);
                } else if (k == 1) {
                    const uint8_t *b = le;
                    while (b > p && py_space(b[-1])) --b;
                    fn(1, idx >> 2, p, b);
                }
                ++idx;
                if (!nl) break;
                p = nl + 1;
            }
        };
        pool.run([&](int t) {
            PieceCount c;
            walk(t, [&](int kind, int64_t rec, const uint8_t *a, const uint8_t *b) {
                if (rec >= n) return;
                if (kind == 0) {
                    if (a == b) {
                        if (c.bad_line < 0) c.bad_line = rec * 4;
                    } else {
                        c.name_bytes += b - a - 1;      // [1:] drops the '@'
                    }
                } else {
                    c.seg_bytes += b - a;
                }
            });
            pc[t] = c;
        });
        int64_t nb = 0, sb = 0;
        std::vector<int64_t> nb0(nt + 1), sb0(nt + 1);
        for (int t = 0; t < nt; ++t) {
            if (pc[t].bad_line >= 0) {
                const long long line = (long long)pc[t].bad_line + 1;
                delete r;
                return fail(SCT_E_FORMAT, "list index out of range (header line %lld of %s is blank)", line, path);
            }
            nb0[t] = nb, sb0[t] = sb;
            nb += pc[t].name_bytes, sb += pc[t].seg_bytes;
        }
        r->n = n;
        r->name_bytes.resize((size_t)nb + 1);
        r->seg_bytes.resize((size_t)sb + 1);
        r->name_off.resize((size_t)n + 1);
        r->seg_off.resize((size_t)n + 1);
        pool.run([&](int t) {
            int64_t no = nb0[t], so = sb0[t];
            walk(t, [&](int kind, int64_t rec, const uint8_t *a, const uint8_t *b) {
                if (rec >= n) return;
                if (kind == 0) {
                    r->name_off[rec] = no;
                    memcpy(r->name_bytes.data() + no, a + 1, (size_t)(b - a - 1));
                    no += b - a - 1;
                } else {
                    r->seg_off[rec] = so;
                    memcpy(r->seg_bytes.data() + so, a, (size_t)(b - a));
                    so += b - a;
                }
            });
        });
        r->name_off[n] = nb;
        r->seg_off[n] = sb;
        std::vector<uint8_t>().swap(r->raw);
    } catch (const std::bad_alloc &) {
        delete r;
        return fail(SCT_E_NOMEM, "out of host memory reading %s", path);
    }
    *out = r;
    return SCT_OK;
}

int sct_io_reads_info(sct_reads *r, int64_t *n, const uint8_t **name_bytes, const int64_t **name_offsets,
                      const uint8_t **seg_bytes, const int64_t **seg_offsets) {
    if (!r) return fail(SCT_E_ARG, "sct_io_reads_info: null handle");
    if (n) *n = r->n;
    if (name_bytes) *name_bytes = r->name_bytes.data();
    if (name_offsets) *name_offsets = r->name_off.data();
    if (seg_bytes) *seg_bytes = r->seg_bytes.data();
    if (seg_offsets) *seg_offsets = r->seg_off.data();
    return SCT_OK;
}

int sct_io_reads_free(sct_reads *r) {
    delete r;
    return SCT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// Writers: rows are formatted in blocks of consecutive reads by the host threads; with gzip every block becomes one gzip
// member (a multi-member file is what `gzip.open`, zcat and the reference's own readers decompress to the concatenation),
// so compression scales with the threads.  Compressed bytes are not comparable between writers anyway (header mtime).
namespace {

inline uint8_t *put_int(uint8_t *p, int64_t v) {
    if (v < 0) {
        *p++ = '-';
        v = -v;
    }
    char tmp[24];
    int k = 0;
    do {
        tmp[k++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (k) *p++ = (uint8_t)tmp[--k];
    return p;
}

// Every member written here carries its own sizes in the gzip header's FEXTRA field (RFC 1952 §2.3.1.1, subfield "SC":
// u32 compressed member size, u32 uncompressed size — the BGZF idea with 32-bit sizes), which any gzip reader skips and
// which lets read_whole_file() below inflate the members of such a file in parallel.
constexpr int GZX_DATA_AT = 16;          // 10 fixed header bytes + XLEN(2) + SI1 SI2 LEN(2)
constexpr int GZX_HEADER = GZX_DATA_AT + 8;

int gzip_member(const std::vector<uint8_t> &in, int level, std::vector<uint8_t> &out) {
    z_stream z;
    memset(&z, 0, sizeof z);
    if (deflateInit2(&z, level, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK) return -1;
    gz_header gh;
    memset(&gh, 0, sizeof gh);
    uint8_t extra[12] = {'S', 'C', 8, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    gh.extra = extra;
    gh.extra_len = sizeof extra;
    gh.os = 255;
    if (deflateSetHeader(&z, &gh) != Z_OK) {
        deflateEnd(&z);
        return -1;
    }
    out.resize(deflateBound(&z, (uLong)in.size()) + 64);
    z.next_in = (Bytef *)in.data();
    z.avail_in = (uInt)in.size();
    z.next_out = out.data();
    z.avail_out = (uInt)out.size();
    const int rc = deflate(&z, Z_FINISH);
    const size_t produced = out.size() - z.avail_out;
    deflateEnd(&z);
    if (rc != Z_STREAM_END || produced < (size_t)GZX_HEADER + 8 || produced > 0xFFFFFFFFull) return -1;
    out.resize(produced);
    const uint32_t cs = (uint32_t)produced, us = (uint32_t)in.size();
    for (int i = 0; i < 4; ++i) {
        out[GZX_DATA_AT + i] = (uint8_t)(cs >> (8 * i));
        out[GZX_DATA_AT + 4 + i] = (uint8_t)(us >> (8 * i));
    }
    return 0;
}

}  // namespace

// A .gz file whose members all carry the "SC" size subfield (written by gzip_member above) is inflated member by member
// on the host threads; anything else (the reference's single-member files, foreign gzip) takes the serial zlib path.
int read_gz_parallel(const char *path, int threads, std::vector<uint8_t> &out, bool *handled) {
    *handled = false;
    std::vector<uint8_t> raw;
    int rc = read_whole_file(path, false, raw);
    if (rc != SCT_OK) return rc;
    struct Member { size_t at, csize, uoff, usize; };
    std::vector<Member> ms;
    size_t pos = 0, utotal = 0;
    while (pos < raw.size()) {
        const uint8_t *p = raw.data() + pos;
        if (raw.size() - pos < (size_t)GZX_HEADER + 8 || p[0] != 0x1f || p[1] != 0x8b || p[2] != 8 || !(p[3] & 4) ||
            p[10] != 12 || p[11] != 0 || p[12] != 'S' || p[13] != 'C' || p[14] != 8 || p[15] != 0)
            return SCT_OK;                                   // not (only) our members: serial path
        uint32_t cs = 0, us = 0;
        for (int i = 0; i < 4; ++i) {
            cs |= (uint32_t)p[GZX_DATA_AT + i] << (8 * i);
            us |= (uint32_t)p[GZX_DATA_AT + 4 + i] << (8 * i);
        }
        if (cs < (uint32_t)GZX_HEADER + 8 || cs > raw.size() - pos) return SCT_OK;
        ms.push_back({pos, cs, utotal, us});
        pos += cs;
        utotal += us;
    }
    if (ms.empty()) return SCT_OK;
    out.resize(utotal);
    std::atomic<size_t> next{0};
    std::atomic<int> bad{0};
    HostPool pool(std::max(1, std::min<int>(threads, 64)));
    pool.run([&](int) {
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= ms.size() || bad.load()) break;
            z_stream z;
            memset(&z, 0, sizeof z);
            if (inflateInit2(&z, 15 + 16) != Z_OK) {
                bad = 1;
                break;
            }
            z.next_in = raw.data() + ms[k].at;
            z.avail_in = (uInt)ms[k].csize;
            z.next_out = out.data() + ms[k].uoff;
            z.avail_out = (uInt)ms[k].usize;
            const int r = inflate(&z, Z_FINISH);
            if (r != Z_STREAM_END || z.avail_out != 0 || z.avail_in != 0) bad = 1;
            inflateEnd(&z);
        }
    });
    if (bad.load()) {
        out.clear();
        return SCT_OK;                                       // sizes lied: let the serial path decide (and report)
    }
    *handled = true;
    return SCT_OK;
}

int read_input_file(const char *path, bool gz, int threads, std::vector<uint8_t> &out) {
    if (gz) {
        bool handled = false;
        const int rc = read_gz_parallel(path, threads, out, &handled);
        if (rc != SCT_OK || handled) return rc;
        out.clear();
    }
    return read_whole_file(path, gz, out);
}

namespace {

// bound(first, last) -> upper bound of the bytes of rows [first, last); emit(first, last, p) -> end pointer, counts rows
template <class Bound, class Emit>
int format_blocks(int64_t n, int threads, int gzip_level, int64_t block_reads, Bound bound, Emit emit, sct_buffer **out) {
    sct_buffer *b = new (std::nothrow) sct_buffer;
    if (!b) return fail(SCT_E_NOMEM, "out of memory");
    const int64_t n_blocks = (n + block_reads - 1) / block_reads;
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(threads, 64), n_blocks));
    std::atomic<int64_t> next{0}, rows{0};
    std::atomic<int> err{0};
    try {
        b->blocks.resize((size_t)n_blocks);
        HostPool pool(nt);
        pool.run([&](int) {
            std::vector<uint8_t> plain;
            try {
                for (;;) {
                    const int64_t k = next.fetch_add(1);
                    if (k >= n_blocks || err.load()) break;
                    const int64_t lo = k * block_reads, hi = std::min(n, lo + block_reads);
                    plain.resize((size_t)bound(lo, hi));
                    int64_t r = 0;
                    uint8_t *e = emit(lo, hi, plain.data(), &r);
                    plain.resize((size_t)(e - plain.data()));
                    rows += r;
                    if (gzip_level > 0) {
                        if (!plain.empty() && gzip_member(plain, gzip_level, b->blocks[(size_t)k]) != 0) err = 1;
                    } else {
                        b->blocks[(size_t)k] = plain;
                    }
                }
            } catch (const std::bad_alloc &) {
                err = 2;
            }
        });
        if (err.load()) {
            delete b;
            return fail(err.load() == 2 ? SCT_E_NOMEM : SCT_E_IO, "formatting the output failed (%s)",
                        err.load() == 2 ? "out of memory" : "zlib");
        }
        std::vector<int64_t> off((size_t)n_blocks + 1, 0);
        for (int64_t k = 0; k < n_blocks; ++k) off[(size_t)k + 1] = off[(size_t)k] + (int64_t)b->blocks[(size_t)k].size();
        int64_t total = off[(size_t)n_blocks];
        if (gzip_level > 0 && total == 0) {                  // an empty gzip file is still one (empty) member
            std::vector<uint8_t> none, member;
            if (gzip_member(none, gzip_level, member) != 0) {
                delete b;
                return fail(SCT_E_IO, "zlib failed");
            }
            b->flat = member;
        } else {
            b->flat.resize((size_t)total);
            next = 0;
            pool.run([&](int) {
                for (;;) {
                    const int64_t k = next.fetch_add(1);
                    if (k >= n_blocks) break;
                    auto &blk = b->blocks[(size_t)k];
                    if (!blk.empty()) memcpy(b->flat.data() + off[(size_t)k], blk.data(), blk.size());
                    std::vector<uint8_t>().swap(blk);
                }
            });
        }
        b->blocks.clear();
        b->rows = rows.load();
    } catch (const std::bad_alloc &) {
        delete b;
        return fail(SCT_E_NOMEM, "out of host memory formatting the output");
    }
    *out = b;
    return SCT_OK;
}

}  // namespace

extern "C" {

int sct_io_format_matches(const uint8_t *name_bytes, const int64_t *name_offsets, const uint8_t *seg_bytes,
                          const int64_t *seg_offsets, int64_t n, const int32_t *min_ed, const int32_t *n_ties,
                          const int64_t *tie_offsets, const int32_t *tie_ids, const uint8_t *barcodes,
                          int32_t n_barcodes, int32_t barcode_length, int32_t gzip_level, int32_t threads,
                          sct_buffer **out) {
    if (!out) return fail(SCT_E_ARG, "sct_io_format_matches: null out");
    *out = nullptr;
    if (n < 0 || n_barcodes < 0 || barcode_length <= 0 || gzip_level < 0 || gzip_level > 9)
        return fail(SCT_E_ARG, "sct_io_format_matches: bad sizes");
    if (n > 0 && (!name_bytes || !name_offsets || !seg_bytes || !seg_offsets || !min_ed || !n_ties || !tie_offsets))
        return fail(SCT_E_ARG, "sct_io_format_matches: null argument");
    const int bl = barcode_length;
    // rev_compl(barcodes[b]) as printed for strand False (scTagger.py:789; identity on non-ACGT, scTagger.py:148-156)
    std::vector<uint8_t> rc((size_t)n_barcodes * bl);
    for (int64_t b = 0; b < n_barcodes; ++b)
        for (int i = 0; i < bl; ++i) rc[(size_t)(b * bl + i)] = comp_char(barcodes[b * bl + (bl - 1 - i)]);
    for (int64_t r = 0; r < n; ++r)
        if (min_ed[r] >= 0)
            for (int64_t t = tie_offsets[r]; t < tie_offsets[r + 1]; ++t)
                if (tie_ids[t] < 0 || (tie_ids[t] >> 1) >= n_barcodes)
                    return fail(SCT_E_ARG, "sct_io_format_matches: tie id %d out of range at read %lld", tie_ids[t], (long long)r);
    auto bound = [&](int64_t lo, int64_t hi) {
        int64_t s = 0;
        for (int64_t r = lo; r < hi; ++r)
            if (min_ed[r] >= 0)
                s += (name_offsets[r + 1] - name_offsets[r]) + (seg_offsets[r + 1] - seg_offsets[r]) + 32 +
                     (tie_offsets[r + 1] - tie_offsets[r]) * (bl + 1);
        return s;
    };
    auto emit = [&](int64_t lo, int64_t hi, uint8_t *p, int64_t *rows) {
        for (int64_t r = lo; r < hi; ++r) {
            if (min_ed[r] < 0) continue;                    // only reads with a match get a row (scTagger.py:777)
            const int64_t nl = name_offsets[r + 1] - name_offsets[r], sl = seg_offsets[r + 1] - seg_offsets[r];
            memcpy(p, name_bytes + name_offsets[r], (size_t)nl);
            p += nl;
            *p++ = '\t';
            p = put_int(p, min_ed[r]);
            *p++ = '\t';
            p = put_int(p, n_ties[r]);
            *p++ = '\t';
            memcpy(p, seg_bytes + seg_offsets[r], (size_t)sl);
            p += sl;
            *p++ = '\t';
            for (int64_t t = tie_offsets[r]; t < tie_offsets[r + 1]; ++t) {
                const int32_t id = tie_ids[t];
                memcpy(p, ((id & 1) ? barcodes : rc.data()) + (int64_t)(id >> 1) * bl, (size_t)bl);
                p += bl;
                *p++ = ',';
            }
            if (tie_offsets[r + 1] > tie_offsets[r]) --p;
            *p++ = '\n';
            ++*rows;
        }
        return p;
    };
    return format_blocks(n, threads, gzip_level, 16384, bound, emit, out);
}

// output_matches of extract_lr_bc (reference scTagger.py:317-320): `rname \t dist \t loc \t seq[s:e or None]`; dist < 0
// marks a read that filter_aln rejected (scTagger.py:287-304): it prints `-1`, `NA` and an empty sequence.
int sct_io_format_lr_bc(const uint8_t *name_bytes, const int64_t *name_offsets, const uint8_t *read_bytes,
                        const int64_t *read_offsets, int64_t n, const int64_t *dist, const int64_t *loc,
                        const int64_t *s, const int64_t *e, int32_t gzip_level, int32_t threads, sct_buffer **out) {
    if (!out) return fail(SCT_E_ARG, "sct_io_format_lr_bc: null out");
    *out = nullptr;
    if (n < 0 || gzip_level < 0 || gzip_level > 9) return fail(SCT_E_ARG, "sct_io_format_lr_bc: bad sizes");
    if (n > 0 && (!name_bytes || !name_offsets || !read_bytes || !read_offsets || !dist || !loc || !s || !e))
        return fail(SCT_E_ARG, "sct_io_format_lr_bc: null argument");
    // Python slice seq[a:b or None] on a sequence of length L
    auto slice = [&](int64_t r, int64_t *a, int64_t *b) {
        const int64_t L = read_offsets[r + 1] - read_offsets[r];
        int64_t x = s[r], y = e[r];
        x = x < 0 ? std::max<int64_t>(0, x + L) : std::min(x, L);
        y = y == 0 ? L : y < 0 ? std::max<int64_t>(0, y + L) : std::min(y, L);
        *a = x;
        *b = std::max(x, y);
    };
    auto bound = [&](int64_t lo, int64_t hi) {
        int64_t t = 0;
        for (int64_t r = lo; r < hi; ++r) {
            t += (name_offsets[r + 1] - name_offsets[r]) + 48;
            if (dist[r] >= 0) {
                int64_t a, b;
                slice(r, &a, &b);
                t += b - a;
            }
        }
        return t;
    };
    auto emit = [&](int64_t lo, int64_t hi, uint8_t *p, int64_t *rows) {
        for (int64_t r = lo; r < hi; ++r) {
            const int64_t nl = name_offsets[r + 1] - name_offsets[r];
            memcpy(p, name_bytes + name_offsets[r], (size_t)nl);
            p += nl;
            *p++ = '\t';
            if (dist[r] < 0) {
                memcpy(p, "-1\tNA\t\n", 7);
                p += 7;
            } else {
                p = put_int(p, dist[r]);
                *p++ = '\t';
                p = put_int(p, loc[r]);
                *p++ = '\t';
                int64_t a, b;
                slice(r, &a, &b);
                memcpy(p, read_bytes + read_offsets[r] + a, (size_t)(b - a));
                p += b - a;
                *p++ = '\n';
            }
            ++*rows;
        }
        return p;
    };
    return format_blocks(n, threads, gzip_level, 2048, bound, emit, out);
}

int sct_io_buffer_info(sct_buffer *b, const uint8_t **data, int64_t *size, int64_t *rows) {
    if (!b) return fail(SCT_E_ARG, "sct_io_buffer_info: null handle");
    if (data) *data = b->flat.data();
    if (size) *size = (int64_t)b->flat.size();
    if (rows) *rows = b->rows;
    return SCT_OK;
}

int sct_io_buffer_free(sct_buffer *b) {
    delete b;
    return SCT_OK;
}

}  // extern "C"
