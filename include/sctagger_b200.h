/*
 * sctagger_b200 — C ABI of the B200-native `match_trie` hot path of vpc-ccg/scTagger.
 *
 * The reference has no FFI: its in-process seam for this path is the Python call
 *     run_get_matches_prefix(max_error, threads) -> dict[rid] -> (e, {(bid, strand)})      reference scTagger.py:688-716
 * with its inputs in module globals (barcodes, lr_segs, lr_segs_idxs; scTagger.py:519-523, filled by
 * read_sr_barcodes scTagger.py:631-640 and read_long_reads scTagger.py:602-628).  The entry points below are what a
 * ctypes binding of that seam needs (see INTEGRATION.md for the stub a maintainer would add):
 *
 *   sct_create   <- the globals `barcodes`/`barcodes_rc` + the (max_error, threads, -bl, -m) arguments
 *                   (scTagger.py:631-640, 758-768); builds the device pattern table once.
 *   sct_match    <- run_get_matches_prefix (scTagger.py:688-716): all segments against all barcodes and their
 *                   reverse complements; per segment the minimum edit distance, the number of tied patterns and the
 *                   tied pattern ids in the order the reference prints them (scTagger.py:789).
 *   sct_destroy  <- process exit.
 *
 * Plain pointers and sizes only; every function returns 0 on success or a negative SCT_E_* code, and never throws.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with SCT_E_CUDA.
 * One handle drives one GPU; calls on different handles may run concurrently from different host threads (ctypes
 * releases the GIL), calls on one handle must be serialised by the caller.
 */
#ifndef SCTAGGER_B200_H
#define SCTAGGER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCT_OK            0
#define SCT_E_ARG        -1   /* invalid argument (the reference would raise AssertionError, scTagger.py:140-143, 760-761) */
#define SCT_E_CUDA       -2   /* CUDA runtime error / no device */
#define SCT_E_NOMEM      -3   /* host or device allocation failed */
#define SCT_E_UNSUPPORTED -4  /* parameter combination outside the implemented range (see sct_create) */
#define SCT_E_STATE      -5   /* call sequence error (e.g. results requested before a run) */
#define SCT_E_IO         -6   /* file cannot be opened / read / decompressed (the reference would raise OSError) */
#define SCT_E_FORMAT     -7   /* malformed input row (the reference would raise IndexError, scTagger.py:613-614) */

typedef struct sct_handle sct_handle;

/* Number of visible CUDA devices (0 if none or if the runtime cannot be initialised). */
int sct_device_count(void);

/* Thread-local message describing the last error returned on this thread. */
const char *sct_last_error(void);

/*
 * barcodes: n_barcodes * barcode_length bytes, file order = barcode id (scTagger.py:638).  Bytes other than upper-case
 *   ACGT are scored as 'A' (map_char, scTagger.py:512-516) and kept as they are by the reverse complement
 *   (rev_compl, scTagger.py:148-156).
 * max_error: -mr.  Requires barcode_length > max_error >= 0 (scTagger.py:143).
 * threads:   the reference's -t.  It fixes k = min{k : 4^k >= threads} (scTagger.py:689-697), which decides where a
 *   match may start in segments holding non-ACGT bytes.  Requires barcode_length-max_error > k.
 * mem_gb:    -m; sizes the streamed chunks (pinned staging + device buffers), > 0.
 * Implemented range: barcode_length <= 32; the bit-parallel kernels cover barcode_length <= 16 and max_error <= 3,
 *   anything else runs on the (slower, exact) dynamic-programming kernel.
 */
int sct_create(int device, const uint8_t *barcodes, int32_t n_barcodes, int32_t barcode_length, int32_t max_error,
               int32_t threads, double mem_gb, sct_handle **out);

int sct_destroy(sct_handle *h);

/*
 * seg_bytes / seg_offsets: the concatenated segment bytes and n+1 offsets (lr_segs / lr_segs_idxs of
 *   scTagger.py:617-626), in HOST memory, owned by the caller for the duration of the call.
 * Outputs (host memory, caller-allocated, n entries unless noted):
 *   min_ed[r]      minimum edit distance, or -1 when no pattern is within max_error (the reference then writes no row)
 *   n_ties[r]      number of (barcode, strand) patterns at that minimum
 *   tie_offsets    n+1 entries; ties of segment r are tie_ids[tie_offsets[r] .. tie_offsets[r+1])
 *   *tie_ids       library-owned buffer, valid until the next sct_match / sct_destroy on this handle;
 *                  id = (barcode_index << 1) | strand, strand 1 = forward, 0 = reverse complement, ascending — the
 *                  sorted((bid, strand)) order of scTagger.py:789
 *   *n_tie_ids     = tie_offsets[n]
 * The call streams the segments through the GPU in chunks sized from mem_gb (pinned double-buffered async copies).
 */
int sct_match(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n,
              int32_t *min_ed, int32_t *n_ties, int64_t *tie_offsets, const int32_t **tie_ids, int64_t *n_tie_ids);

/*
 * Device-resident variant used by bench.py's `value` leg: sct_stage copies one batch (it must fit one chunk) to the
 * GPU, sct_run_staged runs the kernels on the resident batch (results stay on the device) and reports the device
 * time of the whole pass and of the dominant match kernel, measured with CUDA events on the launching stream;
 * sct_fetch_staged copies the results of the last run to the host with the sct_match output conventions.
 */
int sct_stage(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n);
int sct_run_staged(sct_handle *h, float *pass_ms, float *match_kernel_ms);
int sct_fetch_staged(sct_handle *h, int32_t *min_ed, int32_t *n_ties, int64_t *tie_offsets,
                     const int32_t **tie_ids, int64_t *n_tie_ids);

/* Counters of the last sct_match / sct_run_staged on this handle: [0] kernel launches, [1] segments on the one-window
 * seeded (or tile) kernel: clean ACGT, <= 31 bases, [2] segments scored by the all-pairs generic kernel (tie-list
 * overflow re-runs; every non-short segment when the barcodes cannot be seeded), [3] host->device bytes, [4] device->host
 * bytes, [5] chunks, [6] segments on the DP kernel, [7] (barcode, strand) candidates the seeded kernels generated.
 * Segments longer than 31 bases or with non-ACGT bytes run on the WINDOWED seeded kernel: sct_work_counters [4]. */
int sct_stats(sct_handle *h, int64_t stats[8]);

/*
 * Tuning / test hooks (set before the first sct_match / sct_stage):
 *   "chunk_segments", "chunk_bytes"       override the chunk sizing derived from mem_gb
 *   "tie_capacity_per_segment"            initial size of the device tie buffer (it grows and the pass is redone if short)
 *   "path"                                0 = automatic (seeded kernels: one window for clean segments of <= 31 bases,
 *                                         overlapping windows for every other segment), 1 = force the generic
 *                                         bit-parallel kernel, 2 = force the DP kernel, 3 = use the all-pairs tile
 *                                         kernel instead of the seeded one
 *   "stream_segments", "ramp_segments"    segments per streamed chunk of sct_match, and of its first chunk (the size
 *                                         quadruples from chunk to chunk up to stream_segments so the pipeline fills fast)
 *   "chain_kernels"                       1 (default) = a chunk's kernels start behind the previous chunk's match kernel
 *   "tail_segments"                       size of the last streamed chunk of a call (default 65536; 0 = no small last chunk)
 *   "reserve_segments"                    allocate the streaming slots now, for chunks of up to this many segments from
 *                                         pageable caller arrays, instead of inside the first sct_match (the CLI does it
 *                                         while it parses the -lr file)
 *   "seed_ctas_per_sm"                    resident CTAs per SM of the seeded kernel (grid = SMs x this)
 *   "seed_deepening"                      1 (default) = search error bounds 1, 2, .. max_error in turn and stop at the first
 *                                         bound with a hit (only the minimum distance and its ties are reported);
 *                                         0 = search max_error at once.  Same results.
 *   "seed_prefilter"                      -1 (default) = anchored pre-check in the phases whose index buckets are dense
 *                                         enough for it to pay, 1 = wherever possible, 0 = verify every candidate in full.
 *                                         Same results.
 *   "seed_direct"                         -1 (default) = phases with a sparse index probe and verify per lane, 0 = never,
 *                                         1 = every phase.  Same results.
 *   "seed_fast_max"                       longest clean segment of the one-window kernel (default 31; 0 sends every
 *                                         segment through the windowed kernel — a test hook)
 */
int sct_set_option(sct_handle *h, const char *key, int64_t value);

/* Chunk sizing derived from mem_gb (segments / bytes per streamed chunk of sct_match).  sct_stage itself accepts any
 * batch that fits the device. */
int sct_chunk_capacity(sct_handle *h, int64_t *max_segments, int64_t *max_bytes);

/*
 * INT32 ALU roofline denominator: runs a register-only kernel of independent LOP3 / IADD3 chains on
 * every SM and reports lane-operations per second (best of `repeats`).  mode 0 = LOP3 only (ALU pipe),
 * 1 = IADD3 only, 2 = LOP3 + IMAD mix (ALU + FMA pipes).
 */
int sct_measure_int32_peak(int device, int mode, int repeats, double *lane_ops_per_s);

/* Wipe L2 by overwriting a buffer larger than it (used between timed bench iterations). */
int sct_flush_l2(sct_handle *h);

/*
 * Page-locked host memory.  The reference keeps lr_segs / lr_segs_idxs in multiprocessing.RawArray's
 * (scTagger.py:617-626) so that forked workers read them without a copy; the equivalent here is host memory the GPU's
 * copy engines can read and write directly.  Every array handed to sct_match that lies in page-locked memory —
 * allocated with sct_host_alloc, pinned in place with sct_host_register, or by any other cudaHostAlloc /
 * cudaHostRegister — is copied from / to without the staging pass through the library's own pinned buffers: inputs go
 * caller -> device, results device -> caller, and the host threads touch no segment byte.  Pageable arrays keep
 * working (they are staged by the handle's host threads).
 */
int sct_host_alloc(size_t bytes, void **out);
int sct_host_free(void *p);
int sct_host_register(void *p, size_t bytes);
int sct_host_unregister(void *p);

/* Live work counters of the last sct_match / sct_run_staged (device-side, summed over the chunks): [0] index bucket
 * probes, [1] verifications, [2] anchored pre-checks, [3] phases run (windows x error bounds searched: with iterative
 * deepening about 1.3 per segment on the bench input), [4] segments on the windowed seeded kernel (longer than 31 bases
 * or with non-ACGT bytes), [5] LOP3 lane-operations of the automata (steps x (2k+1), k = error bound of the phase),
 * [6..7] reserved.  bench.py derives `roofline.useful` from [5]. */
int sct_work_counters(sct_handle *h, int64_t counters[8]);

/* Name of the kernel that scores the clean segments with the handle's parameters and options (the dominant kernel of a
 * pass; bench.py labels its roofline with it). */
const char *sct_kernel_name(sct_handle *h);

/* ---------------------------------------------------------------------------------------------------------------------
 * extract_lr_bc adapter alignment (BASELINE config 5).  The reference's seam is
 *     get_lr_bc_alns(seqs, threads) -> [(strand, dist, locs)]                         reference scTagger.py:307-314
 * which maps get_alns (scTagger.py:176-196) over the reads: two edlib.align(a, seq, 'HW', 'locations') calls per read,
 * a1 = the short-read adapter (-sa, scTagger.py:41-43) and a2 = rev_compl(a1) (scTagger.py:363-365), characters compared
 * literally.  edlib is a third-party dependency that is not part of the reference tree (env.yaml:9, unpinned); the
 * semantics implemented here are the ones restated in SURVEY.md Appendix B / oracle/align_oracle.c (parity unpinned).
 *
 *   sct_align_create  <- the globals a1/a2 (scTagger.py:363-365) and -t
 *   sct_align         <- get_lr_bc_alns: per read
 *        strand[r]   '+' (a1 strictly better), '-' (a2 strictly better) or 0 (equal distances: the reference's 'NA')
 *        dist[r]     the smaller edit distance, or -1 for 'NA'
 *        locs        '+': every 0-based end position of a1 at that distance (scTagger.py:185)
 *                    '-': start - len(seq) - 1 for every location of a2 (scTagger.py:189), start = edlib's start of the
 *                         location = the leftmost start with the same distance for that end
 *                    locations of read r are (*locs)[loc_offsets[r] .. loc_offsets[r+1]); none for 'NA'
 *   Adapter length <= 32.  Host buffers in, host buffers out; *locs is library-owned (page-locked) until the next call /
 *   destroy.  As with sct_match, arrays that lie in page-locked memory (sct_host_alloc / sct_host_register) are copied
 *   from and to directly; pageable ones are staged by the handle's host threads.  read_offsets are validated on the
 *   device (SCT_E_ARG when they decrease).
 */
typedef struct sct_align_handle sct_align_handle;
int sct_align_create(int device, const uint8_t *adapter, int32_t adapter_length, int32_t threads, double mem_gb,
                     sct_align_handle **out);
int sct_align_destroy(sct_align_handle *h);
int sct_align(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n,
              int8_t *strand, int32_t *dist, int64_t *loc_offsets, const int32_t **locs, int64_t *n_locs);
/* Device-resident variant for timing: one batch (one chunk) staged once, then passes timed with CUDA events. */
int sct_align_stage(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n);
int sct_align_run_staged(sct_align_handle *h, float *pass_ms, float *pass1_kernel_ms);
/* [0] kernel launches, [1] chunks, [2] host->device bytes, [3] device->host bytes, [4] device time of the passes in ns,
 * [5] reads, [6..7] reserved.   Options: "chunk_bytes", "chunk_reads". */
int sct_align_stats(sct_align_handle *h, int64_t stats[8]);
int sct_align_set_option(sct_align_handle *h, const char *key, int64_t value);

/*
 * Allocation-free variant of pass 1 of the adapter alignment: strand[r] ('+', '-', 0 = NA) and dist[r] (-1 = NA) of every
 * read, exactly what sct_align reports in those two arrays.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it): it needs no handle and calls no cudaMalloc / cudaFree / allocating routine.  Its device
 * scratch (the device copy of the reads, the Peq table, the per-read arrays of the pass-1 kernel) is one file-scope static
 * device array of about 2.6 MB that every GPU loading the library holds until exit.
 * Limits (SCT_E_UNSUPPORTED beyond them): n <= 8192 reads, read_offsets[n] - read_offsets[0] <= 2 MiB, every read
 * <= 16384 bases, adapter length <= 32.  Host buffers in, host buffers out, synchronous (legacy default stream).
 * All calls on one device share the scratch array, which keeps the last call's contents (the call clears what must start
 * at zero): calls that overlap on one GPU race — serialise them.
 */
int sct_align_strand_dist_noalloc(int device, const uint8_t *adapter, int32_t adapter_length, const uint8_t *read_bytes,
                                  const int64_t *read_offsets, int64_t n, int8_t *strand, int32_t *dist);

/* ---------------------------------------------------------------------------------------------------------------------
 * Native host I/O of the formats either side of the kernels (§8(f) rank 2).  Host code only
 * (threads + zlib); no GPU is needed and nothing here computes a match.
 *
 *   sct_io_read_long_reads  <- read_long_reads (scTagger.py:602-628): TSV, gzip iff the name ends with ".gz"; per line
 *        column 0 = read name, column 3 = segment (may be empty); a row with fewer than four columns -> SCT_E_FORMAT
 *        (the reference's IndexError).  -> names and segments as concatenated bytes + n+1 offsets, library-owned.
 *   sct_io_read_fastqs      <- read_fastqs (scTagger.py:159-173), one file per call: gzip iff `gzipped` or ".gz"; line 0
 *        of every record: name = first whitespace-separated token minus its first character; line 1: the sequence with
 *        trailing whitespace stripped.
 *   sct_io_format_matches   <- the writer of match_trie (scTagger.py:777-789): one row `name \t e \t n_ties \t segment \t
 *        b1,b2,...` per read with min_ed >= 0, in read order; tie ids as sct_match returns them (forward barcode for
 *        strand 1, its reverse complement for strand 0).  gzip_level 0 = plain text, 1..9 = gzip (one member per block
 *        of rows, compressed in parallel; decompresses to the same bytes as the reference's single member).
 *   sct_io_format_lr_bc     <- output_matches of extract_lr_bc (scTagger.py:317-320): `rname \t dist \t loc \t
 *        seq[s:e or None]` for every read; dist < 0 = rejected by filter_aln: `-1`, `NA`, empty sequence.
 *   The buffers stay valid until the matching *_free call.
 *   sct_io_read_lr_bc_segments <- the -i reader of extract_sr_bc_from_lr (scTagger.py:811-822): `_,_,p,seg = l.split('\t')`,
 *        exactly four columns (else SCT_E_FORMAT, the reference's ValueError); rows with p == "NA" are skipped; only the
 *        segments are kept (no names).
 *   sct_io_read_whitelist   <- the -wl reader (scTagger.py:793-799): every line minus its last character (`l[:-1]`),
 *        returned as the "segments" of the handle, line order = barcode index.
 */
typedef struct sct_reads sct_reads;
typedef struct sct_buffer sct_buffer;
int sct_io_read_long_reads(const char *path, int32_t threads, sct_reads **out);
int sct_io_read_lr_bc_segments(const char *path, int32_t threads, sct_reads **out);
int sct_io_read_whitelist(const char *path, int32_t threads, sct_reads **out);
int sct_io_read_fastqs(const char *path, int32_t gzipped, int32_t threads, sct_reads **out);
int sct_io_reads_info(sct_reads *r, int64_t *n, const uint8_t **name_bytes, const int64_t **name_offsets,
                      const uint8_t **seg_bytes, const int64_t **seg_offsets);
int sct_io_reads_free(sct_reads *r);
int sct_io_format_matches(const uint8_t *name_bytes, const int64_t *name_offsets, const uint8_t *seg_bytes,
                          const int64_t *seg_offsets, int64_t n, const int32_t *min_ed, const int32_t *n_ties,
                          const int64_t *tie_offsets, const int32_t *tie_ids, const uint8_t *barcodes,
                          int32_t n_barcodes, int32_t barcode_length, int32_t gzip_level, int32_t threads,
                          sct_buffer **out);
int sct_io_format_lr_bc(const uint8_t *name_bytes, const int64_t *name_offsets, const uint8_t *read_bytes,
                        const int64_t *read_offsets, int64_t n, const int64_t *dist, const int64_t *loc,
                        const int64_t *s, const int64_t *e, int32_t gzip_level, int32_t threads, sct_buffer **out);
int sct_io_buffer_info(sct_buffer *b, const uint8_t **data, int64_t *size, int64_t *rows);
int sct_io_buffer_free(sct_buffer *b);

/* ---------------------------------------------------------------------------------------------------------------------
 * extract_sr_bc_from_lr: exact whitelist search.  The reference's seam is the loop
 *     A = ahocorasick.Automaton(); A.add_word(bc, idx) ...; A.add_word(rev_compl(bc), -idx) ...     scTagger.py:802-809
 *     for every segment: hits = tuple(A.iter(seg)); if len(hits) == 1: C[abs(hits[0][1])] += 1     scTagger.py:819-827
 * pyahocorasick is a third-party dependency that is not part of the reference tree (env.yaml, unpinned); the semantics
 * implemented are its documented ones — add_word on an existing key replaces the value, iter reports every occurrence
 * of every key in order of its end position — restated in oracle/wl_oracle.py (parity with pyahocorasick unpinned).
 *
 *   sct_wl_create   <- the automaton: words = the whitelist lines as the reference reads them (concatenated bytes +
 *                      n_words+1 offsets, line order = barcode index).  Blank lines add no word; all other words must
 *                      share one length <= 31 (SCT_E_UNSUPPORTED otherwise; the 10x whitelists are all 16-mers).  Bytes
 *                      are compared literally (a word holding a byte outside ACGT only matches that byte).
 *   sct_wl_count    <- the per-segment loop over one batch of segments (HOST buffers; streamed in chunks).  Counts are
 *                      accumulated on the device across calls.  Optional per-segment outputs (may be NULL):
 *                      n_hits[r] = len(hits), hit_word[r] = abs(value) of the first hit or -1.
 *                      Arrays in page-locked memory (sct_host_alloc / sct_host_register) are copied from / to directly.
 *   sct_wl_result   <- the Counter C in insertion order: barcode indices and counts of every barcode counted so far,
 *                      ordered by the first segment that counted it (library-owned until the next call / destroy).
 *                      Only the counted words leave the device (compacted there), not one counter per whitelist word.
 *   sct_wl_reset    clears the counts.   sct_wl_stage / sct_wl_run_staged: device-resident timing of one batch.
 *   sct_wl_stats    [0] kernel launches, [1] chunks of the last call, [2] host->device bytes, [3] device->host bytes,
 *                   [4] distinct packed keys in the table, [5] segments so far, [6] table slots, [7] literal (non-ACGT) words.
 */
typedef struct sct_wl_handle sct_wl_handle;
int sct_wl_create(int device, const uint8_t *word_bytes, const int64_t *word_offsets, int64_t n_words, int32_t threads,
                  sct_wl_handle **out);
int sct_wl_destroy(sct_wl_handle *h);
int sct_wl_count(sct_wl_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n, int32_t *n_hits,
                 int32_t *hit_word);
int sct_wl_result(sct_wl_handle *h, int64_t *n_unique, const int32_t **word_idx, const int64_t **count);
int sct_wl_reset(sct_wl_handle *h);
int sct_wl_stage(sct_wl_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n);
int sct_wl_run_staged(sct_wl_handle *h, float *kernel_ms);
int sct_wl_stats(sct_wl_handle *h, int64_t stats[8]);

#ifdef __cplusplus
}
#endif
#endif
