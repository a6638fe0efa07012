/*
 * TEST INFRASTRUCTURE ONLY — second, independent CPU oracle for scTagger `match_trie`
 * (see trie_oracle.c for the usage rules: tests/, smoke() and bench.py's CPU legs only).
 *
 * Two restatements of the SEMANTICS of the reference, neither of which
 * shares any code or data structure with the trie port or with the CUDA kernel:
 *
 *   sct_oracle_anchored : literal Appendix A.2/A.3 — for every valid start i (prefix rule of
 *                         get_matches_prefix, scTagger.py:677-682) the window seg[i:i+bl+mr]
 *                         (scTagger.py:680), Levenshtein(pattern, window[:l]) for every
 *                         l in [bl-mr, len(window)] (read ids are stored at depths >= bl-mr,
 *                         scTagger.py:563-564; dfs emits when the pattern is exhausted, 567-572),
 *                         minimum over i and l, kept when <= mr.  O(len * bl * (bl+mr)) per pair.
 *   sct_oracle_dp       : one semi-global DP per (segment, pattern) with the modified top row of
 *                         Appendix A.4 (D[0][j] = 0 at valid starts, else D[0][j-1]+1).
 *
 * Both use map_char (scTagger.py:512-516: non-ACGT -> 'A') for scoring, byte-exact upper-case
 * ACGT for the start rule, and emit ties as (bid<<1)|strand ascending (scTagger.py:789).
 * Parity pin: tests/test_oracle.py compares both with the reference outputs in tests/golden/.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int mapc(uint8_t c) { return c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : 0; }
static inline uint8_t compc(uint8_t c) { return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c; }
static inline int is_acgt(uint8_t c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }
#define INF 1000000

static int prefix_k(int threads) { int k = 0; long p = 1; while (threads > p) { p *= 4; k++; } return k; }

/* start i is valid iff the k bytes seg[i:i+k] exist and are upper-case ACGT (and the window is long enough) */
static int valid_start(const uint8_t *s, int64_t len, int64_t i, int k) {
    if (i + k > len) return 0;
    for (int j = 0; j < k; j++) if (!is_acgt(s[i + j])) return 0;
    return 1;
}

static int dist_anchored(const uint8_t *p, int bl, int mr, const uint8_t *s, int64_t len, int k) {
    int best = INF;
    int D[2][80];
    for (int64_t i = 0; i < len; i++) {
        int64_t wl = len - i < bl + mr ? len - i : bl + mr;
        if (wl < bl - mr || !valid_start(s, len, i, k)) continue;
        /* Levenshtein(pattern, window[:l]) for all l via the standard DP, rows = pattern */
        int *prev = D[0], *cur = D[1];
        for (int l = 0; l <= wl; l++) prev[l] = l;
        for (int r = 1; r <= bl; r++) {
            cur[0] = r;
            for (int l = 1; l <= wl; l++) {
                int sub = prev[l - 1] + (mapc(p[r - 1]) != mapc(s[i + l - 1]));
                int del = prev[l] + 1, ins = cur[l - 1] + 1;
                int m = sub < del ? sub : del;
                cur[l] = m < ins ? m : ins;
            }
            int *t = prev; prev = cur; cur = t;
        }
        for (int l = bl - mr; l <= wl; l++) if (prev[l] < best) best = prev[l];
    }
    return best;
}

static int dist_toprow(const uint8_t *p, int bl, const uint8_t *s, int64_t len, int k, int *colbuf) {
    /* column-major: col[r] = D[r][j]; top row per Appendix A.4 */
    int *col = colbuf;
    int top = valid_start(s, len, 0, k) ? 0 : INF;
    col[0] = top;
    for (int r = 1; r <= bl; r++) col[r] = col[r - 1] + 1;
    int best = col[bl];
    for (int64_t j = 1; j <= len; j++) {
        int diag = col[0];
        top = valid_start(s, len, j, k) ? 0 : (top >= INF ? INF : top + 1);
        col[0] = top;
        int c = mapc(s[j - 1]);
        for (int r = 1; r <= bl; r++) {
            int sub = diag + (mapc(p[r - 1]) != c);
            int left = col[r] + 1, up = col[r - 1] + 1;
            diag = col[r];
            int m = sub < left ? sub : left;
            col[r] = m < up ? m : up;
        }
        if (col[bl] < best) best = col[bl];
    }
    return best;
}

static int run(int mode, const uint8_t *barcodes, int B, int bl, int mr, int threads, int workers,
               const uint8_t *seg, const int64_t *off, int64_t n,
               int32_t *min_ed, int32_t *n_ties, int64_t *tie_off, int32_t **tie_ids) {
    int k = prefix_k(threads);
    if (!(bl > mr && mr >= 0) || k >= bl - mr || bl + mr >= 79) return -2;
    uint8_t *rcb = (uint8_t *)malloc((size_t)B * bl + 1);
    for (int b = 0; b < B; b++)
        for (int i = 0; i < bl; i++) rcb[(size_t)b * bl + i] = compc(barcodes[(size_t)b * bl + (bl - 1 - i)]);
    int32_t **lists = (int32_t **)calloc((size_t)(n > 0 ? n : 1), sizeof(int32_t *));
#ifdef _OPENMP
    if (workers > 0) omp_set_num_threads(workers);
#endif
#pragma omp parallel
    {
        int *colbuf = (int *)malloc((size_t)(bl + 2) * sizeof(int));
        int32_t *tmp = (int32_t *)malloc((size_t)(2 * B + 1) * sizeof(int32_t));
#pragma omp for schedule(dynamic, 64)
        for (int64_t r = 0; r < n; r++) {
            const uint8_t *s = seg + off[r];
            int64_t len = off[r + 1] - off[r];
            int best = mr + 1, cnt = 0;
            for (int b = 0; b < B; b++) {
                for (int strand = 0; strand < 2; strand++) { /* 0 = reverse complement first (False < True) */
                    const uint8_t *p = strand ? barcodes + (size_t)b * bl : rcb + (size_t)b * bl;
                    int d = mode == 0 ? dist_anchored(p, bl, mr, s, len, k) : dist_toprow(p, bl, s, len, k, colbuf);
                    if (d > mr || d > best) continue;
                    if (d < best) { best = d; cnt = 0; }
                    tmp[cnt++] = (b << 1) | strand;
                }
            }
            min_ed[r] = cnt ? best : -1;
            n_ties[r] = cnt;
            if (cnt) {
                lists[r] = (int32_t *)malloc((size_t)cnt * sizeof(int32_t));
                memcpy(lists[r], tmp, (size_t)cnt * sizeof(int32_t));
            }
        }
        free(colbuf); free(tmp);
    }
    tie_off[0] = 0;
    for (int64_t r = 0; r < n; r++) tie_off[r + 1] = tie_off[r] + n_ties[r];
    int32_t *ids = (int32_t *)malloc((size_t)(tie_off[n] ? tie_off[n] : 1) * sizeof(int32_t));
    for (int64_t r = 0; r < n; r++)
        if (n_ties[r]) { memcpy(ids + tie_off[r], lists[r], (size_t)n_ties[r] * sizeof(int32_t)); free(lists[r]); }
    free(lists); free(rcb);
    *tie_ids = ids;
    return 0;
}

int sct_oracle_anchored(const uint8_t *barcodes, int B, int bl, int mr, int threads, int workers,
                        const uint8_t *seg, const int64_t *off, int64_t n,
                        int32_t *min_ed, int32_t *n_ties, int64_t *tie_off, int32_t **tie_ids) {
    return run(0, barcodes, B, bl, mr, threads, workers, seg, off, n, min_ed, n_ties, tie_off, tie_ids);
}
int sct_oracle_dp(const uint8_t *barcodes, int B, int bl, int mr, int threads, int workers,
                  const uint8_t *seg, const int64_t *off, int64_t n,
                  int32_t *min_ed, int32_t *n_ties, int64_t *tie_off, int32_t **tie_ids) {
    return run(1, barcodes, B, bl, mr, threads, workers, seg, off, n, min_ed, n_ties, tie_off, tie_ids);
}
