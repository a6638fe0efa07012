/*
 * TEST INFRASTRUCTURE ONLY — CPU oracle for the adapter alignment of scTagger `extract_lr_bc`
 * (only tests/, __graft_entry__.smoke() and bench.py's CPU legs may call this; nothing under sctagger_b200/ does).
 *
 * PARITY UNPINNED: the arithmetic of this path lives in edlib (python-edlib, unpinned in the reference's env.yaml:9;
 * call sites scTagger.py:180-181: edlib.align(a, seq, 'HW', 'locations')), which is neither under /root/reference nor
 * installable here.  This file restates edlib's published semantics for that call:
 *
 *   editDistance = min over j >= 1 of D[m][j] of the semi-global DP (D[0][j] = 0, D[i][0] = i), unit costs,
 *                  characters compared literally (byte equality);  for an empty target: m.
 *   locations    = [(start, end)] for EVERY 0-based end position j-1 with D[m][j] == editDistance, increasing;
 *                  start = the LEFTMOST s with Levenshtein(query, target[s..end]) == editDistance (edlib aligns the
 *                  reversed query to the reversed target prefix in SHW mode and takes the last optimal position).
 *
 * and the decision of get_alns (scTagger.py:176-196): equal distances -> ('NA', -1, ['NA']); a1 better -> '+', locs =
 * ends; a2 better -> '-', locs = start - len(seq) - 1.
 *
 *   sct_oracle_align     : plain O(m n) DP, the checker.
 *   sct_oracle_align_bv  : Myers/Hyyro bit-vector restatement (what edlib itself runs), used as the timed CPU
 *                          baseline ("port") and cross-checked against the DP in tests/test_align_oracle.py.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXM 64

static uint8_t compc(uint8_t c) { return c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c; }

typedef struct { int32_t *v; int64_t n, cap; } ivec;
static void push(ivec *a, int32_t x) {
    if (a->n == a->cap) { a->cap = a->cap ? a->cap * 2 : 16; a->v = (int32_t *)realloc(a->v, (size_t)a->cap * 4); }
    a->v[a->n++] = x;
}

/* HW distance + all end positions at the minimum (plain DP, column by column) */
static int hw_dp(const uint8_t *q, int m, const uint8_t *t, int64_t n, ivec *ends) {
    int col[MAXM + 1];
    ends->n = 0;
    if (n == 0) return m;
    for (int i = 0; i <= m; i++) col[i] = i;
    int best = 1 << 30;
    for (int64_t j = 1; j <= n; j++) {
        int diag = col[0];
        col[0] = 0;
        for (int i = 1; i <= m; i++) {
            int sub = diag + (q[i - 1] != t[j - 1]);
            int up = col[i - 1] + 1, left = col[i] + 1;
            diag = col[i];
            int v = sub < up ? sub : up;
            col[i] = v < left ? v : left;
        }
        if (col[m] < best) { best = col[m]; ends->n = 0; }
        if (col[m] == best) push(ends, (int32_t)(j - 1));
    }
    return best;
}

/* leftmost s with Levenshtein(q, t[s..end]) == d: prefix-anchored DP of reversed q against t[end], t[end-1], ... */
static int64_t leftmost_start(const uint8_t *q, int m, const uint8_t *t, int64_t end, int d) {
    int col[MAXM + 1];
    for (int i = 0; i <= m; i++) col[i] = i;
    int64_t last = -1;
    int64_t maxp = end + 1 < (int64_t)m + d ? end + 1 : (int64_t)m + d;   /* longer substrings differ by > d in length */
    for (int64_t p = 1; p <= maxp; p++) {
        int diag = col[0];
        col[0] = (int)p;
        uint8_t c = t[end - (p - 1)];
        for (int i = 1; i <= m; i++) {
            int sub = diag + (q[m - i] != c);
            int up = col[i - 1] + 1, left = col[i] + 1;
            diag = col[i];
            int v = sub < up ? sub : up;
            col[i] = v < left ? v : left;
        }
        if (col[m] == d) last = p - 1;
    }
    return end - last;
}

/* ---- bit-vector restatement (m <= 64) ---- */
static int hw_bv(const uint64_t *peq, int m, const uint8_t *t, int64_t n, ivec *ends) {
    ends->n = 0;
    if (n == 0) return m;
    uint64_t Pv = m == 64 ? ~0ull : ((1ull << m) - 1), Mv = 0, top = 1ull << (m - 1);
    int score = m, best = 1 << 30;
    for (int64_t j = 0; j < n; j++) {
        uint64_t Eq = peq[t[j]];
        uint64_t Xv = Eq | Mv;
        uint64_t Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
        uint64_t Ph = Mv | ~(Xh | Pv);
        uint64_t Mh = Pv & Xh;
        if (Ph & top) score++;
        if (Mh & top) score--;
        Ph <<= 1;
        Mh <<= 1;
        Pv = Mh | ~(Xv | Ph);
        Mv = Ph & Xv;
        if (score < best) { best = score; ends->n = 0; }
        if (score == best) push(ends, (int32_t)j);
    }
    return best;
}

static int64_t leftmost_start_bv(const uint64_t *rpeq, int m, const uint8_t *t, int64_t end, int d) {
    uint64_t Pv = m == 64 ? ~0ull : ((1ull << m) - 1), Mv = 0, top = 1ull << (m - 1);
    int score = m;
    int64_t last = -1;
    int64_t maxp = end + 1 < (int64_t)m + d ? end + 1 : (int64_t)m + d;
    for (int64_t p = 0; p < maxp; p++) {
        uint64_t Eq = rpeq[t[end - p]];
        uint64_t Xv = Eq | Mv;
        uint64_t Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
        uint64_t Ph = Mv | ~(Xh | Pv);
        uint64_t Mh = Pv & Xh;
        if (Ph & top) score++;
        if (Mh & top) score--;
        Ph = (Ph << 1) | 1;              /* prefix mode: the top row grows by one per column */
        Mh <<= 1;
        Pv = Mh | ~(Xv | Ph);
        Mv = Ph & Xv;
        if (score == d) last = p;
    }
    return end - last;
}

/*
 * adapter: m bytes (a1); a2 = rev_compl(a1) is derived here (scTagger.py:363-365, identity on non-ACGT).
 * reads/off: concatenated read bytes and n+1 offsets.
 * Outputs per read: strand ('+', '-' or 0 for 'NA'), dist (-1 for NA), loc_off[n+1], *locs = malloc'd list
 * (ends for '+', start - len - 1 for '-'; empty for NA — the reference's ['NA']).  Free with sct_oracle_free.
 */
static int align_impl(const uint8_t *adapter, int m, const uint8_t *reads, const int64_t *off, int64_t n, int workers,
                      int use_bv, int8_t *strand, int32_t *dist, int64_t *loc_off, int32_t **locs) {
    if (m < 1 || m > MAXM) return -2;
    uint8_t a1[MAXM], a2[MAXM], r2[MAXM];
    for (int i = 0; i < m; i++) { a1[i] = adapter[i]; a2[i] = compc(adapter[m - 1 - i]); }
    for (int i = 0; i < m; i++) r2[i] = a2[m - 1 - i];
    uint64_t peq1[256], peq2[256], rpeq2[256];
    memset(peq1, 0, sizeof peq1); memset(peq2, 0, sizeof peq2); memset(rpeq2, 0, sizeof rpeq2);
    for (int i = 0; i < m; i++) { peq1[a1[i]] |= 1ull << i; peq2[a2[i]] |= 1ull << i; rpeq2[r2[i]] |= 1ull << i; }
    ivec *per = (ivec *)calloc((size_t)(n > 0 ? n : 1), sizeof(ivec));
#ifdef _OPENMP
    if (workers > 0) omp_set_num_threads(workers);
#endif
#pragma omp parallel
    {
        ivec e1 = {0, 0, 0}, e2 = {0, 0, 0};
#pragma omp for schedule(dynamic, 256)
        for (int64_t r = 0; r < n; r++) {
            const uint8_t *t = reads + off[r];
            int64_t len = off[r + 1] - off[r];
            int d1 = use_bv ? hw_bv(peq1, m, t, len, &e1) : hw_dp(a1, m, t, len, &e1);
            int d2 = use_bv ? hw_bv(peq2, m, t, len, &e2) : hw_dp(a2, m, t, len, &e2);
            ivec out = {0, 0, 0};
            if (d1 == d2) {
                strand[r] = 0; dist[r] = -1;
            } else if (d1 < d2) {
                strand[r] = '+'; dist[r] = d1;
                for (int64_t i = 0; i < e1.n; i++) push(&out, e1.v[i]);
            } else {
                strand[r] = '-'; dist[r] = d2;
                for (int64_t i = 0; i < e2.n; i++) {
                    int64_t s = use_bv ? leftmost_start_bv(rpeq2, m, t, e2.v[i], d2) : leftmost_start(a2, m, t, e2.v[i], d2);
                    push(&out, (int32_t)(s - len - 1));
                }
            }
            per[r] = out;
        }
        free(e1.v); free(e2.v);
    }
    loc_off[0] = 0;
    for (int64_t r = 0; r < n; r++) loc_off[r + 1] = loc_off[r] + per[r].n;
    int32_t *all = (int32_t *)malloc((size_t)(loc_off[n] > 0 ? loc_off[n] : 1) * 4);
    for (int64_t r = 0; r < n; r++) {
        if (per[r].n) memcpy(all + loc_off[r], per[r].v, (size_t)per[r].n * 4);
        free(per[r].v);
    }
    free(per);
    *locs = all;
    return 0;
}

int sct_oracle_align(const uint8_t *adapter, int m, const uint8_t *reads, const int64_t *off, int64_t n, int workers,
                     int8_t *strand, int32_t *dist, int64_t *loc_off, int32_t **locs) {
    return align_impl(adapter, m, reads, off, n, workers, 0, strand, dist, loc_off, locs);
}

int sct_oracle_align_bv(const uint8_t *adapter, int m, const uint8_t *reads, const int64_t *off, int64_t n, int workers,
                        int8_t *strand, int32_t *dist, int64_t *loc_off, int32_t **locs) {
    return align_impl(adapter, m, reads, off, n, workers, 1, strand, dist, loc_off, locs);
}

/* One query against one target, edlib.align(query, target, mode='HW', task='locations') restated: the distance and every
 * (start, end) location at it — used to check the restatement against the vectors edlib publishes in its own README
 * (tests/test_align_oracle.py).  starts / ends: caller-allocated, cap entries; returns the number of locations, the
 * distance in *dist (-1 and 0 locations for an empty target... edlib: distance = query length, no location). */
int sct_oracle_hw_locations(const uint8_t *q, int m, const uint8_t *t, int64_t n, int32_t *dist, int32_t *starts, int32_t *ends,
                            int cap) {
    ivec e = {0, 0, 0};
    const int d = hw_dp(q, m, t, n, &e);
    *dist = d;
    int k = 0;
    for (int64_t i = 0; i < e.n && k < cap; ++i, ++k) {
        ends[k] = e.v[i];
        starts[k] = (int32_t)leftmost_start(q, m, t, e.v[i], d);
    }
    free(e.v);
    return k;
}
