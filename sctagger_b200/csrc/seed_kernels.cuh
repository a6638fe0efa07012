// SEEDED match kernel (sm_100a): pigeonhole seeds from per-piece bucket indexes + exact bit-parallel verification,
// error bounds searched in turn.  Replaces the reference's Trie.dfs / trie_search / merge (scTagger.py:566-599, 643-668,
// 705-715) for every segment — clean ACGT segments of <= 31 bases as one window, longer ones and those with non-ACGT
// bytes as overlapping windows — but scores only the (barcode, strand) pairs that can possibly be within k edits instead
// of all 2B of them:
//
//   * ITERATIVE DEEPENING: phase k = 1, 2, .. max_error looks for every barcode within k edits and the search stops at the
//     first phase that finds one — the reference reports only the minimum distance and the barcodes tied at it, and phase
//     k is complete for every distance <= k (match_seed_kernel below).
//   * The barcode is cut into k+1 disjoint pieces.  k edits destroy at most k pieces, so a barcode within k edits of a
//     substring of the segment T has a piece occurring EXACTLY in T, at a text position j with
//     o_p - k <= j <= n - (bl - o_p) + k.  Every piece has a bucket table (bucket = 2-bit code of the piece) built once
//     per error bound in sct_create; a window's candidates are 2 strands x (k+1) pieces x <= 32 positions bucket ranges.
//   * One warp owns one segment.  Dense phases (seed_phase): the ranges' lengths are prefix-summed (warp scan) into a
//     per-warp slot table in shared memory and the flattened candidate list is walked 32 candidates at a time (no lane
//     idles on a short bucket); an anchored pre-check of the RESPONSIBLE seed (seed_precheck) discards most candidates
//     after a few automaton steps.  Sparse phases (seed_phase_direct): every lane probes one bucket and verifies its
//     entries directly.
//   * Each surviving candidate is scored exactly by a k-error Shift-And automaton with the roles swapped: the state vector
//     runs over the positions of T (one 32-bit register per level), one barcode character is consumed per step, and Eq[c] —
//     which positions of T hold character c — is warp-uniform per strand (pair tables in shared memory):
//         R0' = (R0 << 1) & Eq[c]
//         Rd' = ((Rd << 1) & Eq[c]) | R_{d-1} | (R_{d-1} << 1) | (R'_{d-1} << 1)         match | delete | subst | insert
//     start R_d = all positions (free start in T; the window's slice of the -t prefix rule's start states for segments with
//     non-ACGT bytes); after bl steps the barcode is within d edits iff R_d & mask(n+1) != 0.
//     The shifts are IMADs by a run-time 2 (FMA pipe), so the ALU pipe only sees the LOP3s and the character extract.
//   * Reverse strand: d(revcomp(P), T) = d(P, revcomp(T)): revcomp(T) is scored against the same forward barcodes.
//   * Hits (rare) are folded into the warp's (minimum, tie list); the list is sorted and de-duplicated at the end (a
//     barcode is found once per surviving piece and window) so tie ids come out ascending = the reference's
//     sorted((bid, strand)).
// tests/index_model.py is the Python twin of this file (one-window and windowed variants).
#pragma once
#include "match_kernels.cuh"

namespace sct {

#ifndef SEED_ANCHOR
#define SEED_ANCHOR 1
#endif
#ifndef SEED_MIN_CTAS
#define SEED_MIN_CTAS 5                    // resident CTAs per SM the hot variant is compiled for (48 registers)
#endif
constexpr int SEED_THREADS = 256;
constexpr int SEED_WARPS = SEED_THREADS / 32;
constexpr int SEED_LIST_CAP = 96;          // hits staged per warp before sort + unique
constexpr int SEED_MAX_LEN = 31;           // n + 1 state bits must fit one register
constexpr int SEED_MAX_PIECES = 4;

struct SeedIndex {
    const uint32_t *boff;                  // per piece: 4^len + 1 absolute entry offsets, pieces concatenated
    const uint2 *ewords;                   // entries: .x = barcode as 2-bit codes (character i at bits 2i..2i+1),
                                           //          .y = its reverse complement in the same packing (bl = 16 only)
    const uint32_t *eids;                  // entries: barcode index
    int n_pieces;
    int piece_off[SEED_MAX_PIECES];        // offset of the piece in the barcode
    int piece_len[SEED_MAX_PIECES];
    int boff_base[SEED_MAX_PIECES];        // start of the piece's bucket table inside boff
    uint32_t two;                          // always 2 (run-time so that x*2 stays an FMA-pipe IMAD)
    // anchored pre-check per piece: bits 0-4 = characters to skip (right: piece end; left: bl - piece offset),
    // bit 5 = extend to the left of the seed, bits 9-13 = piece length, bit 14 = test level 1 after 5 characters
    uint32_t pre[SEED_MAX_PIECES];
    int prefilter;                         // 1 = run the anchored pre-check (option "seed_prefilter")
    int first_phase;                       // 1 = this phase builds the segment's Eq tables (the first one that runs)
    int direct;                            // 1 = sparse index: probe and verify per lane, no slot table (seed_phase_direct)
};
#ifndef SEED_BATCH_N
#define SEED_BATCH_N 4
#endif
#ifndef SEED_TIE_CHUNK_N
#define SEED_TIE_CHUNK_N 64
#endif
constexpr int SEED_BATCH = SEED_BATCH_N;   // segments a warp takes per visit of the work cursor
constexpr int SEED_QUEUE_CAP = 64;         // pre-check survivors waiting for a full verification, per warp

__device__ __forceinline__ uint32_t mul_rt(uint32_t x, uint32_t m) {
    uint32_t r;
    asm("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(m));
    return r;
}

// (a & b) | c in one LOP3
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

// Exact level (0..K, or K+1) of one barcode word against one strand of the text.  tab_a is the shared-memory address
// (128-byte aligned) of that strand's PAIR table: entry e = c0 | c1 << 2 holds {Eq[c0], Eq[c1]}, so one LDS.64 feeds
// two steps of the automaton and the character extract is one shift + one LOP3 per two barcode characters.
// INIT: the level-d start state is read from shared memory (v_a: 4 words, bit j = a match may start at text position j
// with d edits already spent — segments with restricted starts) instead of "every position".
template <int K, bool BL16, bool INIT>
__device__ __forceinline__ int seed_verify(uint32_t word, int bl, uint32_t tab_a, uint32_t full, uint32_t two, uint32_t v_a) {
    uint32_t R[K + 1], S[K + 1];
    if (INIT) {
        uint32_t v[4];
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(v_a));
#pragma unroll
        for (int d = 0; d <= K; ++d) {
            R[d] = v[d];
            S[d] = v[d] << 1;
        }
    } else {
#pragma unroll
        for (int d = 0; d <= K; ++d) {
            R[d] = full;
            S[d] = full << 1;
        }
    }
    auto row = [&](uint32_t eq) {
        uint32_t r = S[0] & eq;
        uint32_t s = mul_rt(r, two);
        uint32_t pR = R[0], pS = S[0], nS = s;
        R[0] = r;
        S[0] = s;
#pragma unroll
        for (int d = 1; d <= K; ++d) {
            const uint32_t t = pR | pS | nS;
            r = (S[d] & eq) | t;
            s = mul_rt(r, two);
            pR = R[d];
            pS = S[d];
            nS = s;
            R[d] = r;
            S[d] = s;
        }
    };
    if (BL16) {
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const uint32_t sh = m == 0 ? (word << 3) : (word >> (4 * m - 3));
            uint32_t e0, e1;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(e0), "=r"(e1) : "r"(and_or(sh, 0x78u, tab_a)));
            row(e0);
            row(e1);
        }
    } else {
        for (int i = 0; i < bl; ++i) {
            uint32_t e0;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e0) : "r"(tab_a + (((word >> (2 * i)) & 3u) << 3)));
            row(e0);
        }
    }
    int level = K + 1;
#pragma unroll
    for (int d = K; d >= 0; --d)
        if (R[d] & full) level = d;
    return level;
}

// Anchored pre-check of one candidate: the seed piece matched the text exactly at position j, so a barcode within K edits
// THROUGH this seed has the characters next to the piece within few edits of the text next to the match.  The same
// automaton runs for a few characters from the seed boundary outwards, started at the boundary only (bits b0..b0+d at
// level d: d leading text insertions).  Extending to the left is extending the other strand's text to the right with
// the reverse complement of the barcode, so both directions use the same pair tables.  A candidate that fails cannot be
// the RESPONSIBLE seed of any alignment within K edits, and every barcode within K edits has a responsible seed that
// passes — which seed that is, and what it may assume about its neighbours, depends on the error bound (bl = 16):
//   K = 1, pieces 8+8:      the exact piece's sibling has <= 1 error: 6 characters of it, levels 0..1.
//   K = 2, pieces 5+5+6:    responsible = the first exact piece in the order (piece 2, piece 1, piece 0).  Piece 2 (the
//                           6-mer, fewest candidates) may assume nothing: 6 characters to its left within 2 edits.  Piece
//                           1 is responsible only if piece 2 has an error, so piece 0 — the 5 characters to its left — has
//                           <= 1.  Piece 0 is responsible only if pieces 1 and 2 both have exactly one error: the 5
//                           characters to its right within 1 edit.  One automaton (levels 0..2, 6 steps) serves all three:
//                           level 1 is sampled after the fifth character and a per-slot flag picks the test.
//   K = 3, pieces 4+4+4+4:  of the two halves one has <= 1 error, and in it one piece is exact and its sibling has <= 1:
//                           responsible = that piece; test = its sibling's 4 characters within 1 edit (levels 0..1, 4 steps).
// Measured at -mr 2 with the 6-character / 2-edit test on every slot: 89.8 % of the candidates stop here.
__host__ __device__ constexpr int seed_pre_steps(int K) { return K >= 3 ? 4 : 6; }
__host__ __device__ constexpr int seed_pre_levels(int K) { return K == 2 ? 2 : 1; }   // highest level computed
template <int K>
// aux (per slot, built once per segment): bits 0-5 = right shift of the barcode word, bit 6 = extend left (use the
// reverse-complemented word), bit 7 = which strand's pair table, bit 9 = (K = 2) test level 1 after 5 characters,
// bits 16-31 = byte offset of the start states (32 * b0)
__device__ __forceinline__ bool seed_precheck(uint2 words, uint32_t aux, uint32_t pair_a, uint32_t full, uint32_t two,
                                              uint32_t init_a) {
    constexpr int LV = seed_pre_levels(K);
    constexpr int STEPS = seed_pre_steps(K);
    const uint32_t w = ((aux & 64u) ? words.y : words.x) >> (aux & 63u);
    const uint32_t tab_a = pair_a + (aux & 128u);
    const uint32_t st_a = init_a + (aux >> 16);
    // start state of level d = bits b0 .. b0+d, from a per-CTA table indexed by b0: {R0, S0, R1, S1}, {R2, S2, R3, S3}
    uint32_t R[4], S[4];
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(R[0]), "=r"(S[0]), "=r"(R[1]), "=r"(S[1]) : "r"(st_a));
    if (LV >= 2)
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(R[2]), "=r"(S[2]), "=r"(R[3]), "=r"(S[3]) : "r"(st_a + 16u));
    auto row = [&](uint32_t eq) {
        uint32_t r = S[0] & eq;
        uint32_t s = mul_rt(r, two);
        uint32_t pR = R[0], pS = S[0], nS = s;
        R[0] = r;
        S[0] = s;
#pragma unroll
        for (int d = 1; d <= LV; ++d) {
            const uint32_t t = pR | pS | nS;
            r = (S[d] & eq) | t;
            s = mul_rt(r, two);
            pR = R[d];
            pS = S[d];
            nS = s;
            R[d] = r;
            S[d] = s;
        }
    };
    uint32_t r1_at5 = 0u;
#pragma unroll
    for (int m = 0; m < STEPS / 2; ++m) {
        const uint32_t sh = m == 0 ? (w << 3) : (w >> (4 * m - 3));
        uint32_t e0, e1;
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(e0), "=r"(e1) : "r"(and_or(sh, 0x78u, tab_a)));
        row(e0);
        if (K == 2 && m == 2) r1_at5 = R[1];
        row(e1);
    }
    if (K == 2) return (((aux & 512u) ? r1_at5 : R[2]) & full) != 0u;
    return (R[1] & full) != 0u;
}

// The K = 2 test of the two 5-mer pieces alone (neighbouring 5 characters within ONE edit): levels 0..1, 5 steps.  Used
// for a block of 32 candidates whose slots all carry the level-1 flag — most blocks, since the candidate list is ordered by
// (strand, piece, position) and the 5-mers own 90 % of it; a mixed block runs the general automaton above.
__device__ __forceinline__ bool seed_precheck_l1x5(uint2 words, uint32_t aux, uint32_t pair_a, uint32_t full, uint32_t two,
                                                   uint32_t init_a) {
    const uint32_t w = ((aux & 64u) ? words.y : words.x) >> (aux & 63u);
    const uint32_t tab_a = pair_a + (aux & 128u);
    uint32_t R0, S0, R1, S1;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(R0), "=r"(S0), "=r"(R1), "=r"(S1) : "r"(init_a + (aux >> 16)));
    auto row = [&](uint32_t eq) {
        const uint32_t r0 = S0 & eq;
        const uint32_t s0 = mul_rt(r0, two);
        const uint32_t t = R0 | S0 | s0;
        const uint32_t r1 = (S1 & eq) | t;
        R0 = r0;
        S0 = s0;
        R1 = r1;
        S1 = mul_rt(r1, two);
    };
#pragma unroll
    for (int m = 0; m < 3; ++m) {
        const uint32_t sh = m == 0 ? (w << 3) : (w >> (4 * m - 3));
        uint32_t e0, e1;
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(e0), "=r"(e1) : "r"(and_or(sh, 0x78u, tab_a)));
        row(e0);
        if (m < 2) row(e1);
    }
    return (R1 & full) != 0u;
}

// Sort the warp's hit list ascending and drop duplicates (rank sort through a second buffer).  Returns the new count.
__device__ __forceinline__ int sort_unique(uint32_t *list, uint32_t *tmp, int count, int lane) {
    if (count <= 1) return count;
    for (int e = lane; e < count; e += 32) {
        const uint32_t v = list[e];
        int rank = 0;
        for (int j = 0; j < count; ++j) {
            const uint32_t u = list[j];
            rank += (u < v) || (u == v && j < e);
        }
        tmp[rank] = v;
    }
    __syncwarp();
    int out = 0;
    for (int base = 0; base < count; base += 32) {
        const int e = base + lane;
        bool keep = false;
        uint32_t v = 0;
        if (e < count) {
            v = tmp[e];
            keep = e == 0 || tmp[e - 1] != v;
        }
        const unsigned m = __ballot_sync(FULL, keep);
        if (keep) list[out + __popc(m & ((1u << lane) - 1u))] = v;
        out += __popc(m);
    }
    __syncwarp();
    return out;
}

// Per-warp views of the kernel's shared memory.
struct SeedWarp {
    uint32_t *prefix, *start, *aux, *queue, *map, *list, *tmp, *v;
    uint2 *pair;                 // [2][16]
    unsigned long long *cnt;     // work counters of the warp
    uint32_t pair_a, v_a, init_a, queue_a;
};
constexpr int SEED_MAPW = 128;   // slot-start bitmap: candidates 1..4096 of a window

// The text of one window: packed forward text, its reverse complement, (GEN) the start states of the window, and —
// when asked — the Eq pair tables of both strands in the warp's shared memory.
template <bool GEN>
__device__ __forceinline__ void seed_window(const MatchParams &p, const SeedWarp &w, const int lane, const int seg,
                                            const long long so, const int n_total, const unsigned long long *__restrict__ pw,
                                            const int ws, const bool build_eq, const unsigned long long fw0, int &n,
                                            unsigned long long &fw, unsigned long long &rc) {
    if (!GEN) {
        n = n_total;
        fw = fw0;                                                     // fetched with the batch
    } else {
        n = min(SEED_MAX_LEN, n_total - ws);
        const unsigned long long a = __ldg(pw + (ws >> 5)), b = __ldg(pw + (ws >> 5) + 1);
        const int sh = 2 * (ws & 31);
        fw = sh ? ((a >> sh) | (b << (64 - sh))) : a;
        fw &= (1ull << (2 * n)) - 1ull;
        // start states: bit j of level d = [D0[ws + j] <= d], j = 0..n
        uint32_t vb = 0;
        if (lane <= n) {
            const int pos = ws + lane;
            vb = pos == 0 ? ((p.segflag[seg] & SEG_START0) ? 0xFu : 0u) : (uint32_t)(p.colinfo[so + pos - 1] >> 3);
        }
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            const uint32_t m = __ballot_sync(FULL, (vb >> d) & 1u);
            if (lane == d) w.v[d] = m;
        }
    }
    // reverse complement of the packed text: complement, reverse the 2-bit groups, drop the unused tail
    rc = __brevll(~fw);
    rc = ((rc & 0x5555555555555555ull) << 1) | ((rc >> 1) & 0x5555555555555555ull);
    rc >>= (64 - 2 * n);
    // Eq masks (bit j+1 of Eq[c] = [T[j] == c]) of both strands, stored as pair tables: lane = strand * 16 + entry
    // (a later phase of a one-window segment finds them in place)
    if (build_eq) {
        const uint32_t cj = (uint32_t)(fw >> (2 * lane)) & 3u;
        const bool in = lane < n;
        uint32_t bf[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) bf[c] = __ballot_sync(FULL, in && cj == (uint32_t)c);
        const bool fwd = lane >= 16;                                  // strand 1 = forward text
        auto eq_of = [&](uint32_t c) -> uint32_t {
            const uint32_t cc = fwd ? c : 3u - c;                     // reverse strand: complement ...
            const uint32_t b = cc == 0 ? bf[0] : cc == 1 ? bf[1] : cc == 2 ? bf[2] : bf[3];
            return fwd ? (b << 1) : (__brev(b) >> (31 - n));          // ... and reversed positions
        };
        w.pair[lane] = make_uint2(eq_of(lane & 3u), eq_of((lane >> 2) & 3u));
    }
}

// Fold the exact levels of up to 32 candidates (one per lane, K + 1 = no match) into the warp's running (minimum, tie list).
__device__ __forceinline__ void seed_fold(const int level, const uint32_t entry, const uint32_t strand, const int K,
                                          const uint32_t *__restrict__ eids, uint32_t *list, const uint32_t lt_mask,
                                          int &best, int &count) {
    const unsigned hit = __ballot_sync(FULL, level <= K);
    if (hit) {
        const int m = __reduce_min_sync(FULL, level);
        if (m <= best) {
            if (m < best) {
                best = m;
                count = 0;
            }
            const unsigned mine = __ballot_sync(FULL, level == m);
            if (level == m) {
                const int at = count + __popc(mine & lt_mask);
                if (at < SEED_LIST_CAP) list[at] = (__ldg(eids + entry) << 1) | strand;
            }
            count += __popc(mine);
        }
    }
}

// Exact level (0, 1, or 2 = none) of one candidate of the error-bound-1 phase (pieces 8+8, bl = 16, clean one-window
// segment), ANCHORED at its seed: the piece matched the text exactly, so within one edit through this seed means that
// the other half — 8 characters outwards from the seed boundary, same automaton and start states as seed_precheck — costs
// 0 or 1.  A barcode within one edit has an exact half in its optimal alignment and is found through that seed with its
// true distance; through any other seed occurrence it can only score the same or worse, and the fold keeps the minimum.
// Half the steps of a full verification, and two levels.
__device__ __forceinline__ int seed_anchor_level1(uint2 words, uint32_t aux, uint32_t pair_a, uint32_t full, uint32_t two,
                                                  uint32_t init_a) {
    const uint32_t w = ((aux & 64u) ? words.y : words.x) >> (aux & 63u);
    const uint32_t tab_a = pair_a + (aux & 128u);
    uint32_t R0, S0, R1, S1;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(R0), "=r"(S0), "=r"(R1), "=r"(S1) : "r"(init_a + (aux >> 16)));
    auto row = [&](uint32_t eq) {
        const uint32_t r0 = S0 & eq;
        const uint32_t s0 = mul_rt(r0, two);
        const uint32_t t = R0 | S0 | s0;
        const uint32_t r1 = (S1 & eq) | t;
        R0 = r0;
        S0 = s0;
        R1 = r1;
        S1 = mul_rt(r1, two);
    };
#pragma unroll
    for (int m = 0; m < 4; ++m) {
        const uint32_t sh = m == 0 ? (w << 3) : (w >> (4 * m - 3));
        uint32_t e0, e1;
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(e0), "=r"(e1) : "r"(and_or(sh, 0x78u, tab_a)));
        row(e0);
        row(e1);
    }
    return (R0 & full) ? 0 : ((R1 & full) ? 1 : 2);
}

// A phase whose index is SPARSE (fewer than one barcode per bucket on average: the two 8-mers of error bound 1 at
// B = 25 k): no slot table, no queue, no pre-check.  Every lane probes one (strand, piece, position) bucket and the warp
// verifies bucket entry 0 of every lane, then entry 1, ... — one or two rounds per window.  Error bound 1 on a clean
// one-window segment verifies anchored at the seed (8 steps, 2 levels) instead of over the whole barcode.
template <int KP, bool BL16, bool GEN>
__device__ __forceinline__ void seed_phase_direct(const MatchParams &p, const SeedIndex &ix, const SeedWarp &w, const int lane,
                                                  const int seg, const long long so, const int n_total,
                                                  const unsigned long long *__restrict__ pw, const unsigned long long fw0, const int bl,
        int &best, int &count) {
    constexpr int K = KP;
    constexpr bool ANCH = SEED_ANCHOR && BL16 && !GEN && KP == 1;
    const int P = ix.n_pieces;
    const int stride = 32 - bl - K;
    const uint32_t lt_mask = (1u << lane) - 1u;
    best = K + 1;
    count = 0;
    for (int ws = 0;; ws += stride) {
        int n;
        unsigned long long fw, rc;
        seed_window<GEN>(p, w, lane, seg, so, n_total, pw, ws, GEN || ix.first_phase != 0, fw0, n, fw, rc);
        const uint32_t full = (n >= 31) ? 0xFFFFFFFFu : ((1u << (n + 1)) - 1u);
        int lo[K + 1], cnt[K + 1], per = 0;
#pragma unroll
        for (int q = 0; q <= K; ++q) {
            lo[q] = cnt[q] = 0;
            if (q < P) {
                const int o = ix.piece_off[q], ln = ix.piece_len[q];
                lo[q] = max(0, o - K);
                cnt[q] = max(0, min(n - ln, n - (bl - o) + K) - lo[q] + 1);
                per += cnt[q];
            }
        }
        __syncwarp();                                                  // Eq tables / start states visible
        uint32_t n_ver = 0;
        auto verify = [&](uint32_t entry, uint32_t strand, uint32_t aux, bool valid) {
            int level = K + 1;
            if (valid) {
                if (ANCH) {
                    level = seed_anchor_level1(__ldg(ix.ewords + entry), aux, w.pair_a, full, ix.two, w.init_a);
                } else if (GEN) {
                    const uint2 words = __ldg(ix.ewords + entry);
                    level = seed_verify<K, BL16, true>(strand ? words.x : words.y, bl, w.pair_a + 128u, full, ix.two, w.v_a);
                } else {
                    const uint32_t word = __ldg(&ix.ewords[entry].x);
                    level = seed_verify<K, BL16, false>(word, bl, w.pair_a + strand * 128u, full, ix.two, 0u);
                }
            }
            seed_fold(level, entry, strand, K, ix.eids, w.list, lt_mask, best, count);
        };
        for (int c0 = 0; c0 < 2 * per; c0 += 32) {
            const int c = c0 + lane;
            const uint32_t strand = c >= per ? 1u : 0u;
            int r = c - (strand ? per : 0);
            uint32_t st = 0, en = 0, aux = 0;
            if (c < 2 * per) {
                const unsigned long long text = strand ? fw : rc;
#pragma unroll
                for (int q = 0; q <= K; ++q) {
                    if (q < P) {
                        if (r >= 0 && r < cnt[q]) {
                            const int j = lo[q] + r;
                            const uint32_t h = (uint32_t)(text >> (2 * j)) & ((1u << (2 * ix.piece_len[q])) - 1u);
                            const uint32_t *b = ix.boff + ix.boff_base[q] + h;
                            st = __ldg(b);
                            en = __ldg(b + 1);
                            if (ANCH) {                               // operands of the anchored extension (see seed_precheck)
                                const uint32_t pre = ix.pre[q], left = (pre >> 5) & 1u;
                                const uint32_t b0 = left ? (uint32_t)(n - j) : (uint32_t)j + ((pre >> 9) & 31u);
                                aux = (2u * (pre & 31u)) | (left << 6) | ((strand ^ left) << 7) | (b0 << 21);
                            }
                        }
                        r -= cnt[q];
                    }
                }
            }
            // entry 0 of every lane's bucket, then entry 1, ..: one or two rounds (compacting the few entries into
            // full rounds through shared memory was measured slower: the extra round trip costs more than the idle lanes)
            while (__any_sync(FULL, st < en)) {
                const bool valid = st < en;
                n_ver += (uint32_t)__popc(__ballot_sync(FULL, valid));
                verify(st, strand, aux, valid);
                st += valid ? 1u : 0u;
            }
        }
        if (lane == 0) {
            w.cnt[0] += (unsigned long long)(2 * per);
            w.cnt[1] += n_ver;
            w.cnt[2] += n_ver;
            w.cnt[3] += 1ull;
            w.cnt[4] += (unsigned long long)(n_ver * (uint32_t)(ANCH ? 8 : bl) * (uint32_t)(2 * K + 1));
        }
        __syncwarp();
        if (!GEN || ws + SEED_MAX_LEN >= n_total) break;
    }
}

// One PHASE of the search for one segment: every (barcode, strand) within KP edits, exactly (pigeonhole seeds of the
// phase's index -> anchored pre-check -> full verification), folded into (best, count, list).  best = KP + 1 when there
// is none.
// GEN = false: clean segments of <= 31 bases (fast_list), one window, every start allowed, reverse strand scored as the
//               forward barcode against revcomp(T).
// GEN = true:  any other segment (wide_list): overlapping windows of <= 31 bases, window w starting at base w * stride,
//               stride = 32 - bl - KP, so that every substring of <= bl + KP bases — every alignment within KP edits —
//               lies inside one window; min over windows = min over the segment, ties are the union (sort_unique drops the
//               duplicates of overlapping windows).  Non-ACGT bytes are scored as A (pk) and restrict where a match may
//               START (reference: the -t prefix rule): the level-d start state of the automaton is
//               the window's slice of [D0 <= d] (v bits of colinfo) instead of "everywhere".  A start restriction on T is
//               an END restriction on revcomp(T), so here the reverse strand is scored as revcomp(barcode) (the entry's
//               second word) against the forward text.  Seeds and the anchored pre-check ignore the restriction: they stay
//               necessary conditions.
template <int KP, bool BL16, bool PRE, bool GEN>
__device__ __forceinline__ void seed_phase(const MatchParams &p, const SeedIndex &ix, const SeedWarp &w, const int lane,
                                           const int seg, const long long so, const int n_total,
                                           const unsigned long long *__restrict__ pw, const unsigned long long fw0, const int bl,
        int &best, int &count) {
    constexpr int K = KP;
    const int P = ix.n_pieces;
    const int stride = 32 - bl - K;                                      // GEN: window step (>= 13 for bl <= 16, K <= 3)
    const uint32_t lt_mask = (1u << lane) - 1u;
    const bool pre_on = PRE && ix.prefilter != 0;
    uint32_t *prefix = w.prefix, *start = w.start, *list = w.list;
    const uint32_t queue_a = w.queue_a;
    auto q_st = [&](int i, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(queue_a + 4u * (uint32_t)i), "r"(v) : "memory"); };
    auto q_ld = [&](int i) -> uint32_t {
        uint32_t v;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(queue_a + 4u * (uint32_t)i) : "memory");
        return v;
    };
    const uint32_t pair_a = w.pair_a;
    best = K + 1;
    count = 0;
    for (int ws = 0;; ws += stride) {                                    // windows (exactly one when !GEN)
        int n;
        unsigned long long fw, rc;
        seed_window<GEN>(p, w, lane, seg, so, n_total, pw, ws, GEN || ix.first_phase != 0, fw0, n, fw, rc);
        // bucket ranges of every (strand, piece, position); the non-empty ones become slots, prefix-summed in order
        uint32_t total = 0;
        int n_slots = 0;
        uint32_t n_probe = 0;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const unsigned long long text = s ? fw : rc;               // strand 0 first
#pragma unroll
            for (int q = 0; q <= K; ++q) {
                uint32_t st = 0, len = 0;
                if (q < P) {
                    const int o = ix.piece_off[q], ln = ix.piece_len[q];
                    const int lo = max(0, o - K), hi = min(n - ln, n - (bl - o) + K);
                    n_probe += (uint32_t)max(0, hi - lo + 1);
                    if (lane >= lo && lane <= hi) {
                        const uint32_t h = (uint32_t)(text >> (2 * lane)) & ((1u << (2 * ln)) - 1u);
                        const uint32_t *b = ix.boff + ix.boff_base[q] + h;
                        st = __ldg(b);
                        len = __ldg(b + 1) - st;
                    }
                }
                uint32_t incl = len;
#pragma unroll
                for (int o2 = 1; o2 < 32; o2 <<= 1) {
                    const uint32_t v = __shfl_up_sync(FULL, incl, o2);
                    if (lane >= o2) incl += v;
                }
                const unsigned live = __ballot_sync(FULL, len != 0u);
                if (len) {
                    const int slot = n_slots + __popc(live & lt_mask);
                    const uint32_t first = total + incl - len;
                    prefix[slot] = first;
                    start[slot] = st - first;                        // entry of candidate t = start[slot] + t
                    uint32_t aux = (uint32_t)s << 8;
                    if (PRE) {
                        const uint32_t pre = ix.pre[q], left = (pre >> 5) & 1u;
                        const uint32_t b0 = left ? (uint32_t)(n - lane) : (uint32_t)lane + ((pre >> 9) & 31u);
                        aux |= (2u * (pre & 31u)) | (left << 6) | ((((uint32_t)s) ^ left) << 7) | (((pre >> 14) & 1u) << 9) | (b0 << 21);
                    }
                    w.aux[slot] = aux;
                }
                n_slots += __popc(live);
                total += __shfl_sync(FULL, incl, 31);
            }
        }
        prefix[n_slots + lane] = 0xFFFFFFFFu;
        const bool use_map = total <= 32u * SEED_MAPW;
        if (use_map)
            for (uint32_t x = lane; x < (total + 31u) / 32u; x += 32) w.map[x] = 0u;
        __syncwarp();
        if (use_map) {
            for (int i = 1 + lane; i < n_slots; i += 32) {
                const uint32_t t = prefix[i] - 1u;
                atomicOr(&w.map[t >> 5], 1u << (t & 31u));
            }
            __syncwarp();
        }
        const uint32_t full = (n >= 31) ? 0xFFFFFFFFu : ((1u << (n + 1)) - 1u);
        uint32_t n_ver = 0, n_cheap = 0;
        // exact level of up to 32 queued candidates (lane < cnt), folded into the running (minimum, tie list)
        auto verify = [&](uint32_t item, bool valid) {
            int level = K + 1;
            const uint32_t entry = item & 0x7FFFFFFFu, strand = item >> 31;
            if (valid) {
                if (GEN) {
                    const uint2 words = __ldg(ix.ewords + entry);
                    level = seed_verify<K, BL16, true>(strand ? words.x : words.y, bl, pair_a + 128u, full, ix.two, w.v_a);
                } else {
                    const uint32_t word = __ldg(&ix.ewords[entry].x);
                    level = seed_verify<K, BL16, false>(word, bl, pair_a + strand * 128u, full, ix.two, 0u);
                }
            }
            seed_fold(level, entry, strand, K, ix.eids, list, lt_mask, best, count);
        };
        int qn = 0;
        int cur = 0;                                                    // slot holding candidate t0 (warp-uniform)
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            // slots starting at t0 + r, 1 <= r <= 32, as bit r-1: a lane's slot is cur + #starts at or before it
            uint32_t heads;
            if (use_map) {
                heads = w.map[t0 >> 5];
            } else {
                const uint32_t rel = prefix[cur + 1 + lane] - t0 - 1u;
                heads = __reduce_or_sync(FULL, rel < 32u ? (1u << rel) : 0u);
            }
            const int pos = cur + __popc(heads & lt_mask);
            cur += __popc(heads);
            const uint32_t me = t0 + lane;
            // branch-free: the lanes past the end of the last block run the check on whatever follows the last bucket
            // (the entry array is padded by 32 entries) and are masked out
            const uint32_t entry = start[pos] + me;
            const uint32_t aux = w.aux[pos];
            const uint32_t item = entry | ((aux & 256u) << 23);           // strand in bit 31
            bool pass = me < total;
            if (BL16 && PRE) {
                if (pre_on) {
                    const uint2 words = __ldg(ix.ewords + entry);
                    if (K == 2 && __all_sync(FULL, (aux & 512u) != 0u || !pass)) {
                        pass = seed_precheck_l1x5(words, aux, pair_a, full, ix.two, w.init_a) && pass;
                        n_cheap += min(32u, total - t0);
                    } else
                        pass = seed_precheck<K>(words, aux, pair_a, full, ix.two, w.init_a) && pass;
                }
            }
            const unsigned pm = __ballot_sync(FULL, pass);
            if (pm) {                                                     // warp-uniform: four blocks in ten have no survivor
                if (pass) q_st(qn + __popc(pm & lt_mask), item);
                qn += __popc(pm);
                if (qn >= 32) {
                    __syncwarp();
                    verify(q_ld(lane), true);
                    const uint32_t rest = (32 + lane < qn) ? q_ld(32 + lane) : 0u;
                    __syncwarp();
                    q_st(lane, rest);
                    qn -= 32;
                    n_ver += 32u;
                    __syncwarp();
                }
            }
        }
        __syncwarp();
        if (qn > 0) verify(lane < qn ? q_ld(lane) : 0u, lane < qn);
        n_ver += (uint32_t)qn;
        if (lane == 0) {
            w.cnt[0] += n_probe;
            w.cnt[1] += n_ver;
            w.cnt[2] += total;
            w.cnt[3] += 1ull;
            // automaton steps performed (x (2K+1) LOP3 each): full verifications + anchored pre-checks
            w.cnt[4] += (unsigned long long)((n_ver * (uint32_t)bl + 0u) * (uint32_t)(2 * K + 1)) +
                        (pre_on ? (unsigned long long)(total - n_cheap) * (uint32_t)(seed_pre_steps(K) * (2 * seed_pre_levels(K) + 1)) +
                                      (unsigned long long)n_cheap * 15ull : 0ull);
            if (pre_on) w.cnt[5] += total;
        }
        __syncwarp();
        if (!GEN || ws + SEED_MAX_LEN >= n_total) break;
    }
}

// store_result of the seeded kernels: tie space comes from a per-warp CHUNK of the bump-allocated tie buffer (one atomic
// on the global cursor per SEED_TIE_CHUNK ids instead of one per matched segment; finalize gathers the lists into segment
// order wherever they lie).  The unused tails of the chunks are why the host sizes the buffer with seed_tie_slack().
constexpr int SEED_TIE_CHUNK = SEED_TIE_CHUNK_N;
__device__ __forceinline__ void seed_store_result(const MatchParams &p, int seg, int lane, int K, int best, int count,
                                                  const uint32_t *list, unsigned long long *chunk) {
    if (count > SEED_LIST_CAP) {         // the list is incomplete: the segment is re-run with a full-size list
        if (lane == 0) {
            const unsigned int q = atomicAdd(&p.ctr->n_ovf, 1u);
            p.ovf_list[q] = seg;
            p.min_ed[seg] = best;
            p.n_ties[seg] = 0;
        }
        return;
    }
    unsigned long long pos = 0;
    if (lane == 0) {
        p.min_ed[seg] = (best <= K) ? best : -1;
        p.n_ties[seg] = count;
        if (count > 0) {
            unsigned long long at = chunk[0], left = chunk[1];
            if ((unsigned long long)count > left) {
                left = (unsigned long long)max(count, SEED_TIE_CHUNK);
                at = atomicAdd(&p.ctr->tie_cursor, left);
            }
            pos = at;
            chunk[0] = at + (unsigned long long)count;
            chunk[1] = left - (unsigned long long)count;
            p.tie_start[seg] = (unsigned int)pos;
        }
    }
    if (count > 0) {
        pos = __shfl_sync(FULL, pos, 0);
        if (pos + (unsigned long long)count <= p.tie_cap)
            for (int i = lane; i < count; i += 32) p.tie_buf[pos + i] = (int)list[i];
    }
}

// The seeded match kernel.  One warp per segment; ITERATIVE DEEPENING over the error bound: phase k' looks for every
// barcode within k' edits with the seeds of k' + 1 pieces (k' = 1: two 8-mers, ~12 candidates per segment at B = 25 k;
// k' = 2: 5+5+6, ~1 100; k' = 3: four 4-mers, ~9 300) and the search stops at the first phase that finds something: the
// reference reports only the minimum distance and the barcodes tied at it (scTagger.py:643-668), and phase k' is complete
// for every distance <= k', so a segment with an exact or one-error hit never pays for the wide search.  ixs.ph[k] is the
// index of error bound k; with deepening switched off only the final bound runs.
struct SeedPhases {
    SeedIndex ph[4];       // indexed by the error bound (statically, so that every field is a fixed constant-bank offset)
    int run_mask;          // bit k = bound k is searched (the final bound K always is)
};

template <int K, bool BL16, bool PRE, bool GEN>
__global__ void __launch_bounds__(SEED_THREADS, GEN ? 4 : SEED_MIN_CTAS) match_seed_kernel(const MatchParams p, const SeedPhases ixs) {
    constexpr int NSLOT = 2 * (K + 1) * 32;
    // non-empty slots only, in (strand, piece, position) order; 32 sentinels behind the last one
    __shared__ uint32_t s_prefix[SEED_WARPS][NSLOT + 32];
    __shared__ uint32_t s_start[SEED_WARPS][NSLOT];            // first entry of the bucket - index of its first candidate
    __shared__ uint32_t s_aux[SEED_WARPS][NSLOT];              // strand << 8 | pre-check operands of the slot (see seed_precheck)
    __shared__ uint32_t s_queue[SEED_WARPS][SEED_QUEUE_CAP];   // pre-check survivors: entry | strand << 31
    // bit t-1 = "a slot starts at candidate t": word j is the slot-start mask of candidates 32j+1 .. 32j+32.  Used when
    // the window's candidates fit (always at -mr <= 2 on real data); otherwise the mask is derived from s_prefix.
    __shared__ uint32_t s_map[SEED_WARPS][SEED_MAPW];
    __shared__ __align__(128) uint2 s_pair[SEED_WARPS][2][16]; // per strand: entry c0 | c1 << 2 = {Eq[c0], Eq[c1]}
    __shared__ uint32_t s_list[SEED_WARPS][SEED_LIST_CAP];
    __shared__ uint32_t s_tmp[SEED_WARPS][SEED_LIST_CAP];
    __shared__ __align__(16) uint32_t s_init[33][8];           // pre-check start states by boundary: R0 S0 R1 S1 R2 S2 R3 S3
    __shared__ __align__(16) uint32_t s_v[SEED_WARPS][4];      // GEN: start states of the window per level
    __shared__ unsigned long long s_cnt[SEED_WARPS][8];        // work counters (see the flush at the end)
    __shared__ int4 s_batch[SEED_WARPS][SEED_BATCH];           // the warp's batch: segment, length, byte offset (lo, hi)
    __shared__ unsigned long long s_text[SEED_WARPS][SEED_BATCH];   // ... and its packed text (one-window segments)
    __shared__ int2 s_bpos[SEED_WARPS];                         // ... position inside the batch, its length
    __shared__ unsigned long long s_tie[SEED_WARPS][2];        // the warp's chunk of the tie buffer: next position, entries left
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x < 33) {
        const uint32_t b = threadIdx.x;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
            const uint32_t r = b < 32 ? (((2u << d) - 1u) << b) : 0u;
            s_init[b][2 * d] = r;
            s_init[b][2 * d + 1] = r << 1;
        }
    }
    if (lane < 8) s_cnt[warp][lane] = 0ull;
    if (lane < 2) s_tie[warp][lane] = 0ull;
    __syncthreads();
    SeedWarp w;
    w.prefix = s_prefix[warp];
    w.start = s_start[warp];
    w.aux = s_aux[warp];
    w.queue = s_queue[warp];
    w.map = s_map[warp];
    w.list = s_list[warp];
    w.tmp = s_tmp[warp];
    w.v = s_v[warp];
    w.pair = &s_pair[warp][0][0];
    w.cnt = s_cnt[warp];
    // opaque copies: ptxas otherwise re-derives these shared-memory addresses from SR_CgaCtaId in every loop iteration
    w.init_a = smem_u32(&s_init[0][0]);
    asm volatile("mov.u32 %0, %0;" : "+r"(w.init_a));
    w.pair_a = smem_u32(&s_pair[warp][0][0]);
    w.v_a = smem_u32(&s_v[warp][0]);
    w.queue_a = smem_u32(&s_queue[warp][0]);
    asm volatile("mov.u32 %0, %0;" : "+r"(w.queue_a));
    const unsigned int n_items = GEN ? p.ctr->n_wide : p.ctr->n_fast;
    unsigned int *cursor = GEN ? &p.ctr->cur_wide : &p.ctr->cur_fast;
    const int *items = GEN ? p.wide_list : p.fast_list;
    const int bl = p.bl;

    if (lane == 0) s_bpos[warp] = make_int2(0, 0);
    unsigned int last_base = 0;
    const unsigned int tail_at = gridDim.x * SEED_WARPS * 16u;          // segments left when the batches shrink to one
    __syncwarp();
    for (;;) {
        // a batch of SEED_BATCH segments per visit of the cursor: lanes 0..SEED_BATCH-1 fetch one segment's id, extent
        // and first packed word each, so the chain atomic -> list -> offsets -> text is paid once per batch.  The position
        // inside the batch lives in shared memory: nothing of the batching stays in registers across the phases.
        const int2 bp = s_bpos[warp];                                  // .x = next position, .y = length of the batch
        int g = bp.x;
        int blen = bp.y;
        if (g == blen) {
            // near the end of the list a batch is one segment: a warp stuck with several expensive segments while the
            // others run dry costs more than the batching saves (measured at -mr 3: 5 %)
            blen = (last_base + tail_at < n_items) ? SEED_BATCH : 1;
            unsigned int base = 0;
            if (lane == 0) base = atomicAdd(cursor, (unsigned int)blen);
            base = __shfl_sync(FULL, base, 0);
            last_base = base;
            if (lane < blen) {
                int4 b = make_int4(-1, 0, 0, 0);
                unsigned long long t = 0ull;
                if (base + lane < n_items) {
                    const int sg = items[base + lane];
                    const long long o = p.seg_off[sg] - p.b0;
                    b = make_int4(sg, (int)(p.seg_off[sg + 1] - p.b0 - o), (int)(unsigned int)(o & 0xFFFFFFFFll), (int)(o >> 32));
                    if (!GEN) t = __ldg(p.pk + (o >> 5) + sg);
                }
                s_batch[warp][lane] = b;
                s_text[warp][lane] = t;
            }
            g = 0;
        }
        __syncwarp();
        if (lane == 0) s_bpos[warp] = make_int2(g + 1, blen);
        const int4 bt = s_batch[warp][g];
        const int seg = bt.x;
        if (seg < 0) break;                                              // the work list is exhausted
        const unsigned long long fw0 = s_text[warp][g];
        const int n_total = bt.y;
        const long long so = ((long long)bt.w << 32) | (long long)(unsigned int)bt.z;
        const unsigned long long *pw = p.pk + (so >> 5) + seg;
        int best = K + 1, count = 0;
        if constexpr (K >= 3) {
            if (ixs.run_mask & (1 << (K - 2))) {
                if (ixs.ph[K - 2].direct) seed_phase_direct<K - 2, BL16, GEN>(p, ixs.ph[K - 2], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
                else seed_phase<K - 2, BL16, PRE, GEN>(p, ixs.ph[K - 2], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
                if (best > K - 2) best = K + 1;
            }
        }
        if constexpr (K >= 2) {
            if ((ixs.run_mask & (1 << (K - 1))) && best > K) {
                if (ixs.ph[K - 1].direct) seed_phase_direct<K - 1, BL16, GEN>(p, ixs.ph[K - 1], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
                else seed_phase<K - 1, BL16, PRE, GEN>(p, ixs.ph[K - 1], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
                if (best > K - 1) best = K + 1;
            }
        }
        if (best > K) {
            if (ixs.ph[K].direct) seed_phase_direct<K, BL16, GEN>(p, ixs.ph[K], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
            else seed_phase<K, BL16, PRE, GEN>(p, ixs.ph[K], w, lane, seg, so, n_total, pw, fw0, bl, best, count);
        }
        // more hits than the list holds (in GEN mode possibly only duplicates from overlapping windows): exact re-run
        if (count <= SEED_LIST_CAP) count = sort_unique(w.list, w.tmp, count, lane);
        seed_store_result(p, seg, lane, K, best, count, w.list, s_tie[warp]);
        __syncwarp();
    }
    if (lane == 0) {
        if (s_cnt[warp][2]) atomicAdd(&p.ctr->n_cand, s_cnt[warp][2]);
        if (s_cnt[warp][0]) atomicAdd(&p.ctr->n_probe, s_cnt[warp][0]);
        if (s_cnt[warp][1]) atomicAdd(&p.ctr->n_verify, s_cnt[warp][1]);
        if (s_cnt[warp][3]) atomicAdd(&p.ctr->n_setup, s_cnt[warp][3]);
        if (s_cnt[warp][4]) atomicAdd(&p.ctr->n_lop3, s_cnt[warp][4]);
        if (s_cnt[warp][5]) atomicAdd(&p.ctr->n_precheck, s_cnt[warp][5]);
    }
}

}  // namespace sct
