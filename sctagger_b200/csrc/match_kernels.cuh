// sm_100a kernels of the scTagger `match_trie` hot path (reference scTagger.py:526-599 Trie, 643-668 trie_search,
// 671-685 get_matches_prefix, 705-715 merge).  The reference builds a trie of segment windows and runs an
// error-bounded DFS per barcode; here every (segment, barcode, strand) triple is scored directly by a bit-parallel
// automaton, one warp per segment, lanes over barcodes.  See DESIGN.md for the derivation and data layout.
//
// Arithmetic (checked bit-for-bit on the CPU by tests/wm_model.py, the numpy twin of this file):
//   k-error Shift-And automaton (Wu-Manber), one register R_d per error level, two 16-bit barcode fields interleaved
//   per 32-bit word (barcode 2w = even bits, 2w+1 = odd bits) so a pattern shift is `x*4` and the always-active start
//   state is `+3`:  S_d = 4*R_d + 3*v_d  — a single IMAD on the FMA pipe, leaving the ALU pipe to the LOP3s:
//       R0' = S0 & Eq                          S0' = 4*R0' + 3*v0
//       Rd' = (Sd & Eq) | (R_{d-1} | S_{d-1} | S'_{d-1})      Sd' = 4*Rd' + 3*vd      (ins | sub | del)
//   A barcode matches with <= d edits iff bit (bl-1) of its field is ever set in R_d.
//   v_d[j] = [D0[j] <= d] encodes the reference's rule for where a match may start in segments with non-ACGT bytes
//   (SURVEY.md Appendix A.2/A.4); for clean segments v == 1 and the addend is the immediate 3.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sct {

constexpr int TILE_WORDS = 128;               // words per table tile: 32 lanes x 4 words (= 256 barcodes)
constexpr int TILE_BYTES = TILE_WORDS * 4 * 4; // 4 text codes x 128 words x 4 B = 2048 B, layout [code][lane][4]
constexpr int PLANE_BYTES = 512;               // one text code of one tile
constexpr int REC_BYTES = 64;                  // fast-path segment record: 32 column bytes forward + 32 reverse strand
constexpr int FAST_MAX_LEN = 32;
constexpr int FAST_THREADS = 512;
constexpr int FAST_WARPS = FAST_THREADS / 32;
constexpr int FAST_LIST_CAP = 224;             // tie ids staged per warp in shared memory (fast kernel)
constexpr int GEN_THREADS = 256;
constexpr int GEN_WARPS = GEN_THREADS / 32;
constexpr int GEN_LIST_CAP = 256;
constexpr unsigned FULL = 0xFFFFFFFFu;

// segflag bits written by prep_kernel
constexpr uint8_t SEG_START0 = 1;   // a match may start at position 0
constexpr uint8_t SEG_FAST = 2;     // clean ACGT, <= 32 bases: record in `rec`, on the fast list
constexpr uint8_t SEG_SHORT = 4;    // shorter than bl - mr: can never produce a row

struct Counters {                    // device-side cursors, zeroed before every pass
    unsigned int n_fast, n_gen, n_ovf, n_dp;
    unsigned int cur_fast, cur_gen, cur_ovf, cur_dp;
    unsigned long long tie_cursor;   // bump allocator over tie_buf (may run past tie_cap: the pass is then redone)
    unsigned long long n_cand;       // seeded kernel: candidates generated (statistics)
    unsigned int n_bad;              // segments whose offsets are decreasing or leave the chunk (the call fails)
    unsigned int pad_;
    // live work counters of the seeded kernels (bench.py's roofline.useful): bucket probes, full verifications,
    // pre-checks, segments set up
    unsigned long long n_probe, n_verify, n_precheck, n_setup;
    // seeded mode: segments that are longer than one 31-base window or hold a non-ACGT byte (windowed seeded kernel)
    unsigned int n_wide, cur_wide;
    unsigned long long n_lop3;       // automaton steps x (2k+1): the LOP3 lane-ops of the seeded kernels' automata, per lane
};

struct MatchParams {
    const uint32_t *tabF, *tabR;     // [tile][code][lane][4] interleaved Eq masks, forward / explicit reverse complement
    int n_tiles;
    int n_barcodes;
    int bl;                          // barcode length (<= 16 for the bit-parallel kernels)
    const uint8_t *seg_bytes;        // raw segment bytes of the chunk
    const long long *seg_off;        // n+1 offsets in the CALLER's numbering: byte j of the chunk is offset b0 + j
    long long b0;
    const uint8_t *rec;              // REC_BYTES per segment (fast segments only)
    const uint8_t *colinfo;          // per raw byte: (2*code) | v-bits << 3 (generic segments only)
    const uint8_t *segflag;
    const int *fast_list, *gen_list;
    const int *wide_list;            // seeded mode: segments for the windowed variant of the seeded kernel
    const unsigned long long *pk;    // seeded mode: 2-bit packed text, segment i at word (byte offset >> 5) + i
    int *ovf_list;
    Counters *ctr;
    int *min_ed, *n_ties;
    unsigned int *tie_start;
    int *tie_buf;
    unsigned long long tie_cap;
    uint32_t *ovf_scratch;           // overflow re-run: per-warp tie lists in global memory
    int ovf_scratch_cap;             // entries per warp
    uint32_t four;                   // always 4; a run-time value so that x*4+c stays an FMA-pipe IMAD (see shl2_add)
};

// ---------------------------------------------------------------------------------------------------------------------
// small PTX helpers
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    } while (!done);
}
// TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// x*4 + addend as one integer multiply-add on the FMA pipe (SASS: IMAD), leaving the ALU pipe to the LOP3s.  ptxas
// constant-folds anything it can see (an `asm volatile("mov %0, 4")` included) and then strength-reduces x*4+c into an
// ALU-pipe LEA, so the 4 has to arrive at run time: it is a kernel parameter (MatchParams::four, always 4).
__device__ __forceinline__ uint32_t shl2_add(uint32_t x, uint32_t four, uint32_t addend) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(four), "r"(addend));
    return r;
}
// One text column for one chain (one table word = two barcodes), clean segment: start state always active.
template <int K>
__device__ __forceinline__ void wm_step(uint32_t (&R)[K + 1], uint32_t (&S)[K + 1], uint32_t eq, uint32_t four) {
    uint32_t r = S[0] & eq;
    uint32_t s = shl2_add(r, four, 3u);
    uint32_t pR = R[0], pS = S[0], nS = s;
    R[0] = r;
    S[0] = s;
#pragma unroll
    for (int d = 1; d <= K; ++d) {
        uint32_t t = pR | pS | nS;            // insertion | substitution | deletion: one LOP3
        r = (S[d] & eq) | t;                  // match: one LOP3
        s = shl2_add(r, four, 3u);
        pR = R[d];
        pS = S[d];
        nS = s;
        R[d] = r;
        S[d] = s;
    }
}
// Same with a per-level start-state addend (0 or 3), used for segments with restricted starts.
template <int K>
__device__ __forceinline__ void wm_step_v(uint32_t (&R)[K + 1], uint32_t (&S)[K + 1], uint32_t eq, uint32_t four,
                                          const uint32_t (&add)[K + 1]) {
    uint32_t r = S[0] & eq;
    uint32_t s = shl2_add(r, four, add[0]);
    uint32_t pR = R[0], pS = S[0], nS = s;
    R[0] = r;
    S[0] = s;
#pragma unroll
    for (int d = 1; d <= K; ++d) {
        uint32_t t = pR | pS | nS;
        r = (S[d] & eq) | t;
        s = shl2_add(r, four, add[d]);
        pR = R[d];
        pS = S[d];
        nS = s;
        R[d] = r;
        S[d] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Exact per-tile resolution: both strands of the 256 barcodes of one tile against one segment, with one accumulator
// per error level.  Returns 16 nibbles per lane: nibble i = g*4 + field*2 + strand (strand 0 = reverse complement)
// holds the smallest level at which that pattern matched, 15 if none.  Pattern id = tile*512 + lane*16 + i, i.e.
// (barcode << 1) | strand — ascending ids are the reference's sorted((bid, strand)) order (scTagger.py:789).
//   cols[j]: bits 1-2 = text code of column j, bits 3.. = v_d at position j+1;  allv is OR-ed into the v bits
//   (0xF for clean segments whose record stores codes only);  start0 = a match may start at position 0.
// ---------------------------------------------------------------------------------------------------------------------
template <int K>
__device__ __noinline__ unsigned long long tile_levels(const uint32_t *__restrict__ tabF_tile,
                                                       const uint32_t *__restrict__ tabR_tile,
                                                       const uint8_t *__restrict__ cols, int n, bool start0,
                                                       uint32_t allv, int lane, int topshift, uint32_t four) {
    unsigned long long lv = 0;
#pragma unroll 1
    for (int strand = 0; strand < 2; ++strand) {
        const uint4 *tab = reinterpret_cast<const uint4 *>(strand ? tabF_tile : tabR_tile) + lane;
        uint32_t R[4][K + 1], S[4][K + 1], acc[4][K + 1];
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int d = 0; d <= K; ++d) {
                R[g][d] = start0 ? ((1u << (2 * d)) - 1u) : 0u;
                S[g][d] = start0 ? ((1u << (2 * d + 2)) - 1u) : 0u;
                acc[g][d] = 0u;
            }
        for (int j = 0; j < n; ++j) {
            const uint32_t cb = cols[j];
            const uint32_t vb = (cb >> 3) | allv;
            const uint4 e = __ldg(tab + ((cb >> 1) & 3u) * 32);
            uint32_t add[K + 1];
#pragma unroll
            for (int d = 0; d <= K; ++d) add[d] = ((vb >> d) & 1u) * 3u;
            wm_step_v<K>(R[0], S[0], e.x, four, add);
            wm_step_v<K>(R[1], S[1], e.y, four, add);
            wm_step_v<K>(R[2], S[2], e.z, four, add);
            wm_step_v<K>(R[3], S[3], e.w, four, add);
#pragma unroll
            for (int g = 0; g < 4; ++g)
#pragma unroll
                for (int d = 0; d <= K; ++d) acc[g][d] |= R[g][d];
        }
#pragma unroll
        for (int g = 0; g < 4; ++g)
#pragma unroll
            for (int f = 0; f < 2; ++f) {
                unsigned long long level = 15;
#pragma unroll
                for (int d = K; d >= 0; --d)
                    if ((acc[g][d] >> (topshift + f)) & 1u) level = d;
                lv |= level << (4 * (g * 4 + f * 2 + strand));
            }
    }
    return lv;
}

// Fold one tile's levels into the warp's running (best, tie list).  All values are warp-uniform except lv/lane.
__device__ __forceinline__ void fold_tile(unsigned long long lv, int tile, int lane, int K, int &best, int &count,
                                          uint32_t *list, int cap) {
    int m = 15;
#pragma unroll
    for (int i = 0; i < 16; ++i) m = min(m, (int)((lv >> (4 * i)) & 15ull));
    m = __reduce_min_sync(FULL, m);
    if (m > K || m > best) return;
    if (m < best) {
        best = m;
        count = 0;
    }
    uint32_t hits = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if ((int)((lv >> (4 * i)) & 15ull) == m) hits |= 1u << i;
    const int c = __popc(hits);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
    }
    int pos = count + incl - c;
    const uint32_t pid0 = (uint32_t)tile * 512u + (uint32_t)lane * 16u;
    while (hits) {
        const int i = __ffs(hits) - 1;
        hits &= hits - 1;
        if (pos < cap) list[pos] = pid0 + i;
        ++pos;
    }
    count += __shfl_sync(FULL, incl, 31);
    __syncwarp();
}

// Write one segment's result; ties go to the bump-allocated tie buffer (re-ordered by finalize kernels).
__device__ __forceinline__ void store_result(const MatchParams &p, int seg, int lane, int K, int best, int count,
                                             const uint32_t *list, int cap, bool allow_overflow_list) {
    if (count > cap) {
        // more ties than the staging list holds: the segment is re-run with a full-size list in global memory
        if (lane == 0) {
            if (allow_overflow_list) {
                unsigned int q = atomicAdd(&p.ctr->n_ovf, 1u);
                p.ovf_list[q] = seg;
            }
            p.min_ed[seg] = best;
            p.n_ties[seg] = 0;   // rewritten by the re-run
        }
        return;
    }
    unsigned long long pos = 0;
    if (lane == 0) {
        p.min_ed[seg] = (best <= K) ? best : -1;
        p.n_ties[seg] = count;
        if (count > 0) {
            pos = atomicAdd(&p.ctr->tie_cursor, (unsigned long long)count);
            p.tie_start[seg] = (unsigned int)pos;
        }
    }
    if (count > 0) {
        pos = __shfl_sync(FULL, pos, 0);
        if (pos + (unsigned long long)count <= p.tie_cap)
            for (int i = lane; i < count; i += 32) p.tie_buf[pos + i] = (int)list[i];
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// FAST kernel: clean ACGT segments of <= 32 bases.  One persistent CTA per SM; the forward pattern table (<= ~222 KB)
// is staged once per CTA in shared memory with TMA bulk copies; each warp owns one segment at a time and walks the
// tiles, scoring T (forward strand) and reverse-complement(T) (reverse strand: d(rc(P),T) = d(P, rc(T))) against the
// same table.  Per column and lane: one PRMT (column byte -> plane offset), one LDS.128 (Eq masks of 8 barcodes) and
// four wm_step chains.  Only the top level is accumulated in the hot loop; a tile with any candidate (rare) is
// re-scored exactly by tile_levels().
// ---------------------------------------------------------------------------------------------------------------------
template <int K, bool BL16>
__device__ __forceinline__ uint32_t fast_pass(uint32_t saddr, const uint32_t (&tx)[8], int n, uint32_t four) {
    uint32_t R[4][K + 1], S[4][K + 1];
#pragma unroll
    for (int g = 0; g < 4; ++g)
#pragma unroll
        for (int d = 0; d <= K; ++d) {
            R[g][d] = (1u << (2 * d)) - 1u;
            S[g][d] = (1u << (2 * d + 2)) - 1u;
        }
    uint32_t acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    constexpr int JCHK = BL16 ? (16 - K) : 1;       // segments shorter than bl-K never reach this kernel
    constexpr int JACC = BL16 ? (16 - K - 1) : 0;   // a full match needs >= bl-K text characters
#pragma unroll
    for (int j = 0; j < FAST_MAX_LEN; ++j) {
        if (j >= JCHK) {
            if (j >= n) break;
        }
        const uint32_t off = __byte_perm(tx[j >> 2], 0u, 0x4404u | ((j & 3) << 4));   // (2*code) << 8 = code * 512
        const uint4 e = lds128(saddr + off);
        wm_step<K>(R[0], S[0], e.x, four);
        wm_step<K>(R[1], S[1], e.y, four);
        wm_step<K>(R[2], S[2], e.z, four);
        wm_step<K>(R[3], S[3], e.w, four);
        if (j >= JACC) {
            acc0 |= R[0][K];
            acc1 |= R[1][K];
            acc2 |= R[2][K];
            acc3 |= R[3][K];
        }
    }
    return acc0 | acc1 | acc2 | acc3;
}

template <int K, bool BL16>
__global__ void __launch_bounds__(FAST_THREADS, 1) match_fast_kernel(const MatchParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const uint32_t tab_bytes = (uint32_t)p.n_tiles * TILE_BYTES;
    uint32_t *lists = reinterpret_cast<uint32_t *>(smem + tab_bytes);
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(smem + tab_bytes + FAST_WARPS * FAST_LIST_CAP * 4);
    const uint32_t bar_a = smem_u32(bar);
    const uint32_t tab_a = smem_u32(smem);

    const unsigned int n_fast = p.ctr->n_fast;
    if (n_fast == 0) return;

    if (threadIdx.x == 0) mbar_init(bar_a, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar_a, tab_bytes);
        constexpr uint32_t PIECE = 16 * TILE_BYTES;   // 32 KB per bulk copy
        for (uint32_t o = 0; o < tab_bytes; o += PIECE) {
            const uint32_t sz = min(PIECE, tab_bytes - o);
            tma_bulk_g2s(tab_a + o, reinterpret_cast<const uint8_t *>(p.tabF) + o, sz, bar_a);
        }
    }
    mbar_wait(bar_a, 0);

    uint32_t *list = lists + warp * FAST_LIST_CAP;
    const uint32_t four = p.four;
    const int topshift = 2 * (p.bl - 1);
    const uint32_t topmask = 3u << topshift;

    for (;;) {
        unsigned int idx = 0;
        if (lane == 0) idx = atomicAdd(&p.ctr->cur_fast, 1u);
        idx = __shfl_sync(FULL, idx, 0);
        if (idx >= n_fast) break;
        const int seg = p.fast_list[idx];
        const int n = (int)(p.seg_off[seg + 1] - p.seg_off[seg]);
        const uint4 *r4 = reinterpret_cast<const uint4 *>(p.rec + (size_t)seg * REC_BYTES);
        const uint4 f0 = __ldg(r4), f1 = __ldg(r4 + 1), b0 = __ldg(r4 + 2), b1 = __ldg(r4 + 3);
        int best = K + 1, count = 0;
        for (int t = 0; t < p.n_tiles; ++t) {
            const uint32_t saddr = tab_a + (uint32_t)t * TILE_BYTES + (uint32_t)lane * 16u;
            uint32_t h = 0;
#pragma unroll 1
            for (int s = 0; s < 2; ++s) {
                uint32_t tx[8];
                tx[0] = s ? b0.x : f0.x; tx[1] = s ? b0.y : f0.y; tx[2] = s ? b0.z : f0.z; tx[3] = s ? b0.w : f0.w;
                tx[4] = s ? b1.x : f1.x; tx[5] = s ? b1.y : f1.y; tx[6] = s ? b1.z : f1.z; tx[7] = s ? b1.w : f1.w;
                h |= fast_pass<K, BL16>(saddr, tx, n, four);
            }
            if (__any_sync(FULL, (h & topmask) != 0u)) {
                const unsigned long long lv = tile_levels<K>(p.tabF + (size_t)t * (TILE_BYTES / 4), p.tabR + (size_t)t * (TILE_BYTES / 4),
                                                             p.rec + (size_t)seg * REC_BYTES, n, true, 0xFu, lane, topshift, four);
                fold_tile(lv, t, lane, K, best, count, list, FAST_LIST_CAP);
            }
        }
        __syncwarp();
        store_result(p, seg, lane, K, best, count, list, FAST_LIST_CAP, true);
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// GENERIC kernel: any segment (non-ACGT bytes, > 32 bases, or a re-run after tie-list overflow).  Same automaton,
// tables read from global memory (L1/L2 resident), start-state row from colinfo, every level accumulated.
// mode 0: segments from gen_list, tie list in shared memory;  mode 1: segments from ovf_list, list in global scratch.
// ---------------------------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(GEN_THREADS) match_generic_kernel(const MatchParams p, int mode) {
    __shared__ uint32_t s_lists[GEN_WARPS * GEN_LIST_CAP];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int gwarp = blockIdx.x * GEN_WARPS + warp;
    uint32_t *list;
    int cap;
    if (mode == 0) {
        list = s_lists + warp * GEN_LIST_CAP;
        cap = GEN_LIST_CAP;
    } else {
        list = p.ovf_scratch + (size_t)gwarp * p.ovf_scratch_cap;
        cap = p.ovf_scratch_cap;
    }
    const unsigned int n_items = mode == 0 ? p.ctr->n_gen : p.ctr->n_ovf;
    unsigned int *cursor = mode == 0 ? &p.ctr->cur_gen : &p.ctr->cur_ovf;
    const int *items = mode == 0 ? p.gen_list : p.ovf_list;
    const int topshift = 2 * (p.bl - 1);
    for (;;) {
        unsigned int idx = 0;
        if (lane == 0) idx = atomicAdd(cursor, 1u);
        idx = __shfl_sync(FULL, idx, 0);
        if (idx >= n_items) break;
        const int seg = items[idx];
        const long long o = p.seg_off[seg] - p.b0;
        const int n = (int)(p.seg_off[seg + 1] - p.b0 - o);
        const uint8_t flag = p.segflag[seg];
        const bool is_fast = (flag & SEG_FAST) != 0;
        const uint8_t *cols = is_fast ? p.rec + (size_t)seg * REC_BYTES : p.colinfo + o;
        const uint32_t allv = is_fast ? 0xFu : 0u;
        const bool start0 = (flag & SEG_START0) != 0;
        int best = K + 1, count = 0;
        for (int t = 0; t < p.n_tiles; ++t) {
            const unsigned long long lv = tile_levels<K>(p.tabF + (size_t)t * (TILE_BYTES / 4), p.tabR + (size_t)t * (TILE_BYTES / 4),
                                                         cols, n, start0, allv, lane, topshift, p.four);
            fold_tile(lv, t, lane, K, best, count, list, cap);
        }
        __syncwarp();
        store_result(p, seg, lane, K, best, count, list, cap, mode == 0);
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// PREP kernel: one thread per segment.  Classifies the segment, writes the fast-path record (column bytes 2*code for T
// and for reverse-complement(T)) or the generic column info (code + start-state bits), and
// appends the segment to the matching work list.  Replaces the per-character loop of read_long_reads
// (scTagger.py:621-626) + map_char (512-516) + the prefix rule of get_matches_prefix (677-682).
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int base_code(uint8_t b) {   // -1 for anything but upper-case ACGT
    return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : -1;
}

__global__ void __launch_bounds__(256) prep_kernel(const uint8_t *__restrict__ bytes, const long long *__restrict__ off,
                                                   long long b0, long long nbytes,
                                                   int n_seg, int bl, int K, int kpre, int path /*0 auto,1 generic,2 dp*/,
                                                   int fast_max /* longest segment of the fast path */,
                                                   unsigned long long *__restrict__ pk /* 2-bit packed text, or NULL */,
                                                   uint8_t *__restrict__ rec, uint8_t *__restrict__ colinfo,
                                                   uint8_t *__restrict__ segflag, int *__restrict__ fast_list,
                                                   int *__restrict__ gen_list, int *__restrict__ dp_list,
                                                   Counters *__restrict__ ctr, int *__restrict__ min_ed,
                                                   int *__restrict__ n_ties) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_seg) return;
    const long long o = off[i] - b0;
    long long L = off[i + 1] - b0 - o;
    if (L < 0 || o < 0 || o + L > nbytes) {   // the caller's offsets are not non-decreasing inside the chunk: the call fails
        atomicAdd(&ctr->n_bad, 1u);
        L = 0;
    }
    if (L < bl - K) {   // cannot hold a barcode within K edits: the reference emits no row (scTagger.py:681)
        segflag[i] = SEG_SHORT;
        min_ed[i] = -1;
        n_ties[i] = 0;
        return;
    }
    if (path == 2) {
        segflag[i] = 0;
        dp_list[atomicAdd(&ctr->n_dp, 1u)] = i;
        return;
    }
    bool clean = true;
    if (L <= fast_max && path == 0) {
        uint32_t f[8] = {0, 0, 0, 0, 0, 0, 0, 0}, r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const int n = (int)L;
#pragma unroll
        for (int j = 0; j < FAST_MAX_LEN; ++j) {
            if (j < n) {
                const int c = base_code(bytes[o + j]);
                if (c < 0) clean = false;
                f[j >> 2] |= (uint32_t)(2 * (c & 3)) << (8 * (j & 3));
            }
        }
        if (clean) {
            // reverse strand text: column j holds the complement of T[n-1-j]
#pragma unroll
            for (int j = 0; j < FAST_MAX_LEN; ++j) {
                if (j < n) {
                    const int jj = n - 1 - j;
                    const int c = base_code(bytes[o + jj]);
                    r[j >> 2] |= (uint32_t)(2 * (3 - c)) << (8 * (j & 3));
                }
            }
            uint4 *dst = reinterpret_cast<uint4 *>(rec + (size_t)i * REC_BYTES);
            dst[0] = make_uint4(f[0], f[1], f[2], f[3]);
            dst[1] = make_uint4(f[4], f[5], f[6], f[7]);
            dst[2] = make_uint4(r[0], r[1], r[2], r[3]);
            dst[3] = make_uint4(r[4], r[5], r[6], r[7]);
            segflag[i] = SEG_START0 | SEG_FAST;
            if (pk) {
                unsigned long long w = 0;
#pragma unroll
                for (int j = 0; j < FAST_MAX_LEN; ++j) w |= (unsigned long long)((f[j >> 2] >> (8 * (j & 3) + 1)) & 3u) << (2 * j);
                pk[i] = w;
            }
            fast_list[atomicAdd(&ctr->n_fast, 1u)] = i;
            return;
        }
    }
    // generic: start-state row D0 (0 at a valid start, else previous + 1) folded into v bits per level
    int prev = 1000;
    uint8_t flag = 0;
    for (long long j = 0; j <= L; ++j) {
        bool valid = j + kpre <= L;
        for (int q = 0; valid && q < kpre; ++q) valid = base_code(bytes[o + j + q]) >= 0;
        prev = valid ? 0 : min(prev + 1, 1000);
        uint32_t vb = 0;
#pragma unroll
        for (int d = 0; d < 4; ++d) vb |= (prev <= d ? 1u : 0u) << d;
        if (j == 0) {
            flag = (prev == 0) ? SEG_START0 : 0;
        } else {
            const int c = base_code(bytes[o + j - 1]);
            colinfo[o + j - 1] = (uint8_t)((2 * (c < 0 ? 0 : c)) | (vb << 3));
        }
    }
    segflag[i] = flag;
    gen_list[atomicAdd(&ctr->n_gen, 1u)] = i;
}

// PREP kernel of the seeded mode: one thread per segment, ANY length, ANY bytes.  Writes
//   * pk: the text as 2-bit codes (non-ACGT bytes as A = 0, map_char scTagger.py:512-516), 32 bases per 64-bit word,
//     segment i starting at word (byte offset >> 5) + i (always enough room: the next segment starts at least
//     ceil(L / 32) words later), zero-padded inside its last word;
//   * colinfo: per byte (2 * code) | v-bits << 3 — the start-state row D0 of SURVEY.md Appendix A.4 folded into one bit
//     per error level (bit d = [D0 <= d]); the windowed seeded kernel reads the v bits, the overflow re-run all of it;
//   * the work list: clean segments of <= fast_max bases -> fast_list (one window, every start allowed), the others
//     (longer, or with a non-ACGT byte) -> wide_list (overlapping 31-base windows, start states from colinfo).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) prep_seed_kernel(const uint8_t *__restrict__ bytes, const long long *__restrict__ off,
                                                        long long b0, long long nbytes, int n_seg, int bl, int K, int kpre,
                                                        int fast_max, unsigned long long *__restrict__ pk,
                                                        uint8_t *__restrict__ colinfo, uint8_t *__restrict__ segflag,
                                                        int *__restrict__ fast_list, int *__restrict__ wide_list,
                                                        Counters *__restrict__ ctr, int *__restrict__ min_ed,
                                                        int *__restrict__ n_ties) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_seg) return;
    const long long o = off[i] - b0;
    long long L = off[i + 1] - b0 - o;
    if (L < 0 || o < 0 || o + L > nbytes) {   // the caller's offsets are not non-decreasing inside the chunk: the call fails
        atomicAdd(&ctr->n_bad, 1u);
        L = 0;
    }
    if (L < bl - K) {   // cannot hold a barcode within K edits: the reference emits no row (scTagger.py:681)
        segflag[i] = SEG_SHORT;
        min_ed[i] = -1;
        n_ties[i] = 0;
        return;
    }
    unsigned long long *dst = pk + (o >> 5) + i;
    unsigned long long w = 0;
    bool clean = true;
    int prev = 1000;
    uint8_t flag = 0;
    for (long long j = 0; j <= L; ++j) {
        bool valid = j + kpre <= L;
        for (int q = 0; valid && q < kpre; ++q) valid = base_code(bytes[o + j + q]) >= 0;
        prev = valid ? 0 : min(prev + 1, 1000);
        uint32_t vb = 0;
#pragma unroll
        for (int d = 0; d < 4; ++d) vb |= (prev <= d ? 1u : 0u) << d;
        if (j == 0) {
            flag = (prev == 0) ? SEG_START0 : 0;
        } else {
            const int cc = base_code(bytes[o + j - 1]);
            const int c = cc < 0 ? 0 : cc;
            clean = clean && cc >= 0;
            colinfo[o + j - 1] = (uint8_t)((2 * c) | (vb << 3));
            const int at = (int)((j - 1) & 31);
            w |= (unsigned long long)c << (2 * at);
            if (at == 31) {
                dst[(j - 1) >> 5] = w;
                w = 0;
            }
        }
    }
    if (L & 31) dst[L >> 5] = w;
    segflag[i] = flag;
    if (clean && L <= fast_max) fast_list[atomicAdd(&ctr->n_fast, 1u)] = i;
    else wide_list[atomicAdd(&ctr->n_wide, 1u)] = i;
}

// ---------------------------------------------------------------------------------------------------------------------
// DP kernel: exact fallback for parameter ranges outside the bit-parallel kernels (barcode length 17..32 or
// max_error > 3).  One warp per segment, lanes over barcodes, both strands; plain semi-global DP with the start-state
// top row held in registers.  Two-pass (minimum, then ties) so no list overflow can occur.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int DP_MAX_BL = 32;

__device__ __forceinline__ int dp_distance(const uint8_t *__restrict__ pat /* bl codes */, int bl,
                                           const uint8_t *__restrict__ seg, int n, int kpre) {
    int col[DP_MAX_BL + 1];
    int top = 1000;
    {
        bool valid = kpre <= n;
        for (int q = 0; valid && q < kpre; ++q) valid = base_code(seg[q]) >= 0;
        top = valid ? 0 : 1000;
    }
#pragma unroll
    for (int r = 0; r <= DP_MAX_BL; ++r) col[r] = top + r;
    int best = 1000;
    if (bl <= DP_MAX_BL) best = min(best, top + bl);
    for (int j = 1; j <= n; ++j) {
        bool valid = j + kpre <= n;
        for (int q = 0; valid && q < kpre; ++q) valid = base_code(seg[j + q]) >= 0;
        top = valid ? 0 : min(top + 1, 1000);
        const int cc = base_code(seg[j - 1]);
        const int c = cc < 0 ? 0 : cc;
        int diag = col[0];
        col[0] = top;
#pragma unroll
        for (int r = 1; r <= DP_MAX_BL; ++r) {
            if (r <= bl) {
                const int sub = diag + (pat[r - 1] != c ? 1 : 0);
                const int left = col[r] + 1, up = col[r - 1] + 1;
                diag = col[r];
                col[r] = min(sub, min(left, up));
                if (r == bl) best = min(best, col[r]);
            }
        }
    }
    return best;
}

struct DpParams {
    const uint8_t *patF, *patR;   // n_barcodes x bl text codes, forward / explicit reverse complement
    int n_barcodes, bl, K, kpre;
};

__global__ void __launch_bounds__(128) match_dp_kernel(const MatchParams p, const DpParams q, const int *__restrict__ dp_list) {
    const int lane = threadIdx.x & 31;
    const unsigned int n_items = p.ctr->n_dp;
    for (;;) {
        unsigned int idx = 0;
        if (lane == 0) idx = atomicAdd(&p.ctr->cur_dp, 1u);
        idx = __shfl_sync(FULL, idx, 0);
        if (idx >= n_items) break;
        const int seg = dp_list[idx];
        const long long o = p.seg_off[seg] - p.b0;
        const int n = (int)(p.seg_off[seg + 1] - p.b0 - o);
        const uint8_t *s = p.seg_bytes + o;
        // pass 1: minimum
        int best = q.K + 1;
        for (int b = lane; b < q.n_barcodes; b += 32) {
            const int dr = dp_distance(q.patR + (size_t)b * q.bl, q.bl, s, n, q.kpre);
            const int df = dp_distance(q.patF + (size_t)b * q.bl, q.bl, s, n, q.kpre);
            best = min(best, min(dr, df));
        }
        best = __reduce_min_sync(FULL, best);
        if (best > q.K) {
            if (lane == 0) {
                p.min_ed[seg] = -1;
                p.n_ties[seg] = 0;
            }
            continue;
        }
        // pass 2: count ties, allocate, then write them in id order
        int cnt = 0;
        for (int b = lane; b < q.n_barcodes; b += 32) {
            cnt += dp_distance(q.patR + (size_t)b * q.bl, q.bl, s, n, q.kpre) == best;
            cnt += dp_distance(q.patF + (size_t)b * q.bl, q.bl, s, n, q.kpre) == best;
        }
        const int total = __reduce_add_sync(FULL, cnt);
        unsigned long long pos = 0;
        if (lane == 0) {
            pos = atomicAdd(&p.ctr->tie_cursor, (unsigned long long)total);
            p.min_ed[seg] = best;
            p.n_ties[seg] = total;
            p.tie_start[seg] = (unsigned int)pos;
        }
        pos = __shfl_sync(FULL, pos, 0);
        const bool fits = pos + (unsigned long long)total <= p.tie_cap;
        int written = 0;
        for (int b0 = 0; b0 < q.n_barcodes; b0 += 32) {
            const int b = b0 + lane;
            int hr = 0, hf = 0;
            if (b < q.n_barcodes) {
                hr = dp_distance(q.patR + (size_t)b * q.bl, q.bl, s, n, q.kpre) == best;
                hf = dp_distance(q.patF + (size_t)b * q.bl, q.bl, s, n, q.kpre) == best;
            }
            const int c = hr + hf;
            int incl = c;
#pragma unroll
            for (int sft = 1; sft < 32; sft <<= 1) {
                int v = __shfl_up_sync(FULL, incl, sft);
                if (lane >= sft) incl += v;
            }
            int w = written + incl - c;
            if (fits) {
                if (hr) p.tie_buf[pos + w++] = (b << 1);
                if (hf) p.tie_buf[pos + w++] = (b << 1) | 1;
            }
            written += __shfl_sync(FULL, incl, 31);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// FINALIZE: exclusive scan of n_ties -> tie_offsets (int64), then gather the bump-allocated lists into segment order.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int SCAN_BLOCK = 1024;

__global__ void __launch_bounds__(SCAN_BLOCK) scan_block_sums(const int *__restrict__ n_ties, int n, long long *__restrict__ block_sums) {
    __shared__ long long warp_sums[32];
    const int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    long long v = i < n ? n_ties[i] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        long long w = warp_sums[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_down_sync(FULL, w, o);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = w;
    }
}

// total[0] = ties of this chunk; total[1] = ties of all earlier chunks of the call (`run` is the device-side running sum,
// advanced once per chunk — not again when the pass is repeated after the tie buffer grew), so that tie_offsets leave
// the device already in the call's numbering and the host never touches them.
__global__ void __launch_bounds__(1024) scan_block_prefix(long long *__restrict__ block_sums, int n_blocks, long long *__restrict__ total,
                                                          long long *__restrict__ run, int redo) {
    // single block: serial over chunks of 1024 block sums, exclusive prefix in place
    __shared__ long long sh[1024];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n_blocks; base += 1024) {
        const int i = base + threadIdx.x;
        const long long v = i < n_blocks ? block_sums[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            long long t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        const long long incl = sh[threadIdx.x];
        const long long c = carry;
        if (i < n_blocks) block_sums[i] = c + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = c + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        total[0] = carry;
        if (!redo) {
            const long long b = *run;
            total[1] = b;
            *run = b + carry;
        }
    }
}

__global__ void __launch_bounds__(SCAN_BLOCK) scan_gather(const int *__restrict__ n_ties, int n, const long long *__restrict__ block_prefix,
                                                          const unsigned int *__restrict__ tie_start, const int *__restrict__ tie_buf,
                                                          long long *__restrict__ tie_off, int *__restrict__ tie_out,
                                                          unsigned long long tie_cap, const long long *__restrict__ total) {
    __shared__ long long warp_sums[32];
    const int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = i < n ? n_ties[i] : 0;
    long long incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        long long w = warp_sums[lane];
        long long wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            long long v = __shfl_up_sync(FULL, wi, o);
            if (lane >= o) wi += v;
        }
        warp_sums[lane] = wi - w;
    }
    __syncthreads();
    const long long excl = block_prefix[blockIdx.x] + warp_sums[warp] + incl - c;
    if (i < n) {
        const long long base = total[1];       // tie_out stays chunk-local, tie_off is in the call's numbering
        tie_off[i] = base + excl;
        if (i == n - 1) tie_off[n] = base + excl + c;
        // when the bump allocator ran past tie_cap the host grows the buffers and redoes the pass: write nothing here
        if (c > 0 && (unsigned long long)(excl + c) <= tie_cap) {
            const unsigned int s = tie_start[i];
            if ((unsigned long long)s + (unsigned long long)c <= tie_cap)
                for (int k = 0; k < c; ++k) tie_out[excl + k] = tie_buf[s + k];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// INT32 issue-rate microbenchmark (roofline denominator): 8 independent chains per thread whose
// inputs rotate so that consecutive logic ops cannot be algebraically merged.
// ---------------------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(1024) int32_peak_kernel(uint32_t *out, int iters, uint32_t seed, uint32_t four) {
    uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint32_t x = a[(i + 1) & 7], y = a[(i + 3) & 7];
                if (MODE == 0) {
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(x), "r"(y));
                } else if (MODE == 1) {
                    asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(x));
                } else {
                    if (i & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(four), "r"(x));
                    else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(x), "r"(y));
                }
            }
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s ^= a[i];
    if (s == 0x12345678u) out[0] = s;   // practically never: keeps the chains live
}

}  // namespace sct
