// sm_100a kernels of the `extract_lr_bc` adapter alignment (reference scTagger.py:176-196 get_alns: two
// edlib.align(adapter, read, 'HW', 'locations') calls per read, adapter and its reverse complement).
//
// One thread owns one read and advances BOTH adapters' Myers/Hyyro bit-vectors (pattern <= 32 characters: one register
// each) over the read, one column per base:
//     Xv = Eq | Mv;  Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;  Ph = Mv | ~(Xh | Pv);  Mh = Pv & Xh
//     score += Ph[m-1] - Mh[m-1];  Ph <<= 1;  Mh <<= 1;  Pv = Mh | ~(Xv | Ph);  Mv = Ph & Xv        (free top row: HW)
// Eq comes from a 256-entry table indexed by the raw byte (literal comparison like edlib: N only equals N), both adapters
// interleaved so that one LDS.64 serves a column.  Reads are fetched as aligned 128-bit words.
//   pass 1: minimum distance and number of end positions at the minimum per adapter -> strand, distance, #locations
//   (exclusive scan of #locations)
//   pass 2: the winning adapter again, emitting the locations the reference derives (scTagger.py:184-191):
//           '+' : every end position at the minimum;
//           '-' : for every such end the LEFTMOST start with the same distance (edlib's start location: reversed
//                 adapter against the reversed read prefix, prefix-anchored, last optimal column), as start - len - 1.
// Reads of one chunk are processed in descending length order (counting sort on len/32) so that the 32 reads of a warp
// are of similar length.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sct {

constexpr int ALN_THREADS = 128;
constexpr int ALN_LONG_WORK = 160;        // pass 2: reads with more columns left than this get a warp instead of a thread
constexpr int ALN_LONG_READ = 16384;      // pass 1: longer reads get a warp (aln_pass1_long_kernel) instead of a thread
constexpr int ALN_BUCKETS = 1024;        // length buckets of 32 bases; longer reads share the last bucket

struct AlignParams {
    const uint8_t *bytes;                // read bytes of the chunk; the allocation is 16-byte aligned and padded
    const long long *off;                // n + 1 offsets in the CALLER's numbering: byte j of the chunk is offset b0 + j
    long long b0, nbytes;
    int n;
    int m;                               // adapter length (<= 32)
    const uint2 *peq;                    // [256] {Eq of a1, Eq of a2} by byte value
    const uint32_t *rpeq2;               // [256] Eq of reversed a2 (start search)
    const int *order;                    // read ids, long reads first
    signed char *strand;                 // '+', '-', 0 = NA
    int *dist;                           // -1 = NA
    int *n_locs;
    int2 *span;                          // first and last end position at the minimum of the winning adapter (pass 1 -> 2)
    int *work;                           // columns pass 2 has to scan for the read
    int *long_list;                      // reads whose span is long (work > ALN_LONG_WORK): one warp each in the count / emit kernels
    unsigned int *n_long;
    int *short_list;                     // reads with several end positions in a short span: one thread each, compacted
    unsigned int *n_short;
    int *long_cnt;                       // 32 per long read: end positions at the minimum in each lane's piece of the span
    const long long *loc_off;            // exclusive scan of n_locs (pass 2)
    int *locs;
    unsigned long long loc_cap;
};

// ---- length ordering ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int len_bucket(long long len) {
    const long long b = len >> 5;
    return ALN_BUCKETS - 1 - (int)(b < ALN_BUCKETS - 1 ? b : ALN_BUCKETS - 1);   // descending length
}

// also validates the offsets (non-decreasing, inside the chunk): *bad counts the reads that are not — the call fails
__global__ void __launch_bounds__(256) aln_hist_kernel(const long long *__restrict__ off, int n, long long b0, long long nbytes,
                                                       unsigned int *__restrict__ hist, unsigned int *__restrict__ bad) {
    __shared__ unsigned int sh[ALN_BUCKETS];
    for (int i = threadIdx.x; i < ALN_BUCKETS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const long long a = off[i] - b0, b = off[i + 1] - b0;
        if (a < 0 || b < a || b > nbytes) atomicAdd(bad, 1u);
        atomicAdd(&sh[len_bucket(b > a ? b - a : 0)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ALN_BUCKETS; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// one block of ALN_BUCKETS threads: exclusive scan of the histogram in place -> bucket cursors
__global__ void __launch_bounds__(ALN_BUCKETS) aln_bucket_scan_kernel(unsigned int *__restrict__ hist) {
    __shared__ unsigned int sh[ALN_BUCKETS];
    const unsigned int v = hist[threadIdx.x];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < ALN_BUCKETS; o <<= 1) {
        const unsigned int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0u;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    hist[threadIdx.x] = sh[threadIdx.x] - v;
}

// block-aggregated: ranks inside the block from shared-memory atomics, one global atomic per (block, bucket)
__global__ void __launch_bounds__(1024) aln_scatter_kernel(const long long *__restrict__ off, int n, unsigned int *__restrict__ cursor,
                                                           int *__restrict__ order) {
    __shared__ unsigned int sh[ALN_BUCKETS];
    for (int i = threadIdx.x; i < ALN_BUCKETS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int b = 0;
    unsigned int rank = 0;
    if (i < n) {
        const long long len = off[i + 1] - off[i];
        b = len_bucket(len > 0 ? len : 0);
        rank = atomicAdd(&sh[b], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < ALN_BUCKETS; k += blockDim.x) {
        const unsigned int c = sh[k];
        if (c) sh[k] = atomicAdd(&cursor[k], c);
    }
    __syncthreads();
    if (i < n) order[sh[b] + rank] = i;
}

// extent of read r inside the chunk's byte buffer (clamped: bad offsets fail the call, they must not leave the buffer)
__device__ __forceinline__ void read_extent(const AlignParams &p, int r, long long &beg, long long &end) {
    beg = p.off[r] - p.b0;
    end = p.off[r + 1] - p.b0;
    beg = beg < 0 ? 0 : (beg > p.nbytes ? p.nbytes : beg);
    end = end < beg ? beg : (end > p.nbytes ? p.nbytes : end);
}

// ---- Myers column ---------------------------------------------------------------------------------------------------
struct Myers {
    uint32_t Pv, Mv;
    int score;
    __device__ __forceinline__ void init(int m) {
        Pv = m >= 32 ? 0xFFFFFFFFu : ((1u << m) - 1u);
        Mv = 0u;
        score = m;
    }
    // HW: free top row (horizontal delta 0 enters at the top);  PREFIX: top row grows by one per column (+1 enters)
    template <bool PREFIX>
    __device__ __forceinline__ void step(uint32_t Eq, uint32_t top) {
        const uint32_t Xv = Eq | Mv;
        const uint32_t Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
        uint32_t Ph = Mv | ~(Xh | Pv);
        uint32_t Mh = Pv & Xh;
        score += (Ph & top) ? 1 : 0;
        score -= (Mh & top) ? 1 : 0;
        Ph = (Ph << 1) | (PREFIX ? 1u : 0u);
        Mh <<= 1;
        Pv = Mh | ~(Xv | Ph);
        Mv = Ph & Xv;
    }
};

// Calls f(byte, position) for every byte of [beg, end) using aligned 128-bit loads (the buffer is padded).
template <class F>
__device__ __forceinline__ void for_each_byte(const uint8_t *__restrict__ bytes, long long beg, long long end, F f) {
    long long w = beg & ~15ll;
    for (; w < end; w += 16) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(bytes + w));
        const uint32_t q[4] = {v.x, v.y, v.z, v.w};
        if (w >= beg && w + 16 <= end) {
#pragma unroll
            for (int k = 0; k < 16; ++k) f((q[k >> 2] >> (8 * (k & 3))) & 0xFFu, w + k);
        } else {
#pragma unroll
            for (int k = 0; k < 16; ++k)
                if (w + k >= beg && w + k < end) f((q[k >> 2] >> (8 * (k & 3))) & 0xFFu, w + k);
        }
    }
}

// ---- pass 1 ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ALN_THREADS) aln_pass1_kernel(const AlignParams p) {
    __shared__ uint2 s_peq[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) s_peq[i] = p.peq[i];
    __syncthreads();
    const int t = blockIdx.x * ALN_THREADS + threadIdx.x;
    if (t >= p.n) return;
    const int r = p.order[t];
    long long beg, end;
    read_extent(p, r, beg, end);
    if (end - beg > ALN_LONG_READ) return;                        // aln_pass1_long_kernel
    const int m = p.m;
    const uint32_t top = 1u << (m - 1);
    Myers a, b;
    a.init(m);
    b.init(m);
    int best1 = 1 << 20, best2 = 1 << 20, c1 = 0, c2 = 0, f1 = 0, l1 = 0, f2 = 0, l2 = 0, j = 0;
    for_each_byte(p.bytes, beg, end, [&](uint32_t ch, long long) {
        const uint2 e = s_peq[ch];
        a.step<false>(e.x, top);
        b.step<false>(e.y, top);
        if (a.score < best1) { best1 = a.score; c1 = 0; f1 = j; }
        if (a.score == best1) { ++c1; l1 = j; }
        if (b.score < best2) { best2 = b.score; c2 = 0; f2 = j; }
        if (b.score == best2) { ++c2; l2 = j; }
        ++j;
    });
    if (end == beg) best1 = best2 = m;       // edlib: empty target -> distance = query length (both adapters: a tie)
    signed char s = 0;
    int d = -1, nl = 0;
    int2 span = make_int2(0, -1);
    if (best1 < best2) { s = '+'; d = best1; nl = c1; span = make_int2(f1, l1); }
    else if (best2 < best1) { s = '-'; d = best2; nl = c2; span = make_int2(f2, l2); }
    p.span[r] = span;
    const int work = nl ? (span.x > 2 * m + 2 ? 2 * m + 2 : span.x) + (span.y - span.x + 1) + (s == '-' ? nl * (m + d) : 0) : 0;
    p.work[r] = work;
    if (work > ALN_LONG_WORK) p.long_list[atomicAdd(p.n_long, 1u)] = r;
    else if (s != 0 && span.x != span.y) p.short_list[atomicAdd(p.n_short, 1u)] = r;
    p.strand[r] = s;
    p.dist[r] = d;
    p.n_locs[r] = nl;
}

// ---- pass 1, packed keys (16 <= m <= 27) ---------------------------------------------------------------------------------
// The bookkeeping of the kernel above costs as much as the recurrence: per adapter and column two compares and five
// selects for (minimum, count, first, last).  Here the score is kept in units of U = 2^(m-1) — the weight of the top bit,
// so score += (Ph & top) - (Mh & top) needs no shift — and two packed keys are folded with one add and one min each:
//     score * U + (U - 1 - j)  ->  minimum score, LAST position at it;      score * U + j  ->  minimum score, FIRST position.
// The number of end positions at the minimum is no longer counted here: a read whose first and last optimal end coincide
// (nearly every read with an adapter) has one location; the others are counted by the count kernels over [first, last]
// only.  Positions are relative to a segment of at most U - 1 columns; longer reads fold their segments (state carried).
struct MyersU {                          // Myers column with the score in units of the top bit
    uint32_t Pv, Mv, sk;
    __device__ __forceinline__ void init(int m, uint32_t top) {
        Pv = m >= 32 ? 0xFFFFFFFFu : ((1u << m) - 1u);
        Mv = 0u;
        sk = (uint32_t)m * top;
    }
    __device__ __forceinline__ void step(uint32_t Eq, uint32_t top) {
        const uint32_t Xv = Eq | Mv;
        const uint32_t Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
        uint32_t Ph = Mv | ~(Xh | Pv);
        uint32_t Mh = Pv & Xh;
        sk = sk + (Ph & top) - (Mh & top);
        Ph <<= 1;
        Mh <<= 1;
        Pv = Mh | ~(Xv | Ph);
        Mv = Ph & Xv;
    }
};

__global__ void __launch_bounds__(ALN_THREADS) aln_pass1_fast_kernel(const AlignParams p) {
    __shared__ uint2 s_peq[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) s_peq[i] = p.peq[i];
    __syncthreads();
    const int t = blockIdx.x * ALN_THREADS + threadIdx.x;
    if (t >= p.n) return;
    const int r = p.order[t];
    long long beg, end;
    read_extent(p, r, beg, end);
    if (end - beg > ALN_LONG_READ) return;                        // aln_pass1_long_kernel
    const int m = p.m;
    const uint32_t top = 1u << (m - 1), low = top - 1u;
    MyersU a, b;
    a.init(m, top);
    b.init(m, top);
    int best1 = m, best2 = m, f1 = 0, l1 = -1, f2 = 0, l2 = -1;      // HW scores never exceed m
    bool any1 = false, any2 = false;
    for (long long s0 = beg; s0 < end; s0 += (long long)low) {
        const long long s1 = (s0 + (long long)low < end) ? s0 + (long long)low : end;
        uint32_t j = 0, jr = low, kl1 = 0x7FFFFFFFu, kf1 = 0x7FFFFFFFu, kl2 = 0x7FFFFFFFu, kf2 = 0x7FFFFFFFu;
        for_each_byte(p.bytes, s0, s1, [&](uint32_t ch, long long) {
            const uint2 e = s_peq[ch];
            a.step(e.x, top);
            b.step(e.y, top);
            kl1 = min(kl1, a.sk + jr);
            kf1 = min(kf1, a.sk + j);
            kl2 = min(kl2, b.sk + jr);
            kf2 = min(kf2, b.sk + j);
            ++j;
            --jr;
        });
        const int base = (int)(s0 - beg);
        const int sa = (int)(kl1 >> (m - 1)), sb = (int)(kl2 >> (m - 1));
        if (sa < best1 || !any1) { best1 = sa; f1 = base + (int)(kf1 & low); l1 = base + (int)(low - (kl1 & low)); any1 = true; }
        else if (sa == best1) l1 = base + (int)(low - (kl1 & low));
        if (sb < best2 || !any2) { best2 = sb; f2 = base + (int)(kf2 & low); l2 = base + (int)(low - (kl2 & low)); any2 = true; }
        else if (sb == best2) l2 = base + (int)(low - (kl2 & low));
    }
    // end == beg: edlib's empty target -> distance = query length for both adapters: a tie
    signed char s = 0;
    int d = -1;
    int2 span = make_int2(0, -1);
    if (best1 < best2) { s = '+'; d = best1; span = make_int2(f1, l1); }
    else if (best2 < best1) { s = '-'; d = best2; span = make_int2(f2, l2); }
    const int uniq = s != 0 && span.x == span.y;
    const int work = s ? (span.x > 2 * m + 2 ? 2 * m + 2 : span.x) + (span.y - span.x + 1) + (s == '-' ? m + d : 0) : 0;
    p.span[r] = span;
    p.work[r] = work;
    p.strand[r] = s;
    p.dist[r] = d;
    p.n_locs[r] = uniq;                                                // the others: count kernels
    if (work > ALN_LONG_WORK) p.long_list[atomicAdd(p.n_long, 1u)] = r;
    else if (s != 0 && !uniq) p.short_list[atomicAdd(p.n_short, 1u)] = r;
}

// ---- pass 1 of very long reads: one warp per read ------------------------------------------------------------------------
// One thread per read makes a pass as long as its longest read (a 100 kb nanopore read: 100 k serial columns).  Reads
// longer than ALN_LONG_READ — the front of the length-sorted order — get a warp: lane l scans the l-th 1/32 of the read
// behind 2m + 2 warm-up columns (row i of a semi-global DP column is <= i, so the columns from the piece's start on are
// exact) and tracks (minimum, first, last) of its piece for both adapters; the warp folds them: minimum of the minima,
// first optimal end from the lowest lane that has it, last from the highest.
__global__ void __launch_bounds__(ALN_THREADS) aln_pass1_long_kernel(const AlignParams p) {
    __shared__ uint2 s_peq[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) s_peq[i] = p.peq[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int m = p.m;
    const uint32_t top = 1u << (m - 1);
    for (int w = (blockIdx.x * ALN_THREADS + threadIdx.x) >> 5; w < p.n; w += (gridDim.x * ALN_THREADS) >> 5) {
        const int r = p.order[w];
        long long beg, end;
        read_extent(p, r, beg, end);
        if (end - beg <= ALN_LONG_READ) {
            // the order is descending in len / 32 and every read of the open-ended last bucket is long: behind the first
            // short read of a lower bucket there is no long one — but inside one bucket lengths are unordered, so only a
            // read that could not share a bucket with a long one ends the scan
            if (((end - beg) >> 5) < (ALN_LONG_READ >> 5)) break;
            continue;
        }
        const long long piece = (end - beg + 31) / 32;
        const long long lo = beg + piece * lane, hi = (lo + piece < end) ? lo + piece : end;
        const long long from = (lo - beg > 2 * m + 2) ? lo - (2 * m + 2) : beg;
        int b1 = 1 << 20, b2 = 1 << 20;
        long long f1 = 0, l1 = -1, f2 = 0, l2 = -1;
        if (lo < hi) {
            Myers a, b;
            a.init(m);
            b.init(m);
            for_each_byte(p.bytes, from, hi, [&](uint32_t ch, long long pos) {
                const uint2 e = s_peq[ch];
                a.step<false>(e.x, top);
                b.step<false>(e.y, top);
                if (pos >= lo) {
                    if (a.score < b1) { b1 = a.score; f1 = pos; }
                    if (a.score == b1) l1 = pos;
                    if (b.score < b2) { b2 = b.score; f2 = pos; }
                    if (b.score == b2) l2 = pos;
                }
            });
        }
        const int best1 = __reduce_min_sync(0xFFFFFFFFu, b1), best2 = __reduce_min_sync(0xFFFFFFFFu, b2);
        const bool win1 = best1 < best2, win2 = best2 < best1;
        signed char s = 0;
        int d = -1;
        int2 span = make_int2(0, -1);
        if (win1 || win2) {
            const unsigned has = __ballot_sync(0xFFFFFFFFu, (win1 ? b1 : b2) == (win1 ? best1 : best2));
            const int lf = __ffs(has) - 1, ll = 31 - __clz(has);
            const long long first = __shfl_sync(0xFFFFFFFFu, win1 ? f1 : f2, lf), last = __shfl_sync(0xFFFFFFFFu, win1 ? l1 : l2, ll);
            s = win1 ? '+' : '-';
            d = win1 ? best1 : best2;
            span = make_int2((int)(first - beg), (int)(last - beg));
        }
        if (lane == 0) {
            const int uniq = s != 0 && span.x == span.y;
            const int work = s ? (span.x > 2 * m + 2 ? 2 * m + 2 : span.x) + (span.y - span.x + 1) + (s == '-' ? m + d : 0) : 0;
            p.span[r] = span;
            p.work[r] = work;
            p.strand[r] = s;
            p.dist[r] = d;
            p.n_locs[r] = uniq;
            if (work > ALN_LONG_WORK) p.long_list[atomicAdd(p.n_long, 1u)] = r;
            else if (s != 0 && !uniq) p.short_list[atomicAdd(p.n_short, 1u)] = r;
        }
    }
}

// ---- pass 2 ---------------------------------------------------------------------------------------------------------
// leftmost start s with Levenshtein(a2, read[s..e]) == d: reversed a2 against read[e], read[e-1], ... prefix-anchored
__device__ __noinline__ long long leftmost_start(const uint8_t *__restrict__ bytes, long long beg, long long e,
                                                    const uint32_t *rpeq2 /* shared memory */, int m, int d) {
    Myers q;
    q.init(m);
    const uint32_t top = 1u << (m - 1);
    const long long avail = e - beg + 1;
    const long long maxp = avail < (long long)m + d ? avail : (long long)m + d;
    const long long lo = e - maxp + 1;                             // read[e], read[e-1], .. read[lo]
    long long last_pos = e + 1;
    // aligned 128-bit words, walked backwards (the buffer is padded and 16-byte aligned)
    for (long long w = e & ~15ll; w + 15 >= lo; w -= 16) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(bytes + w));
        const uint32_t qd[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 15; k >= 0; --k) {
            const long long pos = w + k;
            if (pos <= e && pos >= lo) {
                q.step<true>(rpeq2[(qd[k >> 2] >> (8 * (k & 3))) & 0xFFu], top);
                if (q.score == d) last_pos = pos;
            }
        }
        if (w == 0) break;
    }
    return last_pos;
}

// Reads whose first and last optimal end differ, span short (the compacted short_list: full warps of similar work).
// COUNT = true: the number of end positions at the minimum (n_locs); COUNT = false: the locations, behind the scan.
template <bool COUNT>
__global__ void __launch_bounds__(ALN_THREADS) aln_pass2_kernel(const AlignParams p) {
    __shared__ uint2 s_peq[256];
    __shared__ uint32_t s_rpeq2[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) {
        s_peq[i] = p.peq[i];
        s_rpeq2[i] = p.rpeq2[i];
    }
    __syncthreads();
    const unsigned int t = blockIdx.x * ALN_THREADS + threadIdx.x;
    if (t >= *p.n_short) return;
    const int r = p.short_list[t];                                // several end positions inside a short span
    const signed char st = p.strand[r];
    const int2 span = p.span[r];
    const int nl = COUNT ? 0x7FFFFFFF : p.n_locs[r];
    long long beg, end;
    read_extent(p, r, beg, end);
    const long long out0 = COUNT ? 0 : p.loc_off[r];
    if (!COUNT && (unsigned long long)(out0 + nl) > p.loc_cap) return;      // the host grows the buffer and repeats the pass
    const int m = p.m, d = p.dist[r];
    const bool fwd = st == '+';
    const uint32_t top = 1u << (m - 1);
    // Every location lies in [first, last] (pass 1).  Row i of a semi-global DP column is <= i, so an optimal alignment
    // of any adapter prefix ends within 2m text characters of its start: an automaton restarted 2m + 2 characters before
    // `first` reproduces the exact columns from `first` on (and only over-estimates before it), and the scan stops at `last`.
    const long long from = beg + (span.x > 2 * m + 2 ? span.x - (2 * m + 2) : 0), first = beg + span.x, to = beg + span.y + 1;
    Myers a;
    a.init(m);
    int k = 0;
    for_each_byte(p.bytes, from, to < end ? to : end, [&](uint32_t ch, long long pos) {
        const uint2 e = s_peq[ch];
        a.step<false>(fwd ? e.x : e.y, top);
        if (a.score == d && k < nl && pos >= first) {
            if (!COUNT) p.locs[out0 + k] = (int)(pos - beg);      // '+': the end position (scTagger.py:185)
            ++k;
        }
    });
    if (COUNT) p.n_locs[r] = k;
    // '-': start - len(seq) - 1 (scTagger.py:189).  Converted behind the scan so that the lanes of a warp walk their
    // start searches together (inside the scan each lane would call it at a different column: one lane at a time)
    if (!COUNT && !fwd)
        for (int i = 0; i < k; ++i) {
            const long long e = beg + p.locs[out0 + i];
            p.locs[out0 + i] = (int)((leftmost_start(p.bytes, beg, e, s_rpeq2, m, d) - beg) - (end - beg) - 1);
        }
}

// The reads with ONE location (first = last optimal end in pass 1 — nearly every read with an adapter): '+' is that end
// position, '-' the leftmost start of the alignment ending there.
__global__ void __launch_bounds__(ALN_THREADS) aln_emit_uniq_kernel(const AlignParams p) {
    __shared__ uint32_t s_rpeq2[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) s_rpeq2[i] = p.rpeq2[i];
    __syncthreads();
    const int t = blockIdx.x * ALN_THREADS + threadIdx.x;
    if (t >= p.n) return;
    const int r = p.order[t];
    const signed char st = p.strand[r];
    const int2 span = p.span[r];
    if (st == 0 || span.x != span.y) return;
    const long long out0 = p.loc_off[r];
    if ((unsigned long long)(out0 + 1) > p.loc_cap) return;
    long long beg, end;
    read_extent(p, r, beg, end);
    const long long pos = beg + span.x;
    p.locs[out0] = (int)(st == '+' ? (pos - beg) : (leftmost_start(p.bytes, beg, pos, s_rpeq2, p.m, p.dist[r]) - beg) - (end - beg) - 1);
}

// Reads whose span is long (no adapter: equal minima scattered over the whole read) would each be one serial chain of
// thousands of columns — the floor of the whole pass.  Here a warp takes one such read: every lane scans 1/32 of the span
// behind its own 2m + 2 warm-up columns (same exactness argument), first counting its locations, then — after a warp
// prefix sum of the counts — scanning again to write them in order.
template <bool COUNT>
__global__ void __launch_bounds__(ALN_THREADS) aln_pass2_long_kernel(const AlignParams p) {
    __shared__ uint2 s_peq[256];
    __shared__ uint32_t s_rpeq2[256];
    for (int i = threadIdx.x; i < 256; i += ALN_THREADS) {
        s_peq[i] = p.peq[i];
        s_rpeq2[i] = p.rpeq2[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const unsigned int n_long = *p.n_long;
    for (unsigned int w = (blockIdx.x * ALN_THREADS + threadIdx.x) >> 5; w < n_long; w += (gridDim.x * ALN_THREADS) >> 5) {
        const int r = p.long_list[w];
        const int nl = COUNT ? 0x7FFFFFFF : p.n_locs[r];
        long long beg, end;
        read_extent(p, r, beg, end);
        const long long out0 = COUNT ? 0 : p.loc_off[r];
        if (!COUNT && (nl == 0 || (unsigned long long)(out0 + nl) > p.loc_cap)) continue;
        const int m = p.m, d = p.dist[r];
        const bool fwd = p.strand[r] == '+';
        const uint32_t top = 1u << (m - 1);
        const int2 span = p.span[r];
        const long long first = beg + span.x, stop = (beg + span.y + 1 < end) ? beg + span.y + 1 : end;
        const long long piece = (stop - first + 31) / 32;
        const long long lo = first + piece * lane, hi = (lo + piece < stop) ? lo + piece : stop;     // this lane emits [lo, hi)
        const long long from = (lo - beg > 2 * m + 2) ? lo - (2 * m + 2) : beg;
        int cnt = 0;
        if (!COUNT) {
            cnt = p.long_cnt[(size_t)w * 32 + lane];              // counted by the COUNT launch
        } else if (lo < hi) {
            Myers a;
            a.init(m);
            for_each_byte(p.bytes, from, hi, [&](uint32_t ch, long long pos) {
                const uint2 e = s_peq[ch];
                a.step<false>(fwd ? e.x : e.y, top);
                cnt += (a.score == d && pos >= lo);
            });
        }
        if (COUNT) p.long_cnt[(size_t)w * 32 + lane] = cnt;
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += v;
        }
        if (COUNT) {
            if (lane == 31) p.n_locs[r] = incl;
            continue;
        }
        int k = incl - cnt;
        if (cnt > 0) {
            Myers a;
            a.init(m);
            for_each_byte(p.bytes, from, hi, [&](uint32_t ch, long long pos) {
                const uint2 e = s_peq[ch];
                a.step<false>(fwd ? e.x : e.y, top);
                if (a.score == d && pos >= lo && k < nl) {
                    p.locs[out0 + k] = (int)(pos - beg);
                    ++k;
                }
            });
            if (!fwd)                                              // converted behind the scan: see aln_pass2_kernel
                for (int i = incl - cnt; i < k; ++i) {
                    const long long e = beg + p.locs[out0 + i];
                    p.locs[out0 + i] = (int)((leftmost_start(p.bytes, beg, e, s_rpeq2, m, d) - beg) - (end - beg) - 1);
                }
        }
    }
}

// ---- exclusive scan of n_locs (int32 -> int64 offsets, n + 1 entries) --------------------------------------------------
__global__ void __launch_bounds__(1024) aln_block_sums_kernel(const int *__restrict__ n_locs, int n, long long *__restrict__ block_sums) {
    __shared__ long long warp_sums[32];
    const int i = blockIdx.x * 1024 + threadIdx.x;
    long long v = i < n ? n_locs[i] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        long long w = warp_sums[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_down_sync(0xFFFFFFFFu, w, o);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = w;
    }
}

// single block: exclusive prefix of the block sums in place, grand total to *total
// total[0] = locations of this chunk, total[1] = locations of all earlier chunks of the call (`run`, advanced here)
__global__ void __launch_bounds__(1024) aln_block_prefix_kernel(long long *__restrict__ block_sums, int n_blocks, long long *__restrict__ total,
                                                                long long *__restrict__ run) {
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
            const long long t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
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
        const long long b = *run;
        total[0] = carry;
        total[1] = b;
        *run = b + carry;
    }
}

__global__ void __launch_bounds__(1024) aln_scan_apply_kernel(const int *__restrict__ n_locs, int n,
                                                              const long long *__restrict__ block_prefix,
                                                              long long *__restrict__ loc_off /* chunk-local: pass 2 */,
                                                              long long *__restrict__ loc_off_out /* the call's numbering */,
                                                              const long long *__restrict__ total) {
    __shared__ long long warp_sums[32];
    const int i = blockIdx.x * 1024 + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = i < n ? n_locs[i] : 0;
    long long incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const long long v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const long long w = warp_sums[lane];
        long long wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const long long v = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (lane >= o) wi += v;
        }
        warp_sums[lane] = wi - w;
    }
    __syncthreads();
    const long long excl = block_prefix[blockIdx.x] + warp_sums[warp] + incl - c;
    if (i < n) {
        const long long base = total[1];
        loc_off[i] = excl;
        loc_off_out[i] = base + excl;
        if (i == n - 1) {
            loc_off[n] = excl + c;
            loc_off_out[n] = base + excl + c;
        }
    }
}

}  // namespace sct
