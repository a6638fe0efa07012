// Exact whitelist search of `extract_sr_bc_from_lr` (reference scTagger.py:792-844) on sm_100a.
//
// The reference builds an Aho-Corasick automaton over every whitelist barcode and its reverse complement
// (scTagger.py:802-809) and, per long-read segment, lists all occurrences (`A.iter(seg)`, scTagger.py:823); a segment with
// exactly one occurrence adds one to the count of that barcode (scTagger.py:824-827).  All words of a whitelist have one
// length L, so "an occurrence ending at position j" is "the window seg[j-L+1 .. j] is a key": the automaton becomes an
// open-addressing hash table in HBM keyed by the 2-bit packed window (L <= 31 -> one 64-bit key), probed once per window
// with a rolling key.  The table is built on the GPU too: one thread per add_word call; the automaton's "a later add_word
// of the same key replaces the value" is an atomicMax on (call order << 32 | barcode index).
//
// HBM-bound integer work: per window one 16-byte slot read = one 32-byte DRAM sector is the
// algorithmic traffic; the table (2 x 2W keys at load factor <= 0.5, 16 B/slot: 512 MiB for the 6.8 M-barcode 10x
// whitelist) does not fit L2 and a random probe really costs a 128-byte line (ncu: 130 B of DRAM reads per probe), so a
// one-hash bitmap of <= 32 MiB — a quarter of B200's 126 MB L2 — screens the windows first.  The segments stream through once.
// No tensor cores, no collective.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace sctwl {

constexpr unsigned long long WL_EMPTY = ~0ull;
constexpr int WL_THREADS = 256;

struct WlTable {
    ulonglong2 *slots;            // .x = key (WL_EMPTY = free), .y = call order << 32 | barcode index
    unsigned long long mask;      // n_slots - 1
    int L;                        // word length (1..31)
    unsigned int *filter;         // one-hash membership bitmap sized to stay L2-resident (<= 32 MiB): bit set for every key
    unsigned long long fmask;     // n_filter_bits - 1
    const uint8_t *odd;           // literal words: a byte outside ACGT, or a length other than L; concatenated
    const int *odd_off;           // n_odd + 1 offsets into odd
    const int *odd_idx;           // their barcode index
    int n_odd;
    int odd_always;               // some literal word is pure ACGT (other length): every segment is compared
};

__host__ __device__ __forceinline__ unsigned long long wl_hash(unsigned long long k) {
    k ^= k >> 33;
    k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33;
    k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return k;
}

__device__ __forceinline__ int wl_code(uint8_t c) { return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1; }

// One thread per add_word call: calls 0..W-1 are the forward words (value idx), W..2W-1 the reverse complements
// (value -idx; only abs(value) is ever used, scTagger.py:827).  Words with a byte outside ACGT are left to the host
// (they go to the literal list, like the words of another length, whose rows of `words` are zero bytes).
__global__ void __launch_bounds__(WL_THREADS) wl_insert_kernel(WlTable t, const uint8_t *__restrict__ words,
                                                               long long n_words, unsigned int *__restrict__ n_inserted) {
    const long long call = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (call >= 2 * n_words) return;
    const bool rc = call >= n_words;
    const long long w = rc ? call - n_words : call;
    const uint8_t *s = words + w * t.L;
    unsigned long long key = 0;
    for (int i = 0; i < t.L; ++i) {
        const int c = wl_code(s[rc ? t.L - 1 - i : i]);
        if (c < 0) return;
        key = (key << 2) | (unsigned long long)(rc ? 3 - c : c);
    }
    const unsigned long long val = ((unsigned long long)call << 32) | (unsigned long long)w;
    unsigned long long slot = wl_hash(key) & t.mask;
    for (;;) {
        unsigned long long *kp = &t.slots[slot].x;
        const unsigned long long prev = atomicCAS(kp, WL_EMPTY, key);
        if (prev == WL_EMPTY || prev == key) {
            const unsigned long long fb = (wl_hash(key) >> 24) & t.fmask;
            atomicOr(&t.filter[fb >> 5], 1u << (fb & 31));
            atomicMax(&t.slots[slot].y, val);
            if (prev == WL_EMPTY) atomicAdd(n_inserted, 1u);
            return;
        }
        slot = (slot + 1) & t.mask;
    }
}

// A window that is not a whitelist word (~95 % of them) is answered by the filter word in L2; only filter positives
// (true hits + n_keys / n_filter_bits false positives) touch the table in HBM, with streaming loads so that the table's
// lines do not push the filter out of L2.
__device__ __forceinline__ int wl_probe(const WlTable &t, unsigned long long key) {
    const unsigned long long hsh = wl_hash(key);
    const unsigned long long fb = (hsh >> 24) & t.fmask;
    if (!((__ldg(&t.filter[fb >> 5]) >> (fb & 31)) & 1u)) return -1;
    unsigned long long slot = hsh & t.mask;
    for (;;) {
        const ulonglong2 e = __ldcs(&t.slots[slot]);
        if (e.x == key) return (int)(unsigned int)(e.y & 0xFFFFFFFFull);
        if (e.x == WL_EMPTY) return -1;
        slot = (slot + 1) & t.mask;
    }
}

struct WlCountParams {
    const uint8_t *bytes;                 // concatenated segments of the chunk
    const long long *off;                 // n + 1 offsets in the CALLER's numbering: byte j of the chunk is offset b0 + j
    long long b0;
    long long n;
    long long first_rid;                  // index of segment 0 of the chunk in the whole job (first-seen order)
    int *n_hits;                          // per segment: occurrences of whitelist words (nullable)
    int *hit_word;                        // per segment: barcode index of the occurrence when there is exactly one
                                          // (one of them when there are several), -1 if none (nullable)
    unsigned int *counts;                 // per barcode: segments whose only occurrence is this barcode
    unsigned long long *first_seen;       // per barcode: smallest such segment index (Counter insertion order)
};

// One thread per segment, rolling 2-bit key over the bytes; a byte outside ACGT restarts the run (no packed key can
// match across it).  Segments holding such a byte are additionally compared, window by window, with the literal words
// (all segments are when a literal word is pure ACGT).  Real whitelists have no literal words.
__global__ void __launch_bounds__(WL_THREADS) wl_count_kernel(WlTable t, WlCountParams p) {
    const unsigned long long kmask = (t.L >= 32) ? ~0ull : ((1ull << (2 * t.L)) - 1ull);
    for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < p.n; r += (long long)gridDim.x * blockDim.x) {
        const long long o = p.off[r] - p.b0;
        const int n = (int)(p.off[r + 1] - p.b0 - o);
        const uint8_t *s = p.bytes + o;
        unsigned long long key = 0;
        int run = 0, hits = 0, first = -1;
        bool dirty = false;
        for (int j = 0; j < n; ++j) {
            const int c = wl_code(__ldg(s + j));
            if (c < 0) {
                run = 0;
                dirty = true;
                continue;
            }
            key = ((key << 2) | (unsigned long long)c) & kmask;
            if (++run >= t.L) {
                const int w = wl_probe(t, key);
                if (w >= 0) {
                    if (hits == 0) first = w;
                    ++hits;
                }
            }
        }
        if (t.n_odd > 0 && (dirty || t.odd_always)) {
            for (int q = 0; q < t.n_odd; ++q) {
                const int wo = t.odd_off[q], wl = t.odd_off[q + 1] - wo;
                const uint8_t *w = t.odd + wo;
                for (int j = 0; j + wl <= n; ++j) {
                    int i = 0;
                    while (i < wl && __ldg(s + j + i) == w[i]) ++i;
                    if (i == wl) {
                        if (hits == 0) first = t.odd_idx[q];
                        ++hits;
                    }
                }
            }
        }
        if (hits == 1) {
            atomicAdd(&p.counts[first], 1u);
            atomicMin(&p.first_seen[first], (unsigned long long)(p.first_rid + r));
        }
        if (p.n_hits) p.n_hits[r] = hits;
        if (p.hit_word) p.hit_word[r] = first;
    }
}

// Counter read-back: the words that were counted at all (a few thousand cells out of millions of whitelist words) are
// compacted on the device — (first segment that counted it, word index, count) — so that sct_wl_result moves kilobytes
// instead of 12 bytes per whitelist word.
struct WlSeen {
    unsigned long long first;
    int word;
    unsigned int count;
};
__global__ void __launch_bounds__(256) wl_compact_kernel(const unsigned int *__restrict__ counts,
                                                         const unsigned long long *__restrict__ first_seen, long long n_words,
                                                         WlSeen *__restrict__ out, unsigned long long cap,
                                                         unsigned long long *__restrict__ n_out) {
    for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < n_words; w += (long long)gridDim.x * blockDim.x) {
        const unsigned int c = counts[w];
        if (c) {
            const unsigned long long at = atomicAdd(n_out, 1ull);
            if (at < cap) out[at] = WlSeen{first_seen[w], (int)w, c};
        }
    }
}

}  // namespace sctwl
