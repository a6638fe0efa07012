// Host side of libsctagger_b200.so: the C ABI of include/sctagger_b200.h around the kernels of match_kernels.cuh.
// Replaces run_get_matches_prefix (reference scTagger.py:688-716): instead of forking one worker per DNA prefix that
// rebuilds a trie over all reads, the reads are streamed through the GPU in chunks (pinned, double-buffered, one CUDA
// stream per buffer) and every chunk is one pass of: prep -> match (fast / generic / overflow re-run | dp) -> finalize.
// No PyTorch, no CPU fallback: every entry point fails with SCT_E_CUDA when there is no device.
#include "match_kernels.cuh"
#include "seed_kernels.cuh"
#include "../../include/sctagger_b200.h"

#include "host_common.h"

using namespace sct;
using namespace scth;

namespace {

// Chunks in flight per handle.  Three, not two: the small trailing kernels of chunk c (generic, scan: blocks of up to
// 1024 threads) cannot be placed while the persistent match kernel of chunk c+1 fills the SMs, so they run when that
// kernel ends — with two slots the host would sit in "wait for chunk c" until then and enqueue chunk c+2 too late; with
// three it has already queued chunk c+2 and the GPU goes from one match kernel straight into the next.
constexpr int N_SLOTS = 3;

struct Slot {
    int64_t cap_segs = 0, cap_bytes = 0;
    int64_t hcap_segs = 0, hcap_bytes = 0;   // pinned staging (allocated only for pageable caller arrays)
    unsigned long long tie_cap = 0;
    // pinned host staging
    uint8_t *h_bytes = nullptr;
    long long *h_off = nullptr;
    int *h_min = nullptr, *h_nt = nullptr;
    long long *h_toff = nullptr;
    Counters *h_ctr = nullptr;
    long long *h_total = nullptr;      // [0] ties of the chunk, [1] ties of all chunks before it (device-side running sum)
    // device
    uint8_t *d_bytes = nullptr, *d_colinfo = nullptr, *d_rec = nullptr, *d_segflag = nullptr;
    long long *d_off = nullptr;
    int *d_fast = nullptr, *d_gen = nullptr, *d_ovf = nullptr, *d_dp = nullptr, *d_wide = nullptr;
    int64_t pk_words = 0;
    unsigned long long *d_pk = nullptr;
    Counters *d_ctr = nullptr;
    int *d_min = nullptr, *d_nt = nullptr;
    unsigned int *d_tstart = nullptr;
    int *d_tiebuf = nullptr, *d_tieout = nullptr;
    long long *d_toff = nullptr, *d_bsums = nullptr, *d_total = nullptr;   // d_total: [0] chunk total, [1] chunk base
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_k0 = nullptr, ev_k1 = nullptr, ev_end = nullptr;
    // in flight
    int64_t n = 0, nbytes = 0, first = 0;
    long long b0 = 0;               // offset of the chunk's first byte in the caller's numbering (seg_offsets stay absolute)
    bool busy = false;
    bool copies_enqueued = false;   // the fixed-size result copies follow the pass on the stream (sct_match)
    bool redo = false;              // pass repeated after the tie buffer grew: the running tie base is not advanced again
    // where the fixed-size results of the chunk go: the caller's arrays when they are pinned, else the slot's staging
    int *dst_min = nullptr, *dst_nt = nullptr;
    long long *dst_toff = nullptr;
};

}  // namespace

struct sct_handle {
    int device = 0;
    int B = 0, bl = 0, K = 0, threads = 0, kpre = 0;
    double mem_gb = 16.0;
    int sm_count = 0;
    int max_smem_optin = 0;
    uint32_t *d_tabF = nullptr, *d_tabR = nullptr;
    uint8_t *d_patF = nullptr, *d_patR = nullptr;
    int n_tiles = 0;
    bool clean_barcodes = true, use_dp = false, fast_ok = false, seed_ok = false;
    int path_override = 0;       // 0 auto, 1 force generic, 2 force dp, 3 force the tile (all-pairs) kernel
    uint32_t *d_boff[3] = {nullptr, nullptr, nullptr}, *d_eids[3] = {nullptr, nullptr, nullptr};
    uint2 *d_ewords[3] = {nullptr, nullptr, nullptr};
    SeedIndex seed_by_k[4] = {};  // index of error bound k (k = 0 only when -mr 0), built in sct_create for k = 1..K
    bool seed_pre_by_k[4] = {false, false, false, false};   // the anchored pre-check is possible for this bound
    int chain_kernels = 1;        // option: a chunk's kernels start behind the previous chunk's match kernel
    int seed_deepening = 1;       // option: search error bounds 1, 2, .. K in turn and stop at the first hit
    int seed_direct_opt = -1;     // option seed_direct: -1 auto (sparse indexes), 0 never, 1 every phase
    int seed_prefilter_opt = -1;  // option seed_prefilter: -1 auto (per phase, by expected bucket size), 0 off, 1 on where possible
    size_t fast_smem = 0;
    int64_t chunk_segs = 0, chunk_bytes = 0;   // capacity of one slot (sct_stage batch limit), from -m
    int64_t stream_segs = 0;                    // segments per streamed chunk of sct_match (<= chunk_segs)
    int64_t stream_segs_pageable = 1 << 20;     // ... when a caller array is pageable (staged through pinned buffers)
    int64_t tail_segs = 1 << 16;                // size of the last streamed chunk (option tail_segments)
    int64_t ramp_segs = 1 << 18;                // size of the first streamed chunk (x4 per chunk up to stream_segs); 32 k was
                                                // measured 1-5 % slower: every chunk costs ~70 us of kernel tail and launch gaps
    int host_threads = 1;        // worker threads for staging / un-staging (from -t, capped)
    HostPool *pool = nullptr;
    int tie_per_seg = 8;
    int seed_ctas_per_sm = 5;               // = what __launch_bounds__(256, 5) and 35 KB of shared memory let be resident
    int seed_fast_max = SEED_MAX_LEN;       // longest clean segment of the one-window kernel (option: 0 sends everything to the windowed one)
    bool seed_pre_ok = false;
    uint32_t *d_ovf_scratch = nullptr;
    int ovf_blocks = 16, ovf_cap = 0;
    uint8_t *d_flush = nullptr;
    size_t flush_bytes = 256u << 20;
    Slot slot[N_SLOTS];
    int32_t *ties_pinned = nullptr;             // tie ids of the whole call (library-owned, pinned; sct_match's *tie_ids)
    int64_t ties_cap = 0;
    int64_t slot_segs_hint = 0, slot_bytes_hint = 0;   // full-chunk size of the call in progress
    int64_t call_segments = 0;                  // n of the sct_match call in progress
    long long *d_tie_run = nullptr;             // device-side running sum of ties over the chunks of one call
    // which of the caller's arrays are page-locked (cudaHostAlloc / cudaHostRegister / sct_host_alloc): those are
    // copied from / to directly, without the staging pass through the slot's pinned buffers
    bool in_bytes_pinned = false, in_off_pinned = false, out_min_pinned = false, out_nt_pinned = false, out_toff_pinned = false;
    int64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t work[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // live work counters of the seeded kernels (sct_work_counters)
    bool staged = false, staged_ran = false;
};

namespace {

int free_slot(Slot &s) {
    host_free(s.h_bytes); host_free(s.h_off); host_free(s.h_min); host_free(s.h_nt); host_free(s.h_toff);
    host_free(s.h_ctr); host_free(s.h_total);
    dev_free(s.d_bytes); dev_free(s.d_colinfo); dev_free(s.d_rec); dev_free(s.d_segflag); dev_free(s.d_off);
    dev_free(s.d_fast); dev_free(s.d_gen); dev_free(s.d_ovf); dev_free(s.d_dp); dev_free(s.d_ctr); dev_free(s.d_wide);
    dev_free(s.d_pk);
    s.pk_words = 0;
    dev_free(s.d_min); dev_free(s.d_nt); dev_free(s.d_tstart); dev_free(s.d_tiebuf); dev_free(s.d_tieout);
    dev_free(s.d_toff); dev_free(s.d_bsums); dev_free(s.d_total);
    s.cap_segs = s.cap_bytes = s.hcap_segs = s.hcap_bytes = 0;
    s.tie_cap = 0;
    return SCT_OK;
}

// Every warp of the seeded kernels may leave the tail of one chunk of the tie buffer unused (seed_store_result).
unsigned long long seed_tie_slack(const sct_handle *h);

int alloc_ties(Slot &s, unsigned long long cap) {
    dev_free(s.d_tiebuf);
    dev_free(s.d_tieout);
    s.tie_cap = 0;
    CUDA_TRY(cudaMalloc(&s.d_tiebuf, cap * sizeof(int)));
    CUDA_TRY(cudaMalloc(&s.d_tieout, cap * sizeof(int)));
    s.tie_cap = cap;
    return SCT_OK;
}

int ensure_slot(sct_handle *h, Slot &s, int64_t segs, int64_t bytes) {
    if (!s.stream) {
        CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreate(&s.ev_start));
        CUDA_TRY(cudaEventCreate(&s.ev_k0));
        CUDA_TRY(cudaEventCreate(&s.ev_k1));
        CUDA_TRY(cudaEventCreate(&s.ev_end));
        CUDA_TRY(cudaMallocHost(&s.h_ctr, sizeof(Counters)));
        CUDA_TRY(cudaMallocHost(&s.h_total, 2 * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_ctr, sizeof(Counters)));
        CUDA_TRY(cudaMalloc(&s.d_total, 2 * sizeof(long long)));
        if (!h->d_tie_run) {
            CUDA_TRY(cudaMalloc(&h->d_tie_run, sizeof(long long)));
            CUDA_TRY(cudaMemset(h->d_tie_run, 0, sizeof(long long)));
        }
    }
    if (segs > s.cap_segs) {
        const int64_t c = std::max<int64_t>(segs, 1024);
        dev_free(s.d_rec); dev_free(s.d_segflag); dev_free(s.d_off); dev_free(s.d_fast); dev_free(s.d_gen);
        dev_free(s.d_ovf); dev_free(s.d_dp); dev_free(s.d_min); dev_free(s.d_nt); dev_free(s.d_tstart);
        dev_free(s.d_toff); dev_free(s.d_bsums); dev_free(s.d_wide);
        s.cap_segs = 0;
        CUDA_TRY(cudaMalloc(&s.d_rec, (size_t)c * REC_BYTES));
        CUDA_TRY(cudaMalloc(&s.d_segflag, (size_t)c));
        CUDA_TRY(cudaMalloc(&s.d_off, (c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_fast, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_gen, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_ovf, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_dp, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_wide, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_min, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_nt, c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_tstart, c * sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&s.d_toff, (c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_bsums, ((c + SCAN_BLOCK - 1) / SCAN_BLOCK + 1) * sizeof(long long)));
        s.cap_segs = c;
        const unsigned long long want = (unsigned long long)c * h->tie_per_seg + 4096 + seed_tie_slack(h);
        if (want > s.tie_cap) TRY(alloc_ties(s, want));
    }
    if (bytes > s.cap_bytes) {
        const int64_t c = std::max<int64_t>(bytes, 4096);
        dev_free(s.d_bytes);
        dev_free(s.d_colinfo);
        s.cap_bytes = 0;
        CUDA_TRY(cudaMalloc(&s.d_bytes, (size_t)c + 64));
        CUDA_TRY(cudaMalloc(&s.d_colinfo, (size_t)c + 64));
        s.cap_bytes = c;
    }
    // pinned staging for the caller's pageable arrays only (page-locking costs ~1 ms per MB: a fresh handle that is fed
    // page-locked arrays allocates none of it)
    const bool need_h = !(h->in_off_pinned && h->out_min_pinned && h->out_nt_pinned && h->out_toff_pinned);
    if (need_h && segs > s.hcap_segs) {
        const int64_t c = std::max<int64_t>(segs, 1024);
        host_free(s.h_off); host_free(s.h_min); host_free(s.h_nt); host_free(s.h_toff);
        s.hcap_segs = 0;
        CUDA_TRY(cudaMallocHost(&s.h_off, (c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMallocHost(&s.h_min, c * sizeof(int)));
        CUDA_TRY(cudaMallocHost(&s.h_nt, c * sizeof(int)));
        CUDA_TRY(cudaMallocHost(&s.h_toff, (c + 1) * sizeof(long long)));
        s.hcap_segs = c;
    }
    if (!h->in_bytes_pinned && bytes > s.hcap_bytes) {
        const int64_t c = std::max<int64_t>(bytes, 4096);
        host_free(s.h_bytes);
        s.hcap_bytes = 0;
        CUDA_TRY(cudaMallocHost(&s.h_bytes, (size_t)c + 64));
        s.hcap_bytes = c;
    }
    // 2-bit packed text of the seeded mode: segment i starts at word (byte offset >> 5) + i
    const int64_t pkw = s.cap_bytes / 32 + s.cap_segs + 4;
    if (pkw > s.pk_words) {
        dev_free(s.d_pk);
        s.pk_words = 0;
        CUDA_TRY(cudaMalloc(&s.d_pk, (size_t)pkw * sizeof(unsigned long long)));
        s.pk_words = pkw;
    }
    return SCT_OK;
}

unsigned long long seed_tie_slack(const sct_handle *h) {
    return (unsigned long long)h->sm_count * (unsigned long long)(std::max(h->seed_ctas_per_sm, 4) * SEED_WARPS * SEED_TIE_CHUNK) * 2ull;
}

MatchParams make_params(sct_handle *h, Slot &s) {
    MatchParams p;
    p.tabF = h->d_tabF;
    p.tabR = h->d_tabR;
    p.n_tiles = h->n_tiles;
    p.n_barcodes = h->B;
    p.bl = h->bl;
    p.seg_bytes = s.d_bytes;
    p.seg_off = s.d_off;
    p.b0 = s.b0;
    p.rec = s.d_rec;
    p.colinfo = s.d_colinfo;
    p.segflag = s.d_segflag;
    p.fast_list = s.d_fast;
    p.gen_list = s.d_gen;
    p.ovf_list = s.d_ovf;
    p.wide_list = s.d_wide;
    p.pk = s.d_pk;
    p.ctr = s.d_ctr;
    p.min_ed = s.d_min;
    p.n_ties = s.d_nt;
    p.tie_start = s.d_tstart;
    p.tie_buf = s.d_tiebuf;
    p.tie_cap = s.tie_cap;
    p.ovf_scratch = h->d_ovf_scratch;
    p.ovf_scratch_cap = h->ovf_cap;
    p.four = 4u;
    return p;
}

// The phases of one launch: error bounds 1 .. K with iterative deepening (0 only for -mr 0), else K alone.  The anchored
// pre-check runs in a phase when it is possible there and — unless the option forces it — when the phase's buckets hold
// enough candidates for the filter to pay (expected entries per probe >= 2).  -> any phase uses the pre-check
bool seed_phases(const sct_handle *h, SeedPhases &ph) {
    const int K = h->K;
    const int first = (h->seed_deepening && K >= 1) ? 1 : K;
    ph.run_mask = 0;
    bool any = false;
    for (int k = first; k <= K; ++k) {
        ph.run_mask |= 1 << k;
        SeedIndex &ix = ph.ph[k];
        ix = h->seed_by_k[k];
        ix.first_phase = k == first ? 1 : 0;
        int min_len = 99;
        for (int q = 0; q < ix.n_pieces; ++q) min_len = std::min(min_len, ix.piece_len[q]);
        const double per_probe = (double)h->B / (double)(1ull << (2 * min_len));
        bool on = h->seed_pre_by_k[k] && h->bl == 16;
        if (h->seed_prefilter_opt == 0) on = false;
        else if (h->seed_prefilter_opt < 0) on = on && per_probe >= 2.0;
        // a sparse index (less than one barcode per bucket) is probed and verified per lane, without slot table / queue
        ix.direct = (h->seed_direct_opt < 0 ? per_probe < 1.0 : h->seed_direct_opt != 0) ? 1 : 0;
        if (ix.direct) on = false;
        ix.prefilter = on ? 1 : 0;
        any = any || on;
    }
    return any;
}

template <int K>
int launch_match(sct_handle *h, Slot &s, const MatchParams &p, int fast /* 0 none, 1 tile kernel, 2 seeded kernel */,
                 cudaEvent_t after /* everything behind the main kernel waits for this event (may be null) */) {
    if (fast == 2) {
        // seeded kernel: small static shared memory, many resident warps (latency-bound gathers + verification)
        const int grid = h->sm_count * h->seed_ctas_per_sm;
        SeedPhases ph{};
        const bool any_pre = seed_phases(h, ph);
        if (h->bl == 16 && any_pre) match_seed_kernel<K, true, true, false><<<grid, SEED_THREADS, 0, s.stream>>>(p, ph);
        else if (h->bl == 16) match_seed_kernel<K, true, false, false><<<grid, SEED_THREADS, 0, s.stream>>>(p, ph);
        else match_seed_kernel<K, false, false, false><<<grid, SEED_THREADS, 0, s.stream>>>(p, ph);
        CUDA_TRY(cudaGetLastError());
        // the segments the one-window kernel does not take (longer than 31 bases, or with a non-ACGT byte)
        const int wgrid = h->sm_count * 4;
        if (h->bl == 16 && any_pre) match_seed_kernel<K, true, true, true><<<wgrid, SEED_THREADS, 0, s.stream>>>(p, ph);
        else if (h->bl == 16) match_seed_kernel<K, true, false, true><<<wgrid, SEED_THREADS, 0, s.stream>>>(p, ph);
        else match_seed_kernel<K, false, false, true><<<wgrid, SEED_THREADS, 0, s.stream>>>(p, ph);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(s.ev_k1, s.stream));
    } else if (fast == 1) {
        if (h->bl == 16) {
            CUDA_TRY(cudaFuncSetAttribute(match_fast_kernel<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fast_smem));
            match_fast_kernel<K, true><<<h->sm_count, FAST_THREADS, h->fast_smem, s.stream>>>(p);
        } else {
            CUDA_TRY(cudaFuncSetAttribute(match_fast_kernel<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fast_smem));
            match_fast_kernel<K, false><<<h->sm_count, FAST_THREADS, h->fast_smem, s.stream>>>(p);
        }
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(s.ev_k1, s.stream));
    }
    // the generic kernels share the handle's overflow scratch with the other slot's chunk, the finalize advances the
    // running tie base: both stay behind that chunk's finalize
    if (after && fast) CUDA_TRY(cudaStreamWaitEvent(s.stream, after, 0));
    if (fast != 2) {                 // in seeded mode prep_seed_kernel leaves the generic list empty
        match_generic_kernel<K><<<h->sm_count * 4, GEN_THREADS, 0, s.stream>>>(p, 0);
        CUDA_TRY(cudaGetLastError());
    }
    match_generic_kernel<K><<<h->ovf_blocks, GEN_THREADS, 0, s.stream>>>(p, 1);
    CUDA_TRY(cudaGetLastError());
    return SCT_OK;
}

// Which kernel scores the clean short segments: 2 = seeded, 1 = tile (all pairs), 0 = none (generic / dp only).
int fast_kind(const sct_handle *h) {
    if (h->use_dp || h->path_override == 1 || h->path_override == 2) return 0;
    if (h->path_override == 3) return h->fast_ok ? 1 : 0;
    return h->seed_ok ? 2 : (h->fast_ok ? 1 : 0);
}

// Enqueue one full pass over the n segments resident in slot s (no host synchronisation).
int enqueue_pass(sct_handle *h, Slot &s, int64_t n, Slot *prev = nullptr) {
    const int fk = fast_kind(h);
    // Ordering against the chunk in the other slot (sct_match): this chunk's kernels start behind that chunk's MATCH
    // kernel (ev_k1), so its small trailing kernels (generic, scan) overlap this chunk's prep + match kernel instead of
    // leaving the GPU to themselves; only the finalize, which advances the running tie base, waits for that chunk's
    // finalize (ev_end).  Without any ordering the two streams run both chunks' persistent kernels side by side, they
    // finish together, and the GPU idles while the host handles one chunk and prepares the next.
    const bool chained = prev && prev->busy && h->chain_kernels;
    if (chained) CUDA_TRY(cudaStreamWaitEvent(s.stream, fk ? prev->ev_k1 : prev->ev_end, 0));
    const int path = h->use_dp || h->path_override == 2 ? 2 : (fk ? 0 : 1);
    MatchParams p = make_params(h, s);
    CUDA_TRY(cudaEventRecord(s.ev_start, s.stream));
    CUDA_TRY(cudaMemsetAsync(s.d_ctr, 0, sizeof(Counters), s.stream));
    int launches = 0;
    if (n > 0) {
        if (fk == 2 && path == 0)
            prep_seed_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s.stream>>>(s.d_bytes, s.d_off, s.b0, s.nbytes, (int)n, h->bl, h->K, h->kpre,
                                                                              h->seed_fast_max, s.d_pk, s.d_colinfo, s.d_segflag,
                                                                              s.d_fast, s.d_wide, s.d_ctr, s.d_min, s.d_nt);
        else
            prep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s.stream>>>(s.d_bytes, s.d_off, s.b0, s.nbytes, (int)n, h->bl, h->K, h->kpre, path,
                                                                         FAST_MAX_LEN, nullptr, s.d_rec, s.d_colinfo, s.d_segflag, s.d_fast,
                                                                         s.d_gen, s.d_dp, s.d_ctr, s.d_min, s.d_nt);
        CUDA_TRY(cudaGetLastError());
        ++launches;
        CUDA_TRY(cudaEventRecord(s.ev_k0, s.stream));
        if (path == 2) {
            DpParams q;
            q.patF = h->d_patF;
            q.patR = h->d_patR;
            q.n_barcodes = h->B;
            q.bl = h->bl;
            q.K = h->K;
            q.kpre = h->kpre;
            match_dp_kernel<<<h->sm_count * 8, 128, 0, s.stream>>>(p, q, s.d_dp);
            CUDA_TRY(cudaGetLastError());
            CUDA_TRY(cudaEventRecord(s.ev_k1, s.stream));
            ++launches;
        } else {
            switch (h->K) {
                case 0: TRY(launch_match<0>(h, s, p, fk, chained ? prev->ev_end : nullptr)); break;
                case 1: TRY(launch_match<1>(h, s, p, fk, chained ? prev->ev_end : nullptr)); break;
                case 2: TRY(launch_match<2>(h, s, p, fk, chained ? prev->ev_end : nullptr)); break;
                default: TRY(launch_match<3>(h, s, p, fk, chained ? prev->ev_end : nullptr)); break;
            }
            if (!fk) CUDA_TRY(cudaEventRecord(s.ev_k1, s.stream));
            launches += fk == 2 ? 3 : (fk ? 3 : 2);
        }
        if (chained) CUDA_TRY(cudaStreamWaitEvent(s.stream, prev->ev_end, 0));
        const int nb = (int)((n + SCAN_BLOCK - 1) / SCAN_BLOCK);
        scan_block_sums<<<nb, SCAN_BLOCK, 0, s.stream>>>(s.d_nt, (int)n, s.d_bsums);
        CUDA_TRY(cudaGetLastError());
        scan_block_prefix<<<1, 1024, 0, s.stream>>>(s.d_bsums, nb, s.d_total, h->d_tie_run, s.redo ? 1 : 0);
        CUDA_TRY(cudaGetLastError());
        scan_gather<<<nb, SCAN_BLOCK, 0, s.stream>>>(s.d_nt, (int)n, s.d_bsums, s.d_tstart, s.d_tiebuf, s.d_toff, s.d_tieout, s.tie_cap, s.d_total);
        CUDA_TRY(cudaGetLastError());
        launches += 3;
    } else {
        CUDA_TRY(cudaEventRecord(s.ev_k0, s.stream));
        CUDA_TRY(cudaEventRecord(s.ev_k1, s.stream));
        CUDA_TRY(cudaMemsetAsync(s.d_total, 0, 2 * sizeof(long long), s.stream));
    }
    CUDA_TRY(cudaEventRecord(s.ev_end, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.h_ctr, s.d_ctr, sizeof(Counters), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.h_total, s.d_total, 2 * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
    h->stats[0] += launches;
    return SCT_OK;
}

// Fixed-size results (min_ed, n_ties, tie offsets) start their way back right behind the kernels of the pass, straight
// into the caller's arrays when those are page-locked.
int enqueue_result_copies(Slot &s) {
    const int64_t n = s.n;
    if (n <= 0) return SCT_OK;
    CUDA_TRY(cudaMemcpyAsync(s.dst_min, s.d_min, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.dst_nt, s.d_nt, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.dst_toff, s.d_toff, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
    return SCT_OK;
}

// Wait for the pass in slot s; if the bump-allocated tie buffer was too small, grow it and redo the pass.
int complete_pass(sct_handle *h, Slot &s) {
    for (int attempt = 0; attempt < 4; ++attempt) {
        const double t0 = Trace::now();
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        g_tr.t_wait += Trace::now() - t0;
        if (s.h_ctr->n_bad)
            return fail(SCT_E_ARG, "seg_offsets must be non-decreasing (chunk starting at segment %lld: %u bad segments)",
                        (long long)s.first, s.h_ctr->n_bad);
        if (s.h_ctr->tie_cursor > 0xFFFFFFFFull)
            return fail(SCT_E_UNSUPPORTED, "more than 2^32 ties in one chunk of %lld segments: lower -m", (long long)s.n);
        if (s.h_ctr->tie_cursor <= s.tie_cap) {
            h->stats[1] += s.h_ctr->n_fast;
            h->stats[2] += s.h_ctr->n_gen + s.h_ctr->n_ovf;
            h->stats[6] += s.h_ctr->n_dp;
            h->stats[7] += (int64_t)s.h_ctr->n_cand;
            h->work[0] += (int64_t)s.h_ctr->n_probe;
            h->work[1] += (int64_t)s.h_ctr->n_verify;
            h->work[2] += (int64_t)s.h_ctr->n_precheck;
            h->work[3] += (int64_t)s.h_ctr->n_setup;
            h->work[4] += (int64_t)s.h_ctr->n_wide;
            h->work[5] += (int64_t)s.h_ctr->n_lop3;
            s.redo = false;
            return SCT_OK;
        }
        const unsigned long long need = s.h_ctr->tie_cursor + 4096 + seed_tie_slack(h);
        TRY(alloc_ties(s, need));
        s.redo = true;
        TRY(enqueue_pass(h, s, s.n));
        if (s.copies_enqueued) TRY(enqueue_result_copies(s));
    }
    s.redo = false;
    return fail(SCT_E_STATE, "tie buffer did not converge");
}

// The library-owned pinned buffer holding the tie ids of the whole call grows geometrically and keeps its content.
int ensure_ties(sct_handle *h, int64_t need, int64_t keep) {
    if (need <= h->ties_cap) return SCT_OK;
    const int64_t cap = std::max<int64_t>({need, 2 * h->ties_cap, (int64_t)1 << 16});
    int32_t *fresh = nullptr;
    CUDA_TRY(cudaMallocHost(&fresh, (size_t)cap * sizeof(int32_t)));
    if (keep > 0) par_memcpy(h->pool, fresh, h->ties_pinned, (size_t)keep * sizeof(int32_t));
    host_free(h->ties_pinned);
    h->ties_pinned = fresh;
    h->ties_cap = cap;
    return SCT_OK;
}

// Start H2D + pass for chunk [first, first+n) of the caller's arrays.  Page-locked caller arrays are copied from
// directly; pageable ones go through the slot's pinned staging (host threads).  The offsets travel as they are (the
// kernels subtract the chunk's first offset b0), so neither path rewrites them.
// `prev` = the slot holding the chunk before this one (still in flight): see enqueue_pass for the ordering.  The H2D copy
// of this chunk (issued before the wait) and the D2H copies of the previous one overlap the kernels.
int start_chunk(sct_handle *h, Slot &s, Slot *prev, const uint8_t *seg_bytes, const int64_t *off, int64_t first, int64_t n,
                int32_t *min_ed, int32_t *n_ties, int64_t *tie_offsets) {
    const int64_t b0 = off[first], nbytes = off[first + n] - b0;
    if (nbytes < 0) return fail(SCT_E_ARG, "seg_offsets must be non-decreasing (chunk starting at segment %lld)", (long long)first);
    const double ta = Trace::now();
    // a slot is allocated once, at the size of the call's full chunks (the ramp's small chunks would otherwise size it
    // first and every slot would be freed and allocated again two chunks later: 0.2 s on a fresh handle)
    TRY(ensure_slot(h, s, std::max<int64_t>(n, h->slot_segs_hint), std::max<int64_t>(nbytes, h->slot_bytes_hint)));
    g_tr.t_alloc += Trace::now() - ta;
    const double t0 = Trace::now();
    const uint8_t *src_bytes = seg_bytes + b0;
    const long long *src_off = reinterpret_cast<const long long *>(off + first);
    if (!h->in_bytes_pinned && nbytes > 0) {
        par_memcpy(h->pool, s.h_bytes, src_bytes, (size_t)nbytes);
        src_bytes = s.h_bytes;
    }
    if (!h->in_off_pinned) {
        par_memcpy(h->pool, s.h_off, src_off, (size_t)(n + 1) * sizeof(long long));
        src_off = s.h_off;
    }
    g_tr.t_stage += Trace::now() - t0;
    s.n = n;
    s.nbytes = nbytes;
    s.first = first;
    s.b0 = b0;
    s.redo = false;
    s.dst_min = h->out_min_pinned ? min_ed + first : s.h_min;
    s.dst_nt = h->out_nt_pinned ? n_ties + first : s.h_nt;
    s.dst_toff = h->out_toff_pinned ? reinterpret_cast<long long *>(tie_offsets + first) : s.h_toff;
    if (nbytes > 0) CUDA_TRY(cudaMemcpyAsync(s.d_bytes, src_bytes, (size_t)nbytes, cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.d_off, src_off, (size_t)(n + 1) * sizeof(long long), cudaMemcpyHostToDevice, s.stream));
    h->stats[3] += nbytes + (n + 1) * (int64_t)sizeof(long long);
    h->stats[5] += 1;
    TRY(enqueue_pass(h, s, n, prev));
    TRY(enqueue_result_copies(s));
    s.copies_enqueued = true;
    s.busy = true;
    return SCT_OK;
}

// Bring the tie ids of slot s back (into the call-wide pinned buffer, at the chunk's base) and hand the fixed-size
// results that were staged — the caller's array was pageable — over to the caller's arrays.
int fetch_results(sct_handle *h, Slot &s, int32_t *min_ed, int32_t *n_ties, int64_t *tie_offsets, int64_t *n_new_ties) {
    const int64_t n = s.n;
    const long long total = n > 0 ? s.h_total[0] : 0;
    const long long base = n > 0 ? s.h_total[1] : 0;
    if (n > 0) {
        if (!s.copies_enqueued) TRY(enqueue_result_copies(s));
        if (total > 0) {
            // grow once, to the size the call will need if the rest of the segments tie like the ones seen so far
            // (page-locking ~1 ms per MB: doubling from 64 k entries up cost a fresh handle 0.2 s on 10 M segments)
            int64_t want = base + total;
            if (want > h->ties_cap && h->call_segments > 0 && s.first + n > 0)
                want = std::max<int64_t>(want, (int64_t)(1.25 * (double)(base + total) / (double)(s.first + n) * (double)h->call_segments));
            const double ta = Trace::now();
            TRY(ensure_ties(h, want, base));
            g_tr.t_alloc += Trace::now() - ta;
            CUDA_TRY(cudaMemcpyAsync(h->ties_pinned + base, s.d_tieout, (size_t)total * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        }
        double t0 = Trace::now();
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        g_tr.t_wait += Trace::now() - t0;
        t0 = Trace::now();
        h->stats[4] += n * 8 + n * 8 + total * 4;
        if (s.dst_min == s.h_min) par_memcpy(h->pool, min_ed + s.first, s.h_min, (size_t)n * sizeof(int));
        if (s.dst_nt == s.h_nt) par_memcpy(h->pool, n_ties + s.first, s.h_nt, (size_t)n * sizeof(int));
        if (s.dst_toff == s.h_toff) par_memcpy(h->pool, tie_offsets + s.first, s.h_toff, (size_t)n * sizeof(long long));
        g_tr.t_fetch += Trace::now() - t0;
    }
    *n_new_ties = total;
    s.busy = false;
    s.copies_enqueued = false;
    return SCT_OK;
}

}  // namespace

extern "C" {

int sct_device_count(void);
static int ndev_visible() { return sct_device_count(); }
int sct_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char *sct_last_error(void) { return g_err.c_str(); }

int sct_create(int device, const uint8_t *barcodes, int32_t n_barcodes, int32_t barcode_length, int32_t max_error,
               int32_t threads, double mem_gb, sct_handle **out) {
    if (!out) return fail(SCT_E_ARG, "out is NULL");
    *out = nullptr;
    if (!barcodes && n_barcodes > 0) return fail(SCT_E_ARG, "barcodes is NULL");
    if (n_barcodes < 0) return fail(SCT_E_ARG, "n_barcodes < 0");
    if (!(mem_gb > 0)) return fail(SCT_E_ARG, "mem must be > 0 (scTagger.py:141)");
    if (!(barcode_length > 0)) return fail(SCT_E_ARG, "barcode_length must be > 0 (scTagger.py:142)");
    if (!(barcode_length > max_error && max_error >= 0))
        return fail(SCT_E_ARG, "need barcode_length > max_error >= 0 (scTagger.py:143)");
    if (threads < 1) return fail(SCT_E_ARG, "threads must be >= 1");
    int kpre = 0;
    for (long long p = 1; threads > p; p *= 4) ++kpre;   // scTagger.py:689-697
    if (kpre >= barcode_length - max_error)
        return fail(SCT_E_ARG, "barcode_length - max_error must exceed the prefix length k=%d implied by -t %d "
                    "(the reference leaks its -1 sentinel read id otherwise, SURVEY.md A.6)", kpre, threads);
    if (barcode_length > DP_MAX_BL)
        return fail(SCT_E_UNSUPPORTED, "barcode_length %d > %d is not implemented", barcode_length, DP_MAX_BL);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(SCT_E_CUDA, "no CUDA device available (sctagger_b200 has no CPU fallback)");
    }
    if (device < 0 || device >= ndev) return fail(SCT_E_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    CUDA_TRY(cudaSetDevice(device));
    sct_handle *h = new (std::nothrow) sct_handle();
    if (!h) return fail(SCT_E_NOMEM, "out of host memory");
    h->device = device;
    h->B = n_barcodes;
    h->bl = barcode_length;
    h->K = max_error;
    h->threads = threads;
    h->kpre = kpre;
    h->mem_gb = mem_gb;
    // -t host threads, at most 8, and no more than this GPU's share of the host cores (one handle per GPU)
    h->host_threads = std::max(1, std::min({(int)threads, 8, (int)std::thread::hardware_concurrency() / std::max(1, ndev_visible())}));
    h->pool = new (std::nothrow) HostPool(h->host_threads);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    h->sm_count = prop.multiProcessorCount;
    CUDA_TRY(cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
    h->use_dp = barcode_length > 16 || max_error > 3;

    const int B = n_barcodes, bl = barcode_length;
    const size_t W = ((size_t)B + 1) / 2;
    h->n_tiles = (int)std::max<size_t>(1, (W + TILE_WORDS - 1) / TILE_WORDS);
    std::vector<uint8_t> rcb((size_t)B * bl + 1), patF((size_t)B * bl + 1), patR((size_t)B * bl + 1);
    for (int b = 0; b < B; ++b)
        for (int i = 0; i < bl; ++i) {
            const uint8_t c = barcodes[(size_t)b * bl + i];
            if (!is_acgt(c)) h->clean_barcodes = false;
            rcb[(size_t)b * bl + i] = comp_char(barcodes[(size_t)b * bl + (bl - 1 - i)]);
        }
    for (size_t i = 0; i < (size_t)B * bl; ++i) {
        patF[i] = (uint8_t)map_char(barcodes[i]);
        patR[i] = (uint8_t)map_char(rcb[i]);
    }
    CUDA_TRY(cudaMalloc(&h->d_patF, (size_t)B * bl + 1));
    CUDA_TRY(cudaMalloc(&h->d_patR, (size_t)B * bl + 1));
    CUDA_TRY(cudaMemcpy(h->d_patF, patF.data(), (size_t)B * bl + 1, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(h->d_patR, patR.data(), (size_t)B * bl + 1, cudaMemcpyHostToDevice));
    if (!h->use_dp) {
        // interleaved Eq tables, layout [tile][code][lane][4]; word w = tile*128 + lane*4 + q holds barcodes 2w, 2w+1
        const size_t tab_words = (size_t)h->n_tiles * TILE_BYTES / 4;
        std::vector<uint32_t> tF(tab_words, 0u), tR(tab_words, 0u);
        for (int b = 0; b < B; ++b) {
            const size_t w = (size_t)b >> 1;
            const int f = b & 1;
            const size_t tile = w / TILE_WORDS, lane = (w % TILE_WORDS) / 4, q = w % 4;
            for (int i = 0; i < bl; ++i) {
                const size_t baseF = tile * (TILE_BYTES / 4) + (size_t)patF[(size_t)b * bl + i] * 128 + lane * 4 + q;
                const size_t baseR = tile * (TILE_BYTES / 4) + (size_t)patR[(size_t)b * bl + i] * 128 + lane * 4 + q;
                tF[baseF] |= 1u << (2 * i + f);
                tR[baseR] |= 1u << (2 * i + f);
            }
        }
        CUDA_TRY(cudaMalloc(&h->d_tabF, tab_words * 4));
        CUDA_TRY(cudaMalloc(&h->d_tabR, tab_words * 4));
        CUDA_TRY(cudaMemcpy(h->d_tabF, tF.data(), tab_words * 4, cudaMemcpyHostToDevice));
        CUDA_TRY(cudaMemcpy(h->d_tabR, tR.data(), tab_words * 4, cudaMemcpyHostToDevice));
        h->fast_smem = (size_t)h->n_tiles * TILE_BYTES + (size_t)FAST_WARPS * FAST_LIST_CAP * 4 + 16;
        h->fast_ok = h->clean_barcodes && h->fast_smem <= (size_t)h->max_smem_optin;
        // seeded kernel: per error bound k = 1..K (0 for -mr 0) an index of k+1 disjoint pieces (longer pieces last, capped
        // at 8 characters), one bucket table per piece
        {
            std::vector<uint32_t> code((size_t)B), rcode((size_t)B);
            for (int b = 0; b < B; ++b) {
                uint32_t w = 0;
                for (int i = 0; i < bl; ++i) w |= (uint32_t)patF[(size_t)b * bl + i] << (2 * i);
                code[b] = w;
                uint32_t rw = 0;                         // reverse complement, same packing
                for (int i = 0; i < bl; ++i) rw |= (uint32_t)(3 - patF[(size_t)b * bl + (bl - 1 - i)]) << (2 * i);
                rcode[b] = rw;
            }
            h->seed_ok = h->clean_barcodes && B > 0;
            for (int k = (max_error >= 1 ? 1 : 0); k <= max_error && h->seed_ok; ++k) {
                SeedIndex &ix = h->seed_by_k[k];
                const int P = k + 1;
                const int base = bl / P, extra = bl % P;
                int o = 0, min_len = 99;
                size_t boff_words = 0;
                ix.n_pieces = P;
                for (int q = 0; q < P && q < SEED_MAX_PIECES; ++q) {
                    const int ln = base + (q >= P - extra ? 1 : 0);
                    ix.piece_off[q] = o;
                    ix.piece_len[q] = std::min(ln, 8);
                    ix.boff_base[q] = (int)boff_words;
                    boff_words += ((size_t)1 << (2 * ix.piece_len[q])) + 1;
                    min_len = std::min(min_len, ix.piece_len[q]);
                    o += ln;
                }
                // below 4 characters per piece nearly every barcode is a candidate: the all-pairs tile kernel is faster there
                if (!(P <= SEED_MAX_PIECES && min_len >= 4 && (size_t)P * B < ((size_t)1 << 31))) {   // entry index shares a word with the strand bit
                    h->seed_ok = false;
                    break;
                }
                // anchored pre-check (seed_precheck): which side of the piece is tested, and how, follows the responsible-seed
                // rule of the error bound (bl = 16 only: pieces 8+8 / 5+5+6 / 4+4+4+4)
                bool pre_ok = bl == 16 && k >= 1 && k <= 3;
                for (int q = 0; q < P && pre_ok; ++q) {
                    const int po = ix.piece_off[q], ln = ix.piece_len[q];
                    bool go_left;
                    int need;                       // characters the test reads on that side
                    uint32_t lvl1_after5 = 0;
                    if (k == 1) {                   // sibling half within 1 edit: 6 of its characters
                        go_left = q == 1;
                        need = 6;
                    } else if (k == 2) {            // priority piece 2 > piece 1 > piece 0
                        go_left = q >= 1;
                        need = q == 2 ? 6 : 5;
                        lvl1_after5 = q == 2 ? 0u : 1u;
                    } else {                        // sibling piece within 1 edit
                        go_left = (q & 1) != 0;
                        need = 4;
                    }
                    const int avail = go_left ? po : bl - (po + ln);
                    if (avail < need) pre_ok = false;
                    ix.pre[q] = (uint32_t)(go_left ? bl - po : po + ln) | ((go_left ? 1u : 0u) << 5) | ((uint32_t)ln << 9) |
                                (lvl1_after5 << 14);
                }
                h->seed_pre_by_k[k] = pre_ok;
                std::vector<uint32_t> boff(boff_words, 0u), ei((size_t)P * B);
                std::vector<uint2> ew((size_t)P * B + 32, make_uint2(0u, 0u));   // + 32: the kernel's branch-free tail reads past the last bucket
                for (int q = 0; q < P; ++q) {
                    const int po = ix.piece_off[q], ln = ix.piece_len[q];
                    const uint32_t mask = (ln == 16) ? 0xFFFFFFFFu : ((1u << (2 * ln)) - 1u);
                    uint32_t *bo = boff.data() + ix.boff_base[q];
                    const size_t nb = (size_t)1 << (2 * ln);
                    for (int b = 0; b < B; ++b) ++bo[((code[b] >> (2 * po)) & mask) + 1];
                    bo[0] = (uint32_t)((size_t)q * B);
                    for (size_t kk = 1; kk <= nb; ++kk) bo[kk] += bo[kk - 1];
                    std::vector<uint32_t> cur(bo, bo + nb);
                    for (int b = 0; b < B; ++b) {            // ascending barcode id inside a bucket
                        const uint32_t at = cur[(code[b] >> (2 * po)) & mask]++;
                        ew[at] = make_uint2(code[b], rcode[b]);
                        ei[at] = (uint32_t)b;
                    }
                }
                const int di = k == 0 ? 0 : k - 1;
                CUDA_TRY(cudaMalloc(&h->d_boff[di], boff.size() * 4));
                CUDA_TRY(cudaMalloc(&h->d_ewords[di], ew.size() * sizeof(uint2)));
                CUDA_TRY(cudaMalloc(&h->d_eids[di], ei.size() * 4));
                CUDA_TRY(cudaMemcpy(h->d_boff[di], boff.data(), boff.size() * 4, cudaMemcpyHostToDevice));
                CUDA_TRY(cudaMemcpy(h->d_ewords[di], ew.data(), ew.size() * sizeof(uint2), cudaMemcpyHostToDevice));
                CUDA_TRY(cudaMemcpy(h->d_eids[di], ei.data(), ei.size() * 4, cudaMemcpyHostToDevice));
                ix.boff = h->d_boff[di];
                ix.ewords = h->d_ewords[di];
                ix.eids = h->d_eids[di];
                ix.two = 2u;
                ix.prefilter = 0;
                ix.first_phase = 0;
                ix.direct = 0;
            }
            h->seed_pre_ok = h->seed_ok && h->seed_pre_by_k[max_error];
        }
        h->ovf_cap = h->n_tiles * 512;
        CUDA_TRY(cudaMalloc(&h->d_ovf_scratch, (size_t)h->ovf_blocks * GEN_WARPS * h->ovf_cap * sizeof(uint32_t)));
    }
    // chunk sizing from -m: roughly 350 B of pinned + device memory per segment and two slots in flight
    const double budget = mem_gb * (double)(1ull << 30);
    int64_t segs = (int64_t)(budget / (2.0 * 350.0));
    segs = std::min<int64_t>(std::max<int64_t>(segs, 1024), 1 << 21);
    h->chunk_segs = segs;
    h->chunk_bytes = segs * 40;
    // streamed chunks are kept small enough that staging / copies of one chunk overlap the kernels of the other slot
    h->stream_segs = std::min<int64_t>(segs, 1 << 20);
    // the table uploads above are pageable cudaMemcpy's on the legacy stream; the kernels run on non-blocking streams
    CUDA_TRY(cudaDeviceSynchronize());
    *out = h;
    return SCT_OK;
}

int sct_destroy(sct_handle *h) {
    if (!h) return SCT_OK;
    cudaSetDevice(h->device);
    for (Slot &s : h->slot) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        free_slot(s);
        if (s.ev_start) cudaEventDestroy(s.ev_start);
        if (s.ev_k0) cudaEventDestroy(s.ev_k0);
        if (s.ev_k1) cudaEventDestroy(s.ev_k1);
        if (s.ev_end) cudaEventDestroy(s.ev_end);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    dev_free(h->d_tabF); dev_free(h->d_tabR); dev_free(h->d_patF); dev_free(h->d_patR);
    dev_free(h->d_ovf_scratch); dev_free(h->d_flush);
    for (int i = 0; i < 3; ++i) { dev_free(h->d_boff[i]); dev_free(h->d_ewords[i]); dev_free(h->d_eids[i]); }
    dev_free(h->d_tie_run);
    host_free(h->ties_pinned);
    delete h->pool;
    delete h;
    return SCT_OK;
}

int sct_set_option(sct_handle *h, const char *key, int64_t value) {
    if (!h || !key) return fail(SCT_E_ARG, "NULL argument");
    const std::string k(key);
    if (k == "chunk_segments") {
        if (value < 1) return fail(SCT_E_ARG, "chunk_segments must be >= 1");
        h->chunk_segs = value;
        h->stream_segs = value;
        h->stream_segs_pageable = value;
        h->ramp_segs = value;
        h->chunk_bytes = std::max<int64_t>(h->chunk_bytes, 64);
    } else if (k == "chunk_bytes") {
        if (value < 1) return fail(SCT_E_ARG, "chunk_bytes must be >= 1");
        h->chunk_bytes = value;
    } else if (k == "tie_capacity_per_segment") {
        if (value < 0) return fail(SCT_E_ARG, "tie_capacity_per_segment must be >= 0");
        h->tie_per_seg = (int)value;
    } else if (k == "path") {   // 0 auto, 1 generic bit-parallel kernel, 2 DP kernel, 3 tile (all-pairs) kernel instead of the seeded one
        if (value < 0 || value > 3) return fail(SCT_E_ARG, "path must be 0..3");
        if ((value == 1 || value == 3) && h->use_dp)
            return fail(SCT_E_UNSUPPORTED, "bit-parallel kernels need barcode_length <= 16 and max_error <= 3");
        if (value == 3 && !h->fast_ok)
            return fail(SCT_E_UNSUPPORTED, "the tile kernel needs ACGT-only barcodes and a table that fits shared memory");
        h->path_override = (int)value;
    } else if (k == "seed_prefilter") {
        if (value > 0 && !h->seed_pre_ok) return fail(SCT_E_ARG, "seed_prefilter is not available for this barcode length / -mr");
        h->seed_prefilter_opt = value < 0 ? -1 : (value != 0 ? 1 : 0);   // -1 auto, 0 off, 1 on wherever possible
    } else if (k == "chain_kernels") {
        h->chain_kernels = value != 0 ? 1 : 0;
    } else if (k == "seed_direct") {        // -1 auto, 0 slot-table path for every phase, 1 per-lane path for every phase
        h->seed_direct_opt = value < 0 ? -1 : (value != 0 ? 1 : 0);
    } else if (k == "seed_deepening") {     // 1 (default): error bounds 1..K in turn, stop at the first hit; 0: bound K only
        h->seed_deepening = value != 0 ? 1 : 0;
    } else if (k == "ramp_segments") {
        if (value < 1) return fail(SCT_E_ARG, "ramp_segments must be >= 1");
        h->ramp_segs = value;
    } else if (k == "reserve_segments") {
        // allocate the streaming slots now (device buffers and the pinned staging of pageable caller arrays) for chunks of
        // up to `value` segments, instead of inside the first sct_match: the CLI does this while it parses the -lr file
        if (value < 1) return fail(SCT_E_ARG, "reserve_segments must be >= 1");
        CUDA_TRY(cudaSetDevice(h->device));
        const int64_t segs = std::min<int64_t>(value, h->stream_segs);
        const int64_t bytes = std::min<int64_t>(h->chunk_bytes, segs * 32 + 4096);
        h->in_bytes_pinned = h->in_off_pinned = h->out_min_pinned = h->out_nt_pinned = h->out_toff_pinned = false;
        for (Slot &s : h->slot) TRY(ensure_slot(h, s, segs, bytes));
    } else if (k == "tail_segments") {
        if (value < 0) return fail(SCT_E_ARG, "tail_segments must be >= 0");
        h->tail_segs = value;
    } else if (k == "stream_segments") {
        if (value < 1) return fail(SCT_E_ARG, "stream_segments must be >= 1");
        h->stream_segs = std::min<int64_t>(value, h->chunk_segs);
        h->stream_segs_pageable = h->stream_segs;
    } else if (k == "seed_fast_max") {   // test hook: 0 routes every segment through the windowed variant
        if (value < 0 || value > SEED_MAX_LEN) return fail(SCT_E_ARG, "seed_fast_max must be 0..%d", SEED_MAX_LEN);
        h->seed_fast_max = (int)value;
    } else if (k == "seed_ctas_per_sm") {
        if (value < 1 || value > 32) return fail(SCT_E_ARG, "seed_ctas_per_sm must be 1..32");
        h->seed_ctas_per_sm = (int)value;
    } else {
        return fail(SCT_E_ARG, "unknown option '%s'", key);
    }
    return SCT_OK;
}

int sct_stats(sct_handle *h, int64_t stats[8]) {
    if (!h || !stats) return fail(SCT_E_ARG, "NULL argument");
    memcpy(stats, h->stats, sizeof h->stats);
    return SCT_OK;
}

int sct_chunk_capacity(sct_handle *h, int64_t *max_segments, int64_t *max_bytes) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (max_segments) *max_segments = h->chunk_segs;
    if (max_bytes) *max_bytes = h->chunk_bytes;
    return SCT_OK;
}

static int match_impl(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n, int32_t *min_ed,
                      int32_t *n_ties, int64_t *tie_offsets, const int32_t **tie_ids, int64_t *n_tie_ids);

int sct_match(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n, int32_t *min_ed,
              int32_t *n_ties, int64_t *tie_offsets, const int32_t **tie_ids, int64_t *n_tie_ids) {
    const int rc = match_impl(h, seg_bytes, seg_offsets, n, min_ed, n_ties, tie_offsets, tie_ids, n_tie_ids);
    if (rc != SCT_OK && h) {
        // a call that failed half way leaves chunks in flight: drain them so that the handle stays usable (the error
        // message of the failure is kept)
        const std::string keep = g_err;
        for (Slot &s : h->slot) {
            if (s.stream) cudaStreamSynchronize(s.stream);
            s.busy = s.copies_enqueued = s.redo = false;
        }
        cudaGetLastError();
        g_err = keep;
    }
    return rc;
}

static int match_impl(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n, int32_t *min_ed,
                      int32_t *n_ties, int64_t *tie_offsets, const int32_t **tie_ids, int64_t *n_tie_ids) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (n < 0) return fail(SCT_E_ARG, "n < 0");
    if (n > 0x7FFFFFF0ll) return fail(SCT_E_UNSUPPORTED, "more than 2^31 segments in one call");
    if (!seg_offsets || !tie_offsets || !tie_ids || !n_tie_ids || (n > 0 && (!min_ed || !n_ties)))
        return fail(SCT_E_ARG, "NULL argument");
    if (n > 0 && !seg_bytes && seg_offsets[n] > seg_offsets[0]) return fail(SCT_E_ARG, "seg_bytes is NULL");
    const double t_begin = Trace::now();
    g_tr.t_stage = g_tr.t_wait = g_tr.t_fetch = g_tr.t_alloc = 0;
    CUDA_TRY(cudaSetDevice(h->device));
    memset(h->stats, 0, sizeof h->stats);
    memset(h->work, 0, sizeof h->work);
    h->staged = h->staged_ran = false;
    h->call_segments = n;
    // page-locked caller arrays (sct_host_alloc / sct_host_register / cudaHostAlloc) are copied from and to directly
    const int64_t all_bytes = n > 0 ? seg_offsets[n] - seg_offsets[0] : 0;
    h->in_bytes_pinned = n > 0 && all_bytes > 0 && is_pinned(seg_bytes + seg_offsets[0], (size_t)all_bytes);
    h->in_off_pinned = n > 0 && is_pinned(seg_offsets, (size_t)(n + 1) * sizeof(int64_t));
    h->out_min_pinned = n > 0 && is_pinned(min_ed, (size_t)n * sizeof(int32_t));
    h->out_nt_pinned = n > 0 && is_pinned(n_ties, (size_t)n * sizeof(int32_t));
    h->out_toff_pinned = n > 0 && is_pinned(tie_offsets, (size_t)(n + 1) * sizeof(int64_t));
    g_tr.t_check = Trace::now() - t_begin;
    NvtxRange nv_call("sct_match");

    int64_t tie_total = 0;
    int64_t first = 0, chunk = 0;
    auto finish = [&](int si) -> int {
        Slot &s = h->slot[si];
        if (!s.busy) return SCT_OK;
        NvtxRange nv("sct_match: wait + D2H ties");
        TRY(complete_pass(h, s));
        int64_t got = 0;
        TRY(fetch_results(h, s, min_ed, n_ties, tie_offsets, &got));
        tie_total += got;
        return SCT_OK;
    };
    if (n > 0) {
        TRY(ensure_slot(h, h->slot[0], 1, 1));
        CUDA_TRY(cudaMemsetAsync(h->d_tie_run, 0, sizeof(long long), h->slot[0].stream));   // chunk 0 runs on this stream
    }
    // the first chunks are small so that the first kernel starts (and the pipeline fills) after a short copy
    // pageable caller arrays go through pinned staging buffers, whose allocation (page-locking, ~1 ms per MB, three slots)
    // a fresh handle pays in its first call: smaller chunks there (a chunk costs ~70 us, the staging of a 1 Mi-segment
    // chunk ~60 ms of allocation)
    const bool all_pinned = h->in_bytes_pinned && h->in_off_pinned && h->out_min_pinned && h->out_nt_pinned && h->out_toff_pinned;
    const int64_t stream_segs = all_pinned ? h->stream_segs : std::min<int64_t>(h->stream_segs, h->stream_segs_pageable);
    int64_t ramp = std::min<int64_t>(stream_segs, h->ramp_segs);
    h->slot_segs_hint = std::min<int64_t>(stream_segs, n);
    h->slot_bytes_hint = n > 0 ? std::min<int64_t>(h->chunk_bytes, (int64_t)(1.1 * (double)all_bytes / (double)n * (double)h->slot_segs_hint) + 4096) : 0;
    while (first < n) {
        int64_t cn = std::min<int64_t>(std::min(stream_segs, ramp), n - first);
        // the call ends with a SMALL chunk: what follows the last kernel — its result copies, the tie copy, two host
        // synchronisations — is exposed, and it is proportional to the last chunk
        if (cn == n - first && cn > 2 * h->tail_segs) cn -= h->tail_segs;
        ramp = std::min<int64_t>(stream_segs, ramp * 4);   // 32 k, 128 k, 512 k, then full chunks: every chunk costs ~70 us of kernel tail + launch gaps
        const int64_t limit = seg_offsets[first] + h->chunk_bytes;
        if (seg_offsets[first + cn] > limit) {
            const int64_t *e = std::upper_bound(seg_offsets + first, seg_offsets + first + cn + 1, limit);
            cn = std::max<int64_t>(1, (e - (seg_offsets + first)) - 1);
        }
        const int si = (int)(chunk % N_SLOTS);
        TRY(finish(si));                 // the chunk that used this slot N_SLOTS chunks ago
        {
            NvtxRange nv("sct_match: stage + enqueue chunk");
            Slot *prev = chunk > 0 ? &h->slot[(chunk - 1) % N_SLOTS] : nullptr;
            TRY(start_chunk(h, h->slot[si], prev, seg_bytes, seg_offsets, first, cn, min_ed, n_ties, tie_offsets));
        }
        first += cn;
        ++chunk;
    }
    // drain in chunk order
    for (int64_t c = std::max<int64_t>(0, chunk - N_SLOTS); c < chunk; ++c) TRY(finish((int)(c % N_SLOTS)));
    tie_offsets[n] = tie_total;
    *tie_ids = tie_total > 0 ? h->ties_pinned : nullptr;
    *n_tie_ids = tie_total;
    if (g_tr.on)
        fprintf(stderr, "[sct_match] n=%lld chunks=%lld total %.3f ms: check %.3f alloc %.3f stage %.3f wait %.3f fetch %.3f pinned in %d%d out %d%d%d\n",
                (long long)n, (long long)chunk, 1e3 * (Trace::now() - t_begin), 1e3 * g_tr.t_check, 1e3 * g_tr.t_alloc, 1e3 * g_tr.t_stage,
                1e3 * g_tr.t_wait, 1e3 * g_tr.t_fetch, h->in_bytes_pinned, h->in_off_pinned, h->out_min_pinned,
                h->out_nt_pinned, h->out_toff_pinned);
    return SCT_OK;
}

int sct_stage(sct_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n) {
    if (!h || !seg_offsets || n < 0) return fail(SCT_E_ARG, "bad argument");
    if (n > 0x7FFFFFF0ll) return fail(SCT_E_UNSUPPORTED, "more than 2^31 segments in one batch");
    CUDA_TRY(cudaSetDevice(h->device));
    Slot &s = h->slot[0];
    const int64_t nbytes = seg_offsets[n] - seg_offsets[0];
    if (nbytes < 0) return fail(SCT_E_ARG, "seg_offsets must be non-decreasing");
    // the resident batch may be larger than a streamed chunk (bench.py keeps a whole 10 M-segment shard in HBM): it only
    // has to fit the device, which cudaMalloc decides
    // the resident batch is copied straight from / to the caller's arrays: no pinned staging
    h->in_bytes_pinned = h->in_off_pinned = h->out_min_pinned = h->out_nt_pinned = h->out_toff_pinned = true;
    TRY(ensure_slot(h, s, n, nbytes));
    const int64_t b0 = seg_offsets[0];
    if (nbytes > 0) CUDA_TRY(cudaMemcpyAsync(s.d_bytes, seg_bytes + b0, (size_t)nbytes, cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.d_off, seg_offsets, (size_t)(n + 1) * sizeof(long long), cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    s.n = n;
    s.nbytes = nbytes;
    s.first = 0;
    s.b0 = b0;
    s.redo = false;
    h->staged = true;
    h->staged_ran = false;
    return SCT_OK;
}

int sct_run_staged(sct_handle *h, float *pass_ms, float *match_kernel_ms) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (!h->staged) return fail(SCT_E_STATE, "sct_stage has not been called");
    CUDA_TRY(cudaSetDevice(h->device));
    Slot &s = h->slot[0];
    memset(h->stats, 0, sizeof h->stats);
    memset(h->work, 0, sizeof h->work);
    CUDA_TRY(cudaMemsetAsync(h->d_tie_run, 0, sizeof(long long), s.stream));
    s.redo = false;
    s.copies_enqueued = false;
    TRY(enqueue_pass(h, s, s.n));
    TRY(complete_pass(h, s));
    float a = 0.f, b = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&a, s.ev_start, s.ev_end));
    CUDA_TRY(cudaEventElapsedTime(&b, s.ev_k0, s.ev_k1));
    if (pass_ms) *pass_ms = a;
    if (match_kernel_ms) *match_kernel_ms = b;
    h->staged_ran = true;
    return SCT_OK;
}

int sct_fetch_staged(sct_handle *h, int32_t *min_ed, int32_t *n_ties, int64_t *tie_offsets, const int32_t **tie_ids,
                     int64_t *n_tie_ids) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (!h->staged_ran) return fail(SCT_E_STATE, "sct_run_staged has not been called");
    if (!tie_offsets || !tie_ids || !n_tie_ids) return fail(SCT_E_ARG, "NULL argument");
    CUDA_TRY(cudaSetDevice(h->device));
    Slot &s = h->slot[0];
    const int64_t n = s.n;
    int64_t got = 0;
    if (n > 0) {
        // straight into the caller's arrays (cudaMemcpy handles pageable destinations)
        CUDA_TRY(cudaMemcpyAsync(min_ed, s.d_min, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaMemcpyAsync(n_ties, s.d_nt, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaMemcpyAsync(tie_offsets, s.d_toff, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
        got = s.h_total[0];
        if (got > 0) {
            TRY(ensure_ties(h, got, 0));
            CUDA_TRY(cudaMemcpyAsync(h->ties_pinned, s.d_tieout, (size_t)got * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        }
        CUDA_TRY(cudaStreamSynchronize(s.stream));
    }
    tie_offsets[n] = got;
    *tie_ids = got > 0 ? h->ties_pinned : nullptr;
    *n_tie_ids = got;
    return SCT_OK;
}

int sct_host_alloc(size_t bytes, void **out) {
    if (!out) return fail(SCT_E_ARG, "out is NULL");
    *out = nullptr;
    CUDA_TRY(cudaHostAlloc(out, std::max<size_t>(bytes, 64), cudaHostAllocPortable));
    return SCT_OK;
}

int sct_host_free(void *p) {
    if (p) CUDA_TRY(cudaFreeHost(p));
    return SCT_OK;
}

int sct_host_register(void *p, size_t bytes) {
    if (!p || bytes == 0) return fail(SCT_E_ARG, "bad argument");
    CUDA_TRY(cudaHostRegister(p, bytes, cudaHostRegisterPortable));
    return SCT_OK;
}

int sct_host_unregister(void *p) {
    if (p) CUDA_TRY(cudaHostUnregister(p));
    return SCT_OK;
}

const char *sct_kernel_name(sct_handle *h) {
    if (!h) return "";
    if (h->use_dp || h->path_override == 2) return "match_dp_kernel";
    switch (fast_kind(h)) {
        case 2: return "match_seed_kernel";
        case 1: return "match_fast_kernel";
        default: return "match_generic_kernel";
    }
}

int sct_work_counters(sct_handle *h, int64_t counters[8]) {
    if (!h || !counters) return fail(SCT_E_ARG, "NULL argument");
    memcpy(counters, h->work, sizeof h->work);
    return SCT_OK;
}

int sct_flush_l2(sct_handle *h) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    CUDA_TRY(cudaSetDevice(h->device));
    if (!h->d_flush) CUDA_TRY(cudaMalloc(&h->d_flush, h->flush_bytes));
    Slot &s = h->slot[0];
    if (!s.stream) TRY(ensure_slot(h, s, 1, 1));
    CUDA_TRY(cudaMemsetAsync(h->d_flush, 0x5a, h->flush_bytes, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    return SCT_OK;
}

int sct_measure_int32_peak(int device, int mode, int repeats, double *lane_ops_per_s) {
    if (!lane_ops_per_s || mode < 0 || mode > 2 || repeats < 1) return fail(SCT_E_ARG, "bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(SCT_E_CUDA, "no CUDA device available");
    }
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    uint32_t *d_out = nullptr;
    CUDA_TRY(cudaMalloc(&d_out, 64));
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    const int blocks = prop.multiProcessorCount * 2, threads = 1024, iters = 4096;
    double best = 0.0;
    for (int r = 0; r < repeats + 1; ++r) {
        CUDA_TRY(cudaEventRecord(e0));
        if (mode == 0) int32_peak_kernel<0><<<blocks, threads>>>(d_out, iters, 12345u + r, 4u);
        else if (mode == 1) int32_peak_kernel<1><<<blocks, threads>>>(d_out, iters, 12345u + r, 4u);
        else int32_peak_kernel<2><<<blocks, threads>>>(d_out, iters, 12346u + 2 * r, 4u);
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaEventRecord(e1));
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        const double ops = (double)blocks * threads * (double)iters * 16.0 * 8.0;
        if (r > 0) best = std::max(best, ops / (ms * 1e-3));   // r == 0 is the warm-up
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    *lane_ops_per_s = best;
    return SCT_OK;
}

}  // extern "C"
