// Host side of the exact whitelist search of `extract_sr_bc_from_lr` (C ABI: sct_wl_* in include/sctagger_b200.h) around
// the kernels of wl_kernels.cuh.  Replaces the Aho-Corasick automaton and the per-line Counter loop of the reference
// (scTagger.py:802-827): the table is built on the GPU once, the segments are streamed through in chunks (pinned,
// double-buffered, one stream per buffer) and the per-barcode counts stay on the device until sct_wl_result.
// No CPU fallback.
#include "wl_kernels.cuh"
#include "host_common.h"

#include <map>

using namespace sctwl;
using namespace scth;

namespace {

struct WSlot {
    int64_t cap_segs = 0, cap_bytes = 0;
    uint8_t *h_bytes = nullptr, *d_bytes = nullptr;
    long long *h_off = nullptr, *d_off = nullptr;
    int *h_hits = nullptr, *h_word = nullptr, *d_hits = nullptr, *d_word = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int64_t n = 0, first = 0;
    long long b0 = 0;                       // offset of the chunk's first byte in the caller's numbering
    int *dst_hits = nullptr, *dst_word = nullptr;   // the caller's arrays when they are page-locked, else the staging
    bool busy = false;
};

}  // namespace

struct sct_wl_handle {
    int device = 0, sm_count = 148, host_threads = 1;
    WlTable tab{};
    int64_t n_words = 0, n_slots = 0, n_keys = 0, n_filter_bits = 0;
    uint8_t *d_odd = nullptr;
    int *d_odd_idx = nullptr, *d_odd_off = nullptr;
    unsigned int *d_counts = nullptr;
    unsigned long long *d_first = nullptr;
    WlSeen *d_seen = nullptr;               // sct_wl_result: compacted (first segment, word, count) of the counted words
    unsigned long long *d_nseen = nullptr;
    int64_t seen_cap = 0;
    int64_t chunk_segs = (int64_t)1 << 21, chunk_bytes = (int64_t)64 << 20;
    int64_t next_rid = 0;
    HostPool *pool = nullptr;
    WSlot slot[2];
    std::vector<int32_t> res_idx;
    std::vector<int64_t> res_cnt;
    int64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool staged = false;
};

namespace {

void free_wslot(WSlot &s) {
    host_free(s.h_bytes); host_free(s.h_off); host_free(s.h_hits); host_free(s.h_word);
    dev_free(s.d_bytes); dev_free(s.d_off); dev_free(s.d_hits); dev_free(s.d_word);
    if (s.ev0) cudaEventDestroy(s.ev0);
    if (s.ev1) cudaEventDestroy(s.ev1);
    if (s.stream) cudaStreamDestroy(s.stream);
    s = WSlot{};
}

int ensure_wslot(WSlot &s, int64_t segs, int64_t bytes) {
    if (!s.stream) {
        CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreate(&s.ev0));
        CUDA_TRY(cudaEventCreate(&s.ev1));
    }
    if (segs > s.cap_segs) {
        const int64_t c = std::max<int64_t>(segs, 1024);
        host_free(s.h_off); host_free(s.h_hits); host_free(s.h_word);
        dev_free(s.d_off); dev_free(s.d_hits); dev_free(s.d_word);
        s.cap_segs = 0;
        CUDA_TRY(cudaMallocHost(&s.h_off, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMallocHost(&s.h_hits, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMallocHost(&s.h_word, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_off, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_hits, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_word, (size_t)c * sizeof(int)));
        s.cap_segs = c;
    }
    if (bytes > s.cap_bytes) {
        const int64_t c = std::max<int64_t>(bytes, 4096);
        host_free(s.h_bytes);
        dev_free(s.d_bytes);
        s.cap_bytes = 0;
        CUDA_TRY(cudaMallocHost(&s.h_bytes, (size_t)c + 64));
        CUDA_TRY(cudaMalloc(&s.d_bytes, (size_t)c + 64));
        s.cap_bytes = c;
    }
    return SCT_OK;
}

int launch_count(sct_wl_handle *h, WSlot &s, bool per_segment) {
    WlCountParams p;
    p.bytes = s.d_bytes;
    p.off = s.d_off;
    p.b0 = s.b0;
    p.n = s.n;
    p.first_rid = s.first;
    p.n_hits = per_segment ? s.d_hits : nullptr;
    p.hit_word = per_segment ? s.d_word : nullptr;
    p.counts = h->d_counts;
    p.first_seen = h->d_first;
    if (s.n > 0) {
        const int64_t want = (s.n + WL_THREADS - 1) / WL_THREADS;
        const int grid = (int)std::min<int64_t>(want, (int64_t)h->sm_count * 8 * 4);   // 8 resident CTAs/SM, 4 waves
        wl_count_kernel<<<grid, WL_THREADS, 0, s.stream>>>(h->tab, p);
        CUDA_TRY(cudaGetLastError());
        h->stats[0] += 1;
    }
    return SCT_OK;
}

int reset_counts(sct_wl_handle *h) {
    if (h->n_words > 0) {
        CUDA_TRY(cudaMemset(h->d_counts, 0, (size_t)h->n_words * sizeof(unsigned int)));
        CUDA_TRY(cudaMemset(h->d_first, 0xFF, (size_t)h->n_words * sizeof(unsigned long long)));
        // the memsets run on the legacy stream, the count kernels on non-blocking streams that do not order behind it
        CUDA_TRY(cudaDeviceSynchronize());
    }
    h->next_rid = 0;
    return SCT_OK;
}

}  // namespace

extern "C" {

int sct_wl_destroy(sct_wl_handle *h) {
    if (!h) return SCT_OK;
    cudaSetDevice(h->device);
    for (auto &s : h->slot) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        free_wslot(s);
    }
    dev_free(h->tab.slots);
    dev_free(h->tab.filter);
    dev_free(h->d_odd);
    dev_free(h->d_odd_idx);
    dev_free(h->d_odd_off);
    dev_free(h->d_counts);
    dev_free(h->d_first);
    dev_free(h->d_seen);
    dev_free(h->d_nseen);
    delete h->pool;
    delete h;
    return SCT_OK;
}

namespace {
__global__ void wl_init_kernel(ulonglong2 *slots, unsigned long long n) {
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x)
        slots[i] = make_ulonglong2(WL_EMPTY, 0ull);
}
}  // namespace

int sct_wl_create(int device, const uint8_t *word_bytes, const int64_t *word_offsets, int64_t n_words, int32_t threads,
                  sct_wl_handle **out) {
    if (!out) return fail(SCT_E_ARG, "sct_wl_create: null out");
    *out = nullptr;
    if (n_words < 0 || n_words > 0x7fffffffLL || (n_words > 0 && (!word_bytes || !word_offsets)))
        return fail(SCT_E_ARG, "sct_wl_create: bad whitelist arguments");
    // L = the most frequent word length (packed keys); blank lines add no word; every other word is a literal word
    int L = 0;
    {
        std::map<int64_t, int64_t> hist;
        for (int64_t w = 0; w < n_words; ++w) {
            const int64_t len = word_offsets[w + 1] - word_offsets[w];
            if (len < 0) return fail(SCT_E_ARG, "sct_wl_create: offsets must be non-decreasing");
            if (len > 0) ++hist[len];
        }
        int64_t best = 0;
        for (auto &kv : hist)
            if (kv.second > best) best = kv.second, L = (int)std::min<int64_t>(kv.first, 1 << 20);
    }
    if (L > 31) return fail(SCT_E_UNSUPPORTED, "whitelist word length %d > 31 (one 64-bit 2-bit-packed key)", L);
    if (L == 0) L = 1;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1)
        return fail(SCT_E_CUDA, "no CUDA device: sctagger_b200 has no CPU fallback");
    if (device < 0 || device >= n_dev) return fail(SCT_E_ARG, "device %d out of range (0..%d)", device, n_dev - 1);
    CUDA_TRY(cudaSetDevice(device));
    sct_wl_handle *h = new (std::nothrow) sct_wl_handle;
    if (!h) return fail(SCT_E_NOMEM, "out of memory");
    h->device = device;
    h->n_words = n_words;
    h->host_threads = std::max(1, std::min<int>(threads, 16));
    h->pool = new HostPool(h->host_threads);
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    auto bail = [&](int rc) {
        std::string keep = g_err;
        sct_wl_destroy(h);
        g_err = keep;
        return rc;
    };
    // dense W x L copy (blank lines become a row of zero bytes, which no text can match) + the words with a byte outside
    // ACGT, which are compared literally: last add_word of a string wins (forward calls first, then the RC calls)
    std::vector<uint8_t> dense;
    std::map<std::string, std::pair<int64_t, int>> odd;
    bool odd_always = false;
    try {
        dense.assign((size_t)std::max<int64_t>(1, n_words) * L, 0);
        std::vector<uint8_t> is_odd((size_t)std::max<int64_t>(1, n_words), 0);
        par_ranges(h->pool, n_words, 1 << 14, [&](int64_t a, int64_t b) {
            for (int64_t w = a; w < b; ++w) {
                const int64_t len = word_offsets[w + 1] - word_offsets[w];
                if (len == 0) continue;
                const uint8_t *s = word_bytes + word_offsets[w];
                if (len != L) {
                    is_odd[(size_t)w] = 2;
                    continue;
                }
                memcpy(dense.data() + w * L, s, (size_t)L);
                for (int i = 0; i < L; ++i)
                    if (!is_acgt(s[i])) is_odd[(size_t)w] = 1;
                if (is_odd[(size_t)w]) memset(dense.data() + w * L, 0, (size_t)L);
            }
        });
        for (int pass = 0; pass < 2; ++pass)
            for (int64_t w = 0; w < n_words; ++w) {
                if (!is_odd[(size_t)w]) continue;
                std::string s((const char *)word_bytes + word_offsets[w], (size_t)(word_offsets[w + 1] - word_offsets[w]));
                bool pure = true;
                for (char c : s) pure = pure && is_acgt((uint8_t)c);
                odd_always = odd_always || pure;
                if (pass == 1) {
                    std::string r(s.rbegin(), s.rend());
                    for (auto &c : r) c = (char)comp_char((uint8_t)c);
                    s = r;
                }
                odd[s] = std::make_pair(pass * n_words + w, (int)w);      // later calls overwrite
            }
        if (odd.size() > 65536)
            return bail(fail(SCT_E_UNSUPPORTED, "%zu whitelist words are not ACGT-only words of the common length %d",
                             odd.size(), L));
    } catch (const std::bad_alloc &) {
        return bail(fail(SCT_E_NOMEM, "out of host memory building the whitelist table"));
    }
    int64_t slots = 1024;
    while (slots < 4 * n_words) slots <<= 1;
    h->n_slots = slots;
    h->tab.mask = (unsigned long long)slots - 1;
    int64_t fbits = 1 << 16;                              // up to 32 bits per key, at most 32 MiB (a quarter of L2)
    int cap_log2 = 28;                                    // 32 MiB: measured best on B200 (2^26..2^29: 0.51, 0.43, 0.40, 0.44 ms)
    if (const char *e = getenv("SCT_WL_FILTER_LOG2")) cap_log2 = std::max(16, std::min(33, atoi(e)));
    while (fbits < 64 * n_words && fbits < ((int64_t)1 << cap_log2)) fbits <<= 1;
    h->n_filter_bits = fbits;
    h->tab.fmask = (unsigned long long)fbits - 1;
    h->tab.L = L;
    h->tab.n_odd = (int)odd.size();
    uint8_t *d_words = nullptr;
    unsigned int *d_n = nullptr;
    auto build = [&]() -> int {
        CUDA_TRY(cudaMalloc(&h->tab.slots, (size_t)slots * sizeof(ulonglong2)));
        CUDA_TRY(cudaMalloc(&h->tab.filter, (size_t)(fbits / 8)));
        CUDA_TRY(cudaMemset(h->tab.filter, 0, (size_t)(fbits / 8)));
        CUDA_TRY(cudaMalloc(&h->d_counts, (size_t)std::max<int64_t>(1, n_words) * sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&h->d_first, (size_t)std::max<int64_t>(1, n_words) * sizeof(unsigned long long)));
        CUDA_TRY(cudaMalloc(&d_words, dense.size()));
        CUDA_TRY(cudaMalloc(&d_n, sizeof(unsigned int)));
        CUDA_TRY(cudaMemset(d_n, 0, sizeof(unsigned int)));
        CUDA_TRY(cudaMemcpy(d_words, dense.data(), dense.size(), cudaMemcpyHostToDevice));
        wl_init_kernel<<<h->sm_count * 8, 256>>>(h->tab.slots, (unsigned long long)slots);
        CUDA_TRY(cudaGetLastError());
        if (n_words > 0) {
            const long long calls = 2 * (long long)n_words;
            wl_insert_kernel<<<(unsigned)((calls + WL_THREADS - 1) / WL_THREADS), WL_THREADS>>>(h->tab, d_words, n_words, d_n);
            CUDA_TRY(cudaGetLastError());
        }
        unsigned int n_keys = 0;
        CUDA_TRY(cudaMemcpy(&n_keys, d_n, sizeof n_keys, cudaMemcpyDeviceToHost));
        h->n_keys = n_keys;
        if (!odd.empty()) {
            std::vector<uint8_t> ob;
            std::vector<int> oi, oo(1, 0);
            for (auto &kv : odd) {
                ob.insert(ob.end(), kv.first.begin(), kv.first.end());
                oi.push_back(kv.second.second);
                oo.push_back((int)ob.size());
            }
            CUDA_TRY(cudaMalloc(&h->d_odd, ob.size()));
            CUDA_TRY(cudaMalloc(&h->d_odd_idx, oi.size() * sizeof(int)));
            CUDA_TRY(cudaMalloc(&h->d_odd_off, oo.size() * sizeof(int)));
            CUDA_TRY(cudaMemcpy(h->d_odd, ob.data(), ob.size(), cudaMemcpyHostToDevice));
            CUDA_TRY(cudaMemcpy(h->d_odd_idx, oi.data(), oi.size() * sizeof(int), cudaMemcpyHostToDevice));
            CUDA_TRY(cudaMemcpy(h->d_odd_off, oo.data(), oo.size() * sizeof(int), cudaMemcpyHostToDevice));
            h->tab.odd = h->d_odd;
            h->tab.odd_idx = h->d_odd_idx;
            h->tab.odd_off = h->d_odd_off;
            h->tab.odd_always = odd_always ? 1 : 0;
        }
        TRY(reset_counts(h));
        CUDA_TRY(cudaDeviceSynchronize());
        return SCT_OK;
    };
    const int rc = build();
    if (d_words) cudaFree(d_words);
    if (d_n) cudaFree(d_n);
    if (rc != SCT_OK) return bail(rc);
    *out = h;
    return SCT_OK;
}

static int wl_finish_slot(sct_wl_handle *h, WSlot &s, int32_t *n_hits, int32_t *hit_word) {
    if (!s.busy) return SCT_OK;
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    s.busy = false;
    if (n_hits && s.n > 0 && s.dst_hits == s.h_hits) par_memcpy(h->pool, n_hits + s.first, s.h_hits, (size_t)s.n * sizeof(int));
    if (hit_word && s.n > 0 && s.dst_word == s.h_word) par_memcpy(h->pool, hit_word + s.first, s.h_word, (size_t)s.n * sizeof(int));
    return SCT_OK;
}

int sct_wl_count(sct_wl_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n, int32_t *n_hits,
                 int32_t *hit_word) {
    if (!h) return fail(SCT_E_ARG, "sct_wl_count: null handle");
    if (n < 0 || (n > 0 && (!seg_bytes || !seg_offsets))) return fail(SCT_E_ARG, "sct_wl_count: bad arguments");
    CUDA_TRY(cudaSetDevice(h->device));
    const bool per_seg = n_hits || hit_word;
    const int64_t all_bytes = n > 0 ? seg_offsets[n] - seg_offsets[0] : 0;
    const bool bytes_pinned = all_bytes > 0 && is_pinned(seg_bytes + seg_offsets[0], (size_t)all_bytes);
    const bool off_pinned = n > 0 && is_pinned(seg_offsets, (size_t)(n + 1) * sizeof(int64_t));
    const bool hits_pinned = n > 0 && n_hits && is_pinned(n_hits, (size_t)n * sizeof(int32_t));
    const bool word_pinned = n > 0 && hit_word && is_pinned(hit_word, (size_t)n * sizeof(int32_t));
    const int64_t rid0 = h->next_rid;
    h->staged = false;
    h->stats[2] = h->stats[3] = h->stats[1] = 0;
    int64_t a = 0;
    int k = 0;
    // per-segment outputs are indexed from the start of THIS call; first-seen order runs over the whole job
    while (a < n) {
        int64_t b = a;
        const int64_t base = seg_offsets[a];
        while (b < n && b - a < h->chunk_segs && seg_offsets[b + 1] - base <= h->chunk_bytes) ++b;
        if (b == a) {
            if (seg_offsets[a + 1] < seg_offsets[a]) return fail(SCT_E_ARG, "sct_wl_count: offsets must be non-decreasing");
            b = a + 1;                                   // a single segment larger than a chunk: the buffers grow
        }
        WSlot &s = h->slot[k & 1];
        TRY(wl_finish_slot(h, s, n_hits, hit_word));
        const int64_t cn = b - a, cb = seg_offsets[b] - base;
        if (cb < 0) return fail(SCT_E_ARG, "sct_wl_count: offsets must be non-decreasing");
        TRY(ensure_wslot(s, cn, cb));
        // page-locked caller arrays are copied from / to directly (as in sct_match); the offsets travel in the caller's
        // numbering (the kernel subtracts the chunk's first offset) and are only READ here, to validate them
        std::atomic<bool> bad(false);
        par_ranges(h->pool, cn, 1 << 16, [&](int64_t x, int64_t y) {
            bool mine = false;
            for (int64_t i = x; i < y; ++i) mine |= seg_offsets[a + i + 1] < seg_offsets[a + i];
            if (mine) bad = true;
        });
        if (bad) return fail(SCT_E_ARG, "sct_wl_count: offsets must be non-decreasing");
        const uint8_t *src_bytes = seg_bytes + base;
        const long long *src_off = reinterpret_cast<const long long *>(seg_offsets + a);
        if (!bytes_pinned && cb > 0) {
            par_memcpy(h->pool, s.h_bytes, src_bytes, (size_t)cb);
            src_bytes = s.h_bytes;
        }
        if (!off_pinned) {
            par_memcpy(h->pool, s.h_off, src_off, (size_t)(cn + 1) * sizeof(long long));
            src_off = s.h_off;
        }
        s.n = cn;
        s.first = a;
        s.b0 = base;
        s.dst_hits = hits_pinned ? n_hits + a : s.h_hits;
        s.dst_word = word_pinned ? hit_word + a : s.h_word;
        if (cb > 0) CUDA_TRY(cudaMemcpyAsync(s.d_bytes, src_bytes, (size_t)cb, cudaMemcpyHostToDevice, s.stream));
        CUDA_TRY(cudaMemcpyAsync(s.d_off, src_off, (size_t)(cn + 1) * sizeof(long long), cudaMemcpyHostToDevice, s.stream));
        // first_rid of the kernel = job-wide index; s.first stays call-relative for the un-staging
        const int64_t keep = s.first;
        s.first = rid0 + a;
        TRY(launch_count(h, s, per_seg));
        s.first = keep;
        if (per_seg) {
            CUDA_TRY(cudaMemcpyAsync(s.dst_hits, s.d_hits, (size_t)cn * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
            CUDA_TRY(cudaMemcpyAsync(s.dst_word, s.d_word, (size_t)cn * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
            h->stats[3] += cn * 8;
        }
        s.busy = true;
        h->stats[2] += cb + (cn + 1) * 8;
        h->stats[1] += 1;
        a = b;
        ++k;
    }
    TRY(wl_finish_slot(h, h->slot[0], n_hits, hit_word));
    TRY(wl_finish_slot(h, h->slot[1], n_hits, hit_word));
    h->next_rid = rid0 + n;
    h->stats[5] += n;
    return SCT_OK;
}

int sct_wl_result(sct_wl_handle *h, int64_t *n_unique, const int32_t **word_idx, const int64_t **count) {
    if (!h || !n_unique || !word_idx || !count) return fail(SCT_E_ARG, "sct_wl_result: null argument");
    CUDA_TRY(cudaSetDevice(h->device));
    try {
        std::vector<WlSeen> seen;
        if (h->n_words > 0) {
            if (!h->d_nseen) CUDA_TRY(cudaMalloc(&h->d_nseen, sizeof(unsigned long long)));
            unsigned long long n_seen = 0;
            for (int attempt = 0; attempt < 2; ++attempt) {
                if (h->seen_cap == 0) {
                    const int64_t cap = std::min<int64_t>(h->n_words, (int64_t)1 << 18);
                    CUDA_TRY(cudaMalloc(&h->d_seen, (size_t)cap * sizeof(WlSeen)));
                    h->seen_cap = cap;
                }
                CUDA_TRY(cudaMemset(h->d_nseen, 0, sizeof(unsigned long long)));
                const int grid = (int)std::min<int64_t>((h->n_words + 255) / 256, (int64_t)h->sm_count * 16);
                wl_compact_kernel<<<grid, 256>>>(h->d_counts, h->d_first, h->n_words, h->d_seen, (unsigned long long)h->seen_cap, h->d_nseen);
                CUDA_TRY(cudaGetLastError());
                CUDA_TRY(cudaMemcpy(&n_seen, h->d_nseen, sizeof n_seen, cudaMemcpyDeviceToHost));
                h->stats[0] += 1;
                if ((int64_t)n_seen <= h->seen_cap) break;
                dev_free(h->d_seen);                       // more counted words than the buffer holds: once more, at size
                h->seen_cap = 0;
                CUDA_TRY(cudaMalloc(&h->d_seen, (size_t)n_seen * sizeof(WlSeen)));
                h->seen_cap = (int64_t)n_seen;
            }
            seen.resize((size_t)n_seen);
            if (n_seen) CUDA_TRY(cudaMemcpy(seen.data(), h->d_seen, (size_t)n_seen * sizeof(WlSeen), cudaMemcpyDeviceToHost));
            h->stats[3] += (int64_t)n_seen * (int64_t)sizeof(WlSeen) + 8;
        }
        // Counter insertion order = first segment that counted the word (ties cannot occur: a segment counts one word)
        std::sort(seen.begin(), seen.end(), [](const WlSeen &x, const WlSeen &y) {
            return x.first != y.first ? x.first < y.first : x.word < y.word;
        });
        h->res_idx.resize(seen.size());
        h->res_cnt.resize(seen.size());
        for (size_t i = 0; i < seen.size(); ++i) {
            h->res_idx[i] = seen[i].word;
            h->res_cnt[i] = seen[i].count;
        }
    } catch (const std::bad_alloc &) {
        return fail(SCT_E_NOMEM, "out of host memory collecting the counts");
    }
    *n_unique = (int64_t)h->res_idx.size();
    *word_idx = h->res_idx.data();
    *count = h->res_cnt.data();
    return SCT_OK;
}

int sct_wl_reset(sct_wl_handle *h) {
    if (!h) return fail(SCT_E_ARG, "sct_wl_reset: null handle");
    CUDA_TRY(cudaSetDevice(h->device));
    return reset_counts(h);
}

int sct_wl_stage(sct_wl_handle *h, const uint8_t *seg_bytes, const int64_t *seg_offsets, int64_t n) {
    if (!h) return fail(SCT_E_ARG, "sct_wl_stage: null handle");
    if (n <= 0 || !seg_bytes || !seg_offsets) return fail(SCT_E_ARG, "sct_wl_stage: bad arguments");
    CUDA_TRY(cudaSetDevice(h->device));
    WSlot &s = h->slot[0];
    const int64_t base = seg_offsets[0], cb = seg_offsets[n] - base;
    TRY(ensure_wslot(s, n, cb));
    memcpy(s.h_bytes, seg_bytes + base, (size_t)cb);
    for (int64_t i = 0; i <= n; ++i) s.h_off[i] = seg_offsets[i] - base;
    s.n = n;
    s.first = 0;
    s.b0 = 0;                                   // the staged offsets are chunk-relative
    CUDA_TRY(cudaMemcpyAsync(s.d_bytes, s.h_bytes, (size_t)cb, cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.d_off, s.h_off, (size_t)(n + 1) * sizeof(long long), cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    h->staged = true;
    return SCT_OK;
}

int sct_wl_run_staged(sct_wl_handle *h, float *kernel_ms) {
    if (!h || !h->staged) return fail(SCT_E_STATE, "sct_wl_run_staged: nothing staged");
    CUDA_TRY(cudaSetDevice(h->device));
    WSlot &s = h->slot[0];
    TRY(reset_counts(h));
    CUDA_TRY(cudaEventRecord(s.ev0, s.stream));
    TRY(launch_count(h, s, false));
    CUDA_TRY(cudaEventRecord(s.ev1, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, s.ev0, s.ev1));
    if (kernel_ms) *kernel_ms = ms;
    return SCT_OK;
}

int sct_wl_stats(sct_wl_handle *h, int64_t stats[8]) {
    if (!h || !stats) return fail(SCT_E_ARG, "sct_wl_stats: null argument");
    memcpy(stats, h->stats, sizeof h->stats);
    stats[4] = h->n_keys;
    stats[6] = h->n_slots;
    stats[7] = h->tab.n_odd;
    return SCT_OK;
}

}  // extern "C"
