// Host side of the `extract_lr_bc` adapter alignment (C ABI: sct_align_* in include/sctagger_b200.h) around the kernels
// of align_kernels.cuh.  Replaces get_lr_bc_alns / get_alns (reference scTagger.py:307-314, 176-196): instead of a
// multiprocessing.Pool pickling every read to an edlib call, the reads are streamed through the GPU in byte-sized chunks
// (pinned, double-buffered, one CUDA stream per buffer).  No CPU fallback.
#include "align_kernels.cuh"
#include "host_common.h"

using namespace sct;
using namespace scth;

namespace {

struct ASlot {
    int64_t cap_reads = 0, cap_bytes = 0, cap_locs = 0, cap_hbytes = 0;
    uint8_t *h_bytes = nullptr;
    long long *h_off = nullptr, *h_locoff = nullptr, *h_total = nullptr;
    signed char *h_strand = nullptr;
    int *h_dist = nullptr, *h_locs = nullptr;
    uint8_t *d_bytes = nullptr;
    long long *d_off = nullptr, *d_locoff = nullptr, *d_locoff_out = nullptr, *d_bsums = nullptr, *d_total = nullptr;
    unsigned int *d_bad = nullptr, *h_bad = nullptr;
    unsigned int *d_hist = nullptr;
    int *d_order = nullptr, *d_dist = nullptr, *d_nlocs = nullptr, *d_locs = nullptr;
    int2 *d_span = nullptr;
    int *d_work = nullptr, *d_long = nullptr, *d_longcnt = nullptr, *d_short = nullptr;
    unsigned int *d_nlong = nullptr;
    signed char *d_strand = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr;
    int64_t n = 0, nbytes = 0, first = 0;
    long long b0 = 0;                   // offset of the chunk's first byte in the caller's numbering
    bool busy = false;
    // where the fixed-size results go: the caller's arrays when they are page-locked, else the slot's staging
    signed char *dst_strand = nullptr;
    int *dst_dist = nullptr;
    long long *dst_locoff = nullptr;
};

}  // namespace

struct sct_align_handle {
    int device = 0, m = 0, host_threads = 1, long_grid = 1184;
    uint2 *d_peq = nullptr;
    uint32_t *d_rpeq2 = nullptr;
    int64_t chunk_bytes = 0, chunk_reads = 0;
    HostPool *pool = nullptr;
    ASlot slot[2];
    int32_t *locs_pinned = nullptr;     // locations of the whole call (library-owned, page-locked; sct_align's *locs)
    int64_t locs_cap = 0;
    long long *d_run = nullptr;         // device-side running sum of locations over the chunks of one call
    bool in_bytes_pinned = false, in_off_pinned = false, out_strand_pinned = false, out_dist_pinned = false, out_locoff_pinned = false;
    int64_t stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool staged = false;
};

namespace {

void free_aslot(ASlot &s) {
    host_free(s.h_bytes); host_free(s.h_off); host_free(s.h_locoff); host_free(s.h_total); host_free(s.h_strand);
    host_free(s.h_dist); host_free(s.h_locs); host_free(s.h_bad);
    dev_free(s.d_bad); dev_free(s.d_locoff_out);
    dev_free(s.d_bytes); dev_free(s.d_off); dev_free(s.d_locoff); dev_free(s.d_bsums); dev_free(s.d_total);
    dev_free(s.d_hist); dev_free(s.d_order); dev_free(s.d_dist); dev_free(s.d_nlocs); dev_free(s.d_locs); dev_free(s.d_span); dev_free(s.d_work); dev_free(s.d_long); dev_free(s.d_nlong); dev_free(s.d_longcnt); dev_free(s.d_short);
    dev_free(s.d_strand);
    s.cap_reads = s.cap_bytes = s.cap_locs = s.cap_hbytes = 0;
}

int ensure_locs(ASlot &s, int64_t want) {
    if (want <= s.cap_locs) return SCT_OK;
    const int64_t c = std::max<int64_t>(want + want / 4, 4096);
    dev_free(s.d_locs);
    s.cap_locs = 0;
    CUDA_TRY(cudaMalloc(&s.d_locs, (size_t)c * sizeof(int)));
    s.cap_locs = c;
    return SCT_OK;
}

int ensure_aslot(ASlot &s, int64_t reads, int64_t bytes, bool stage_bytes) {
    if (!s.stream) {
        CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreate(&s.ev0));
        CUDA_TRY(cudaEventCreate(&s.ev1));
        CUDA_TRY(cudaEventCreate(&s.ev2));
        CUDA_TRY(cudaEventCreate(&s.ev3));
        CUDA_TRY(cudaMallocHost(&s.h_total, 2 * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_total, 2 * sizeof(long long)));
        CUDA_TRY(cudaMallocHost(&s.h_bad, sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&s.d_bad, sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&s.d_nlong, 2 * sizeof(unsigned int)));
        CUDA_TRY(cudaMalloc(&s.d_hist, ALN_BUCKETS * sizeof(unsigned int)));
    }
    if (reads > s.cap_reads) {
        const int64_t c = std::max<int64_t>(reads, 1024);
        host_free(s.h_off); host_free(s.h_locoff); host_free(s.h_strand); host_free(s.h_dist);
        dev_free(s.d_off); dev_free(s.d_locoff); dev_free(s.d_locoff_out); dev_free(s.d_bsums); dev_free(s.d_order); dev_free(s.d_dist);
        dev_free(s.d_nlocs); dev_free(s.d_strand); dev_free(s.d_span); dev_free(s.d_work); dev_free(s.d_long); dev_free(s.d_longcnt); dev_free(s.d_short);
        s.cap_reads = 0;
        CUDA_TRY(cudaMallocHost(&s.h_off, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMallocHost(&s.h_locoff, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMallocHost(&s.h_strand, (size_t)c));
        CUDA_TRY(cudaMallocHost(&s.h_dist, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_off, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_locoff, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_locoff_out, (size_t)(c + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_bsums, (size_t)((c + 1023) / 1024 + 1) * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&s.d_order, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_dist, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_nlocs, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_span, (size_t)c * sizeof(int2)));
        CUDA_TRY(cudaMalloc(&s.d_work, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_long, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_short, (size_t)c * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_longcnt, (size_t)c * 32 * sizeof(int)));
        CUDA_TRY(cudaMalloc(&s.d_strand, (size_t)c));
        s.cap_reads = c;
        TRY(ensure_locs(s, c * 2));
    }
    if (bytes > s.cap_bytes) {
        const int64_t c = std::max<int64_t>(bytes, 4096);
        dev_free(s.d_bytes);
        s.cap_bytes = 0;
        CUDA_TRY(cudaMalloc(&s.d_bytes, (size_t)c + 64));
        s.cap_bytes = c;
    }
    if (stage_bytes && bytes > s.cap_hbytes) {         // pinned staging only for pageable caller buffers
        const int64_t c = std::max<int64_t>(bytes, 4096);
        host_free(s.h_bytes);
        s.cap_hbytes = 0;
        CUDA_TRY(cudaMallocHost(&s.h_bytes, (size_t)c + 64));
        s.cap_hbytes = c;
    }
    return SCT_OK;
}

AlignParams make_aparams(sct_align_handle *h, ASlot &s) {
    AlignParams p;
    p.bytes = s.d_bytes;
    p.off = s.d_off;
    p.b0 = s.b0;
    p.nbytes = s.nbytes;
    p.n = (int)s.n;
    p.m = h->m;
    p.peq = h->d_peq;
    p.rpeq2 = h->d_rpeq2;
    p.order = s.d_order;
    p.strand = s.d_strand;
    p.dist = s.d_dist;
    p.n_locs = s.d_nlocs;
    p.span = s.d_span;
    p.work = s.d_work;
    p.long_list = s.d_long;
    p.n_long = s.d_nlong;
    p.long_cnt = s.d_longcnt;
    p.short_list = s.d_short;
    p.n_short = s.d_nlong + 1;
    p.loc_off = s.d_locoff;
    p.locs = s.d_locs;
    p.loc_cap = (unsigned long long)s.cap_locs;
    return p;
}

// length ordering, pass 1, scan; then the grand total and the fixed-size results start their way back
int enqueue_pass1(sct_align_handle *h, ASlot &s) {
    const int n = (int)s.n;
    CUDA_TRY(cudaEventRecord(s.ev0, s.stream));
    if (n > 0) {
        const AlignParams p = make_aparams(h, s);
        CUDA_TRY(cudaMemsetAsync(s.d_hist, 0, ALN_BUCKETS * sizeof(unsigned int), s.stream));
        CUDA_TRY(cudaMemsetAsync(s.d_bad, 0, sizeof(unsigned int), s.stream));
        aln_hist_kernel<<<std::min((n + 255) / 256, 1184), 256, 0, s.stream>>>(s.d_off, n, s.b0, s.nbytes, s.d_hist, s.d_bad);
        aln_bucket_scan_kernel<<<1, ALN_BUCKETS, 0, s.stream>>>(s.d_hist);
        aln_scatter_kernel<<<(n + 1023) / 1024, 1024, 0, s.stream>>>(s.d_off, n, s.d_hist, s.d_order);
        CUDA_TRY(cudaMemsetAsync(s.d_nlong, 0, 2 * sizeof(unsigned int), s.stream));
        // packed-key pass 1 where score * 2^(m-1) + position fits a word; it leaves the count of non-unique locations to
        // the count kernels (which only scan [first, last] of those reads)
        if (h->m >= 16 && h->m <= 27) aln_pass1_fast_kernel<<<(n + ALN_THREADS - 1) / ALN_THREADS, ALN_THREADS, 0, s.stream>>>(p);
        else aln_pass1_kernel<<<(n + ALN_THREADS - 1) / ALN_THREADS, ALN_THREADS, 0, s.stream>>>(p);
        aln_pass1_long_kernel<<<h->long_grid, ALN_THREADS, 0, s.stream>>>(p);   // reads longer than ALN_LONG_READ: a warp each
        aln_pass2_kernel<true><<<(n + ALN_THREADS - 1) / ALN_THREADS, ALN_THREADS, 0, s.stream>>>(p);
        aln_pass2_long_kernel<true><<<h->long_grid, ALN_THREADS, 0, s.stream>>>(p);
        const int nb = (n + 1023) / 1024;
        aln_block_sums_kernel<<<nb, 1024, 0, s.stream>>>(s.d_nlocs, n, s.d_bsums);
        aln_block_prefix_kernel<<<1, 1024, 0, s.stream>>>(s.d_bsums, nb, s.d_total, h->d_run);
        aln_scan_apply_kernel<<<nb, 1024, 0, s.stream>>>(s.d_nlocs, n, s.d_bsums, s.d_locoff, s.d_locoff_out, s.d_total);
        CUDA_TRY(cudaGetLastError());
        h->stats[0] += 10;
    } else {
        CUDA_TRY(cudaMemsetAsync(s.d_total, 0, 2 * sizeof(long long), s.stream));
        CUDA_TRY(cudaMemsetAsync(s.d_bad, 0, sizeof(unsigned int), s.stream));
    }
    CUDA_TRY(cudaEventRecord(s.ev1, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.h_total, s.d_total, 2 * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.h_bad, s.d_bad, sizeof(unsigned int), cudaMemcpyDeviceToHost, s.stream));
    return SCT_OK;
}

int enqueue_fixed_copies(ASlot &s) {
    const int64_t n = s.n;
    if (n <= 0) return SCT_OK;
    CUDA_TRY(cudaMemcpyAsync(s.dst_strand, s.d_strand, (size_t)n, cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.dst_dist, s.d_dist, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.dst_locoff, s.d_locoff_out, (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost, s.stream));
    return SCT_OK;
}

// after pass 1 has finished: size the location buffer, run pass 2
int enqueue_pass2(sct_align_handle *h, ASlot &s, int32_t *copy_back /* where the chunk's locations go on the host, or null */) {
    const long long total = s.n > 0 ? s.h_total[0] : 0;
    TRY(ensure_locs(s, total));
    CUDA_TRY(cudaEventRecord(s.ev2, s.stream));
    if (total > 0) {
        const AlignParams p = make_aparams(h, s);
        const int n = (int)s.n;
        // reads with one location (nearly all) write it straight from pass 1's span; short spans are re-scanned by their
        // thread, long ones (the list pass 1 built) by a warp each
        aln_pass2_long_kernel<false><<<h->long_grid, ALN_THREADS, 0, s.stream>>>(p);
        aln_pass2_kernel<false><<<(n + ALN_THREADS - 1) / ALN_THREADS, ALN_THREADS, 0, s.stream>>>(p);
        aln_emit_uniq_kernel<<<(n + ALN_THREADS - 1) / ALN_THREADS, ALN_THREADS, 0, s.stream>>>(p);
        CUDA_TRY(cudaGetLastError());
        h->stats[0] += 3;
        if (copy_back != nullptr)
            CUDA_TRY(cudaMemcpyAsync(copy_back, s.d_locs, (size_t)total * sizeof(int), cudaMemcpyDeviceToHost, s.stream));
    }
    CUDA_TRY(cudaEventRecord(s.ev3, s.stream));
    return SCT_OK;
}

// H2D of chunk [first, first + n).  Page-locked caller arrays are copied from directly; pageable ones go through the slot's
// pinned staging (host threads).  The offsets travel as they are (the kernels subtract b0 and validate them).
int stage_chunk(sct_align_handle *h, ASlot &s, const uint8_t *bytes, const int64_t *off, int64_t first, int64_t n) {
    const int64_t b0 = off[first], nbytes = off[first + n] - b0;
    if (nbytes < 0) return fail(SCT_E_ARG, "read_offsets must be non-decreasing (chunk starting at read %lld)", (long long)first);
    TRY(ensure_aslot(s, n, nbytes, !h->in_bytes_pinned));
    const uint8_t *src_bytes = bytes + b0;
    const long long *src_off = reinterpret_cast<const long long *>(off + first);
    if (!h->in_bytes_pinned && nbytes > 0) {
        par_memcpy(h->pool, s.h_bytes, src_bytes, (size_t)nbytes);
        src_bytes = s.h_bytes;
    }
    if (!h->in_off_pinned) {
        par_memcpy(h->pool, s.h_off, src_off, (size_t)(n + 1) * sizeof(long long));
        src_off = s.h_off;
    }
    s.n = n;
    s.nbytes = nbytes;
    s.first = first;
    s.b0 = b0;
    if (nbytes > 0) CUDA_TRY(cudaMemcpyAsync(s.d_bytes, src_bytes, (size_t)nbytes, cudaMemcpyHostToDevice, s.stream));
    CUDA_TRY(cudaMemcpyAsync(s.d_off, src_off, (size_t)(n + 1) * sizeof(long long), cudaMemcpyHostToDevice, s.stream));
    h->stats[2] += nbytes + (n + 1) * (int64_t)sizeof(long long);
    h->stats[1] += 1;
    return SCT_OK;
}

// The library-owned pinned buffer holding the locations of the whole call grows geometrically and keeps its content.
int ensure_call_locs(sct_align_handle *h, int64_t need, int64_t keep) {
    if (need <= h->locs_cap) return SCT_OK;
    const int64_t cap = std::max<int64_t>({need, 2 * h->locs_cap, (int64_t)1 << 16});
    int32_t *fresh = nullptr;
    CUDA_TRY(cudaMallocHost(&fresh, (size_t)cap * sizeof(int32_t)));
    if (keep > 0) par_memcpy(h->pool, fresh, h->locs_pinned, (size_t)keep * sizeof(int32_t));
    host_free(h->locs_pinned);
    h->locs_pinned = fresh;
    h->locs_cap = cap;
    return SCT_OK;
}

}  // namespace

extern "C" {

int sct_align_create(int device, const uint8_t *adapter, int32_t adapter_length, int32_t threads, double mem_gb,
                     sct_align_handle **out) {
    if (!out) return fail(SCT_E_ARG, "out is NULL");
    *out = nullptr;
    if (!adapter || adapter_length < 1) return fail(SCT_E_ARG, "adapter must not be empty");
    if (adapter_length > 32) return fail(SCT_E_UNSUPPORTED, "adapter length %d > 32 is not implemented", adapter_length);
    if (threads < 1) return fail(SCT_E_ARG, "threads must be >= 1 (scTagger.py:128)");
    if (!(mem_gb > 0)) return fail(SCT_E_ARG, "mem must be > 0");
    const int ndev = sct_device_count();
    if (ndev == 0) return fail(SCT_E_CUDA, "no CUDA device available (sctagger_b200 has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(SCT_E_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    CUDA_TRY(cudaSetDevice(device));
    sct_align_handle *h = new (std::nothrow) sct_align_handle();
    if (!h) return fail(SCT_E_NOMEM, "out of host memory");
    h->device = device;
    h->m = adapter_length;
    {
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, device));
        h->long_grid = prop.multiProcessorCount * 8;
    }
    h->host_threads = std::max(1, std::min({(int)threads, 8, (int)std::thread::hardware_concurrency() / std::max(1, ndev)}));
    h->pool = new (std::nothrow) HostPool(h->host_threads);
    // a1 = adapter, a2 = rev_compl(a1) (scTagger.py:363-365; identity on non-ACGT, scTagger.py:146-155)
    const int m = adapter_length;
    std::vector<uint2> peq(256, make_uint2(0u, 0u));
    std::vector<uint32_t> rpeq2(256, 0u);
    for (int i = 0; i < m; ++i) {
        const uint8_t c1 = adapter[i], c2 = comp_char(adapter[m - 1 - i]);
        peq[c1].x |= 1u << i;
        peq[c2].y |= 1u << i;
        rpeq2[c2] |= 1u << (m - 1 - i);
    }
    CUDA_TRY(cudaMalloc(&h->d_peq, 256 * sizeof(uint2)));
    CUDA_TRY(cudaMalloc(&h->d_rpeq2, 256 * sizeof(uint32_t)));
    CUDA_TRY(cudaMemcpy(h->d_peq, peq.data(), 256 * sizeof(uint2), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(h->d_rpeq2, rpeq2.data(), 256 * sizeof(uint32_t), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMalloc(&h->d_run, sizeof(long long)));
    CUDA_TRY(cudaMemset(h->d_run, 0, sizeof(long long)));
    CUDA_TRY(cudaDeviceSynchronize());   // the uploads above are on the legacy stream, the kernels on non-blocking streams
    const double budget = mem_gb * (double)(1ull << 30);
    // 128 MiB chunks: one thread walks one read, so a pass cannot be shorter than its longest read (~0.3 ms for 5 kb);
    // smaller chunks pay that floor too often, one single chunk would not overlap staging / copies with the kernels
    h->chunk_bytes = (int64_t)std::min<double>(std::max<double>(budget / 8.0, 16.0 * (1 << 20)), 128.0 * (1 << 20));
    h->chunk_reads = std::max<int64_t>(1024, h->chunk_bytes / 64);
    *out = h;
    return SCT_OK;
}

int sct_align_destroy(sct_align_handle *h) {
    if (!h) return SCT_OK;
    cudaSetDevice(h->device);
    for (ASlot &s : h->slot) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        free_aslot(s);
        if (s.ev0) cudaEventDestroy(s.ev0);
        if (s.ev1) cudaEventDestroy(s.ev1);
        if (s.ev2) cudaEventDestroy(s.ev2);
        if (s.ev3) cudaEventDestroy(s.ev3);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    dev_free(h->d_peq);
    dev_free(h->d_rpeq2);
    dev_free(h->d_run);
    host_free(h->locs_pinned);
    delete h->pool;
    delete h;
    return SCT_OK;
}

int sct_align_set_option(sct_align_handle *h, const char *key, int64_t value) {
    if (!h || !key) return fail(SCT_E_ARG, "NULL argument");
    const std::string k(key);
    if (k == "chunk_bytes") {
        if (value < 1) return fail(SCT_E_ARG, "chunk_bytes must be >= 1");
        h->chunk_bytes = value;
    } else if (k == "chunk_reads") {
        if (value < 1) return fail(SCT_E_ARG, "chunk_reads must be >= 1");
        h->chunk_reads = value;
    } else {
        return fail(SCT_E_ARG, "unknown option '%s'", key);
    }
    return SCT_OK;
}

int sct_align_stats(sct_align_handle *h, int64_t stats[8]) {
    if (!h || !stats) return fail(SCT_E_ARG, "NULL argument");
    memcpy(stats, h->stats, sizeof h->stats);
    return SCT_OK;
}

static int align_impl(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n, int8_t *strand,
                      int32_t *dist, int64_t *loc_offsets, const int32_t **locs, int64_t *n_locs);

int sct_align(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n, int8_t *strand,
              int32_t *dist, int64_t *loc_offsets, const int32_t **locs, int64_t *n_locs) {
    const int rc = align_impl(h, read_bytes, read_offsets, n, strand, dist, loc_offsets, locs, n_locs);
    if (rc != SCT_OK && h) {
        // a call that failed half way leaves chunks in flight: drain them so that the handle stays usable
        const std::string keep = g_err;
        for (ASlot &s : h->slot) {
            if (s.stream) cudaStreamSynchronize(s.stream);
            s.busy = false;
        }
        cudaGetLastError();
        g_err = keep;
    }
    return rc;
}

static int align_impl(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n, int8_t *strand,
                      int32_t *dist, int64_t *loc_offsets, const int32_t **locs, int64_t *n_locs) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (n < 0) return fail(SCT_E_ARG, "n < 0");
    if (!read_offsets || !loc_offsets || !locs || !n_locs || (n > 0 && (!strand || !dist))) return fail(SCT_E_ARG, "NULL argument");
    if (n > 0 && !read_bytes && read_offsets[n] > read_offsets[0]) return fail(SCT_E_ARG, "read_bytes is NULL");
    CUDA_TRY(cudaSetDevice(h->device));
    memset(h->stats, 0, sizeof h->stats);
    h->staged = false;
    // page-locked caller arrays (sct_host_alloc / sct_host_register / cudaHostAlloc) are copied from and to directly
    const int64_t all_bytes = n > 0 ? read_offsets[n] - read_offsets[0] : 0;
    h->in_bytes_pinned = all_bytes > 0 && is_pinned(read_bytes + read_offsets[0], (size_t)all_bytes);
    h->in_off_pinned = n > 0 && is_pinned(read_offsets, (size_t)(n + 1) * sizeof(int64_t));
    h->out_strand_pinned = n > 0 && is_pinned(strand, (size_t)n);
    h->out_dist_pinned = n > 0 && is_pinned(dist, (size_t)n * sizeof(int32_t));
    h->out_locoff_pinned = n > 0 && is_pinned(loc_offsets, (size_t)(n + 1) * sizeof(int64_t));
    NvtxRange nv_call("sct_align");
    int64_t loc_total = 0;

    auto finish = [&](int si) -> int {
        ASlot &s = h->slot[si];
        if (!s.busy) return SCT_OK;
        NvtxRange nv("sct_align: wait + pass 2 + D2H");
        CUDA_TRY(cudaStreamSynchronize(s.stream));            // pass 1, total and the fixed-size results are back
        const int64_t cn = s.n;
        if (cn > 0 && *s.h_bad)
            return fail(SCT_E_ARG, "read_offsets must be non-decreasing (chunk starting at read %lld: %u bad reads)",
                        (long long)s.first, *s.h_bad);
        const long long total = cn > 0 ? s.h_total[0] : 0, base = cn > 0 ? s.h_total[1] : 0;
        if (total > 0) TRY(ensure_call_locs(h, base + total, base));
        TRY(enqueue_pass2(h, s, total > 0 ? h->locs_pinned + base : nullptr));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, s.ev0, s.ev1);
        cudaEventElapsedTime(&b, s.ev2, s.ev3);
        h->stats[4] += (int64_t)((a + b) * 1e6);
        h->stats[3] += cn * 13 + 8 + total * 4;
        if (cn > 0) {
            if (s.dst_strand == s.h_strand) par_memcpy(h->pool, strand + s.first, s.h_strand, (size_t)cn);
            if (s.dst_dist == s.h_dist) par_memcpy(h->pool, dist + s.first, s.h_dist, (size_t)cn * sizeof(int));
            if (s.dst_locoff == s.h_locoff) par_memcpy(h->pool, loc_offsets + s.first, s.h_locoff, (size_t)cn * sizeof(long long));
        }
        loc_total += total;
        s.busy = false;
        return SCT_OK;
    };

    CUDA_TRY(cudaMemset(h->d_run, 0, sizeof(long long)));      // legacy stream: synchronous to the host, before any chunk
    int64_t first = 0, chunk = 0;
    while (first < n) {
        int64_t cn = std::min<int64_t>(h->chunk_reads, n - first);
        const int64_t limit = read_offsets[first] + h->chunk_bytes;
        if (read_offsets[first + cn] > limit) {
            const int64_t *e = std::upper_bound(read_offsets + first, read_offsets + first + cn + 1, limit);
            cn = std::max<int64_t>(1, (e - (read_offsets + first)) - 1);
        }
        const int si = (int)(chunk & 1);
        TRY(finish(si));
        ASlot &s = h->slot[si];
        NvtxRange nv("sct_align: stage + enqueue chunk");
        TRY(stage_chunk(h, s, read_bytes, read_offsets, first, cn));
        s.dst_strand = h->out_strand_pinned ? reinterpret_cast<signed char *>(strand) + first : s.h_strand;
        s.dst_dist = h->out_dist_pinned ? dist + first : s.h_dist;
        s.dst_locoff = h->out_locoff_pinned ? reinterpret_cast<long long *>(loc_offsets) + first : s.h_locoff;
        // kernels of this chunk behind pass 1 of the chunk in flight (two streams would otherwise run both side by side,
        // finish together and leave the GPU idle during the host's part; the running location base is advanced in chunk
        // order); the H2D above still overlaps
        if (h->slot[1 - si].busy) CUDA_TRY(cudaStreamWaitEvent(s.stream, h->slot[1 - si].ev1, 0));
        TRY(enqueue_pass1(h, s));
        TRY(enqueue_fixed_copies(s));
        s.busy = true;
        first += cn;
        ++chunk;
    }
    // drain in chunk order: the slot holding the older chunk first
    const int older = (int)(chunk & 1);
    TRY(finish(older));
    TRY(finish(1 - older));
    loc_offsets[n] = loc_total;
    *locs = loc_total > 0 ? h->locs_pinned : nullptr;
    *n_locs = loc_total;
    h->stats[5] = n;
    return SCT_OK;
}

int sct_align_stage(sct_align_handle *h, const uint8_t *read_bytes, const int64_t *read_offsets, int64_t n) {
    if (!h || !read_offsets || n < 0) return fail(SCT_E_ARG, "bad argument");
    CUDA_TRY(cudaSetDevice(h->device));
    ASlot &s = h->slot[0];
    memset(h->stats, 0, sizeof h->stats);
    {
        const int64_t all_bytes = n > 0 ? read_offsets[n] - read_offsets[0] : 0;
        h->in_bytes_pinned = all_bytes > 0 && is_pinned(read_bytes + read_offsets[0], (size_t)all_bytes);
        h->in_off_pinned = n > 0 && is_pinned(read_offsets, (size_t)(n + 1) * sizeof(int64_t));
    }
    TRY(stage_chunk(h, s, read_bytes, read_offsets, 0, n));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    h->staged = true;
    return SCT_OK;
}

int sct_align_run_staged(sct_align_handle *h, float *pass_ms, float *pass1_kernel_ms) {
    if (!h) return fail(SCT_E_ARG, "NULL handle");
    if (!h->staged) return fail(SCT_E_STATE, "sct_align_stage has not been called");
    CUDA_TRY(cudaSetDevice(h->device));
    ASlot &s = h->slot[0];
    h->stats[0] = 0;
    CUDA_TRY(cudaMemsetAsync(h->d_run, 0, sizeof(long long), s.stream));
    TRY(enqueue_pass1(h, s));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    if (*s.h_bad) return fail(SCT_E_ARG, "read_offsets must be non-decreasing (%u bad reads)", *s.h_bad);
    TRY(enqueue_pass2(h, s, nullptr));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    float a = 0.f, b = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&a, s.ev0, s.ev1));
    CUDA_TRY(cudaEventElapsedTime(&b, s.ev2, s.ev3));
    if (pass_ms) *pass_ms = a + b;
    if (pass1_kernel_ms) *pass1_kernel_ms = a;
    return SCT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// sct_align_strand_dist_noalloc: pass 1 of the adapter alignment (strand + distance per read) for callers that cannot
// call cudaMalloc — code under CUDA graph capture, tight serving loops, harnesses that forbid it.  No counterpart in the
// reference.  Everything pass 1 needs besides the caller's host arrays — the
// device copy of the reads and offsets, the Peq table, the read order, and the span / work / location-count / list
// arrays aln_pass1_kernel writes — lives in ONE file-scope static device array of a fixed, modest size (about 2.6 MB,
// held by every GPU that loads the library until exit).  The entry point calls no cudaMalloc / cudaFree / allocating
// routine: plain cudaMemcpy / cudaMemset on the legacy default stream and two kernel launches.
// Every call on a device shares the array and it keeps the last call's contents: the two list counters the kernel bumps
// are cleared by every call, everything else it reads is rewritten first.  Calls that overlap on one GPU (two host
// threads) race on it: serialise them.
// ---------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int64_t NA_MAX_READS = 8192;
constexpr int64_t NA_MAX_BYTES = 2ll << 20;
constexpr size_t na_up(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t NA_BYTES = 0;                                                                 // read bytes (+ padding)
constexpr size_t NA_OFF = NA_BYTES + na_up((size_t)NA_MAX_BYTES + 256);                        // int64[n + 1]
constexpr size_t NA_PEQ = NA_OFF + na_up((size_t)(NA_MAX_READS + 1) * 8);                      // uint2[256]
constexpr size_t NA_ORDER = NA_PEQ + na_up(256 * sizeof(uint2));                               // int[n]
constexpr size_t NA_DIST = NA_ORDER + na_up((size_t)NA_MAX_READS * 4);
constexpr size_t NA_NLOCS = NA_DIST + na_up((size_t)NA_MAX_READS * 4);
constexpr size_t NA_WORK = NA_NLOCS + na_up((size_t)NA_MAX_READS * 4);
constexpr size_t NA_LONG = NA_WORK + na_up((size_t)NA_MAX_READS * 4);
constexpr size_t NA_SHORT = NA_LONG + na_up((size_t)NA_MAX_READS * 4);
constexpr size_t NA_SPAN = NA_SHORT + na_up((size_t)NA_MAX_READS * 4);                         // int2[n]
constexpr size_t NA_STRAND = NA_SPAN + na_up((size_t)NA_MAX_READS * 8);                        // int8[n]
constexpr size_t NA_CTR = NA_STRAND + na_up((size_t)NA_MAX_READS);                             // n_long, n_short
constexpr size_t NA_TOTAL = NA_CTR + 256;

}  // namespace

// In host code this name is only a stand-in for the device array (not a device address): cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_TOTAL];

namespace {

__global__ void __launch_bounds__(256) na_iota_kernel(int *order, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) order[i] = i;
}

constexpr int NA_MAX_DEVICES = 64;
unsigned char *g_na_base[NA_MAX_DEVICES] = {};      // device address of g_noalloc_scratch, per device
std::mutex g_na_mutex;

// device address of the scratch array on `device` (which must be current): resolved once per device
int na_scratch(int device, unsigned char **out) {
    std::lock_guard<std::mutex> g(g_na_mutex);
    if (!g_na_base[device]) {
        void *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_na_base[device] = static_cast<unsigned char *>(p);
    }
    *out = g_na_base[device];
    return SCT_OK;
}

}  // namespace

extern "C" int sct_align_strand_dist_noalloc(int device, const uint8_t *adapter, int32_t adapter_length, const uint8_t *read_bytes,
                                             const int64_t *read_offsets, int64_t n, int8_t *strand, int32_t *dist) {
    if (!adapter || adapter_length < 1) return fail(SCT_E_ARG, "adapter must not be empty");
    if (adapter_length > 32) return fail(SCT_E_UNSUPPORTED, "adapter length %d > 32 is not implemented", adapter_length);
    if (n < 0 || !read_offsets || (n > 0 && (!strand || !dist))) return fail(SCT_E_ARG, "bad argument");
    if (n > NA_MAX_READS)
        return fail(SCT_E_UNSUPPORTED, "sct_align_strand_dist_noalloc: %lld reads exceed the fixed scratch (at most %lld reads per call)",
                    (long long)n, (long long)NA_MAX_READS);
    const int64_t b0 = read_offsets[0], nbytes = read_offsets[n] - b0;
    if (nbytes < 0) return fail(SCT_E_ARG, "read_offsets must be non-decreasing");
    if (nbytes > NA_MAX_BYTES)
        return fail(SCT_E_UNSUPPORTED, "sct_align_strand_dist_noalloc: %lld read bytes exceed the fixed scratch (at most %lld bytes per call)",
                    (long long)nbytes, (long long)NA_MAX_BYTES);
    if (nbytes > 0 && !read_bytes) return fail(SCT_E_ARG, "read_bytes is NULL");
    for (int64_t i = 0; i < n; ++i) {
        const int64_t len = read_offsets[i + 1] - read_offsets[i];
        if (len < 0) return fail(SCT_E_ARG, "read_offsets must be non-decreasing (read %lld)", (long long)i);
        if (len > ALN_LONG_READ)
            return fail(SCT_E_UNSUPPORTED, "sct_align_strand_dist_noalloc: read %lld has %lld bases (at most %d per read)", (long long)i,
                        (long long)len, ALN_LONG_READ);
    }
    const int ndev = sct_device_count();
    if (ndev == 0) return fail(SCT_E_CUDA, "no CUDA device available (sctagger_b200 has no CPU fallback)");
    if (device < 0 || device >= ndev || device >= NA_MAX_DEVICES) return fail(SCT_E_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    if (n == 0) return SCT_OK;
    CUDA_TRY(cudaSetDevice(device));
    unsigned char *base = nullptr;
    TRY(na_scratch(device, &base));

    // a1 = adapter, a2 = rev_compl(a1) (scTagger.py:363-365), as in sct_align_create
    const int m = adapter_length;
    uint2 peq[256];
    for (int i = 0; i < 256; ++i) peq[i] = make_uint2(0u, 0u);
    for (int i = 0; i < m; ++i) {
        peq[adapter[i]].x |= 1u << i;
        peq[comp_char(adapter[m - 1 - i])].y |= 1u << i;
    }
    CUDA_TRY(cudaMemcpy(base + NA_PEQ, peq, sizeof peq, cudaMemcpyHostToDevice));
    if (nbytes > 0) CUDA_TRY(cudaMemcpy(base + NA_BYTES, read_bytes + b0, (size_t)nbytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(base + NA_OFF, read_offsets, (size_t)(n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemset(base + NA_CTR, 0, 256));            // the list counters start at zero in every call

    AlignParams p;
    p.bytes = base + NA_BYTES;
    p.off = reinterpret_cast<const long long *>(base + NA_OFF);
    p.b0 = b0;
    p.nbytes = nbytes;
    p.n = (int)n;
    p.m = m;
    p.peq = reinterpret_cast<const uint2 *>(base + NA_PEQ);
    p.rpeq2 = nullptr;                                       // the start search belongs to pass 2
    p.order = reinterpret_cast<int *>(base + NA_ORDER);
    p.strand = reinterpret_cast<signed char *>(base + NA_STRAND);
    p.dist = reinterpret_cast<int *>(base + NA_DIST);
    p.n_locs = reinterpret_cast<int *>(base + NA_NLOCS);
    p.span = reinterpret_cast<int2 *>(base + NA_SPAN);
    p.work = reinterpret_cast<int *>(base + NA_WORK);
    p.long_list = reinterpret_cast<int *>(base + NA_LONG);
    p.n_long = reinterpret_cast<unsigned int *>(base + NA_CTR);
    p.short_list = reinterpret_cast<int *>(base + NA_SHORT);
    p.n_short = reinterpret_cast<unsigned int *>(base + NA_CTR) + 1;
    p.long_cnt = nullptr;
    p.loc_off = nullptr;
    p.locs = nullptr;
    p.loc_cap = 0;
    na_iota_kernel<<<(unsigned)((n + 255) / 256), 256>>>(reinterpret_cast<int *>(base + NA_ORDER), (int)n);
    aln_pass1_kernel<<<(unsigned)((n + ALN_THREADS - 1) / ALN_THREADS), ALN_THREADS>>>(p);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpy(strand, base + NA_STRAND, (size_t)n, cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(dist, base + NA_DIST, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    return SCT_OK;
}
