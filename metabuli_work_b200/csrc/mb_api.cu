// mb_api.cu — the C ABI of libmetabuli_b200.so (include/metabuli_b200.h): context, HBM-resident
// database / taxonomy, and the launch sequence of the classification hot path.
//
// Replaces the body of Classifier::startClassify's per-batch loop (src/commons/Classifier.cpp:81-150):
//   extractQueryKmers -> matchKmers -> sortMatches -> assignTaxonomy.
// Everything runs on one CUDA stream; the only host synchronisations per batch are four 8-byte
// read-backs (k-mer count, match count, longest per-read match list, taxCnt size) that size the next
// launch.  There is no CPU fallback anywhere in this file.
#include <chrono>
#include <cmath>

#include "mb_common.cuh"
#ifndef MB_RADIX_V2_DEFAULT
#define MB_RADIX_V2_DEFAULT true
#endif
#include "mb_extract.cuh"
#include "mb_radix.cuh"
#include "mb_match.cuh"
#include "mb_score.cuh"
#include "mb_segsort.cuh"
#include "mb_fastq.cuh"
#include "mb_mask.cuh"
#include "mb_build.cuh"

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace mb;

namespace {

thread_local std::string g_global_err;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; return e; }
        cap = want;
        return cudaSuccess;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
    void swap(DevBuf &o) { std::swap(p, o.p); std::swap(cap, o.cap); }
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
};

enum { EV_START, EV_H2D, EV_EXTRACT, EV_SORTK, EV_MATCH, EV_SORTM, EV_SCORE, EV_D2H, EV_COUNT };

struct Counters {                 // device-side scalars, one allocation
    unsigned long long match_count;
    unsigned long long tc_count;
    unsigned long long kmer_count;
    unsigned long long slow_matches;     // matches of the reads listed for the general scorer (compact scratch offsets)
    unsigned long long seg_huge_words;   // key scratch the segments beyond SS_CAP need (k_sort_segments_huge)
    u32 error_flag;
    u32 max_seg;
    u32 score_ticket;
    u32 radix_ticket;
    u32 match_ticket;
    u32 seg_ticket;
    u32 seg_too_long;
    u32 seg_long_count;
    u32 seg_long_ticket;
    u32 slow_count;
    u32 slow_ticket;
    u32 finish_ticket;
    u32 fast_overflow[8];     // reads the compact scorer left to the general kernel, by reason
    u32 lr_n_grp, lr_n_sp, lr_ticket[3];   // long-read scorer: list lengths and kernel tickets
    u32 trivial[RS_MAX_PASSES];
};

// kernels of this library, for the per-kernel timing table (mb_get_kernel_times)
enum { KID_TILE_MATE, KID_EXTRACT_COUNT, KID_EXTRACT_SCAN, KID_EXTRACT_WRITE, KID_KSORT_HIST, KID_KSORT_SCAN, KID_KSORT_PASS,
       KID_MATCH_PARTITION, KID_MATCH, KID_MSORT_HIST, KID_MSORT_SCAN, KID_MSORT_PASS, KID_SEGMENT, KID_SEGMENT_MAX, KID_SEGSORT, KID_SEGSORT_LONG, KID_SCORE_FAST, KID_SCORE_FINISH, KID_SCORE,
       KID_LR_HEADS, KID_LR_PATHS, KID_LR_SPECIES, KID_LR_BEST, KID_MASK, KID_SCORE_ORDER,
       KID_COUNT };
const char *const KID_NAMES[KID_COUNT] = {"k_tile_first_mate", "k_extract<count>", "k_scan_counts", "k_extract",
                                          "k_radix_hist<kmer>", "k_radix_scan_hist<kmer>", "k_radix_pass<kmer>",
                                          "k_match_partition", "k_match", "k_radix_hist<match>", "k_radix_scan_hist<match>",
                                          "k_radix_pass<match>", "k_segment_matches", "k_segment_max", "k_sort_segments", "k_sort_segments_long", "k_score_fast", "k_score_finish", "k_score",
                                          "k_lr_heads", "k_lr_paths", "k_lr_species", "k_lr_best", "k_mask_mates", "k_order_reads"};
struct KSample { int kid; cudaEvent_t a, b; };

} // namespace

struct mb_ctx {
    mb_params par;
    cudaStream_t stream = nullptr;
    std::string err;
    DevBuf tables;
    // database
    DevBuf words, info, units, raw_kidx, raw_vbefore;
    u64 n_words = 0, n_db_kmers = 0, n_raw = 0;
    u32 n_units = 0;
    SortT units_t{-1, 0};             // the T function the unit index was cut for (rebuilt when the sort key changes)
    // per-context (= per-device) launch facts: function attributes and occupancy are per device, so nothing of this is static
    int sms = 0;                      // multiprocessors of the context's device
    int match_ctas = 0;               // resident CTAs of k_match (0 = not asked yet)
    int seg_ctas = 0, fast_ctas = 0;
    bool match_attr_set = false, ex_attr_set = false, fast_attr_set = false;
    unsigned radix_attr_mask = 0;     // bit per radix instantiation whose shared-memory opt-in was done (radix_slot<>)
    // taxonomy
    DevBuf parent, depth, species_of, species_rank_of, tax_node;
    DevBuf fq_text[2], fq_tile_cnt, fq_tile_off, fq_lines[2], fq_src, fq_name_off, fq_name_len, fq_info, fq_sums, fq_sum_off;   // mb_stage_fastq
    u32 fq_n_reads = 0;
    DevBuf fq_pf_text[2];             // text of the NEXT chunk, copied on copy_stream (mb_prefetch_fastq)
    const void *fq_pf_host[2] = {nullptr, nullptr};
    size_t fq_pf_n[2] = {0, 0};
    bool fq_pf_pending = false;
    cudaEvent_t fq_pf_done = nullptr;
    DevBuf seg_long;                  // reads whose segment goes to k_sort_segments_long
    int32_t max_taxid = -1, node0 = 0, euk = 0;
    // batch
    DevBuf bases, offsets, lens, stage_info;
    // mb_prefetch_reads: the next batch, copied on a second stream while the current one is being classified
    DevBuf pf_bases, pf_offsets, pf_lens, pf_info;
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t pf_done = nullptr;
    cudaEvent_t fetch_ready = nullptr, fetch_done = nullptr;   // mb_fetch_results_async: scoring finished / results on the host
    bool fetch_pending = false;
    bool res_valid = false;            // the result arrays hold a scored batch (until the next mb_score_reads starts): what
    u32 res_n_reads = 0;               // mb_fetch_results_async copies, also after the NEXT batch has been staged
    u64 res_tc_used = 0;
    cudaStream_t side_stream = nullptr;          // the leftover scorer runs beside k_score_finish
    cudaEvent_t side_fork = nullptr, side_join = nullptr;
    bool pf_pending = false;
    const void *pf_host[3] = {nullptr, nullptr, nullptr};
    u64 pf_n_bases = 0;
    u32 pf_n_reads = 0;
    u64 n_bases = 0;
    u32 n_reads = 0;
    u32 seq_base = 0;
    SortT sort_t{0, 0};               // the function of the amino-acid part the resident k-mers are ordered by
    u64 kmer_bound = 0;
    u32 max_query_len = 0;
    DevBuf tile_mate, tile_counts, tile_offsets;
    DevBuf kmers_a, kmers_b;
    u64 n_kmers = 0;
    DevBuf hist, tile_state;
    DevBuf qstart, matches_a, matches_b;
    u64 n_matches = 0;
    DevBuf seg_begin, seg_end, results, tc_id, tc_cnt, arena;
    DevBuf slow_list, slow_off, best_recs;
    bool score_full_scratch = false;  // the compact scratch of the general scorer was too small once: size it by the batch
    DevBuf sc_local, sc_paths, sc_comb, sc_linked, sc_spid, sc_spscore;   // per-match scoring scratch
    DevBuf lr_grp_np, lr_grp_end, lr_sp_end, lr_grp_heads, lr_sp_heads;      // long-read scorer (k_lr_*)
    u64 tc_used = 0;
    DevBuf counters;
    void *pinned = nullptr;       // 256-byte pinned scratch for read-backs
    cudaEvent_t ev[EV_COUNT];
    bool ev_set[EV_COUNT];
    mb_phase_times times;
    u32 launches = 0;
    bool profile = false;             // mb_set_profile: bracket every kernel launch with an event pair
    std::vector<KSample> samples;     // event pairs of the current batch
    size_t n_samples = 0;
    bool reads_masked = false;         // --mask 1 has been applied to the staged bases
    bool fine_key = false, key_decided = false;    // seven-residue sort key: decided by the first large batch after mb_load_db
    DevBuf mask_probs, mask_scales;
    DevBuf read_order, order_hist;     // short-read scorers: reads by match count
    DevBuf seg_huge, seg_huge_keys, seg_huge_idx;      // segments beyond SS_CAP: list and (key, index) scratch
    bool have_reads = false, have_kmers = false, kmers_sorted = false, have_matches = false, matches_sorted = false, have_results = false;
    bool segments_valid = false;      // seg_begin/seg_end describe matches_a
    // multi-GPU routing (mb_peer_*): this context's receive side and the routing table of all ranks
    DevBuf peer_kmers, peer_matches, peer_cnt, route_dev, sent_cnt;
    u64 peer_kmer_cap = 0, peer_match_cap = 0;
    Route route;
    unsigned long long *peer_counters[MB_MAX_RANKS] = {};   // every rank's arrival counters (peer addresses)
    bool routed = false;
    SegIn kmers_seg = {}, matches_seg = {};    // n > 0: the resident k-mers / matches still lie in the receive regions

    Counters *dcnt() const { return counters.as<Counters>(); }
};

#define CK(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            char buf_[512];                                                                          \
            snprintf(buf_, sizeof(buf_), "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            ctx->err = buf_;                                                                         \
            return MB_E_CUDA;                                                                        \
        }                                                                                            \
    } while (0)

// one kernel launch: bracketed by profiling events, counted, checked
#define KL_ON(kid, strm, ...)                                                                        \
    do {                                                                                             \
        { KTimer kt_(ctx, kid, strm); __VA_ARGS__; }                                                 \
        ++ctx->launches;                                                                             \
        CK(cudaGetLastError());                                                                      \
    } while (0)
#define KL(kid, ...)                                                                                 \
    do {                                                                                             \
        { KTimer kt_(ctx, kid); __VA_ARGS__; }                                                       \
        ++ctx->launches;                                                                             \
        CK(cudaGetLastError());                                                                      \
    } while (0)

// Brackets one kernel launch with events on the launching stream when profiling is on.
struct KTimer {
    mb_ctx *c;
    KSample *s = nullptr;
    cudaStream_t st;
    KTimer(mb_ctx *ctx, int kid, cudaStream_t on = nullptr) : c(ctx), st(on ? on : ctx->stream) {
        if (!c->profile || kid < 0) return;
        if (c->n_samples == c->samples.size()) {
            KSample n;
            n.kid = kid;
            if (cudaEventCreate(&n.a) != cudaSuccess || cudaEventCreate(&n.b) != cudaSuccess) return;
            c->samples.push_back(n);
        }
        s = &c->samples[c->n_samples++];
        s->kid = kid;
        cudaEventRecord(s->a, st);
    }
    ~KTimer() { if (s) cudaEventRecord(s->b, st); }
};

static int fail(mb_ctx *ctx, int code, const char *msg) { ctx->err = msg; return code; }

static inline int bits_for(u64 v) { int b = 1; while (b < 63 && (v >> b)) ++b; return b; }

static void mark(mb_ctx *ctx, int ev) {
    cudaEventRecord(ctx->ev[ev], ctx->stream);
    ctx->ev_set[ev] = true;
}

__global__ void k_read_back(u32 *__restrict__ host_dst, const u32 *__restrict__ src, u32 n_words) {
    if (threadIdx.x < n_words) host_dst[threadIdx.x] = src[threadIdx.x];
    __threadfence_system();
}

template <class T>
static int read_back(mb_ctx *ctx, const T *dptr, T *out) {
    // A few bytes from the device, synchronously.  Not a cudaMemcpy: a device->host memcpy queues on the DMA engine behind a
    // result transfer that mb_fetch_results_async has in flight (0.9 ms at configs[1], measured in mb_stage_reads); a
    // one-warp kernel writes the words straight into the page-locked host buffer instead (unified addressing).
    if constexpr (sizeof(T) % 4 == 0 && sizeof(T) <= 256) {
        k_read_back<<<1, 64, 0, ctx->stream>>>(reinterpret_cast<u32 *>(ctx->pinned), reinterpret_cast<const u32 *>(dptr), (u32) (sizeof(T) / 4));
        CK(cudaGetLastError());
    } else {
        CK(cudaMemcpyAsync(ctx->pinned, dptr, sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    memcpy(out, ctx->pinned, sizeof(T));
    return MB_OK;
}

// ---- radix sort driver --------------------------------------------------------------------------------
// one bit of mb_ctx::radix_attr_mask per instantiation of the pass kernel (numbered in order of first use, process wide)
static std::atomic<int> g_radix_slots{0};
template <typename Rec, typename Key, int ITEMS>
static int radix_slot() { static const int slot = g_radix_slots++; return slot; }

template <typename Rec, typename Key, int ITEMS>
static int radix_sort(mb_ctx *ctx, DevBuf &a, DevBuf &b, u64 n, const Key &key, int n_passes, int kid_base, SegIn *seg_in = nullptr) {
    // seg_in (n > 0): the records lie in the regions of a peer receive buffer, outside the ping-pong pair; the histogram
    // and the first pass read them there, the result ends up in `a` like always and *seg_in is cleared
    SegIn seg = {};
    if (seg_in && seg_in->n) { seg = *seg_in; seg_in->n = 0; }
    if (n == 0) return MB_OK;
    if (n_passes == 0) {
        if (seg.n) {
            CK(a.ensure(n * sizeof(Rec)));
            for (u32 r = 0; r < seg.n; ++r)
                if (seg.prefix[r + 1] > seg.prefix[r])
                    CK(cudaMemcpyAsync(a.as<Rec>() + seg.prefix[r], seg.base[r], (seg.prefix[r + 1] - seg.prefix[r]) * sizeof(Rec),
                                       cudaMemcpyDeviceToDevice, ctx->stream));
        }
        return MB_OK;
    }
    if (n_passes > RS_MAX_PASSES) return fail(ctx, MB_E_ARG, "too many radix passes");
    if (seg.n) CK(a.ensure(n * sizeof(Rec)));
    constexpr u64 TILE = (u64) RS_THREADS * ITEMS;
    const u64 tiles = (n + TILE - 1) / TILE;
    // the pass with the tile kept in shared memory (k_radix_pass2: half the items per thread)
    constexpr bool HAS_V2 = ITEMS % 2 == 0 && ITEMS >= 8;
    constexpr int ITEMS2 = HAS_V2 ? ITEMS / 2 : ITEMS;
    constexpr u64 TILE2 = (u64) R2_THREADS * ITEMS2;
    const u64 tiles2 = (n + TILE2 - 1) / TILE2;
    CK(b.ensure(n * sizeof(Rec)));
    CK(ctx->hist.ensure(sizeof(u64) * RS_BINS * RS_MAX_PASSES));
    CK(ctx->tile_state.ensure(std::max(tiles, tiles2) * RS_BINS * sizeof(u64)));
    CK(cudaMemsetAsync(ctx->hist.p, 0, sizeof(u64) * RS_BINS * n_passes, ctx->stream));
    CK(cudaMemsetAsync(&ctx->dcnt()->trivial, 0, sizeof(u32) * RS_MAX_PASSES, ctx->stream));
    const u64 hist_ctas = std::min<u64>((n + RS_THREADS - 1) / RS_THREADS, (u64) ctx->sms * 8);
    if (seg.n)
        KL(kid_base + 0, k_radix_hist<Rec, Key, true><<<(unsigned) hist_ctas, RS_THREADS, n_passes * RS_BINS * sizeof(u32), ctx->stream>>>(
            a.as<Rec>(), n, key, n_passes, ctx->hist.as<u64>(), seg));
    else
        KL(kid_base + 0, k_radix_hist<Rec, Key, false><<<(unsigned) hist_ctas, RS_THREADS, n_passes * RS_BINS * sizeof(u32), ctx->stream>>>(
            a.as<Rec>(), n, key, n_passes, ctx->hist.as<u64>(), seg));
    KL(kid_base + 1, k_radix_scan_hist<<<n_passes, RS_BINS, 0, ctx->stream>>>(ctx->hist.as<u64>(), n, ctx->dcnt()->trivial));
    constexpr size_t smem = radix_pass_smem<Rec, ITEMS>();
    static const int radix_dbg = getenv("MB_RADIX_DBG") ? atoi(getenv("MB_RADIX_DBG")) : 0;     // timing experiments only
    const unsigned slot_bit = 1u << (radix_slot<Rec, Key, ITEMS>() & 31);
    static const bool radix_v2 = getenv("MB_RADIX_V2") ? atoi(getenv("MB_RADIX_V2")) != 0 : MB_RADIX_V2_DEFAULT;
    if (!(ctx->radix_attr_mask & slot_bit)) {
        CK(cudaFuncSetAttribute(k_radix_pass<Rec, Key, ITEMS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        CK(cudaFuncSetAttribute(k_radix_pass<Rec, Key, ITEMS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        if constexpr (HAS_V2) {
            CK(cudaFuncSetAttribute(k_radix_pass2<Rec, Key, ITEMS2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) radix_pass2_smem<Rec, ITEMS2>()));
            CK(cudaFuncSetAttribute(k_radix_pass2<Rec, Key, ITEMS2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) radix_pass2_smem<Rec, ITEMS2>()));
        }
        ctx->radix_attr_mask |= slot_bit;
    }
    for (int p = 0; p < n_passes; ++p) {
        const bool v2 = HAS_V2 && radix_v2 && !radix_dbg;
        CK(cudaMemsetAsync(ctx->tile_state.p, 0, (v2 ? tiles2 : tiles) * RS_BINS * sizeof(u64), ctx->stream));
        CK(cudaMemsetAsync(&ctx->dcnt()->radix_ticket, 0, sizeof(u32), ctx->stream));
        if constexpr (HAS_V2) {
            if (v2) {
                if (seg.n)
                    KL(kid_base + 2, k_radix_pass2<Rec, Key, ITEMS2, true><<<(unsigned) tiles2, R2_THREADS, radix_pass2_smem<Rec, ITEMS2>(), ctx->stream>>>(
                        a.as<Rec>(), b.as<Rec>(), n, key, p, ctx->hist.as<u64>() + (size_t) p * RS_BINS, ctx->tile_state.as<u64>(),
                        &ctx->dcnt()->radix_ticket, seg));
                else
                    KL(kid_base + 2, k_radix_pass2<Rec, Key, ITEMS2, false><<<(unsigned) tiles2, R2_THREADS, radix_pass2_smem<Rec, ITEMS2>(), ctx->stream>>>(
                        a.as<Rec>(), b.as<Rec>(), n, key, p, ctx->hist.as<u64>() + (size_t) p * RS_BINS, ctx->tile_state.as<u64>(),
                        &ctx->dcnt()->radix_ticket, seg));
                seg.n = 0;
                std::swap(a.p, b.p);
                std::swap(a.cap, b.cap);
                continue;
            }
        }
        if (seg.n)
            KL(kid_base + 2, k_radix_pass<Rec, Key, ITEMS, true><<<(unsigned) tiles, RS_THREADS, smem, ctx->stream>>>(
                a.as<Rec>(), b.as<Rec>(), n, key, p, ctx->hist.as<u64>() + (size_t) p * RS_BINS, ctx->tile_state.as<u64>(),
                &ctx->dcnt()->radix_ticket, radix_dbg, seg));
        else
            KL(kid_base + 2, k_radix_pass<Rec, Key, ITEMS, false><<<(unsigned) tiles, RS_THREADS, smem, ctx->stream>>>(
                a.as<Rec>(), b.as<Rec>(), n, key, p, ctx->hist.as<u64>() + (size_t) p * RS_BINS, ctx->tile_state.as<u64>(),
                &ctx->dcnt()->radix_ticket, radix_dbg, seg));
        seg.n = 0;
        std::swap(a.p, b.p);
        std::swap(a.cap, b.cap);
    }
    return MB_OK;
}

// stable multi-way partition = one radix pass whose digit is the destination rank
template <typename Rec, typename Key, int ITEMS>
static int partition_by(mb_ctx *ctx, DevBuf &a, DevBuf &b, u64 n, const Key &key, int n_parts, uint64_t *counts, SegIn *src) {
    for (int i = 0; i < n_parts; ++i) counts[i] = 0;
    if (n == 0) return MB_OK;
    if (int rc = radix_sort<Rec, Key, ITEMS>(ctx, a, b, n, key, 1, -100, src)) return rc;
    // digit_base (exclusive scan of the histogram) is still in ctx->hist
    std::vector<u64> base(RS_BINS);
    CK(cudaMemcpyAsync(base.data(), ctx->hist.p, RS_BINS * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < n_parts; ++i) counts[i] = (i + 1 < RS_BINS ? base[i + 1] : n) - base[i];
    counts[n_parts - 1] = n - base[n_parts - 1];
    return MB_OK;
}

// records that still lie in the regions of a receive buffer are gathered into the ordinary resident buffer
template <typename Rec>
static int materialize(mb_ctx *ctx, SegIn &seg, DevBuf &a, u64 n) {
    if (!seg.n) return MB_OK;
    CK(a.ensure((n + 1) * sizeof(Rec)));
    for (u32 r = 0; r < seg.n; ++r)
        if (seg.prefix[r + 1] > seg.prefix[r])
            CK(cudaMemcpyAsync(a.as<Rec>() + seg.prefix[r], seg.base[r], (seg.prefix[r + 1] - seg.prefix[r]) * sizeof(Rec),
                               cudaMemcpyDeviceToDevice, ctx->stream));
    seg.n = 0;
    return MB_OK;
}
#define MATERIALIZE_K() do { if (int rc_ = materialize<mb_kmer>(ctx, ctx->kmers_seg, ctx->kmers_a, ctx->n_kmers)) return rc_; } while (0)
#define MATERIALIZE_M() do { if (int rc_ = materialize<mb_match>(ctx, ctx->matches_seg, ctx->matches_a, ctx->n_matches)) return rc_; } while (0)

extern "C" {

static int publish_sent(mb_ctx *ctx, int which);

void mb_default_params(mb_params *p) {
    memset(p, 0, sizeof(*p));
    p->device = 0; p->kmer_format = 1; p->syncmer = 0; p->smer_len = 5; p->seq_mode = 2;
    p->min_cons_cnt = 4; p->min_cons_cnt_euk = 9; p->min_score = 0.f; p->min_sp_score = 0.f; p->tie_ratio = 0.95f;
    p->skip_redundancy = 0; p->accession_level = 0; p->last_shard = 1;
    p->mask_mode = 0; p->mask_prob = 0.9f;
}

const char *mb_global_error(void) { return g_global_err.c_str(); }
const char *mb_last_error(const mb_ctx *ctx) { return ctx ? ctx->err.c_str() : g_global_err.c_str(); }

static int check_params(const mb_params *p, std::string &err) {
    if (p->kmer_format != 1 && p->kmer_format != 2) { err = "kmer_format must be 1 or 2"; return MB_E_ARG; }
    if (p->seq_mode < 1 || p->seq_mode > 3) { err = "seq_mode must be 1, 2 or 3"; return MB_E_ARG; }
    if (p->syncmer && (p->smer_len < 1 || p->smer_len > 7)) { err = "smer_len must be in 1..7"; return MB_E_ARG; }
    if (p->kmer_sort_bits != 0 && p->kmer_sort_bits != 21 && p->kmer_sort_bits != 31 && p->kmer_sort_bits != 24 && p->kmer_sort_bits != 32 &&
        p->kmer_sort_bits != 40) {
        err = "kmer_sort_bits must be 0, 21, 31, 24, 32 or 40"; return MB_E_ARG;
    }
    if ((p->kmer_sort_bits == 21 || p->kmer_sort_bits == 31) && p->kmer_format != 2) { err = "kmer_sort_bits 21 / 31 (re-coded keys) need kmer_format 2"; return MB_E_ARG; }
    if (p->accession_level != 0) { err = "accession_level != 0 is not on this path"; return MB_E_ARG; }
    if (p->mask_mode != 0 && p->mask_mode != 1) { err = "mask_mode must be 0 or 1"; return MB_E_ARG; }
    if (p->mask_mode && !(p->mask_prob > 0.f && p->mask_prob <= 1.f)) { err = "mask_prob must be in (0, 1]"; return MB_E_ARG; }
    return MB_OK;
}

mb_ctx *mb_create(const mb_params *p) {
    if (!p) { g_global_err = "null params"; return nullptr; }
    if (check_params(p, g_global_err) != MB_OK) return nullptr;
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        g_global_err = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libmetabuli_b200 has no CPU fallback)";
        return nullptr;
    }
    if (p->device < 0 || p->device >= n_dev) { g_global_err = "device ordinal out of range"; return nullptr; }
    if ((e = cudaSetDevice(p->device)) != cudaSuccess) { g_global_err = cudaGetErrorString(e); return nullptr; }
    mb_ctx *ctx = new mb_ctx();
    ctx->par = *p;
    memset(&ctx->times, 0, sizeof(ctx->times));
    for (int i = 0; i < EV_COUNT; ++i) ctx->ev_set[i] = false;
    bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaDeviceGetAttribute(&ctx->sms, cudaDevAttrMultiProcessorCount, p->device) == cudaSuccess && ctx->sms > 0;
    for (int i = 0; ok && i < EV_COUNT; ++i) ok = cudaEventCreate(&ctx->ev[i]) == cudaSuccess;
    ok = ok && cudaHostAlloc(&ctx->pinned, 256, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess;
    ok = ok && ctx->tables.ensure(sizeof(Tables)) == cudaSuccess && ctx->counters.ensure(sizeof(Counters)) == cudaSuccess;
    if (ok) {
        Tables t;
        build_tables(t);
        ok = cudaMemcpy(ctx->tables.p, &t, sizeof(t), cudaMemcpyHostToDevice) == cudaSuccess &&
             cudaMemset(ctx->counters.p, 0, sizeof(Counters)) == cudaSuccess;
    }
    if (!ok) {
        g_global_err = std::string("context setup failed: ") + cudaGetErrorString(cudaGetLastError());
        delete ctx;
        return nullptr;
    }
    return ctx;
}

void mb_destroy(mb_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->par.device);
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < EV_COUNT; ++i) cudaEventDestroy(ctx->ev[i]);
    for (auto &k : ctx->samples) { cudaEventDestroy(k.a); cudaEventDestroy(k.b); }
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->pf_done) cudaEventDestroy(ctx->pf_done);
    if (ctx->fetch_done) { cudaEventSynchronize(ctx->fetch_done); cudaEventDestroy(ctx->fetch_done); }
    if (ctx->fetch_ready) cudaEventDestroy(ctx->fetch_ready);
    if (ctx->fq_pf_done) cudaEventDestroy(ctx->fq_pf_done);
    if (ctx->side_stream) { cudaStreamSynchronize(ctx->side_stream); cudaStreamDestroy(ctx->side_stream); }
    if (ctx->side_fork) cudaEventDestroy(ctx->side_fork);
    if (ctx->side_join) cudaEventDestroy(ctx->side_join);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int mb_set_params(mb_ctx *ctx, const mb_params *p) {
    if (!ctx || !p) return MB_E_ARG;
    if (int rc = check_params(p, ctx->err)) return rc;
    if (p->device != ctx->par.device) return fail(ctx, MB_E_ARG, "device cannot change");
    const bool reshard = p->last_shard != ctx->par.last_shard;
    if (reshard && ctx->n_words) return fail(ctx, MB_E_ARG, "last_shard cannot change after mb_load_db");
    ctx->par = *p;
    ctx->have_kmers = ctx->have_matches = ctx->have_results = false;
    return MB_OK;
}

void *mb_stream(mb_ctx *ctx) { return ctx ? (void *) ctx->stream : nullptr; }

// ---- database ---------------------------------------------------------------------------------------
static DbView db_view(const mb_ctx *ctx, bool for_index) {
    DbView v;
    v.words = ctx->words.as<u16>();
    v.n_words = ctx->n_words;
    v.info = ctx->info.as<u32>();
    v.n_kmers_eff = for_index ? ~0ull : (ctx->par.last_shard ? (ctx->n_db_kmers ? ctx->n_db_kmers - 1 : 0) : ctx->n_db_kmers);
    v.info_mask = ~(((u32) (ctx->par.skip_redundancy == 0)) << 31);
    v.species_of = ctx->species_of.as<int32_t>();
    v.max_taxid = ctx->max_taxid;
    return v;
}

// The T function the query sort will order by (mb_sort_kmers) for the context's current parameters.
static SortT planned_sort_t(const mb_ctx *ctx) {
    static const int env_bits = getenv("MB_KSORT_BITS") ? atoi(getenv("MB_KSORT_BITS")) : 0;      // tuning knob
    int bits = env_bits ? env_bits : ctx->par.kmer_sort_bits;
    if ((bits == 21 || bits == 31) && ctx->par.kmer_format != 2) bits = 0;
    // default: format 2 takes a re-coded key — five residues (three passes), or seven (four passes) when the database
    // stream is long against the batch (fine_key, below): the better the queries are ordered, the faster the join runs
    // (measured: k_match 33.0 -> 25.1 ms at configs[2] with 1.32 G DB k-mers for 3.4 ms of extra pass, step 94.9 -> 90.3 ms;
    // 8.85 -> 7.85 ms at configs[1], where the extra pass costs 2.8 ms and the five-residue key stays)
    if (bits == 0) bits = ctx->par.kmer_format == 2 ? (ctx->fine_key ? 31 : 21) : 32;
    SortT t;
    t.mode = bits == 21 ? 1 : bits == 31 ? 2 : 0;
    t.shift = t.mode ? 0 : 40 - bits;
    return t;
}

// Cut the resident stream into units at the T-bucket starts of sort key `st` (mb_match.cuh): pass B over the raw blocks,
// compaction of the owning blocks on the host (a load-time step over a few MB).
static int build_units(mb_ctx *ctx, const SortT &st) {
    const u64 n_raw = ctx->n_raw;
    ctx->n_units = 0;
    if (n_raw == 0) return MB_OK;
    DevBuf raw_units;
    CK(raw_units.ensure(n_raw * sizeof(Unit)));
    const DbView v = db_view(ctx, true);
    KL(-1, k_db_raw_own<<<(unsigned) ((n_raw + 7) / 8), 256, 0, ctx->stream>>>(v, n_raw, ctx->raw_kidx.as<u64>(), ctx->raw_vbefore.as<u64>(), st,
                                                                          raw_units.as<Unit>()));
    std::vector<Unit> ru(n_raw), units;
    CK(cudaMemcpyAsync(ru.data(), raw_units.p, n_raw * sizeof(Unit), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    units.reserve(n_raw + 1);
    for (u64 b = 0; b < n_raw; ++b)
        if (ru[b].word != ~0ull) units.push_back(ru[b]);
    if (units.empty() || units[0].word != 0 || units[0].kidx != 0) return fail(ctx, MB_E_DATA, "internal: the unit index does not start at the stream start");
    if (units.size() > 0xFFFFFFF0ull) return fail(ctx, MB_E_ARG, "database shard too large for one context (unit index)");
    Unit end;
    end.word = ctx->n_words; end.vbefore = 0; end.kidx = ctx->n_db_kmers; end.t_lo = ~0ull;
    units.push_back(end);
    CK(ctx->units.ensure(units.size() * sizeof(Unit)));
    CK(cudaMemcpyAsync(ctx->units.p, units.data(), units.size() * sizeof(Unit), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_units = (u32) (units.size() - 1);
    ctx->units_t = st;
    return MB_OK;
}

int mb_load_db(mb_ctx *ctx, const uint16_t *diffIdx, size_t n_words, const uint32_t *info, size_t n_kmers) {
    if (!ctx || (!diffIdx && n_words) || (!info && n_kmers)) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    ctx->n_words = 0;
    ctx->n_db_kmers = 0;
    ctx->n_units = 0;
    ctx->n_raw = 0;
    CK(ctx->words.ensure((n_words + DB_PAD) * sizeof(u16)));
    CK(ctx->info.ensure((n_kmers + 16) * sizeof(u32)));
    CK(cudaMemsetAsync(ctx->words.as<u16>() + n_words, 0, DB_PAD * sizeof(u16), ctx->stream));
    CK(cudaMemcpyAsync(ctx->words.p, diffIdx, n_words * sizeof(u16), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->info.p, info, n_kmers * sizeof(u32), cudaMemcpyHostToDevice, ctx->stream));
    ctx->n_words = n_words;
    ctx->n_db_kmers = n_kmers;
    const u64 n_raw = (n_words + UN_W - 1) / UN_W;
    ctx->n_raw = n_raw;
    if (n_raw == 0) { CK(cudaStreamSynchronize(ctx->stream)); return MB_OK; }
    CK(ctx->raw_kidx.ensure(n_raw * sizeof(u64)));
    CK(ctx->raw_vbefore.ensure(n_raw * sizeof(u64)));
    DevBuf raw_cnt, raw_sum;
    CK(raw_cnt.ensure(n_raw * sizeof(u32)));
    CK(raw_sum.ensure(n_raw * sizeof(u64)));
    const DbView v = db_view(ctx, true);
    KL(-1, k_db_raw_summary<<<(unsigned) ((n_raw + 7) / 8), 256, 0, ctx->stream>>>(v, n_raw, raw_cnt.as<u32>(), raw_sum.as<u64>()));
    // The scan over raw blocks is a load-time step on a few MB: done on the host.
    std::vector<u32> cnt(n_raw);
    std::vector<u64> sum(n_raw), kidx(n_raw), vbefore(n_raw);
    CK(cudaMemcpyAsync(cnt.data(), raw_cnt.p, n_raw * sizeof(u32), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(sum.data(), raw_sum.p, n_raw * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    u64 k = 0, sv = 0;
    for (u64 b = 0; b < n_raw; ++b) { kidx[b] = k; vbefore[b] = sv; k += cnt[b]; sv += sum[b]; }
    if (k != n_kmers) {
        char msg[256];
        snprintf(msg, sizeof(msg), "diffIdx holds %llu k-mers but info holds %llu entries", (unsigned long long) k, (unsigned long long) n_kmers);
        ctx->n_words = 0; ctx->n_db_kmers = 0; ctx->n_raw = 0;
        return fail(ctx, MB_E_DATA, msg);
    }
    CK(cudaMemcpyAsync(ctx->raw_kidx.p, kidx.data(), n_raw * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->raw_vbefore.p, vbefore.data(), n_raw * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->fine_key = ctx->key_decided = false;
    if (int rc = build_units(ctx, planned_sort_t(ctx))) return rc;
    ctx->times.db_words = n_words;
    ctx->times.db_kmers = n_kmers;
    return MB_OK;
}

int mb_decode_db(mb_ctx *ctx, uint64_t *values_out, size_t cap, size_t *n_out) {
    if (!ctx || !n_out) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    *n_out = ctx->n_db_kmers;
    if (cap < ctx->n_db_kmers) return fail(ctx, MB_E_CAPACITY, "values_out too small");
    if (ctx->n_db_kmers == 0) return MB_OK;
    DevBuf vals;
    CK(vals.ensure(ctx->n_db_kmers * sizeof(u64)));
    const size_t smem = MT_WARPS * sizeof(WarpSmem);
    CK(cudaFuncSetAttribute(k_db_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const unsigned grid = (unsigned) std::min<u64>(((u64) ctx->n_units + MT_WARPS - 1) / MT_WARPS, 4096);
    KL(-1, k_db_decode<<<grid, MT_THREADS, smem, ctx->stream>>>(db_view(ctx, true), ctx->units.as<Unit>(), ctx->n_units, vals.as<u64>()));
    CK(cudaMemcpyAsync(values_out, vals.p, ctx->n_db_kmers * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

// ---- taxonomy ---------------------------------------------------------------------------------------
int mb_load_taxonomy(mb_ctx *ctx, const int32_t *parent, const int32_t *species_of, const int32_t *species_rank_of,
                     int32_t max_taxid, int32_t node0_taxid, int32_t eukaryota_taxid) {
    if (!ctx || !parent || !species_of || !species_rank_of || max_taxid < 0) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    const size_t n = (size_t) max_taxid + 1;
    // depth by memoised parent walk (root: parent == self)
    std::vector<int32_t> depth(n, -1);
    std::vector<int32_t> chain;
    for (size_t t = 0; t < n; ++t) {
        if (parent[t] < 0 || depth[t] >= 0) continue;
        chain.clear();
        int32_t x = (int32_t) t;
        while (depth[x] < 0) {
            const int32_t p = parent[x];
            if (p < 0 || p > max_taxid || parent[p] < 0) return fail(ctx, MB_E_DATA, "taxonomy: parent of an existing taxon does not exist");
            chain.push_back(x);
            if (p == x) { depth[x] = 0; chain.pop_back(); break; }
            if (chain.size() > n) return fail(ctx, MB_E_DATA, "taxonomy: cycle in the parent array");
            x = p;
        }
        int32_t d = depth[x];
        for (size_t k = chain.size(); k-- > 0;) depth[chain[k]] = ++d;
    }
    CK(ctx->parent.ensure(n * 4)); CK(ctx->depth.ensure(n * 4));
    CK(ctx->species_of.ensure(n * 4)); CK(ctx->species_rank_of.ensure(n * 4));
    CK(cudaMemcpyAsync(ctx->parent.p, parent, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->depth.p, depth.data(), n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->species_of.p, species_of, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->species_rank_of.p, species_rank_of, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    // packed (parent, depth) records for tax_lca
    std::vector<u64> node(n);
    for (size_t t = 0; t < n; ++t) node[t] = (u64) (u32) parent[t] | ((u64) (u32) depth[t] << 32);
    CK(ctx->tax_node.ensure(n * 8));
    CK(cudaMemcpyAsync(ctx->tax_node.p, node.data(), n * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->max_taxid = max_taxid;
    ctx->node0 = node0_taxid;
    ctx->euk = eukaryota_taxid;
    return MB_OK;
}

// ---- stages -----------------------------------------------------------------------------------------
// copies + device-side validation of one batch into (bases, offsets, lens, info) on stream `st`
static int enqueue_batch(mb_ctx *ctx, cudaStream_t st, DevBuf &d_bases, DevBuf &d_off, DevBuf &d_len, DevBuf &d_info,
                         const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens, uint32_t n_reads) {
    CK(d_bases.ensure(n_bases + 128));
    CK(d_off.ensure((size_t) n_reads * 2 * sizeof(u64) + 16));
    CK(d_len.ensure((size_t) n_reads * 2 * sizeof(u32) + 16));
    CK(d_info.ensure(sizeof(StageInfo)));
    if (n_bases) CK(cudaMemcpyAsync(d_bases.p, bases, n_bases, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(d_bases.as<u8>() + n_bases, 0, 128, st));
    CK(cudaMemsetAsync(d_info.p, 0, sizeof(StageInfo), st));
    if (n_reads) {
        CK(cudaMemcpyAsync(d_off.p, offsets, (size_t) n_reads * 2 * sizeof(u64), cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(d_len.p, lens, (size_t) n_reads * 2 * sizeof(u32), cudaMemcpyHostToDevice, st));
        k_validate_reads<<<(n_reads + 255) / 256, 256, 0, st>>>(d_off.as<u64>(), d_len.as<u32>(), n_reads, n_bases, ctx->par.seq_mode == 2,
                                                                 d_info.as<StageInfo>());
        CK(cudaGetLastError());
    }
    return MB_OK;
}

int mb_prefetch_reads(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                      uint32_t n_reads) {
    if (!ctx || (!bases && n_bases) || ((!offsets || !lens) && n_reads)) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->pf_done, cudaEventDisableTiming));
    }
    if (ctx->pf_pending) CK(cudaStreamSynchronize(ctx->copy_stream));          // an unclaimed prefetch is overwritten
    ctx->pf_pending = false;
    if (int rc = enqueue_batch(ctx, ctx->copy_stream, ctx->pf_bases, ctx->pf_offsets, ctx->pf_lens, ctx->pf_info, bases, n_bases, offsets, lens, n_reads))
        return rc;
    CK(cudaEventRecord(ctx->pf_done, ctx->copy_stream));
    ctx->pf_host[0] = bases; ctx->pf_host[1] = offsets; ctx->pf_host[2] = lens;
    ctx->pf_n_bases = n_bases; ctx->pf_n_reads = n_reads;
    ctx->pf_pending = true;
    return MB_OK;
}

int mb_stage_reads(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                   uint32_t n_reads) {
    if (!ctx || (!bases && n_bases) || ((!offsets || !lens) && n_reads)) return MB_E_ARG;
    if ((uint64_t) n_reads + ctx->seq_base >= (1u << 29)) return fail(ctx, MB_E_ARG, "at most 2^29-1 reads per batch (29-bit sequenceID, Kmer.h:11-16)");
    CK(cudaSetDevice(ctx->par.device));
    ctx->have_reads = ctx->have_kmers = ctx->have_matches = ctx->have_results = false;
    mark(ctx, EV_START);
    if (ctx->pf_pending && ctx->pf_host[0] == bases && ctx->pf_host[1] == offsets && ctx->pf_host[2] == lens &&
        ctx->pf_n_bases == n_bases && ctx->pf_n_reads == n_reads) {
        // this batch was prefetched: adopt its buffers, order the work stream behind the copy
        ctx->bases.swap(ctx->pf_bases); ctx->offsets.swap(ctx->pf_offsets);
        ctx->lens.swap(ctx->pf_lens); ctx->stage_info.swap(ctx->pf_info);
        CK(cudaStreamWaitEvent(ctx->stream, ctx->pf_done, 0));
        ctx->pf_pending = false;
    } else {
        if (int rc = enqueue_batch(ctx, ctx->stream, ctx->bases, ctx->offsets, ctx->lens, ctx->stage_info, bases, n_bases, offsets, lens, n_reads))
            return rc;
    }
    mark(ctx, EV_H2D);
    // layout check, k-mer slot bound (QueryIndexer.cpp:62-67) and longest query were computed on the device
    StageInfo info;
    if (int rc = read_back(ctx, ctx->stage_info.as<StageInfo>(), &info)) return rc;
    if (info.bad) return fail(ctx, MB_E_ARG, "mates must be packed back to back in increasing offset order");
    ctx->n_bases = n_bases;
    ctx->n_reads = n_reads;
    ctx->kmer_bound = info.bound;
    ctx->max_query_len = info.max_q + 8;
    ctx->have_reads = true;
    ctx->reads_masked = false;
    ctx->launches = 0;
    ctx->n_samples = 0;
    ctx->times.match_retries = 0;
    return MB_OK;
}

// ---- input pipeline on the device: FASTQ text -> staged batch (mb_fastq.cuh) ---------------------------------
// lines of one text: line_start[0 .. n_lines]; returns the number of COMPLETE lines (terminated, or closed by the end of
// the file when final_chunk)
static int fq_index_lines(mb_ctx *ctx, int f, const char *text, size_t n, int final_chunk, u64 *n_lines) {
    *n_lines = 0;
    u8 *d_text;
    if (ctx->fq_pf_pending && ctx->fq_pf_host[f] == text && ctx->fq_pf_n[f] == n) {
        // this text was prefetched (mb_prefetch_fastq): adopt its buffer, order the work stream behind the copy
        ctx->fq_text[f].swap(ctx->fq_pf_text[f]);
        ctx->fq_pf_host[f] = nullptr;
        CK(cudaStreamWaitEvent(ctx->stream, ctx->fq_pf_done, 0));
        d_text = ctx->fq_text[f].as<u8>();
    } else {
        CK(ctx->fq_text[f].ensure(n + 64));
        d_text = ctx->fq_text[f].as<u8>();
        if (n) CK(cudaMemcpyAsync(d_text, text, n, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemsetAsync(d_text + n, 0, 64, ctx->stream));
    }
    const u64 tiles = (n + FQ_TILE - 1) / FQ_TILE;
    if (tiles > 0xFFFFFFFFull) return fail(ctx, MB_E_ARG, "FASTQ chunk too large");
    CK(ctx->fq_tile_cnt.ensure((tiles + 1) * sizeof(u32)));
    CK(ctx->fq_tile_off.ensure((tiles + 2) * sizeof(u64)));
    u64 newlines = 0;
    if (tiles) {
        k_fq_count<<<(unsigned) tiles, FQ_THREADS, 0, ctx->stream>>>(d_text, n, ctx->fq_tile_cnt.as<u32>());
        k_scan_counts<<<1, 1024, 0, ctx->stream>>>(ctx->fq_tile_cnt.as<u32>(), ctx->fq_tile_off.as<u64>(), (u32) tiles);
        CK(cudaGetLastError());
        ctx->launches += 2;
        if (int rc = read_back(ctx, ctx->fq_tile_off.as<u64>() + tiles, &newlines)) return rc;
    }
    CK(ctx->fq_lines[f].ensure((newlines + 3) * sizeof(u64)));
    CK(cudaMemsetAsync(ctx->fq_lines[f].p, 0, sizeof(u64), ctx->stream));
    if (tiles) {
        k_fq_lines<<<(unsigned) tiles, FQ_THREADS, 0, ctx->stream>>>(d_text, n, ctx->fq_tile_off.as<u64>(), ctx->fq_lines[f].as<u64>());
        CK(cudaGetLastError());
        ++ctx->launches;
    }
    u64 lines = newlines;
    if (final_chunk && n && text[n - 1] != '\n') {               // the last line of the file has no newline: closed by the end
        const u64 end = (u64) n + 1;
        CK(cudaMemcpyAsync(ctx->fq_lines[f].as<u64>() + newlines + 1, &end, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        ++lines;
    }
    *n_lines = lines;
    return MB_OK;
}

int mb_stage_fastq(mb_ctx *ctx, const char *text1, size_t n1, const char *text2, size_t n2, uint32_t max_reads, int final_chunk,
                   mb_fastq_index *out) {
    if (!ctx || !out || (!text1 && n1) || (!text2 && n2)) return MB_E_ARG;
    memset(out, 0, sizeof(*out));
    CK(cudaSetDevice(ctx->par.device));
    ctx->have_reads = ctx->have_kmers = ctx->have_matches = ctx->have_results = false;
    ctx->fq_n_reads = 0;
    ctx->launches = 0;
    ctx->n_samples = 0;
    mark(ctx, EV_START);
    const bool paired = text2 != nullptr;
    u64 lines1 = 0, lines2 = 0;
    if (int rc = fq_index_lines(ctx, 0, text1, n1, final_chunk, &lines1)) return rc;
    if (paired) { if (int rc = fq_index_lines(ctx, 1, text2, n2, final_chunk, &lines2)) return rc; }
    ctx->fq_pf_pending = false;
    u64 nr64 = lines1 / 4;
    if (paired) nr64 = std::min<u64>(nr64, lines2 / 4);
    nr64 = std::min<u64>(nr64, max_reads);
    if (nr64 + ctx->seq_base >= (1u << 29)) return fail(ctx, MB_E_ARG, "at most 2^29-1 reads per batch (29-bit sequenceID, Kmer.h:11-16)");
    const u32 nr = (u32) nr64;
    // bytes of each text that belong to these records
    u64 c1 = 0, c2 = 0;
    if (nr) {
        if (int rc = read_back(ctx, ctx->fq_lines[0].as<u64>() + 4 * (u64) nr, &c1)) return rc;
        if (paired) { if (int rc = read_back(ctx, ctx->fq_lines[1].as<u64>() + 4 * (u64) nr, &c2)) return rc; }
    }
    out->consumed1 = std::min<u64>(c1, n1);
    out->consumed2 = std::min<u64>(c2, n2);
    const u64 nm = 2 * (u64) nr;
    CK(ctx->offsets.ensure(nm * sizeof(u64) + 16));
    CK(ctx->lens.ensure(nm * sizeof(u32) + 16));
    CK(ctx->fq_src.ensure(nm * sizeof(u64) + 16));
    CK(ctx->fq_name_off.ensure((u64) nr * sizeof(u64) + 16));
    CK(ctx->fq_name_len.ensure((u64) nr * sizeof(u32) + 16));
    CK(ctx->fq_info.ensure(sizeof(FqInfo)));
    CK(ctx->stage_info.ensure(sizeof(StageInfo)));
    CK(cudaMemsetAsync(ctx->fq_info.p, 0, sizeof(FqInfo), ctx->stream));
    CK(cudaMemsetAsync(ctx->stage_info.p, 0, sizeof(StageInfo), ctx->stream));
    u64 n_bases = 0;
    if (nr) {
        k_fq_records<<<(nr + 255) / 256, 256, 0, ctx->stream>>>(ctx->fq_text[0].as<u8>(), ctx->fq_lines[0].as<u64>(), nr, 0, paired ? 2 : 1,
                                                                 ctx->lens.as<u32>(), ctx->fq_src.as<u64>(), ctx->fq_name_off.as<u64>(),
                                                                 ctx->fq_name_len.as<u32>(), ctx->fq_info.as<FqInfo>());
        if (paired)
            k_fq_records<<<(nr + 255) / 256, 256, 0, ctx->stream>>>(ctx->fq_text[1].as<u8>(), ctx->fq_lines[1].as<u64>(), nr, 1, 2,
                                                                     ctx->lens.as<u32>(), ctx->fq_src.as<u64>(), nullptr, nullptr,
                                                                     ctx->fq_info.as<FqInfo>());
        // packed offsets = exclusive scan of the mate lengths
        const u64 st = (nm + SCAN_TILE - 1) / SCAN_TILE;
        CK(ctx->fq_sums.ensure((st + 1) * sizeof(u32)));
        CK(ctx->fq_sum_off.ensure((st + 2) * sizeof(u64)));
        k_scan_tile_sums<<<(unsigned) st, 256, 0, ctx->stream>>>(ctx->lens.as<u32>(), nm, ctx->fq_sums.as<u32>());
        k_scan_counts<<<1, 1024, 0, ctx->stream>>>(ctx->fq_sums.as<u32>(), ctx->fq_sum_off.as<u64>(), (u32) st);
        k_scan_tile_apply<<<(unsigned) st, 256, 0, ctx->stream>>>(ctx->lens.as<u32>(), nm, ctx->fq_sum_off.as<u64>(), ctx->offsets.as<u64>());
        CK(cudaGetLastError());
        ctx->launches += paired ? 5 : 4;
        if (int rc = read_back(ctx, ctx->fq_sum_off.as<u64>() + st, &n_bases)) return rc;
        FqInfo fi;
        if (int rc = read_back(ctx, ctx->fq_info.as<FqInfo>(), &fi)) return rc;
        if (fi.bad) {
            char msg[160];
            snprintf(msg, sizeof(msg), "FASTQ record %llu of this chunk is malformed (%s)", fi.bad_record,
                     fi.bad == 1 ? "header does not start with '@'" : fi.bad == 2 ? "separator line does not start with '+'"
                                                                                   : "sequence and quality differ in length");
            return fail(ctx, MB_E_DATA, msg);
        }
    }
    CK(ctx->bases.ensure(n_bases + 128));
    if (nr) {
        k_fq_pack<<<(unsigned) ((nm * 32 + 255) / 256), 256, 0, ctx->stream>>>(ctx->fq_text[0].as<u8>(), paired ? ctx->fq_text[1].as<u8>() : nullptr,
                                                                               ctx->fq_src.as<u64>(), ctx->lens.as<u32>(), ctx->offsets.as<u64>(),
                                                                               (u32) nm, ctx->bases.as<u8>());
        k_validate_reads<<<(nr + 255) / 256, 256, 0, ctx->stream>>>(ctx->offsets.as<u64>(), ctx->lens.as<u32>(), nr, n_bases, ctx->par.seq_mode == 2,
                                                                     ctx->stage_info.as<StageInfo>());
        CK(cudaGetLastError());
        ctx->launches += 2;
    }
    CK(cudaMemsetAsync(ctx->bases.as<u8>() + n_bases, 0, 128, ctx->stream));
    mark(ctx, EV_H2D);
    StageInfo info;
    if (int rc = read_back(ctx, ctx->stage_info.as<StageInfo>(), &info)) return rc;
    if (info.bad) return fail(ctx, MB_E_DATA, "internal: packed FASTQ batch failed the layout check");
    ctx->n_bases = n_bases;
    ctx->n_reads = nr;
    ctx->fq_n_reads = nr;
    ctx->kmer_bound = info.bound;
    ctx->max_query_len = info.max_q + 8;
    ctx->have_reads = true;
    ctx->reads_masked = false;
    ctx->times.match_retries = 0;
    out->n_reads = nr;
    out->n_bases = n_bases;
    return MB_OK;
}

int mb_prefetch_fastq(mb_ctx *ctx, const char *text1, size_t n1, const char *text2, size_t n2) {
    if (!ctx || (!text1 && n1) || (!text2 && n2)) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->pf_done, cudaEventDisableTiming));
    }
    if (!ctx->fq_pf_done) CK(cudaEventCreateWithFlags(&ctx->fq_pf_done, cudaEventDisableTiming));
    if (ctx->fq_pf_pending) CK(cudaStreamSynchronize(ctx->copy_stream));       // an unclaimed prefetch is overwritten
    ctx->fq_pf_pending = false;
    const char *t[2] = {text1, text2};
    const size_t n[2] = {n1, n2};
    for (int f = 0; f < 2; ++f) {
        ctx->fq_pf_host[f] = nullptr;
        if (!t[f]) continue;
        CK(ctx->fq_pf_text[f].ensure(n[f] + 64));
        if (n[f]) CK(cudaMemcpyAsync(ctx->fq_pf_text[f].p, t[f], n[f], cudaMemcpyHostToDevice, ctx->copy_stream));
        CK(cudaMemsetAsync(ctx->fq_pf_text[f].as<u8>() + n[f], 0, 64, ctx->copy_stream));
        ctx->fq_pf_host[f] = t[f];
        ctx->fq_pf_n[f] = n[f];
    }
    CK(cudaEventRecord(ctx->fq_pf_done, ctx->copy_stream));
    ctx->fq_pf_pending = true;
    return MB_OK;
}

int mb_mem_info(mb_ctx *ctx, size_t *free_bytes, size_t *total_bytes) {
    if (!ctx || !free_bytes || !total_bytes) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    CK(cudaMemGetInfo(free_bytes, total_bytes));
    return MB_OK;
}

void *mb_alloc_host(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void mb_free_host(void *p) { if (p) cudaFreeHost(p); }

int mb_fetch_fastq_names(mb_ctx *ctx, uint64_t *name_off, uint32_t *name_len) {
    if (!ctx || !name_off || !name_len) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->fq_n_reads) {
        CK(cudaMemcpyAsync(name_off, ctx->fq_name_off.p, (size_t) ctx->fq_n_reads * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaMemcpyAsync(name_len, ctx->fq_name_len.p, (size_t) ctx->fq_n_reads * sizeof(u32), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
    }
    return MB_OK;
}

int mb_fetch_staged_reads(mb_ctx *ctx, char *bases, uint64_t *offsets, uint32_t *lens) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_reads) return fail(ctx, MB_E_ARG, "no batch staged");
    CK(cudaSetDevice(ctx->par.device));
    if (bases && ctx->n_bases) CK(cudaMemcpyAsync(bases, ctx->bases.p, ctx->n_bases, cudaMemcpyDeviceToHost, ctx->stream));
    if (offsets && ctx->n_reads) CK(cudaMemcpyAsync(offsets, ctx->offsets.p, (size_t) ctx->n_reads * 2 * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    if (lens && ctx->n_reads) CK(cudaMemcpyAsync(lens, ctx->lens.p, (size_t) ctx->n_reads * 2 * sizeof(u32), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

// --mask 1 (mb_mask.cuh): every mate of the staged batch masked in place, once per batch
static int mask_impl(mb_ctx *ctx) {
    if (ctx->reads_masked || !ctx->n_reads || !ctx->n_bases) { ctx->reads_masked = true; return MB_OK; }
    // Tantan::Tantan (lib/mmseqs/lib/tantan/tantan.cpp:81-135) for the arguments of SeqIterator.cpp:161-170; the two
    // likelihood ratios are what ProbabilityMatrix(NucleotideMatrix("nucleotide.out", 1.0, 0.0)) holds at run time
    // (dumped from the reference classes: tests/golden/data/mask_lr.bin, checked by the tests)
    MaskModel M;
    const double repeatProb = 0.005, repeatEndProb = 0.05, decay = 0.9;
    M.b2b = 1 - repeatProb;
    M.f2b = repeatEndProb;
    M.f2f0 = 1 - repeatEndProb;
    M.lr_match = 0x1.c6a258ad6ec6ep+1;
    M.lr_mismatch = 0x1.31f371483f33ap-3;
    M.min_prob = (double) ctx->par.mask_prob;
    double p = repeatProb * ((1 - decay) / (1 - std::pow(decay, MK_OFF)));     // firstRepeatOffsetProb, tantan.cpp:26-31
    for (int i = 0; i < MK_OFF; ++i) { M.b2f[i] = p; p *= decay; }
    const u32 n_mates = ctx->n_reads * 2;
    CK(ctx->mask_probs.ensure((ctx->n_bases + 16) * sizeof(float)));
    CK(ctx->mask_scales.ensure(((ctx->n_bases >> 4) + n_mates + 2) * sizeof(double)));
    KL(KID_MASK, k_mask_mates<<<(n_mates + MK_THREADS - 1) / MK_THREADS, MK_THREADS, 0, ctx->stream>>>(
                     ctx->bases.as<u8>(), ctx->offsets.as<u64>(), ctx->lens.as<u32>(), n_mates, M, ctx->mask_probs.as<float>(),
                     ctx->mask_scales.as<double>()));
    ctx->reads_masked = true;
    return MB_OK;
}

int mb_mask_reads(mb_ctx *ctx) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_reads) return fail(ctx, MB_E_ARG, "mb_stage_reads first");
    CK(cudaSetDevice(ctx->par.device));
    if (int rc = mask_impl(ctx)) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

// exclusive scan of u32 -> u64 (offsets[n] = total) with the three scan kernels of the FASTQ parser
static int scan_u32(mb_ctx *ctx, const u32 *v, u64 n, u64 *out, DevBuf &sums, DevBuf &sum_off, u64 *total) {
    const u64 st = (n + SCAN_TILE - 1) / SCAN_TILE;
    CK(sums.ensure((st + 1) * sizeof(u32)));
    CK(sum_off.ensure((st + 2) * sizeof(u64)));
    k_scan_tile_sums<<<(unsigned) st, 256, 0, ctx->stream>>>(v, n, sums.as<u32>());
    k_scan_counts<<<1, 1024, 0, ctx->stream>>>(sums.as<u32>(), sum_off.as<u64>(), (u32) st);
    k_scan_tile_apply<<<(unsigned) st, 256, 0, ctx->stream>>>(v, n, sum_off.as<u64>(), out);
    CK(cudaGetLastError());
    ctx->launches += 3;
    return read_back(ctx, sum_off.as<u64>() + st, total);
}

int mb_build_index(mb_ctx *ctx, const uint64_t *values, const int32_t *taxids, const int32_t *species, size_t n, int split_num,
                   uint16_t *diffIdx_out, size_t cap_words, size_t *n_words, uint32_t *info_out, size_t cap_kmers, size_t *n_kmers,
                   mb_split *splits_out) {
    if (!ctx || !values || !taxids || !species || !n_words || !n_kmers) return MB_E_ARG;
    if (split_num < 2) return fail(ctx, MB_E_ARG, "split_num must be at least 2");
    if (!ctx->parent.p) return fail(ctx, MB_E_ARG, "mb_load_taxonomy first (the duplicates' taxa are folded with LCA)");
    if (n >= (1ull << 32) * SCAN_TILE) return fail(ctx, MB_E_ARG, "too many records for one buffer");
    CK(cudaSetDevice(ctx->par.device));
    *n_words = *n_kmers = 0;
    if (splits_out) memset(splits_out, 0, sizeof(mb_split) * (size_t) split_num);
    if (n == 0) return MB_OK;
    DevBuf in_v, in_t, in_s, rec_a, rec_b, head, pos, out_v, out_t, cnt, woff, words, sums, sum_off, probes;
    CK(in_v.ensure(n * 8)); CK(in_t.ensure(n * 4)); CK(in_s.ensure(n * 4));
    CK(cudaMemcpyAsync(in_v.p, values, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(in_t.p, taxids, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(in_s.p, species, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(rec_a.ensure((n + 1) * sizeof(mb_kmer)));
    const unsigned g = (unsigned) ((n + 255) / 256);
    KL(-100, k_build_pack<<<g, 256, 0, ctx->stream>>>(in_v.as<u64>(), in_t.as<int32_t>(), in_s.as<int32_t>(), n, rec_a.as<mb_kmer>()));
    // compareTargetKmer orders by (value, speciesId, taxId); the taxId order inside a run of duplicates does not reach the
    // output (the run keeps its value and species, and LCA does not depend on the order), so (value, speciesId) is sorted
    TargetKey key;
    int sp_bits = 1;
    while ((1ll << sp_bits) <= (long long) ctx->max_taxid) ++sp_bits;
    key.sp_passes = (sp_bits + 7) / 8;
    if (int rc = radix_sort<mb_kmer, TargetKey, 16>(ctx, rec_a, rec_b, n, key, key.sp_passes + 8, -100)) return rc;
    CK(head.ensure(n * sizeof(u32)));
    CK(pos.ensure((n + 1) * sizeof(u64)));
    KL(-100, k_build_heads<<<g, 256, 0, ctx->stream>>>(rec_a.as<mb_kmer>(), n, head.as<u32>()));
    u64 uniq = 0;
    if (int rc = scan_u32(ctx, head.as<u32>(), n, pos.as<u64>(), sums, sum_off, &uniq)) return rc;
    *n_kmers = (size_t) uniq;
    TaxView tv;
    tv.parent = ctx->parent.as<int32_t>(); tv.depth = ctx->depth.as<int32_t>(); tv.species_of = ctx->species_of.as<int32_t>();
    tv.species_rank_of = ctx->species_rank_of.as<int32_t>(); tv.max_taxid = ctx->max_taxid; tv.node0_taxid = ctx->node0;
    tv.eukaryota = ctx->euk;
    tv.node = ctx->tax_node.as<u64>();
    CK(out_v.ensure((uniq + 1) * 8)); CK(out_t.ensure((uniq + 1) * 4));
    KL(-100, k_build_reduce<<<g, 256, 0, ctx->stream>>>(rec_a.as<mb_kmer>(), n, head.as<u32>(), pos.as<u64>(), tv, out_v.as<u64>(), out_t.as<u32>()));
    const unsigned gu = (unsigned) ((uniq + 255) / 256);
    CK(cnt.ensure(uniq * sizeof(u32)));
    CK(woff.ensure((uniq + 2) * sizeof(u64)));
    KL(-100, k_build_wordcount<<<gu, 256, 0, ctx->stream>>>(out_v.as<u64>(), uniq, cnt.as<u32>()));
    u64 nw = 0;
    if (int rc = scan_u32(ctx, cnt.as<u32>(), uniq, woff.as<u64>(), sums, sum_off, &nw)) return rc;
    CK(cudaMemcpyAsync(woff.as<u64>() + uniq, &nw, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));     // woff[uniq] = total
    *n_words = (size_t) nw;
    if (uniq > cap_kmers || nw > cap_words) return fail(ctx, MB_E_CAPACITY, "output buffers too small (sizes returned in n_words / n_kmers)");
    if (!diffIdx_out || !info_out) return fail(ctx, MB_E_ARG, "null output");
    CK(words.ensure((nw + 8) * sizeof(u16)));
    KL(-100, k_build_encode<<<gu, 256, 0, ctx->stream>>>(out_v.as<u64>(), uniq, woff.as<u64>(), words.as<u16>()));
    CK(cudaMemcpyAsync(diffIdx_out, words.p, nw * sizeof(u16), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(info_out, out_t.p, uniq * sizeof(u32), cudaMemcpyDeviceToHost, ctx->stream));
    if (splits_out) {
        const u64 size_of_split = uniq / (u64) (split_num - 1);
        CK(probes.ensure((size_t) split_num * sizeof(SplitProbe)));
        KL(-100, k_build_splits<<<(split_num + 255) / 256, 256, 0, ctx->stream>>>(out_v.as<u64>(), uniq, woff.as<u64>(), size_of_split, split_num,
                                                                                   probes.as<SplitProbe>()));
        std::vector<SplitProbe> pr((size_t) split_num);
        CK(cudaMemcpyAsync(pr.data(), probes.p, pr.size() * sizeof(SplitProbe), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        // the reference's state machine has ONE pending flag: a threshold reached before the previous one found its
        // amino-acid change merges with it, so equal answers of consecutive thresholds are one split
        int idx = 1;
        u64 last_j = ~0ull;
        for (int os = 1; os < split_num; ++os) {
            if (pr[(size_t) os].j == ~0ull || pr[(size_t) os].j == last_j) continue;
            last_j = pr[(size_t) os].j;
            splits_out[idx].ADkmer = pr[(size_t) os].adkmer;
            splits_out[idx].diffIdxOffset = pr[(size_t) os].words_through;
            splits_out[idx].infoIdxOffset = pr[(size_t) os].j + 1;
            ++idx;
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

static int extract_impl(mb_ctx *ctx, size_t *n_kmers, const bool routed) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_reads) return fail(ctx, MB_E_ARG, "mb_stage_reads first");
    if (routed && !ctx->routed) return fail(ctx, MB_E_ARG, "mb_peer_attach first");
    CK(cudaSetDevice(ctx->par.device));
    ctx->have_kmers = ctx->kmers_sorted = ctx->have_matches = ctx->have_results = false;
    ctx->n_kmers = 0;
    ctx->kmers_seg.n = 0;
    ctx->n_samples = 0;               // the per-kernel table covers one pass of the device stages
    ctx->launches = 0;
    const u32 n_tiles = (u32) ((ctx->n_bases + EX_TILE - 1) / EX_TILE);
    if (ctx->par.mask_mode) { if (int rc = mask_impl(ctx)) return rc; }
    if (n_tiles && ctx->n_reads) {
        ExtractArgs a;
        a.bases = ctx->bases.as<u8>(); a.n_bases = ctx->n_bases; a.offsets = ctx->offsets.as<u64>(); a.lens = ctx->lens.as<u32>();
        a.n_mates = ctx->n_reads * 2; a.seq_mode = ctx->par.seq_mode; a.format = ctx->par.kmer_format;
        a.syncmer = ctx->par.syncmer; a.smer_len = ctx->par.smer_len; a.seq_base = ctx->seq_base;
        CK(ctx->tile_mate.ensure(((size_t) n_tiles + 1) * sizeof(u32)));
        if (!routed) CK(ctx->kmers_a.ensure((ctx->kmer_bound + 1) * sizeof(mb_kmer)));
        else CK(cudaMemsetAsync(ctx->route.sent_k, 0, MB_MAX_RANKS * sizeof(unsigned long long), ctx->stream));
        CK(cudaMemsetAsync(&ctx->dcnt()->kmer_count, 0, sizeof(unsigned long long), ctx->stream));
        KL(KID_TILE_MATE, k_tile_first_mate<<<(n_tiles + 1 + 255) / 256, 256, 0, ctx->stream>>>(a, n_tiles, ctx->tile_mate.as<u32>()));
        if (!ctx->ex_attr_set) {
            CK(cudaFuncSetAttribute(k_extract<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) extract_smem_bytes()));
            CK(cudaFuncSetAttribute(k_extract<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) extract_smem_bytes()));
            CK(cudaFuncSetAttribute(k_extract<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) extract_smem_bytes()));
            CK(cudaFuncSetAttribute(k_extract<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) extract_smem_bytes()));
            ctx->ex_attr_set = true;
        }
        mb_kmer *out = routed ? nullptr : ctx->kmers_a.as<mb_kmer>();
        const Route *rt = routed ? ctx->route_dev.as<Route>() : nullptr;
        const u32 *tm = ctx->tile_mate.as<u32>();
        const Tables *tb = ctx->tables.as<Tables>();
        unsigned long long *cnt = &ctx->dcnt()->kmer_count;
        const size_t sm = extract_smem_bytes();
        if (a.format == 2 && !routed) KL(KID_EXTRACT_WRITE, k_extract<true, false><<<n_tiles, EX_THREADS, sm, ctx->stream>>>(a, tm, tb, out, cnt, ctx->kmer_bound, rt));
        else if (a.format == 2) KL(KID_EXTRACT_WRITE, k_extract<true, true><<<n_tiles, EX_THREADS, sm, ctx->stream>>>(a, tm, tb, out, cnt, ctx->kmer_bound, rt));
        else if (!routed) KL(KID_EXTRACT_WRITE, k_extract<false, false><<<n_tiles, EX_THREADS, sm, ctx->stream>>>(a, tm, tb, out, cnt, ctx->kmer_bound, rt));
        else KL(KID_EXTRACT_WRITE, k_extract<false, true><<<n_tiles, EX_THREADS, sm, ctx->stream>>>(a, tm, tb, out, cnt, ctx->kmer_bound, rt));
        if (routed) { if (int rc = publish_sent(ctx, 0)) return rc; }
        unsigned long long total = 0;
        if (int rc = read_back(ctx, &ctx->dcnt()->kmer_count, &total)) return rc;
        if (total > ctx->kmer_bound) return fail(ctx, MB_E_DATA, "internal: extracted more k-mers than the slot bound");
        ctx->n_kmers = total;
    }
    mark(ctx, EV_EXTRACT);
    if (n_kmers) *n_kmers = ctx->n_kmers;
    if (routed) ctx->n_kmers = 0;            // they left for their owners; mb_take_routed_kmers brings this rank's share
    else ctx->have_kmers = true;
    return MB_OK;
}
int mb_extract_kmers(mb_ctx *ctx, size_t *n_kmers) { return extract_impl(ctx, n_kmers, false); }
int mb_extract_kmers_routed(mb_ctx *ctx, size_t *n_emitted) { return extract_impl(ctx, n_emitted, true); }

int mb_sort_kmers(mb_ctx *ctx) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_kmers) return fail(ctx, MB_E_ARG, "no k-mers resident");
    CK(cudaSetDevice(ctx->par.device));
    // The join needs the queries of one DB block contiguous, not ordered (every query looks its group up in the decoded
    // block): ordering the leading `bits` of the 40 amino-acid bits takes bits/8 passes instead of 5, and for format 2
    // the re-coded 24-bit key (SortT mode 1, mb_common.cuh; kmer_sort_bits 21) resolves 5.6 residues in 3 passes.
    // the first batch of at least a million k-mers after mb_load_db decides the default key for this database (kept, so
    // that the unit index is not rebuilt from batch to batch): seven residues when the DB stream has more than four words
    // per query k-mer
    if (!ctx->key_decided && ctx->n_kmers >= (1u << 20)) {
        ctx->key_decided = true;
        ctx->fine_key = ctx->n_words > 4 * ctx->n_kmers;
    }
    ctx->sort_t = planned_sort_t(ctx);
    const int mode = ctx->sort_t.mode;
    const int bits = mode ? 0 : 40 - ctx->sort_t.shift;
    if (!mode && bits != 24 && bits != 32 && bits != 40) return fail(ctx, MB_E_ARG, "MB_KSORT_BITS must be 21, 31, 24, 32 or 40");
    KmerKey key;
    key.first_bit = 24 + ctx->sort_t.shift;
    const int passes = mode == 1 ? 3 : mode == 2 ? 4 : bits / 8;
    static const int k_items = getenv("MB_KSORT_ITEMS") ? atoi(getenv("MB_KSORT_ITEMS")) : 16;     // tuning knob
    int rc = 0;
    if (mode == 1) {
        KmerKeyPacked pk;
        if (k_items == 8) rc = radix_sort<mb_kmer, KmerKeyPacked, 8>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, pk, passes, KID_KSORT_HIST, &ctx->kmers_seg);
        else if (k_items == 12) rc = radix_sort<mb_kmer, KmerKeyPacked, 12>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, pk, passes, KID_KSORT_HIST, &ctx->kmers_seg);
        else rc = radix_sort<mb_kmer, KmerKeyPacked, 16>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, pk, passes, KID_KSORT_HIST, &ctx->kmers_seg);
    }
    else if (mode == 2) {
        KmerKeyPacked7 pk;
        rc = radix_sort<mb_kmer, KmerKeyPacked7, 16>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, pk, passes, KID_KSORT_HIST, &ctx->kmers_seg);
    }
    else if (k_items == 8) rc = radix_sort<mb_kmer, KmerKey, 8>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, key, passes, KID_KSORT_HIST, &ctx->kmers_seg);
    else if (k_items == 12) rc = radix_sort<mb_kmer, KmerKey, 12>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, key, passes, KID_KSORT_HIST, &ctx->kmers_seg);
    else rc = radix_sort<mb_kmer, KmerKey, 16>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, key, passes, KID_KSORT_HIST, &ctx->kmers_seg);
    if (rc) return rc;
    mark(ctx, EV_SORTK);
    ctx->kmers_sorted = true;
    return MB_OK;
}

static int match_impl(mb_ctx *ctx, size_t *n_matches, const bool routed) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_kmers || !ctx->kmers_sorted) return fail(ctx, MB_E_ARG, "sorted k-mers required");
    if (routed && !ctx->routed) return fail(ctx, MB_E_ARG, "mb_peer_attach first");
    ctx->matches_seg.n = 0;
    if (ctx->max_taxid < 0) return fail(ctx, MB_E_ARG, "mb_load_taxonomy first");
    CK(cudaSetDevice(ctx->par.device));
    ctx->have_matches = ctx->matches_sorted = ctx->have_results = false;
    ctx->segments_valid = false;
    ctx->n_matches = 0;
    if (ctx->n_units && ctx->n_kmers) {
        // the unit index follows the key the resident queries are ordered by
        if (ctx->units_t.mode != ctx->sort_t.mode || ctx->units_t.shift != ctx->sort_t.shift)
            if (int rc = build_units(ctx, ctx->sort_t)) return rc;
        const u32 nu = ctx->n_units;
        CK(ctx->qstart.ensure(((size_t) nu + 2) * sizeof(u64)));
        u64 *qstart = ctx->qstart.as<u64>();
        KL(KID_MATCH_PARTITION, k_match_partition<<<(nu + 1 + 255) / 256, 256, 0, ctx->stream>>>(ctx->units.as<Unit>(), nu, ctx->kmers_a.as<mb_kmer>(),
                                                                   ctx->n_kmers, ctx->sort_t, qstart));
        if (!ctx->match_attr_set) {
            CK(cudaFuncSetAttribute(k_match<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) match_smem_bytes()));
            CK(cudaFuncSetAttribute(k_match<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) match_smem_bytes()));
            ctx->match_attr_set = true;
        }
        size_t cap = std::max<size_t>(ctx->matches_a.cap / sizeof(mb_match), std::max<size_t>(1u << 16, ctx->n_kmers / 2));
        for (int attempt = 0; attempt < 3; ++attempt) {
            if (!routed) {
                CK(ctx->matches_a.ensure(cap * sizeof(mb_match)));
                cap = ctx->matches_a.cap / sizeof(mb_match);
            }
            CK(cudaMemsetAsync(&ctx->dcnt()->match_count, 0, sizeof(unsigned long long), ctx->stream));
            CK(cudaMemsetAsync(&ctx->dcnt()->error_flag, 0, sizeof(u32), ctx->stream));
            MatchOut mo;
            mo.out = routed ? nullptr : ctx->matches_a.as<mb_match>(); mo.cap = routed ? 0 : cap;
            mo.counter = &ctx->dcnt()->match_count; mo.error_flag = &ctx->dcnt()->error_flag; mo.route = nullptr;
            CK(cudaMemsetAsync(&ctx->dcnt()->match_ticket, 0, sizeof(u32), ctx->stream));
            if (routed) CK(cudaMemsetAsync(ctx->route.sent_m, 0, MB_MAX_RANKS * sizeof(unsigned long long), ctx->stream));
            // persistent: as many CTAs as are resident at once (3 per SM: 66 KB of shared memory each); every warp draws units
            if (!ctx->match_ctas) {
                int per_sm = 0;
                CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_match<false>, MT_THREADS, match_smem_bytes()));
                static const int env_ctas = getenv("MB_MATCH_CTAS_PER_SM") ? atoi(getenv("MB_MATCH_CTAS_PER_SM")) : 0;     // tuning knob
                if (env_ctas > 0 && env_ctas < per_sm) per_sm = env_ctas;
                ctx->match_ctas = std::max(1, per_sm) * ctx->sms;
            }
            const unsigned grid = (unsigned) std::min<u64>(((u64) nu + MT_WARPS - 1) / MT_WARPS, (u64) ctx->match_ctas);
#define MB_LAUNCH_MATCH(R, ROUTE) KL(KID_MATCH, (k_match<R><<<grid, MT_THREADS, match_smem_bytes(), ctx->stream>>>( \
                db_view(ctx, false), ctx->units.as<Unit>(), nu, ctx->kmers_a.as<mb_kmer>(), qstart, mo, \
                ctx->tables.as<Tables>(), ctx->par.kmer_format, &ctx->dcnt()->match_ticket, ROUTE)))
            if (!routed) MB_LAUNCH_MATCH(false, nullptr);
            else MB_LAUNCH_MATCH(true, ctx->route_dev.as<Route>());
#undef MB_LAUNCH_MATCH
            if (routed) { if (int rc = publish_sent(ctx, 1)) return rc; }
            unsigned long long total = 0;
            if (int rc = read_back(ctx, &ctx->dcnt()->match_count, &total)) return rc;
            if (routed || total <= cap) { ctx->n_matches = total; break; }
            // capacity miss: the kernel counted everything but wrote only what fits; size exactly and rerun
            ++ctx->times.match_retries;
            cap = (size_t) total;
            if (attempt == 2) return fail(ctx, MB_E_CAPACITY, "match buffer could not be sized");
        }
        u32 flag = 0;
        if (int rc = read_back(ctx, &ctx->dcnt()->error_flag, &flag)) return rc;
        if (flag) return fail(ctx, MB_E_DATA, "a DB k-mer carries a tax ID without species mapping (KmerMatcher.cpp:433-441)");
    }
    mark(ctx, EV_MATCH);
    if (n_matches) *n_matches = ctx->n_matches;
    if (routed) ctx->n_matches = 0;          // they left for their home ranks; mb_take_routed_matches brings this rank's
    else ctx->have_matches = true;
    return MB_OK;
}
int mb_match_kmers(mb_ctx *ctx, size_t *n_matches) { return match_impl(ctx, n_matches, false); }
int mb_match_kmers_routed(mb_ctx *ctx, size_t *n_emitted) { return match_impl(ctx, n_emitted, true); }

static MatchKey match_key(const mb_ctx *ctx) {
    MatchKey k;
    // pos < queryLength + 3 (mate 2 offset, KmerExtractor.cpp:339-347)
    k.pos_bits = bits_for((u64) ctx->max_query_len + 8);
    k.sp_bits = bits_for((u64) std::max(ctx->max_taxid, 1));
    k.seq_bits = bits_for((u64) ctx->n_reads + 1);
    k.seq_base = ctx->seq_base;
    return k;
}

// per-read segments [seg_begin, seg_end) of the match list (which must be grouped by read): Classifier.cpp:174-185
static int compute_segments(mb_ctx *ctx) {
    const u32 nr = ctx->n_reads;
    CK(ctx->seg_begin.ensure((size_t) std::max<u32>(nr, 1) * sizeof(u64)));
    CK(ctx->seg_end.ensure((size_t) std::max<u32>(nr, 1) * sizeof(u64)));
    if (nr) {
        CK(cudaMemsetAsync(ctx->seg_begin.p, 0, (size_t) nr * sizeof(u64), ctx->stream));
        CK(cudaMemsetAsync(ctx->seg_end.p, 0, (size_t) nr * sizeof(u64), ctx->stream));
    }
    if (ctx->n_matches && nr)
        KL(KID_SEGMENT, k_segment_matches<<<(unsigned) ((ctx->n_matches + 255) / 256), 256, 0, ctx->stream>>>(
            ctx->matches_a.as<mb_match>(), ctx->n_matches, nr, ctx->seq_base, ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>()));
    ctx->segments_valid = true;
    return MB_OK;
}

int mb_sort_matches(mb_ctx *ctx) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_matches) return fail(ctx, MB_E_ARG, "no matches resident");
    CK(cudaSetDevice(ctx->par.device));
    const MatchKey key = match_key(ctx);
    ctx->segments_valid = false;
    bool done = false;
    SegKey sk;
    sk.pos_bits = key.pos_bits; sk.sp_bits = key.sp_bits;
    // Short reads: group by read with radix passes on the sequence ID alone, then order every read's segment in
    // shared memory (mb_segsort.cuh).  Long reads have segments of thousands of matches: full-key radix sort.
    if (ctx->par.seq_mode != 3 && sk.total_bits() <= 64 && ctx->have_reads && ctx->n_matches && ctx->n_reads) {
        MatchSeqKey qk;
        qk.seq_base = ctx->seq_base;
        const int seq_passes = (key.seq_bits + 7) / 8;
        static const int m_items = getenv("MB_MSORT_ITEMS") ? atoi(getenv("MB_MSORT_ITEMS")) : 12;    // tuning knob
        int rcs = 0;
        if (m_items == 4) rcs = radix_sort<mb_match, MatchSeqKey, 4>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, qk, seq_passes, KID_MSORT_HIST, &ctx->matches_seg);
        else if (m_items == 6) rcs = radix_sort<mb_match, MatchSeqKey, 6>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, qk, seq_passes, KID_MSORT_HIST, &ctx->matches_seg);
        else if (m_items == 12) rcs = radix_sort<mb_match, MatchSeqKey, 12>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, qk, seq_passes, KID_MSORT_HIST, &ctx->matches_seg);
        else rcs = radix_sort<mb_match, MatchSeqKey, 8>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, qk, seq_passes, KID_MSORT_HIST, &ctx->matches_seg);
        if (rcs) return rcs;
        if (int rc = compute_segments(ctx)) return rc;
        CK(ctx->matches_b.ensure(ctx->n_matches * sizeof(mb_match)));
        CK(cudaMemsetAsync(&ctx->dcnt()->seg_ticket, 0, 4 * sizeof(u32), ctx->stream));
        CK(cudaMemsetAsync(&ctx->dcnt()->seg_huge_words, 0, sizeof(unsigned long long), ctx->stream));
        CK(ctx->seg_long.ensure((size_t) ctx->n_reads * sizeof(u32)));
        CK(ctx->seg_huge.ensure((size_t) SS_HUGE_MAX * sizeof(HugeSeg)));
        if (!ctx->seg_ctas) {
            CK(cudaFuncSetAttribute(k_sort_segments_long, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) segsort_smem_bytes()));
            int per_sm = 0;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sort_segments<true>, SS_THREADS, 0));
            ctx->seg_ctas = std::max(1, per_sm) * ctx->sms;
        }
        const unsigned grid = (unsigned) std::min<u64>((u64) ctx->seg_ctas, ((u64) ctx->n_reads + 63) / 64);
        if (sk.total_bits() <= 55)
            KL(KID_SEGSORT, k_sort_segments<true><<<grid, SS_THREADS, 0, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->matches_b.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->n_reads, sk,
                &ctx->dcnt()->seg_ticket, &ctx->dcnt()->seg_too_long, ctx->seg_long.as<u32>(), &ctx->dcnt()->seg_long_count,
                ctx->seg_huge.as<HugeSeg>(), &ctx->dcnt()->seg_huge_words));
        else
            KL(KID_SEGSORT, k_sort_segments<false><<<grid, SS_THREADS, 0, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->matches_b.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->n_reads, sk,
                &ctx->dcnt()->seg_ticket, &ctx->dcnt()->seg_too_long, ctx->seg_long.as<u32>(), &ctx->dcnt()->seg_long_count,
                ctx->seg_huge.as<HugeSeg>(), &ctx->dcnt()->seg_huge_words));
        // segments of 513..1024 matches (rare with short reads): shared-memory network, one warp each
        KL(KID_SEGSORT_LONG, k_sort_segments_long<<<(unsigned) ctx->sms, SS_THREADS, segsort_smem_bytes(), ctx->stream>>>(
            ctx->matches_a.as<mb_match>(), ctx->matches_b.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), sk,
            &ctx->dcnt()->seg_long_ticket, ctx->seg_long.as<u32>(), &ctx->dcnt()->seg_long_count));
        u32 too_long = 0;
        if (int rc = read_back(ctx, &ctx->dcnt()->seg_too_long, &too_long)) return rc;
        if (too_long && too_long <= SS_HUGE_MAX) {
            // a few reads with more than SS_CAP matches: one CTA each over a slice of global key scratch
            unsigned long long words = 0;
            if (int rc = read_back(ctx, &ctx->dcnt()->seg_huge_words, &words)) return rc;
            CK(ctx->seg_huge_keys.ensure((size_t) words * sizeof(u64)));
            CK(ctx->seg_huge_idx.ensure((size_t) words * sizeof(u32)));
            KL(KID_SEGSORT_LONG, k_sort_segments_huge<<<std::min<unsigned>(too_long, (unsigned) ctx->sms * 4), SS_THREADS, 0, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->matches_b.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), sk,
                ctx->seg_huge.as<HugeSeg>(), too_long, ctx->seg_huge_keys.as<u64>(), ctx->seg_huge_idx.as<u32>()));
            too_long = 0;
        }
        done = too_long == 0;      // otherwise more reads beyond SS_CAP than the list holds: the full-key sort below, from the
                                   // untouched input (the segments beyond SS_CAP were not copied)
        if (done) {
            std::swap(ctx->matches_a.p, ctx->matches_b.p);
            std::swap(ctx->matches_a.cap, ctx->matches_b.cap);
        }
    }
    if (!done) {
        ctx->segments_valid = false;
        // full compareMatches key (long reads, contigs): the 24 DNA bits stay out of the radix passes — records that tie above
        // them are a handful at most and are ordered in place afterwards (10 -> 7 passes at configs[3])
        static const bool whole_key = getenv("MB_MSORT_WHOLE_KEY") != nullptr;          // tuning knob: the 24 DNA bits as passes too
        MatchKey mk = key;
        mk.skip_bits = whole_key ? 0 : 24;
        const int passes = (mk.total_bits() - mk.skip_bits + 7) / 8;
        if (int rc = radix_sort<mb_match, MatchKey, 8>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, mk, passes, KID_MSORT_HIST, &ctx->matches_seg)) return rc;
        if (mk.skip_bits && ctx->n_matches)
            KL(KID_SEGSORT_LONG, k_fix_dna_ties<<<(unsigned) ((ctx->n_matches + 255) / 256), 256, 0, ctx->stream>>>(ctx->matches_a.as<mb_match>(), ctx->n_matches, mk));
    }
    mark(ctx, EV_SORTM);
    ctx->matches_sorted = true;
    return MB_OK;
}

int mb_score_reads(mb_ctx *ctx, size_t *taxcnt_used) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->have_matches || !ctx->matches_sorted) return fail(ctx, MB_E_ARG, "sorted matches required");
    if (!ctx->have_reads) return fail(ctx, MB_E_ARG, "reads must be staged (lengths are needed)");
    CK(cudaSetDevice(ctx->par.device));
    // results of the previous batch still on their way to the host (mb_fetch_results_async): they leave before the result
    // arrays are written again (the copy was enqueued a whole batch ago, so this does not wait in practice)
    if (ctx->fetch_pending) { CK(cudaEventSynchronize(ctx->fetch_done)); ctx->fetch_pending = false; }
    ctx->res_valid = false;
    const u32 nr = ctx->n_reads;
    ctx->tc_used = 0;
    if (nr) {
        CK(ctx->results.ensure((size_t) nr * sizeof(mb_read_result)));
        CK(cudaMemsetAsync(&ctx->dcnt()->tc_count, 0, sizeof(unsigned long long), ctx->stream));
        CK(cudaMemsetAsync(&ctx->dcnt()->score_ticket, 0, sizeof(u32), ctx->stream));
        CK(cudaMemsetAsync(&ctx->dcnt()->error_flag, 0, sizeof(u32), ctx->stream));
        if (!ctx->segments_valid)
            if (int rc = compute_segments(ctx)) return rc;
        ScoreCfg cfg;
        cfg.format = ctx->par.kmer_format;
        if (ctx->par.syncmer) { cfg.dna_shift = (8 - ctx->par.smer_len) * 3; cfg.max_codon_shift = 8 - ctx->par.smer_len; }
        else { cfg.dna_shift = 3; cfg.max_codon_shift = 1; }
        cfg.denominator = (ctx->par.seq_mode == 1 || ctx->par.seq_mode == 2) ? 100 : 1000;
        cfg.min_cons_cnt = ctx->par.min_cons_cnt; cfg.min_cons_cnt_euk = ctx->par.min_cons_cnt_euk;
        cfg.min_score = ctx->par.min_score; cfg.min_sp_score = ctx->par.min_sp_score; cfg.tie_ratio = ctx->par.tie_ratio;
        TaxView tv;
        tv.parent = ctx->parent.as<int32_t>(); tv.depth = ctx->depth.as<int32_t>(); tv.species_of = ctx->species_of.as<int32_t>();
        tv.species_rank_of = ctx->species_rank_of.as<int32_t>(); tv.max_taxid = ctx->max_taxid; tv.node0_taxid = ctx->node0;
        tv.eukaryota = ctx->euk;
        tv.node = ctx->tax_node.as<u64>();
        ScoreArena arena;
        arena.cap_b = (int) ((ctx->max_query_len + 3) / (u32) cfg.dna_shift + 2);
        arena.cap_c = 4 * arena.cap_b + 64;
        arena.slice_bytes = arena_slice_bytes(arena.cap_b, arena.cap_c);
        // one thread per read; resident threads fill the machine, bounded by a 4-GB arena of per-thread slices
        const u64 budget = 4ull << 30;
        u64 resident = (u64) ctx->sms * 2048;
        if (const char *ev = getenv("MB_SCORE_THREADS_PER_SM")) resident = (u64) ctx->sms * (u64) std::max(128, atoi(ev));   // tuning knob
        u64 threads = std::min<u64>(resident, std::max<u64>(budget / arena.slice_bytes, SC_THREADS));
        threads = std::min<u64>(threads, (u64) nr + SC_THREADS);
        const unsigned grid = (unsigned) std::max<u64>(1, threads / SC_THREADS);
        // short reads: the leftover scorer runs beside k_score_finish on a second stream, with its own slices behind the
        // first `grid` blocks' worth
        const unsigned slow_grid = std::min<unsigned>(grid, (unsigned) ctx->sms * 8);
        CK(ctx->arena.ensure((size_t) (grid + slow_grid) * SC_THREADS * arena.slice_bytes));
        arena.base = ctx->arena.as<unsigned char>();
        const size_t nm = std::max<size_t>(ctx->n_matches, 1);
        // Per-match scratch.  Long reads: every match may carry a path; `combined` (species phase) reuses `local` (paths
        // phase).  Short reads: only the few reads the streaming scorer leaves over use it, at compact offsets — an eighth of
        // the batch is reserved, and the whole batch once that has proved too small (error 6 below).
        const bool long_mode = ctx->par.seq_mode == 3;
        const size_t sc_cap = (long_mode || ctx->score_full_scratch) ? nm : std::min<size_t>(nm, std::max<size_t>(nm / 8, 1u << 20));
        CK(ctx->sc_local.ensure(sc_cap * sizeof(PathRec)));
        CK(ctx->sc_paths.ensure(sc_cap * sizeof(PathRec)));
        if (!long_mode) CK(ctx->sc_comb.ensure(sc_cap * sizeof(PathRec)));
        CK(ctx->sc_linked.ensure(sc_cap));
        if (!long_mode) CK(ctx->sc_spid.ensure(sc_cap * sizeof(int32_t)));
        CK(ctx->sc_spscore.ensure(sc_cap * sizeof(float)));
        arena.local = ctx->sc_local.as<PathRec>(); arena.paths = ctx->sc_paths.as<PathRec>();
        arena.combined = long_mode ? arena.local : ctx->sc_comb.as<PathRec>();
        arena.linked = ctx->sc_linked.as<u8>(); arena.sp_id = ctx->sc_spid.as<int32_t>(); arena.sp_score = ctx->sc_spscore.as<float>();
        ScoreArena arena_slow = arena;
        arena_slow.base = arena.base + (size_t) grid * SC_THREADS * arena.slice_bytes;
        const size_t tc_cap = std::max<size_t>(ctx->n_matches, 16);
        CK(ctx->tc_id.ensure(tc_cap * sizeof(int32_t)));
        CK(ctx->tc_cnt.ensure(tc_cap * sizeof(int32_t)));
        const int paired = ctx->par.seq_mode == 2 ? 1 : 0;
        if (ctx->par.seq_mode != 3) {
            // short reads: streaming scorer first; what it cannot hold goes to the general kernel through slow_list
            CK(ctx->slow_list.ensure((size_t) nr * sizeof(u32)));
            CK(ctx->slow_off.ensure((size_t) nr * sizeof(u64)));
            CK(cudaMemsetAsync(&ctx->dcnt()->slow_matches, 0, sizeof(unsigned long long), ctx->stream));
            CK(ctx->best_recs.ensure((size_t) nr * sizeof(BestRec)));
            CK(cudaMemsetAsync(&ctx->dcnt()->slow_count, 0, 11 * sizeof(u32), ctx->stream));
            const size_t fast_smem = (size_t) SC_THREADS * SC_GSTRIDE * sizeof(u32);
            // buckets of k_score_finish: cap_b taxa + cap_b bytes per thread, odd word stride
            int bucket_stride = arena.cap_b + (arena.cap_b + 3) / 4;
            bucket_stride |= 1;
            const size_t fin_smem = (size_t) SC_THREADS * bucket_stride * sizeof(u32);
            if (fin_smem > 200 * 1024) return fail(ctx, MB_E_ARG, "reads too long for seq-mode 1/2 scoring: use --seq-mode 3");
            if (!ctx->fast_attr_set) {
                CK(cudaFuncSetAttribute(k_score_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) fast_smem));
                CK(cudaFuncSetAttribute(k_score_finish, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                ctx->fast_attr_set = true;
            }
            // persistent: every resident slot (the DP state is 100 B of shared memory per thread; registers set the limit)
            if (!ctx->fast_ctas) {
                int per_sm = 0;
                CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_score_fast, SC_THREADS, fast_smem));
                static const int env_ctas = getenv("MB_SCORE_CTAS_PER_SM") ? atoi(getenv("MB_SCORE_CTAS_PER_SM")) : 0;      // tuning knob
                if (env_ctas > 0 && env_ctas < per_sm) per_sm = env_ctas;
                ctx->fast_ctas = std::max(1, per_sm) * ctx->sms;
            }
            const unsigned fast_grid = (unsigned) std::min<u64>((u64) ctx->fast_ctas, ((u64) nr + SC_THREADS - 1) / SC_THREADS);
            // reads by match count, so that the lanes of a warp run loops of similar length (MB_SCORE_ORDER=0: batch order);
            // measured at configs[1]: k_score_fast 3.13 -> 2.73 ms for 0.05 ms of ordering
            static const bool by_count = !(getenv("MB_SCORE_ORDER") && atoi(getenv("MB_SCORE_ORDER")) == 0);        // tuning knob
            const u32 *order = nullptr;
            if (by_count) {
                CK(ctx->read_order.ensure((size_t) nr * sizeof(u32)));
                CK(ctx->order_hist.ensure(256 * sizeof(u32)));
                CK(cudaMemsetAsync(ctx->order_hist.p, 0, 256 * sizeof(u32), ctx->stream));
                const unsigned og = (unsigned) std::min<u64>(((u64) nr + 255) / 256, (u64) ctx->sms * 8);
                KL(KID_SCORE_ORDER, k_order_count<<<og, 256, 0, ctx->stream>>>(ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), nr, ctx->order_hist.as<u32>()));
                KL(KID_SCORE_ORDER, k_order_scan<<<1, 256, 0, ctx->stream>>>(ctx->order_hist.as<u32>()));
                KL(KID_SCORE_ORDER, k_order_scatter<<<(nr + 255) / 256, 256, 0, ctx->stream>>>(ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), nr,
                                                                                              ctx->order_hist.as<u32>(), ctx->read_order.as<u32>()));
                order = ctx->read_order.as<u32>();
            }
            KL(KID_SCORE_FAST, k_score_fast<<<fast_grid, SC_THREADS, fast_smem, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->lens.as<u32>(), nr, paired, cfg, tv,
                ctx->best_recs.as<BestRec>(), &ctx->dcnt()->score_ticket, ctx->slow_list.as<u32>(), &ctx->dcnt()->slow_count,
                ctx->dcnt()->fast_overflow, ctx->slow_off.as<u64>(), &ctx->dcnt()->slow_matches, order));
            if (!ctx->side_stream) {
                CK(cudaStreamCreateWithFlags(&ctx->side_stream, cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&ctx->side_fork, cudaEventDisableTiming));
                CK(cudaEventCreateWithFlags(&ctx->side_join, cudaEventDisableTiming));
            }
            // the few reads the streaming scorer left over are long single-thread jobs; they can run on a side stream while
            // k_score_finish finishes everybody else (disjoint reads, disjoint scratch, shared atomic counters)
            // (measured on B200: beside k_score_finish the leftover kernel slows from 0.95 to 4.8 ms — its scratch is
            // evicted from L1 by the streaming kernel — so the two are serialised by default; MB_SCORE_SIDE_STREAM=1 overlaps)
            static const bool side = getenv("MB_SCORE_SIDE_STREAM") != nullptr;                // tuning knob
            cudaStream_t slow_stream = side ? ctx->side_stream : ctx->stream;
            CK(cudaEventRecord(ctx->side_fork, ctx->stream));
            CK(cudaStreamWaitEvent(slow_stream, ctx->side_fork, 0));
            KL_ON(KID_SCORE, slow_stream, k_score<<<slow_grid, SC_THREADS, 0, slow_stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->lens.as<u32>(), nr, paired, cfg, tv, arena_slow,
                ctx->results.as<mb_read_result>(), ctx->tc_id.as<int32_t>(), ctx->tc_cnt.as<int32_t>(), (u64) tc_cap, &ctx->dcnt()->tc_count,
                &ctx->dcnt()->slow_ticket, &ctx->dcnt()->error_flag, ctx->slow_list.as<u32>(), &ctx->dcnt()->slow_count,
                ctx->slow_off.as<u64>(), (u64) sc_cap));
            CK(cudaEventRecord(ctx->side_join, slow_stream));
            KL(KID_SCORE_FINISH, k_score_finish<<<grid, SC_THREADS, fin_smem, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->lens.as<u32>(), nr, paired, cfg, tv, arena,
                ctx->best_recs.as<BestRec>(), bucket_stride, ctx->results.as<mb_read_result>(), ctx->tc_id.as<int32_t>(),
                ctx->tc_cnt.as<int32_t>(), (u64) tc_cap, &ctx->dcnt()->tc_count, &ctx->dcnt()->finish_ticket, &ctx->dcnt()->error_flag,
                nullptr));      // batch order: measured, the count order costs this latency-bound kernel 0.19 ms (scattered records)
            CK(cudaStreamWaitEvent(ctx->stream, ctx->side_join, 0));
        } else {
            // long reads / contigs: paths per (read, species, frame) run, species scores per (read, species) run, then one
            // thread per read (mb_score.cuh, k_lr_*)
            if (ctx->n_matches >= 0xFFFFFFF0ull) return fail(ctx, MB_E_ARG, "too many matches in one long-read batch (use smaller batches)");
            CK(ctx->lr_grp_np.ensure(nm * sizeof(int32_t)));
            CK(ctx->lr_grp_end.ensure(nm * sizeof(int32_t)));
            CK(ctx->lr_sp_end.ensure(nm * sizeof(int32_t)));
            CK(ctx->lr_grp_heads.ensure(nm * sizeof(u32)));
            CK(ctx->lr_sp_heads.ensure(nm * sizeof(u32)));
            LongScratch ls;
            ls.local = arena.local; ls.paths = arena.paths; ls.combined = arena.combined; ls.linked = arena.linked;
            ls.grp_np = ctx->lr_grp_np.as<int32_t>(); ls.grp_end = ctx->lr_grp_end.as<int32_t>();
            ls.sp_score = arena.sp_score; ls.sp_end = ctx->lr_sp_end.as<int32_t>();
            ls.grp_heads = ctx->lr_grp_heads.as<u32>(); ls.sp_heads = ctx->lr_sp_heads.as<u32>();
            ls.n_grp = &ctx->dcnt()->lr_n_grp; ls.n_sp = &ctx->dcnt()->lr_n_sp;
            CK(cudaMemsetAsync(&ctx->dcnt()->lr_n_grp, 0, 5 * sizeof(u32), ctx->stream));
            const unsigned wide = (unsigned) ctx->sms * 16;          // 16 CTAs of 128 threads per SM
            if (ctx->n_matches) {
                KL(KID_LR_HEADS, k_lr_heads<<<(unsigned) ((ctx->n_matches + 255) / 256), 256, 0, ctx->stream>>>(
                    ctx->matches_a.as<mb_match>(), ctx->n_matches, ls, 3 * cfg.max_codon_shift));
                KL(KID_LR_PATHS, k_lr_paths<<<wide, SC_THREADS, 0, ctx->stream>>>(
                    ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->seq_base, cfg, tv, ls,
                    &ctx->dcnt()->lr_ticket[0]));
                KL(KID_LR_SPECIES, k_lr_species<<<wide, SC_THREADS, 0, ctx->stream>>>(
                    ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->lens.as<u32>(), ctx->seq_base, cfg,
                    ls, &ctx->dcnt()->lr_ticket[1]));
            }
            // one thread block per read at a time; an arena slice per block
            const unsigned best_grid = (unsigned) std::max<u64>(1, std::min<u64>((u64) grid * SC_THREADS, std::min<u64>((u64) nr, (u64) wide)));
            KL(KID_LR_BEST, k_lr_best<<<best_grid, SC_THREADS, 0, ctx->stream>>>(
                ctx->matches_a.as<mb_match>(), ctx->seg_begin.as<u64>(), ctx->seg_end.as<u64>(), ctx->lens.as<u32>(), nr, cfg, tv, arena, ls,
                ctx->results.as<mb_read_result>(), ctx->tc_id.as<int32_t>(), ctx->tc_cnt.as<int32_t>(), (u64) tc_cap, &ctx->dcnt()->tc_count,
                &ctx->dcnt()->lr_ticket[2], &ctx->dcnt()->error_flag));
        }
        unsigned long long used = 0;
        if (int rc = read_back(ctx, &ctx->dcnt()->tc_count, &used)) return rc;
        if (getenv("MB_DEBUG_SCORE") && ctx->par.seq_mode != 3) {
            u32 why[8];
            CK(cudaMemcpyAsync(why, ctx->dcnt()->fast_overflow, sizeof(why), cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            fprintf(stderr, "[mb] compact scorer left over: group %u, paths/read %u, species/read %u, paths/species %u, field %u, size %u (of %u reads)\n",
                    why[1], why[2], why[3], why[4], why[5], why[6], nr);
        }
        u32 flag = 0;
        if (int rc = read_back(ctx, &ctx->dcnt()->error_flag, &flag)) return rc;
        if (flag == 6 && !ctx->score_full_scratch) {       // the leftover reads hold more matches than the compact scratch: redo with all of it
            ctx->score_full_scratch = true;
            return mb_score_reads(ctx, taxcnt_used);
        }
        if (flag) {
            char msg[128];
            snprintf(msg, sizeof(msg), "scoring kernel reported internal error %u", flag);
            return fail(ctx, MB_E_DATA, msg);
        }
        ctx->tc_used = used;
    }
    mark(ctx, EV_SCORE);
    ctx->have_results = true;
    ctx->res_valid = true; ctx->res_n_reads = ctx->n_reads; ctx->res_tc_used = ctx->tc_used;
    if (taxcnt_used) *taxcnt_used = ctx->tc_used;
    return MB_OK;
}

int mb_fetch_results(mb_ctx *ctx, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n, size_t taxcnt_cap) {
    if (!ctx || (!results && ctx->n_reads)) return MB_E_ARG;
    if (!ctx->have_results) return fail(ctx, MB_E_ARG, "no results resident");
    if (taxcnt_cap < ctx->tc_used) return fail(ctx, MB_E_CAPACITY, "taxcnt arrays too small");
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->n_reads) CK(cudaMemcpyAsync(results, ctx->results.p, (size_t) ctx->n_reads * sizeof(mb_read_result), cudaMemcpyDeviceToHost, ctx->stream));
    if (ctx->tc_used) {
        CK(cudaMemcpyAsync(taxcnt_id, ctx->tc_id.p, ctx->tc_used * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaMemcpyAsync(taxcnt_n, ctx->tc_cnt.p, ctx->tc_used * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    }
    mark(ctx, EV_D2H);
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

// The device->host copy of the results on the copy stream: the caller goes on with the next batch (stage, extract, sort,
// join) while the results travel; mb_fetch_wait before the host arrays are read.
int mb_fetch_results_async(mb_ctx *ctx, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n, size_t taxcnt_cap) {
    if (!ctx || (!results && ctx->res_n_reads)) return MB_E_ARG;
    // the last scored batch: still resident after the next batch has been staged and extracted (only the next scoring
    // overwrites it), so the transfer can be started wherever it disturbs the next batch least
    if (!ctx->res_valid) return fail(ctx, MB_E_ARG, "no results resident");
    if (taxcnt_cap < ctx->res_tc_used) return fail(ctx, MB_E_CAPACITY, "taxcnt arrays too small");
    CK(cudaSetDevice(ctx->par.device));
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->pf_done, cudaEventDisableTiming));
    }
    if (!ctx->fetch_done) {
        CK(cudaEventCreateWithFlags(&ctx->fetch_ready, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&ctx->fetch_done, cudaEventDisableTiming));
    }
    if (ctx->fetch_pending) CK(cudaEventSynchronize(ctx->fetch_done));         // one transfer in flight at a time
    CK(cudaEventRecord(ctx->fetch_ready, ctx->stream));
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->fetch_ready, 0));
    if (ctx->res_n_reads) CK(cudaMemcpyAsync(results, ctx->results.p, (size_t) ctx->res_n_reads * sizeof(mb_read_result), cudaMemcpyDeviceToHost, ctx->copy_stream));
    if (ctx->res_tc_used) {
        CK(cudaMemcpyAsync(taxcnt_id, ctx->tc_id.p, ctx->res_tc_used * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->copy_stream));
        CK(cudaMemcpyAsync(taxcnt_n, ctx->tc_cnt.p, ctx->res_tc_used * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->copy_stream));
    }
    CK(cudaEventRecord(ctx->fetch_done, ctx->copy_stream));
    ctx->fetch_pending = true;
    return MB_OK;
}
int mb_fetch_wait(mb_ctx *ctx) {
    if (!ctx) return MB_E_ARG;
    if (ctx->fetch_pending) {
        CK(cudaSetDevice(ctx->par.device));
        CK(cudaEventSynchronize(ctx->fetch_done));
        ctx->fetch_pending = false;
    }
    return MB_OK;
}

int mb_classify_staged(mb_ctx *ctx, size_t *taxcnt_used) {
    if (int rc = mb_extract_kmers(ctx, nullptr)) return rc;
    if (int rc = mb_sort_kmers(ctx)) return rc;
    if (int rc = mb_match_kmers(ctx, nullptr)) return rc;
    if (int rc = mb_sort_matches(ctx)) return rc;
    return mb_score_reads(ctx, taxcnt_used);
}

int mb_classify_batch(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                      uint32_t n_reads, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n,
                      size_t taxcnt_cap, size_t *taxcnt_used) {
    if (int rc = mb_stage_reads(ctx, bases, n_bases, offsets, lens, n_reads)) return rc;
    size_t used = 0;
    if (int rc = mb_classify_staged(ctx, &used)) return rc;
    if (taxcnt_used) *taxcnt_used = used;
    if (used > taxcnt_cap) return fail(ctx, MB_E_CAPACITY, "taxcnt arrays too small; *taxcnt_used holds the required size");
    return mb_fetch_results(ctx, results, taxcnt_id, taxcnt_n, taxcnt_cap);
}

// ---- taps / exchange ----------------------------------------------------------------------------------
int mb_copy_kmers(mb_ctx *ctx, mb_kmer *out, size_t cap, size_t *n) {
    if (!ctx || !n) return MB_E_ARG;
    if (!ctx->have_kmers) return fail(ctx, MB_E_ARG, "no k-mers resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_K();
    *n = ctx->n_kmers;
    if (cap < ctx->n_kmers) return fail(ctx, MB_E_CAPACITY, "k-mer output too small");
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->n_kmers) CK(cudaMemcpyAsync(out, ctx->kmers_a.p, ctx->n_kmers * sizeof(mb_kmer), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

int mb_copy_matches(mb_ctx *ctx, mb_match *out, size_t cap, size_t *n) {
    if (!ctx || !n) return MB_E_ARG;
    if (!ctx->have_matches) return fail(ctx, MB_E_ARG, "no matches resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_M();
    *n = ctx->n_matches;
    if (cap < ctx->n_matches) return fail(ctx, MB_E_CAPACITY, "match output too small");
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->n_matches) CK(cudaMemcpyAsync(out, ctx->matches_a.p, ctx->n_matches * sizeof(mb_match), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}

static int set_kmers(mb_ctx *ctx, const void *src, size_t n, cudaMemcpyKind kind) {
    CK(cudaSetDevice(ctx->par.device));
    CK(ctx->kmers_a.ensure((n + 1) * sizeof(mb_kmer)));
    if (n) CK(cudaMemcpyAsync(ctx->kmers_a.p, src, n * sizeof(mb_kmer), kind, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_kmers = n;
    ctx->kmers_seg.n = 0;
    ctx->have_kmers = true;
    ctx->kmers_sorted = false;
    ctx->have_matches = ctx->have_results = false;
    return MB_OK;
}
static int set_matches(mb_ctx *ctx, const void *src, size_t n, cudaMemcpyKind kind) {
    CK(cudaSetDevice(ctx->par.device));
    CK(ctx->matches_a.ensure((n + 1) * sizeof(mb_match)));
    if (n) CK(cudaMemcpyAsync(ctx->matches_a.p, src, n * sizeof(mb_match), kind, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->n_matches = n;
    ctx->matches_seg.n = 0;
    ctx->have_matches = true;
    ctx->matches_sorted = false;
    ctx->segments_valid = false;
    ctx->have_results = false;
    return MB_OK;
}
int mb_set_kmers(mb_ctx *ctx, const mb_kmer *k, size_t n) { return (!ctx || (!k && n)) ? MB_E_ARG : set_kmers(ctx, k, n, cudaMemcpyHostToDevice); }
int mb_set_matches(mb_ctx *ctx, const mb_match *m, size_t n) { return (!ctx || (!m && n)) ? MB_E_ARG : set_matches(ctx, m, n, cudaMemcpyHostToDevice); }
int mb_set_kmers_device(mb_ctx *ctx, const void *d, size_t n) { return (!ctx || (!d && n)) ? MB_E_ARG : set_kmers(ctx, d, n, cudaMemcpyDeviceToDevice); }
int mb_set_matches_device(mb_ctx *ctx, const void *d, size_t n) { return (!ctx || (!d && n)) ? MB_E_ARG : set_matches(ctx, d, n, cudaMemcpyDeviceToDevice); }

int mb_export_kmers_device(mb_ctx *ctx, void *dst, size_t cap, size_t *n) {
    if (!ctx || !n) return MB_E_ARG;
    if (!ctx->have_kmers) return fail(ctx, MB_E_ARG, "no k-mers resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_K();
    *n = ctx->n_kmers;
    if (cap < ctx->n_kmers) return fail(ctx, MB_E_CAPACITY, "k-mer output too small");
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->n_kmers) CK(cudaMemcpyAsync(dst, ctx->kmers_a.p, ctx->n_kmers * sizeof(mb_kmer), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}
int mb_export_matches_device(mb_ctx *ctx, void *dst, size_t cap, size_t *n) {
    if (!ctx || !n) return MB_E_ARG;
    if (!ctx->have_matches) return fail(ctx, MB_E_ARG, "no matches resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_M();
    *n = ctx->n_matches;
    if (cap < ctx->n_matches) return fail(ctx, MB_E_CAPACITY, "match output too small");
    CK(cudaSetDevice(ctx->par.device));
    if (ctx->n_matches) CK(cudaMemcpyAsync(dst, ctx->matches_a.p, ctx->n_matches * sizeof(mb_match), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return MB_OK;
}
int mb_set_read_id_base(mb_ctx *ctx, uint32_t base) {
    if (!ctx) return MB_E_ARG;
    if ((uint64_t) base + ctx->n_reads >= (1u << 29)) return fail(ctx, MB_E_ARG, "read id base + batch size exceeds the 29-bit sequenceID");
    ctx->seq_base = base;
    return MB_OK;
}

int mb_kmers_device(mb_ctx *ctx, void **dptr, size_t *n) {
    if (!ctx || !dptr || !n) return MB_E_ARG;
    if (!ctx->have_kmers) return fail(ctx, MB_E_ARG, "no k-mers resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_K();
    CK(cudaStreamSynchronize(ctx->stream));
    *dptr = ctx->kmers_a.p; *n = ctx->n_kmers;
    return MB_OK;
}
int mb_matches_device(mb_ctx *ctx, void **dptr, size_t *n) {
    if (!ctx || !dptr || !n) return MB_E_ARG;
    if (!ctx->have_matches) return fail(ctx, MB_E_ARG, "no matches resident");
    CK(cudaSetDevice(ctx->par.device));
    MATERIALIZE_M();
    CK(cudaStreamSynchronize(ctx->stream));
    *dptr = ctx->matches_a.p; *n = ctx->n_matches;
    return MB_OK;
}

int mb_partition_kmers(mb_ctx *ctx, const uint64_t *bounds, int n_shards, uint64_t *counts) {
    if (!ctx || !counts || n_shards < 1 || n_shards > 9 || (!bounds && n_shards > 1)) return MB_E_ARG;
    if (!ctx->have_kmers) return fail(ctx, MB_E_ARG, "no k-mers resident");
    CK(cudaSetDevice(ctx->par.device));
    KmerOwnerKey key;
    key.n_bounds = n_shards - 1;
    for (int i = 0; i < 8; ++i) key.bounds[i] = i < n_shards - 1 ? bounds[i] : ~0ull;
    return partition_by<mb_kmer, KmerOwnerKey, 16>(ctx, ctx->kmers_a, ctx->kmers_b, ctx->n_kmers, key, n_shards, counts, &ctx->kmers_seg);
}

int mb_partition_matches(mb_ctx *ctx, uint32_t reads_per_rank, int n_ranks, uint64_t *counts) {
    if (!ctx || !counts || n_ranks < 1 || n_ranks > 255 || reads_per_rank == 0) return MB_E_ARG;
    if (!ctx->have_matches) return fail(ctx, MB_E_ARG, "no matches resident");
    CK(cudaSetDevice(ctx->par.device));
    MatchHomeKey key;
    key.reads_per_rank = reads_per_rank;
    return partition_by<mb_match, MatchHomeKey, 8>(ctx, ctx->matches_a, ctx->matches_b, ctx->n_matches, key, n_ranks, counts, &ctx->matches_seg);
}

// ---- multi-GPU: peer-routed producers --------------------------------------------------------------------
static cudaError_t alloc_fixed(DevBuf &b, size_t bytes) {
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) { b.p = nullptr; return e; }
    b.cap = bytes;
    return cudaSuccess;
}

int mb_peer_alloc(mb_ctx *ctx, size_t kmer_cap, size_t match_cap, mb_peer_mem *out) {
    if (!ctx || !out || !kmer_cap || !match_cap) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->routed = false;
    ctx->kmers_seg.n = 0; ctx->matches_seg.n = 0;
    CK(alloc_fixed(ctx->peer_kmers, kmer_cap * sizeof(mb_kmer)));
    CK(alloc_fixed(ctx->peer_matches, match_cap * sizeof(mb_match)));
    CK(alloc_fixed(ctx->peer_cnt, 2 * MB_MAX_RANKS * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(ctx->peer_cnt.p, 0, ctx->peer_cnt.cap, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->peer_kmer_cap = kmer_cap; ctx->peer_match_cap = match_cap;
    out->kmers = ctx->peer_kmers.p; out->matches = ctx->peer_matches.p; out->counters = ctx->peer_cnt.p;
    out->kmer_cap = kmer_cap; out->match_cap = match_cap;
    return MB_OK;
}

int mb_ipc_export(mb_ctx *ctx, const void *dptr, unsigned char handle[64]) {
    if (!ctx || !dptr || !handle) return MB_E_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    CK(cudaSetDevice(ctx->par.device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, const_cast<void *>(dptr)));
    memcpy(handle, &h, 64);
    return MB_OK;
}
int mb_ipc_open(mb_ctx *ctx, const unsigned char handle[64], void **dptr) {
    if (!ctx || !dptr || !handle) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    CK(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return MB_OK;
}
int mb_ipc_close(mb_ctx *ctx, void *dptr) {
    if (!ctx || !dptr) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaIpcCloseMemHandle(dptr));
    return MB_OK;
}

int mb_enable_peer_access(mb_ctx *ctx, int peer_device) {
    if (!ctx) return MB_E_ARG;
    if (peer_device == ctx->par.device) return MB_OK;
    CK(cudaSetDevice(ctx->par.device));
    int can = 0;
    CK(cudaDeviceCanAccessPeer(&can, ctx->par.device, peer_device));
    if (!can) return fail(ctx, MB_E_ARG, "the GPUs cannot access each other's memory (no NVLink / PCIe peer path)");
    const cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return MB_OK; }
    CK(e);
    return MB_OK;
}

// Receive-buffer geometry: rank d's buffer of `cap` records is cut into n_ranks equal regions, region s is written by
// source rank s only, so a source reserves slots with its own local counters and no remote atomic is ever needed.
int mb_peer_attach(mb_ctx *ctx, uint32_t rank, uint32_t n_ranks, const mb_peer_mem *peers, const uint64_t *bounds,
                   uint32_t reads_per_rank) {
    if (!ctx || !peers || n_ranks < 1 || n_ranks > MB_MAX_RANKS || rank >= n_ranks || (!bounds && n_ranks > 1) || !reads_per_rank)
        return MB_E_ARG;
    if (!ctx->peer_kmers.p || peers[rank].kmers != ctx->peer_kmers.p || peers[rank].matches != ctx->peer_matches.p ||
        peers[rank].counters != ctx->peer_cnt.p)
        return fail(ctx, MB_E_ARG, "peers[rank] must be this context's own mb_peer_alloc result");
    CK(cudaSetDevice(ctx->par.device));
    CK(ctx->sent_cnt.ensure(2 * MB_MAX_RANKS * sizeof(unsigned long long)));
    Route &r = ctx->route;
    memset(&r, 0, sizeof(r));
    r.n = n_ranks; r.rank = rank; r.reads_per_rank = reads_per_rank; r.inv_reads_per_rank = 1.0f / (float) reads_per_rank;
    for (uint32_t i = 0; i < MB_MAX_RANKS; ++i) r.bounds[i] = (i + 1 < n_ranks) ? bounds[i] : ~0ull;
    for (uint32_t i = 0; i < n_ranks; ++i) {
        if (!peers[i].kmers || !peers[i].matches || !peers[i].counters) return fail(ctx, MB_E_ARG, "peer receive side missing");
        const u64 kreg = peers[i].kmer_cap / n_ranks, mreg = peers[i].match_cap / n_ranks;
        r.kmers[i] = reinterpret_cast<mb_kmer *>(peers[i].kmers) + (u64) rank * kreg;
        r.matches[i] = reinterpret_cast<mb_match *>(peers[i].matches) + (u64) rank * mreg;
        r.kmer_cap[i] = kreg; r.match_cap[i] = mreg;
        ctx->peer_counters[i] = reinterpret_cast<unsigned long long *>(peers[i].counters);
    }
    r.sent_k = ctx->sent_cnt.as<unsigned long long>();
    r.sent_m = r.sent_k + MB_MAX_RANKS;
    CK(ctx->route_dev.ensure(sizeof(Route)));
    CK(cudaMemcpyAsync(ctx->route_dev.p, &r, sizeof(Route), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->routed = true;
    return MB_OK;
}

// after a routed producer kernel: tell every destination how many records this rank put into its region there
// (`which`: 0 k-mers, 1 matches); stream ordered behind the kernel, 8-byte peer copies
static int publish_sent(mb_ctx *ctx, int which) {
    const Route &r = ctx->route;
    const unsigned long long *sent = which ? r.sent_m : r.sent_k;
    for (u32 d = 0; d < r.n; ++d)
        CK(cudaMemcpyAsync(ctx->peer_counters[d] + which * MB_MAX_RANKS + r.rank, sent + d, sizeof(unsigned long long), cudaMemcpyDefault,
                           ctx->stream));
    return MB_OK;
}

struct Arrived { unsigned long long v[MB_MAX_RANKS]; };

// read this rank's arrival counters `which` (0 k-mers, 1 matches), clear them for the next batch and describe the
// filled part of the regions; *total = records that arrived, or (overflow) n_ranks x the fullest region's demand
static int take_regions(mb_ctx *ctx, int which, size_t rec_bytes, void *buf, u64 cap, SegIn &seg, unsigned long long *total, bool *over) {
    Arrived a;
    Arrived *dev = reinterpret_cast<Arrived *>(ctx->peer_cnt.as<unsigned long long>() + which * MB_MAX_RANKS);
    if (int rc = read_back(ctx, dev, &a)) return rc;
    CK(cudaMemsetAsync(dev, 0, sizeof(Arrived), ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const u32 n = ctx->route.n;
    const u64 region = cap / n;
    unsigned long long sum = 0, worst = 0;
    seg.n = n;
    for (u32 s = 0; s < n; ++s) {
        seg.prefix[s] = sum;
        seg.base[s] = reinterpret_cast<const unsigned char *>(buf) + (size_t) s * region * rec_bytes;
        sum += a.v[s];
        worst = std::max(worst, a.v[s]);
    }
    seg.prefix[n] = sum;
    *over = worst > region;
    *total = *over ? worst * n : sum;
    if (*over) seg.n = 0;
    return MB_OK;
}

int mb_take_routed_kmers(mb_ctx *ctx, size_t *n) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->routed) return fail(ctx, MB_E_ARG, "mb_peer_attach first");
    CK(cudaSetDevice(ctx->par.device));
    unsigned long long got = 0;
    bool over = false;
    ctx->have_kmers = ctx->kmers_sorted = ctx->have_matches = ctx->have_results = false;
    if (int rc = take_regions(ctx, 0, sizeof(mb_kmer), ctx->peer_kmers.p, ctx->peer_kmer_cap, ctx->kmers_seg, &got, &over)) return rc;
    if (n) *n = (size_t) got;
    if (over) return fail(ctx, MB_E_CAPACITY, "more k-mers arrived from one rank than its region of the receive buffer holds");
    ctx->n_kmers = got;
    if (got == 0) ctx->kmers_seg.n = 0;
    ctx->have_kmers = true;
    return MB_OK;
}

int mb_take_routed_matches(mb_ctx *ctx, size_t *n) {
    if (!ctx) return MB_E_ARG;
    if (!ctx->routed) return fail(ctx, MB_E_ARG, "mb_peer_attach first");
    CK(cudaSetDevice(ctx->par.device));
    unsigned long long got = 0;
    bool over = false;
    ctx->have_matches = ctx->matches_sorted = ctx->have_results = false;
    ctx->segments_valid = false;
    if (int rc = take_regions(ctx, 1, sizeof(mb_match), ctx->peer_matches.p, ctx->peer_match_cap, ctx->matches_seg, &got, &over)) return rc;
    if (n) *n = (size_t) got;
    if (over) return fail(ctx, MB_E_CAPACITY, "more matches arrived from one rank than its region of the receive buffer holds");
    ctx->n_matches = got;
    if (got == 0) ctx->matches_seg.n = 0;
    ctx->have_matches = true;
    return MB_OK;
}

int mb_get_timings(mb_ctx *ctx, mb_phase_times *t) {
    if (!ctx || !t) return MB_E_ARG;
    CK(cudaSetDevice(ctx->par.device));
    CK(cudaStreamSynchronize(ctx->stream));
    mb_phase_times &o = ctx->times;
    auto span = [&](int a, int b) -> float {
        if (!ctx->ev_set[a] || !ctx->ev_set[b]) return 0.f;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->ev[a], ctx->ev[b]) != cudaSuccess) { cudaGetLastError(); return 0.f; }
        return ms;
    };
    o.h2d_ms = span(EV_START, EV_H2D);
    o.extract_ms = span(EV_H2D, EV_EXTRACT);
    o.sort_kmers_ms = span(EV_EXTRACT, EV_SORTK);
    o.match_ms = span(EV_SORTK, EV_MATCH);
    o.sort_matches_ms = span(EV_MATCH, EV_SORTM);
    o.score_ms = span(EV_SORTM, EV_SCORE);
    o.d2h_ms = span(EV_SCORE, EV_D2H);
    o.total_ms = span(EV_START, ctx->ev_set[EV_D2H] ? EV_D2H : EV_SCORE);
    o.n_bases = ctx->n_bases; o.n_kmers = ctx->n_kmers; o.n_matches = ctx->n_matches; o.n_reads = ctx->n_reads;
    o.db_words = ctx->n_words; o.db_kmers = ctx->n_db_kmers;
    o.kernel_launches = ctx->launches;
    *t = o;
    return MB_OK;
}

int mb_set_profile(mb_ctx *ctx, int on) {
    if (!ctx) return MB_E_ARG;
    ctx->profile = on != 0;
    ctx->n_samples = 0;
    return MB_OK;
}

int mb_get_kernel_times(mb_ctx *ctx, mb_kernel_time *out, size_t cap, size_t *n) {
    if (!ctx || !n || (!out && cap)) return MB_E_ARG;
    *n = KID_COUNT;
    if (cap < (size_t) KID_COUNT) return fail(ctx, MB_E_CAPACITY, "kernel time table needs KID_COUNT entries");
    CK(cudaSetDevice(ctx->par.device));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < KID_COUNT; ++k) {
        memset(&out[k], 0, sizeof(out[k]));
        strncpy(out[k].name, KID_NAMES[k], sizeof(out[k].name) - 1);
    }
    for (size_t i = 0; i < ctx->n_samples; ++i) {
        const KSample &s = ctx->samples[i];
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.a, s.b) != cudaSuccess) { cudaGetLastError(); continue; }
        out[s.kid].ms += ms;
        out[s.kid].launches += 1;
    }
    return MB_OK;
}

} // extern "C"

// ---- allocation-free query k-mer sort (mb_sort_kmers_noalloc) ---------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and the same key as
// mb_sort_kmers, but the scratch of the one-sweep radix sort — the digit histograms of every pass, the per-tile look-back
// states and the tile ticket — lives in the file-scope device array below instead of cudaMalloc'ed buffers.  Every GPU that
// loads the library holds the array until exit.  All calls (of every context on a device) share it and it keeps the last
// call's contents: the parts that must start at zero are cleared on the stream before use, and two calls that overlap
// on one GPU race on it.
namespace {
constexpr int NA_ITEMS2 = 8;                                              // records per thread of k_radix_pass2, as mb_sort_kmers
constexpr u64 NA_TILE = (u64) R2_THREADS * NA_ITEMS2;
constexpr u64 NA_MAX_TILES = (MB_NOALLOC_MAX_KMERS + NA_TILE - 1) / NA_TILE;
struct NoallocScratch {
    u64 hist[RS_MAX_PASSES * RS_BINS];                                    // 32 KB
    u64 tile_state[NA_MAX_TILES * RS_BINS];                               // 2 MB at 4 Mi records
    u32 trivial[RS_MAX_PASSES];
    u32 ticket;
};
}
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoallocScratch)];

namespace {
constexpr int NA_MAX_DEVICES = 64;
struct NoallocDev { bool ready = false; NoallocScratch *scratch = nullptr; };
NoallocDev g_noalloc_dev[NA_MAX_DEVICES];

// once per device, with that device current: the DEVICE address of the array (in host code its name is only a host-side
// stand-in) and the shared-memory opt-in of the pass kernels (function attributes are per device)
int noalloc_setup(mb_ctx *ctx, NoallocScratch **out) {
    const int dev = ctx->par.device;
    if (dev < 0 || dev >= NA_MAX_DEVICES) return fail(ctx, MB_E_ARG, "device ordinal beyond the no-alloc set-up table");
    NoallocDev &d = g_noalloc_dev[dev];
    if (!d.ready) {
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        constexpr int smem = (int) radix_pass2_smem<mb_kmer, NA_ITEMS2>();
        CK(cudaFuncSetAttribute(k_radix_pass2<mb_kmer, KmerKey, NA_ITEMS2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CK(cudaFuncSetAttribute(k_radix_pass2<mb_kmer, KmerKeyPacked, NA_ITEMS2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CK(cudaFuncSetAttribute(k_radix_pass2<mb_kmer, KmerKeyPacked7, NA_ITEMS2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        d.scratch = reinterpret_cast<NoallocScratch *>(p);
        d.ready = true;
    }
    *out = d.scratch;
    return MB_OK;
}

template <typename Key>
int noalloc_sort(mb_ctx *ctx, cudaStream_t st, NoallocScratch *sc, mb_kmer *a, mb_kmer *b, u64 n, const Key &key, int n_passes) {
    const u64 tiles = (n + NA_TILE - 1) / NA_TILE;
    constexpr size_t smem = radix_pass2_smem<mb_kmer, NA_ITEMS2>();
    const SegIn seg = {};
    CK(cudaMemsetAsync(sc->hist, 0, sizeof(u64) * RS_BINS * n_passes, st));
    CK(cudaMemsetAsync(sc->trivial, 0, sizeof(sc->trivial), st));
    const u64 hist_ctas = std::min<u64>((n + RS_THREADS - 1) / RS_THREADS, (u64) ctx->sms * 8);
    k_radix_hist<mb_kmer, Key, false><<<(unsigned) hist_ctas, RS_THREADS, n_passes * RS_BINS * sizeof(u32), st>>>(a, n, key, n_passes, sc->hist, seg);
    CK(cudaGetLastError());
    k_radix_scan_hist<<<n_passes, RS_BINS, 0, st>>>(sc->hist, n, sc->trivial);
    CK(cudaGetLastError());
    for (int p = 0; p < n_passes; ++p) {
        CK(cudaMemsetAsync(sc->tile_state, 0, tiles * RS_BINS * sizeof(u64), st));
        CK(cudaMemsetAsync(&sc->ticket, 0, sizeof(u32), st));
        k_radix_pass2<mb_kmer, Key, NA_ITEMS2, false><<<(unsigned) tiles, R2_THREADS, smem, st>>>(a, b, n, key, p, sc->hist + (size_t) p * RS_BINS,
                                                                                                sc->tile_state, &sc->ticket, seg);
        CK(cudaGetLastError());
        std::swap(a, b);
    }
    // after an odd number of passes the result sits in the caller's tmp buffer (now `a`): bring it home (now `b`)
    if (n_passes & 1) CK(cudaMemcpyAsync(b, a, n * sizeof(mb_kmer), cudaMemcpyDeviceToDevice, st));
    return MB_OK;
}
} // namespace

extern "C" int mb_sort_kmers_noalloc(mb_ctx *ctx, void *d_kmers, void *d_tmp, size_t n, void *stream) {
    if (!ctx) return MB_E_ARG;
    if (n > (size_t) MB_NOALLOC_MAX_KMERS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "mb_sort_kmers_noalloc: %zu k-mers exceed the fixed scratch (MB_NOALLOC_MAX_KMERS = %u); use mb_sort_kmers",
                 n, (unsigned) MB_NOALLOC_MAX_KMERS);
        return fail(ctx, MB_E_CAPACITY, msg);
    }
    if (n == 0) return MB_OK;
    if (!d_kmers || !d_tmp || d_kmers == d_tmp || ((uintptr_t) d_kmers & 15) || ((uintptr_t) d_tmp & 15))
        return fail(ctx, MB_E_ARG, "mb_sort_kmers_noalloc: d_kmers / d_tmp must be two distinct 16-byte aligned device buffers of n records");
    CK(cudaSetDevice(ctx->par.device));
    NoallocScratch *sc = nullptr;
    if (int rc = noalloc_setup(ctx, &sc)) return rc;
    cudaStream_t st = stream ? (cudaStream_t) stream : ctx->stream;
    // the key of mb_sort_kmers (kmer_sort_bits, database size; the MB_KSORT_BITS knob is honoured there and here alike)
    const SortT plan = planned_sort_t(ctx);
    const int bits = plan.mode ? 0 : 40 - plan.shift;
    if (!plan.mode && bits != 24 && bits != 32 && bits != 40) return fail(ctx, MB_E_ARG, "MB_KSORT_BITS must be 21, 31, 24, 32 or 40");
    mb_kmer *a = reinterpret_cast<mb_kmer *>(d_kmers), *b = reinterpret_cast<mb_kmer *>(d_tmp);
    int rc;
    if (plan.mode == 1) {
        KmerKeyPacked pk;
        rc = noalloc_sort(ctx, st, sc, a, b, (u64) n, pk, 3);
    } else if (plan.mode == 2) {
        KmerKeyPacked7 pk;
        rc = noalloc_sort(ctx, st, sc, a, b, (u64) n, pk, 4);
    } else {
        KmerKey key;
        key.first_bit = 24 + plan.shift;
        rc = noalloc_sort(ctx, st, sc, a, b, (u64) n, key, bits / 8);
    }
    if (rc) return rc;
    if (!stream) CK(cudaStreamSynchronize(st));
    return MB_OK;
}
