// mb_match.cuh — K3+K4: unit-parallel diffIdx decode feeding the amino-acid join with the DNA Hamming filter and
// match compaction.
//
// Replaces KmerMatcher::matchKmers (src/commons/KmerMatcher.cpp:123-481) with getNextTargetKmer
// (KmerMatcher.h:282-297) and compareDna (KmerMatcher.cpp:1117-1146).
//
// What the reference computes, stated without its thread ranges: a query k-mer q is compared with
// ALL DB k-mers whose amino-acid part (value >> 24) equals q's — except the very last k-mer of the
// DB, which the `diffIdxPos != numOfDiffIdx` guards never offer (KmerMatcher.cpp:363-406) — and
// every candidate whose codon-id Hamming sum is <= min(2*min_sum, 7) yields a Match.  The result
// per query depends on nothing else, so the join parallelises over pieces of the DB.
//
// B200 design (round 2: warp-autonomous units).  The compressed stream stays in HBM as it is on disk.  The query sort
// orders a monotone function T of the amino-acid part (SortT, mb_common.cuh), so a T BUCKET (all k-mers with one T value)
// is contiguous on both sides.  At load time the stream is cut into UNITS at T-bucket starts, one cut per 512-word raw
// block (the first bucket start inside the block): a unit is a whole number of T buckets — hence of amino-acid groups —
// of about 512 words / 240 k-mers, described by its decode cursor (first word, value before, k-mer ordinal) and its first
// T value.  Per batch one binary search per unit (k_match_partition) finds where the unit's T range starts in the sorted
// query array; the query slices of consecutive units are disjoint and exact, so no query is looked up twice.
//
// k_match is persistent and every WARP works on its own: it draws a unit ticket (one ahead), and for that unit
//   1. waits for the unit's words, which ONE lane fetched into the warp's shared-memory buffer with a bulk asynchronous
//      copy (cp.async.bulk + mbarrier) while the warp was still joining the previous unit (double buffer);
//   2. decodes them: every lane assembles the deltas ending in its 20 words (two sweeps over registers loaded with five
//      64-bit shared loads: count/sum, warp scan of (count, 64-bit sum), then absolute values), marks the amino-acid group
//      heads in a bit mask;
//   3. joins the unit's query slice 32 queries at a time: branch-free binary search of the query's amino-acid key in
//      the decoded values (<= 9 steps), group end from the head bit mask, then ALL (query, candidate) pairs of the 32
//      queries are spread evenly over the lanes (prefix sum of the group sizes), one Hamming evaluation per pair through a
//      two-codon table in shared memory, per-query minimum by shared atomicMin, filter d <= min(2*min, 7);
//   4. stages the surviving matches in the warp's buffer and flushes it with one atomic reservation and coalesced
//      8-byte stores.
// There is no block-wide barrier anywhere in the loop.  Units longer than the buffers (no bucket start for a while, or
// many one-word deltas) are joined in several pieces, each ending on a group boundary; a single amino-acid group longer
// than a piece takes a streaming two-pass path.  Capacity misses are counted, never written out of bounds; the host
// grows the buffer and reruns (the reference's retry, Classifier.cpp:127-131).
#pragma once
#include "mb_common.cuh"
#include <cstddef>

namespace mb {

#ifndef MB_UN_W
#define MB_UN_W 448
#endif
constexpr int UN_W = MB_UN_W;               // raw block of the load-time index: at most one unit starts per raw block (multiple of 32)
constexpr int MT_THREADS = 256;
constexpr int MT_WARPS = MT_THREADS / 32;
constexpr int MT_BUF_W = 640;              // words of one word buffer (32 lanes x 20 words)
constexpr int MT_R = MT_BUF_W / 32;        // words per lane
constexpr int MT_CAP_W = MT_BUF_W - 8;     // words of one piece (the buffer starts at an 8-word boundary of the stream)
constexpr int MT_CAP_K = 256;              // k-mers of one piece
constexpr int MT_HB = MT_CAP_K / 32 + 1;   // head-bit words (one extra bit: the sentinel behind the last k-mer)
#ifndef MB_MT_STAGE
#define MB_MT_STAGE 64
#endif
#ifndef MB_MT_MIN_CTAS
#define MB_MT_MIN_CTAS 4
#endif
#ifndef MB_MT_INFO_PF
#define MB_MT_INFO_PF 1
#endif
constexpr int MT_STAGE = MB_MT_STAGE;      // matches staged per warp before a flush
constexpr int DB_PAD = MT_BUF_W + 64;      // zero words the device stream is padded with

static_assert(MT_R * 2 % 8 == 0, "a lane's run must be whole 64-bit words");

// One unit of the load-time index (see above).  units[n_units] is a sentinel {n_words, last value, n_kmers, ~0}.
struct Unit {
    u64 word;            // first word of the unit's first k-mer
    u64 vbefore;         // value of the preceding k-mer (0 at the stream start)
    u64 kidx;            // ordinal of the unit's first k-mer
    u64 t_lo;            // T of its first k-mer (sort_t)
};

struct DbView {
    const u16 *words;        // padded with DB_PAD zero words
    u64 n_words;
    const u32 *info;
    u64 n_kmers_eff;         // candidates are ordinals < n_kmers_eff (drops the final k-mer on the last shard)
    u32 info_mask;           // KmerMatcher.cpp:204-205
    const int32_t *species_of;
    int32_t max_taxid;
};

struct DbCursor { u64 word, vbefore, kidx; };

// per-warp shared memory
constexpr int MT_KEYS = 256;               // key array: MT_CAP_K entries padded with 0xFFFFFFFF to a power of two (bound-free search)
static_assert(MT_KEYS >= MT_CAP_K && (MT_KEYS & (MT_KEYS - 1)) == 0, "key array: power of two that holds a piece");
struct __align__(128) WarpSmem {
    u16 w[MT_BUF_W];                       // word buffer (bulk-copy destination); free again as soon as a piece is decoded
    u32 key[MT_KEYS];                      // amino-acid part of entry j minus aa0 (entries >= n: 0xFFFFFFFF)
    u32 dna[MT_CAP_K];                     // low 24 bits (DNA part) of entry j; directly behind `key` (decode_piece relies on it)
    u64 stage[MT_STAGE * 3];               // staged match records
    u32 hbits[MT_HB + 3];                  // bit j: k-mer j starts an amino-acid group; bit n: sentinel
    u32 qmin[32];                          // per query of the current 32-query slice: smallest distance in its group
    u64 mbar;                              // mbarrier of the word buffer
    u64 aa0;                               // amino-acid part of the piece's first entry
};
static_assert(offsetof(WarpSmem, dna) == offsetof(WarpSmem, key) + MT_KEYS * sizeof(u32), "dna follows key");
// value of entry j of the decoded piece
__device__ __forceinline__ u64 piece_value(const WarpSmem &ws, u32 j) { return ((ws.aa0 + ws.key[j]) << 24) | ws.dna[j]; }

// the part of the shared memory that the warps share (read only after set-up)
struct MatchTables {
    u8 lut2[4096];
    u8 lutp[4096];
};

constexpr size_t match_smem_bytes() { return sizeof(MatchTables) + MT_WARPS * sizeof(WarpSmem); }

// (prefetch_l2, mbarrier and bulk-copy helpers: mb_common.cuh)

// One lane asks for the words of a piece that starts at stream word `word` and has `wl` words (<= MT_CAP_W): the copy
// starts at the 8-word boundary below `word`.  Returns nothing; the warp later waits on the buffer's mbarrier.
__device__ __forceinline__ void fetch_words(const DbView &db, WarpSmem &ws, u64 word, u32 wl) {
    const u64 wbase = word & ~7ull;
    const u32 pad = (u32) (word - wbase);
    const u32 bytes = ((pad + wl + 7u) >> 3) << 4;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // the lanes' reads of the buffer precede the asynchronous write
    mbar_expect_tx(&ws.mbar, bytes);
    bulk_g2s(ws.w, db.words + wbase, bytes, &ws.mbar);
}

// ---- warp-level decode of one piece -----------------------------------------------------------------------------
// The piece's words lie in buf[pad, pad + wl) (buffer coordinates).  Decodes the deltas that END inside the piece
// (getNextTargetKmer, KmerMatcher.h:282-297).  Of the first min(n_total, MT_CAP_K) k-mers it leaves the amino-acid part
// relative to the first one's (ws.aa0) in ws.key, the DNA part in ws.dna and the amino-acid group heads in ws.hbits (bit 0
// is always set: a piece starts at a group start); ws.key is padded with 0xFFFFFFFF up to MT_KEYS.  *n_fit = entries whose
// relative amino-acid part fits 32 bits (the piece is cut before the first one that does not).  Returns n_total.
// Warp-synchronous; every lane must call it.
__device__ __forceinline__ u32 decode_piece(WarpSmem &ws, const u16 *buf, u32 pad, u32 wl, u64 vbefore, u32 *n_fit) {
    const u32 lane = threadIdx.x & 31;
    const u32 l0 = lane * MT_R, end = pad + wl;
    // my 20 words in five 64-bit registers; words outside the piece become zero words (no terminator, no payload), which
    // leaves the sweeps free of range checks: leading zeros keep the empty delta empty, trailing zeros only shift a delta
    // that nothing finishes
    u64 r[MT_R / 4];
    const u64 *b64 = reinterpret_cast<const u64 *>(buf + l0);
#pragma unroll
    for (int i = 0; i < MT_R / 4; ++i) r[i] = b64[i];
    if (l0 < pad || l0 + MT_R > end) {
#pragma unroll
        for (int i = 0; i < MT_R; ++i) {
            const u32 idx = l0 + i;
            if (idx < pad || idx >= end) r[i >> 2] &= ~(0xFFFFull << (16 * (i & 3)));
        }
    }
#pragma unroll
    for (int i = 0; i < MT_KEYS / 32; ++i) ws.key[lane + 32 * i] = 0xFFFFFFFFu;
    // words of an unfinished delta in front of my run (at most 4 for a valid stream), not before the piece start
    u64 carry = 0;
    if (l0 > pad && l0 < end) {
        u32 st = l0;
        while (st > pad && l0 - st < 4 && !(buf[st - 1] & 0x8000u)) --st;
        for (u32 k = st; k < l0; ++k) carry = (carry | buf[k]) << 15;
    }
    // the piece's first k-mer ends within five words of its start, inside lane 0's run: its amino-acid part is the base of the keys
    u64 aa0 = 0;
    if (lane == 0) {
        u64 d = 0;
        for (u32 k = pad; k < end && k < pad + 5; ++k) {
            const u32 frag = buf[k];
            d = (d << 15) | (frag & 0x7FFFu);
            if (frag & 0x8000u) break;
        }
        aa0 = (vbefore + d) >> 24;
    }
    aa0 = __shfl_sync(0xFFFFFFFFu, aa0, 0);
    // sweep 1: count and sum the deltas terminating in my run (branch-free)
    u32 cnt = 0;
    u64 sum = 0, diff = carry >> 15;                     // `carry` is kept pre-shifted; the sweeps shift on entry
#pragma unroll
    for (int i = 0; i < MT_R; ++i) {
        const u32 frag = (u32) (r[i >> 2] >> (16 * (i & 3))) & 0xFFFFu;
        const bool t = (frag & 0x8000u) != 0;
        diff = (diff << 15) | (frag & 0x7FFFu);
        sum += t ? diff : 0ull;
        cnt += t ? 1u : 0u;
        diff = t ? 0ull : diff;
    }
    // warp inclusive scan of (cnt, sum)
    u32 icnt = cnt;
    u64 isum = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 tc = __shfl_up_sync(0xFFFFFFFFu, icnt, d);
        const u64 ts = __shfl_up_sync(0xFFFFFFFFu, isum, d);
        if (lane >= (u32) d) { icnt += tc; isum += ts; }
    }
    const u32 n_total = __shfl_sync(0xFFFFFFFFu, icnt, 31);
    __syncwarp();                                        // key array cleared
    // sweep 2: keys and DNA parts of my k-mers.  The key is the low 32 bits of (amino-acid part - aa0); whether every key of
    // the piece fits is checked once, on the largest one (the last k-mer decoded), below.
    const u32 pos0 = icnt - cnt;
    u64 run = vbefore + (isum - sum);
    u32 *kp = ws.key + pos0;                             // ws.dna lies MT_KEYS words behind ws.key
    u32 *const kend_p = ws.key + MT_CAP_K;
    const u32 aa0_lo = (u32) aa0;
    diff = carry >> 15;
#pragma unroll
    for (int i = 0; i < MT_R; ++i) {
        const u32 frag = (u32) (r[i >> 2] >> (16 * (i & 3))) & 0xFFFFu;
        diff = (diff << 15) | (frag & 0x7FFFu);
        if (frag & 0x8000u) {
            run += diff;
            if (kp < kend_p) {
                kp[0] = (u32) (run >> 24) - aa0_lo;
                kp[MT_KEYS] = (u32) run & 0xFFFFFFu;
            }
            ++kp;
            diff = 0;
        }
    }
    // amino-acid parts relative to aa0 that do not fit 32 bits (a sparse region): the piece is cut before the first such
    // entry.  The values are ascending, so the last k-mer any lane decoded tells whether that can happen at all.
    u32 fit = 0xFFFFFFFFu;
    if (__any_sync(0xFFFFFFFFu, cnt != 0 && ((run >> 24) - aa0) >= 0xFFFFFFFFull)) {
        u32 pos = pos0;
        u64 v = vbefore + (isum - sum);
        diff = carry >> 15;
        for (int i = 0; i < MT_R; ++i) {
            const u32 frag = (u32) (r[i >> 2] >> (16 * (i & 3))) & 0xFFFFu;
            diff = (diff << 15) | (frag & 0x7FFFu);
            if (frag & 0x8000u) {
                v += diff;
                if (((v >> 24) - aa0) >= 0xFFFFFFFFull) {
                    fit = fit < pos ? fit : pos;
                    if (pos < (u32) MT_CAP_K) ws.key[pos] = 0xFFFFFFFFu;
                }
                ++pos;
                diff = 0;
            }
        }
        fit = __reduce_min_sync(0xFFFFFFFFu, fit);
    }
    if (lane == 0) ws.aa0 = aa0;
    __syncwarp();
    // group heads: entry j opens a group when its key differs from its predecessor's; the padding (0xFFFFFFFF) behind the
    // last entry makes the first padded slot a head too
#pragma unroll
    for (int i = 0; i < MT_HB; ++i) {
        const u32 j = 32 * i + lane;
        const u32 k = j < (u32) MT_KEYS ? ws.key[j] : 0xFFFFFFFFu;
        const u32 kprev = j == 0 ? ~k : (j - 1 < (u32) MT_KEYS ? ws.key[j - 1] : 0xFFFFFFFFu);
        const u32 m = __ballot_sync(0xFFFFFFFFu, k != kprev);
        if (lane == 0) ws.hbits[i] = m;
    }
    *n_fit = fit;
    __syncwarp();
    return n_total;
}

// index of the last group head in [0, n) (n >= 1; bit 0 is always a head).  Warp-uniform result.
__device__ __forceinline__ u32 last_head_before(const WarpSmem &ws, u32 n) {
    const u32 lane = threadIdx.x & 31;
    u32 m = 0;
    if (lane < (u32) MT_HB) {
        m = ws.hbits[lane];
        const u32 lo = lane * 32;
        if (lo >= n) m = 0;
        else if (n - lo < 32) m &= (1u << (n - lo)) - 1u;
    }
    const u32 any = __ballot_sync(0xFFFFFFFFu, m != 0);
    const int top = 31 - __clz((int) any);               // any != 0: bit 0 is set
    const u32 mt = __shfl_sync(0xFFFFFFFFu, m, top);
    return (u32) top * 32 + (31 - __clz((int) mt));
}

// buffer word index just behind the terminator of the piece's k-th k-mer (k >= 1), i.e. the first word of k-mer k.
// Warp-uniform result; the piece's words lie in buf[pad, pad + wl).
__device__ __forceinline__ u32 word_after_kmer(const u16 *buf, u32 pad, u32 wl, u32 k) {
    const u32 lane = threadIdx.x & 31;
    u32 seen = 0;
    for (u32 base = pad; base < pad + wl; base += 32) {
        const u32 idx = base + lane;
        const bool term = idx < pad + wl && (buf[idx] & 0x8000u);
        const u32 m = __ballot_sync(0xFFFFFFFFu, term);
        const u32 c = (u32) __popc(m);
        if (seen + c >= k) {
            // the (k - seen)-th set bit of m
            u32 mm = m;
            for (u32 i = 1; i < k - seen; ++i) mm &= mm - 1;
            return base + (u32) __ffs((int) mm);          // ffs is 1-based: index behind the terminator
        }
        seen += c;
    }
    return pad + wl;
}

struct MatchOut {
    mb_match *out;
    u64 cap;
    unsigned long long *counter;   // total matches produced (may exceed cap)
    u32 *error_flag;               // set when a DB tax ID has no species (KmerMatcher.cpp:433-441)
    const Route *route;            // multi-GPU: matches go to their read's home rank (shared-memory copy), else nullptr
};

// One match record (Match.h:9-25 payload) from query q and DB k-mer ordinal `kidx` with value t, distance d.
__device__ __forceinline__ void make_match(const DbView &db, const MatchOut &mo, const u8 *lutp, const mb_kmer &q, u64 t, u64 kidx,
                                           u32 d, int format, u64 (&rec)[3]) {
    const int32_t tax = (int32_t) (__ldg(db.info + kidx) & db.info_mask);
    int32_t sp = 0;
    if (tax >= 0 && tax <= db.max_taxid) sp = __ldg(db.species_of + tax);
    if (sp == 0 || tax == 0) *mo.error_flag = 1;
    // KmerMatcher.cpp:1137-1143: getHammings for (frame<3) == (format==2), else getHammings_reverse
    const bool rev = ((qi_frame(q.qinfo) < 3) != (format == 2));
    const u32 pack = d ? hamming_pack_lut(lutp, q.value, t, rev) : 0u;       // distance 0 <=> every field 0
    rec[0] = q.qinfo;
    rec[1] = (u64) (u32) tax | ((u64) (u32) sp << 32);
    rec[2] = (t & 0xFFFFFFull) | ((u64) pack << 32) | ((u64) d << 48);
}

// Routed flush (mb_match_kmers_routed): the staged matches are grouped by home rank inside the staging buffer, one
// reservation per home on this rank's LOCAL per-home counter (lanes 0..n-1 in parallel; every source owns a fixed region
// of every home's receive buffer), then each home's run leaves with coalesced 8-byte peer stores.  Warp-synchronous.
__device__ __noinline__ void warp_flush_routed(const MatchOut &mo, u64 *stage, u32 fill) {
    const Route &r = *mo.route;
    const u32 lane = threadIdx.x & 31;
    constexpr int PER = (MT_STAGE + 31) / 32;
    u64 w0[PER], w1[PER], w2[PER];
    u32 d[PER], rank[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const u32 i = lane + 32 * j;
        d[j] = 0xFFFFFFFFu; rank[j] = 0;
        if (i < fill) { w0[j] = stage[i * 3]; w1[j] = stage[i * 3 + 1]; w2[j] = stage[i * 3 + 2]; d[j] = route_home(r, qi_seq(w0[j])); }
    }
    u32 mycnt = 0;                                  // lane dd: matches of this flush for home dd
    for (u32 dd = 0; dd < r.n; ++dd) {
        u32 run = 0;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const u32 mask = __ballot_sync(0xFFFFFFFFu, d[j] == dd);
            if (d[j] == dd) rank[j] = run + __popc(mask & ((1u << lane) - 1u));
            run += __popc(mask);
        }
        if (lane == dd) mycnt = run;
    }
    u32 incl = mycnt;
#pragma unroll
    for (int k = 1; k < 32; k <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, k);
        if (lane >= (u32) k) incl += t;
    }
    const u32 excl = incl - mycnt;
    unsigned long long base = 0;
    if (lane < r.n && mycnt) base = atomicAdd(r.sent_m + lane, (unsigned long long) mycnt);
    if (lane == 0) atomicAdd(mo.counter, (unsigned long long) fill);             // local statistic: matches this rank produced
    __syncwarp();
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const u32 e = __shfl_sync(0xFFFFFFFFu, excl, d[j] & 31u);
        if (d[j] != 0xFFFFFFFFu) { const u32 o = (e + rank[j]) * 3; stage[o] = w0[j]; stage[o + 1] = w1[j]; stage[o + 2] = w2[j]; }
    }
    __syncwarp();
    for (u32 dd = 0; dd < r.n; ++dd) {
        const u32 cnt = __shfl_sync(0xFFFFFFFFu, mycnt, dd);
        const u32 ex = __shfl_sync(0xFFFFFFFFu, excl, dd);
        const unsigned long long b = __shfl_sync(0xFFFFFFFFu, base, dd);
        if (!cnt) continue;
        u64 *dst = reinterpret_cast<u64 *>(r.matches[dd]) + b * 3;
        const u64 *src = stage + ex * 3;
        const u64 cap = r.match_cap[dd];
        if (b + cnt <= cap) {
            for (u32 i = lane; i < cnt * 3; i += 32) dst[i] = src[i];
        } else {
            for (u32 i = lane; i < cnt * 3; i += 32)
                if (b + i / 3 < cap) dst[i] = src[i];
        }
    }
}

// Flush the warp's staged matches: one reservation, coalesced 8-byte stores.  Warp-synchronous.
template <bool ROUTED>
__device__ __forceinline__ void warp_flush(const MatchOut &mo, u64 *stage, u32 &fill) {
    const u32 lane = threadIdx.x & 31;
    __syncwarp();
    if (ROUTED && fill) {
        warp_flush_routed(mo, stage, fill);
    } else if (fill) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(mo.counter, (unsigned long long) fill);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        u64 *dst = reinterpret_cast<u64 *>(mo.out) + base * 3;
        const u32 words = fill * 3;
        if (base + fill <= mo.cap) {
            for (u32 i = lane; i < words; i += 32) dst[i] = stage[i];
        } else {
            for (u32 i = lane; i < words; i += 32)
                if (base + i / 3 < mo.cap) dst[i] = stage[i];
        }
    }
    fill = 0;
    __syncwarp();
}

// Warp-level emission for the streaming path: lane l adds the entries [0, ge) of the decoded piece whose distance to its
// query is <= maxh (cnt of them).  kbase: ordinal of entry 0.
template <bool ROUTED>
__device__ __noinline__ void warp_emit_own(const DbView &db, u64 kbase, WarpSmem &ws, const MatchTables &tb, const MatchOut &mo, bool active,
                                           const mb_kmer &q, u32 ge, u32 maxh, u32 cnt, int format, u32 &fill) {
    const u32 lane = threadIdx.x & 31;
    u32 incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (u32) d) incl += t;
    }
    const u32 total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (total == 0) return;
    // reserve and write straight to HBM (rare path: a group longer than a piece)
    warp_flush<ROUTED>(mo, ws.stage, fill);
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(mo.counter, (unsigned long long) total);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (active && cnt) {
        u64 o = base + incl - cnt;
        for (u32 j = 0; j < ge; ++j) {
            const u64 t = ws.dna[j];
            const u32 d = hamming_sum_lut(tb.lut2, q.value, t);
            if (d > maxh) continue;
            u64 rec[3];
            make_match(db, mo, tb.lutp, q, t, kbase + j, d, format, rec);
            if (ROUTED) {                                     // one reservation per record at its home
                const u32 h = route_home(*mo.route, qi_seq(rec[0]));
                const unsigned long long slot = atomicAdd(mo.route->sent_m + h, 1ull);
                if (slot < mo.route->match_cap[h]) {
                    u64 *p = reinterpret_cast<u64 *>(mo.route->matches[h]) + slot * 3;
                    p[0] = rec[0]; p[1] = rec[1]; p[2] = rec[2];
                }
            } else if (o < mo.cap) {
                u64 *p = reinterpret_cast<u64 *>(mo.out + o);
                p[0] = rec[0]; p[1] = rec[1]; p[2] = rec[2];
            }
            ++o;
        }
    }
}

// Fetch (unless the words are already in flight) + wait + decode of the piece at cursor g in word buffer `buf`.
// Returns n = entries of the piece that can be joined now (decoded, keys fit, candidates below kend) and sets
// *complete = every one of them belongs to a complete amino-acid group (the piece reaches the end of the unit or of the
// candidates).  `parity` is the mbarrier phase of the word buffer.
__device__ __forceinline__ u32 open_piece(const DbView &db, WarpSmem &ws, u32 &parity, const DbCursor &g, u64 wlimit, u64 kend,
                                          bool in_flight, u32 *wl_out, bool *complete) {
    const u32 lane = threadIdx.x & 31;
    const u32 wl = (u32) (wlimit - g.word < (u64) MT_CAP_W ? wlimit - g.word : (u64) MT_CAP_W);
    if (!in_flight) {
        __syncwarp();
        if (lane == 0) fetch_words(db, ws, g.word, wl);
    }
    mbar_wait(&ws.mbar, parity);
    parity ^= 1u;
    u32 n_fit;
    const u32 n_total = decode_piece(ws, ws.w, (u32) (g.word & 7ull), wl, g.vbefore, &n_fit);
    u32 n = n_total < (u32) MT_CAP_K ? n_total : (u32) MT_CAP_K;
    bool all = g.word + wl >= wlimit && n_total <= (u32) MT_CAP_K;
    if (n_fit < n) { n = n_fit; all = false; }
    if ((u64) n >= kend - g.kidx) { n = (u32) (kend - g.kidx); all = true; }
    *wl_out = wl;
    *complete = all;
    return n;
}

// cursor of entry `used` (>= 1) of the piece that was decoded at cursor g (wl words; its words are still in the buffer)
__device__ __forceinline__ DbCursor cursor_after(const WarpSmem &ws, const DbCursor &g, u32 wl, u32 used) {
    DbCursor nx;
    nx.word = (g.word & ~7ull) + word_after_kmer(ws.w, (u32) (g.word & 7ull), wl, used);
    nx.vbefore = piece_value(ws, used - 1);
    nx.kidx = g.kidx + used;
    return nx;
}

// Streaming two-pass join for an amino-acid group that does not fit one piece.  `c` sits on the group's first k-mer;
// the group ends at the first k-mer with another amino-acid part, at ordinal kend or at stream word wlimit.  For every
// 32 queries of the slice that hold queries of that group the group is decoded twice (minimum, then emit).  On return
// `c` is the cursor of the first k-mer behind the group.
template <bool ROUTED>
__device__ __noinline__ void giant_group(const DbView &db, DbCursor &c, u64 kend, u64 wlimit, WarpSmem &ws, const MatchTables &tb,
                                         const MatchOut &mo, u32 &parity, const mb_kmer *__restrict__ q, u64 qs, u64 qe,
                                         int format, u32 &fill) {
    const u32 lane = threadIdx.x & 31;
    u32 wl;
    bool complete;
    open_piece(db, ws, parity, c, wlimit, kend, false, &wl, &complete);
    const u64 gaa = ws.aa0;                              // the group's amino-acid part: that of its first k-mer
    DbCursor gend = c;                                   // becomes the cursor behind the group
    bool have_end = false;
    for (u64 qb = qs; qb < qe || !have_end; qb += 32) {
        const u64 qi = qb + lane;
        mb_kmer qk;
        qk.value = 0; qk.qinfo = 0;
        bool mine = false;
        if (qi < qe) { qk = q[qi]; mine = (qk.value >> 24) == gaa; }
        const bool any = __any_sync(0xFFFFFFFFu, mine);
        if (!any && have_end) continue;                  // (the first pass also finds the group's end, queries or not)
        u32 mind = 255, maxh = 0;
        for (int phase = 0; phase < 2; ++phase) {
            if (phase == 1 && !any) break;
            DbCursor g = c;
            while (true) {
                const u32 n = open_piece(db, ws, parity, g, wlimit, kend, false, &wl, &complete);
                if (n == 0) { if (!have_end) { gend = g; have_end = true; } break; }
                // first entry with another amino-acid part (keys are relative to THIS piece's first entry)
                const u64 aa_p = ws.aa0;
                u32 ge = n;
                for (u32 j0 = 0; j0 < n; j0 += 32) {
                    const u32 j = j0 + lane;
                    const u32 m = __ballot_sync(0xFFFFFFFFu, j < n && aa_p + ws.key[j] != gaa);
                    if (m) { ge = j0 + (u32) __ffs((int) m) - 1; break; }
                }
                const bool finished = ge < n || complete;
                if (phase == 0) {
                    if (mine)
                        for (u32 j = 0; j < ge; ++j) {
                            const u32 d = hamming_sum_lut(tb.lut2, qk.value, (u64) ws.dna[j]);
                            mind = d < mind ? d : mind;
                        }
                } else {
                    u32 cnt = 0;
                    if (mine)
                        for (u32 j = 0; j < ge; ++j) cnt += hamming_sum_lut(tb.lut2, qk.value, (u64) ws.dna[j]) <= maxh ? 1u : 0u;
                    warp_emit_own<ROUTED>(db, g.kidx, ws, tb, mo, mine, qk, ge, maxh, cnt, format, fill);
                }
                // cursor behind the entries consumed from this piece
                const u32 used = finished ? ge : n;
                DbCursor nx = g;
                if (used) nx = cursor_after(ws, g, wl, used);
                __syncwarp();
                if (finished) { if (!have_end) { gend = nx; have_end = true; } break; }
                g = nx;
            }
            if (phase == 0) maxh = mind * 2 < 7 ? mind * 2 : 7;
        }
    }
    c = gend;
}

// ---- the join kernel ---------------------------------------------------------------------------------------------
template <bool ROUTED>
__global__ void __launch_bounds__(MT_THREADS, MB_MT_MIN_CTAS)
k_match(const __grid_constant__ DbView db, const Unit *__restrict__ units, u32 n_units, const mb_kmer *__restrict__ q,
        const u64 *__restrict__ qstart, const __grid_constant__ MatchOut mo_in, const Tables *__restrict__ tables, int format,
        u32 *unit_ticket, const Route *__restrict__ route_g) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ Route s_route;
    __shared__ MatchOut s_mo;                        // routed form only: mo_in with the shared-memory copy of the route
    MatchTables &tb = *reinterpret_cast<MatchTables *>(smem_raw);
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    WarpSmem &ws = *reinterpret_cast<WarpSmem *>(smem_raw + sizeof(MatchTables) + (size_t) warp * sizeof(WarpSmem));
    if (ROUTED) {                                   // multi-GPU: matches leave for their home rank (warp_flush_routed)
        for (u32 i = tid; i < sizeof(Route) / 8; i += MT_THREADS) reinterpret_cast<u64 *>(&s_route)[i] = reinterpret_cast<const u64 *>(route_g)[i];
        if (tid == 0) { s_mo = mo_in; s_mo.route = &s_route; }
    }
    const MatchOut &mo = ROUTED ? s_mo : mo_in;
    {
        const uint4 *src2 = reinterpret_cast<const uint4 *>(tables->lut2);
        const uint4 *srcp = reinterpret_cast<const uint4 *>(tables->lutp);
        uint4 *d2 = reinterpret_cast<uint4 *>(tb.lut2);
        uint4 *dp = reinterpret_cast<uint4 *>(tb.lutp);
        for (u32 i = tid; i < 4096 / 16; i += MT_THREADS) { d2[i] = __ldg(src2 + i); dp[i] = __ldg(srcp + i); }
    }
    if (lane == 0) mbar_init(&ws.mbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();                                // tables, route and barriers are ready; the only block barrier

    u32 fill = 0;                                   // matches staged by this warp (warp-uniform)
    u32 parity = 0;                                 // mbarrier phase of the word buffer

    // Asks for the words of unit u when the unit has queries and candidates; warms L2 with its query slice and info
    // entries.  All lanes compute the same facts from the same loads; nothing is kept in registers.
    auto prefetch_unit = [&](u32 u) {
        if (u >= n_units) return;
        const u64 qs = qstart[u], qe = qstart[u + 1];
        const u64 w0 = units[u].word, k0 = units[u].kidx, w1 = units[u + 1].word;
        u64 k1 = units[u + 1].kidx;
        if (k1 > db.n_kmers_eff) k1 = db.n_kmers_eff;
        if (qs >= qe || k0 >= k1) return;
        __syncwarp();
        if (lane == 0) fetch_words(db, ws, w0, (u32) (w1 - w0 < (u64) MT_CAP_W ? w1 - w0 : (u64) MT_CAP_W));
        u64 qn = qe - qs;
        if (qn > 2048) qn = 2048;
        const char *qp = reinterpret_cast<const char *>(q + qs);
        for (u64 o = (u64) lane * 128; o < qn * 16; o += 32 * 128) prefetch_l2(qp + o);
        u64 nk = k1 - k0;
        if (nk > (u64) MT_CAP_K) nk = MT_CAP_K;
        const char *inf = reinterpret_cast<const char *>(db.info + k0);
        for (u64 o = (u64) lane * 128; o < nk * 4; o += 32 * 128) prefetch_l2(inf + o);
    };

    // tickets are drawn one unit ahead: `cur` is joined now; the words of `nxt` are asked for as soon as the last piece of
    // `cur` is decoded (the decode keeps nothing in the word buffer), so they arrive while `cur` is being joined
    u32 cur = 0, nxt = 0;
    if (lane == 0) { cur = atomicAdd(unit_ticket, 1u); nxt = atomicAdd(unit_ticket, 1u); }
    cur = __shfl_sync(0xFFFFFFFFu, cur, 0);
    nxt = __shfl_sync(0xFFFFFFFFu, nxt, 0);
    prefetch_unit(cur);

    while (cur < n_units) {
        u32 drawn = 0;
        if (lane == 0) drawn = atomicAdd(unit_ticket, 1u);
        bool asked = false;                          // the words of `nxt` have been asked for
        if (lane == 0 && nxt + 2 * gridDim.x * MT_WARPS < n_units) { prefetch_l2(units + nxt + 2 * gridDim.x * MT_WARPS); prefetch_l2(qstart + nxt + 2 * gridDim.x * MT_WARPS); }

        // the current unit (the same facts prefetch_unit saw)
        const u64 qs = qstart[cur], qe = qstart[cur + 1];
        DbCursor c{units[cur].word, units[cur].vbefore, units[cur].kidx};
        const u64 wend = units[cur + 1].word;
        u64 kend = units[cur + 1].kidx;
        if (kend > db.n_kmers_eff) kend = db.n_kmers_eff;
        if (qs < qe && c.kidx < kend) {
            bool in_flight = true;
            while (c.kidx < kend) {
                u32 wl;
                bool complete;
                const u32 n = open_piece(db, ws, parity, c, wend, kend, in_flight, &wl, &complete);
                in_flight = false;
                if (n == 0) break;
                if (complete && !asked) { prefetch_unit(nxt); asked = true; }       // no further piece of this unit: the buffer is free
                // a piece that does not reach the end of the unit leaves its last group, which may continue, to the next piece
                u32 proc = n;
                if (!complete) {
                    proc = last_head_before(ws, n);
                    if (proc == 0) {
                        giant_group<ROUTED>(db, c, kend, wend, ws, tb, mo, parity, q, qs, qe, format, fill);
                        continue;
                    }
                }
                if (lane == 0) atomicOr(&ws.hbits[proc >> 5], 1u << (proc & 31));      // sentinel head behind the last entry
                __syncwarp();

                // ---- join the query slice against entries [0, proc) ----
                const u64 aa0 = ws.aa0;
                const u32 key_hi = ws.key[proc - 1];
                ulonglong2 v_next = make_ulonglong2(0, 0);
                if (qs + lane < qe) v_next = __ldg(reinterpret_cast<const ulonglong2 *>(q + qs + lane));
                for (u64 qb = qs; qb < qe; qb += 32) {
                    const u64 qvalue = v_next.x, qinfo = v_next.y;
                    const bool mine = qb + lane < qe;
                    if (qb + 32 + lane < qe) v_next = __ldg(reinterpret_cast<const ulonglong2 *>(q + qb + 32 + lane));
                    // (1) every lane: its query's amino-acid group [lo, lo + ncand) in the piece
                    u32 lo = 0, ncand = 0;
                    const u64 rel = (qvalue >> 24) - aa0;              // (wraps to a huge value below the piece)
                    if (mine && rel <= (u64) key_hi) {
                        // bound-free lower bound in the padded key array: number of entries below the key
                        const u32 key = (u32) rel;
                        const u32 *kp = ws.key;
#pragma unroll
                        for (int step = MT_KEYS / 2; step; step >>= 1)
                            if (kp[step - 1] < key) kp += step;
                        const u32 pos = (u32) (kp - ws.key);
                        if (*kp == key && pos < proc) {
                            // pos is the group's head; the group ends at the next head (the sentinel bounds the scan)
                            lo = pos;
                            const u32 b = pos + 1;
                            u32 m = ws.hbits[b >> 5] >> (b & 31);
                            u32 hi;
                            if (m) hi = b + (u32) __ffs((int) m) - 1;
                            else {
                                u32 wi = (b >> 5) + 1;
                                while ((m = ws.hbits[wi]) == 0) ++wi;
                                hi = wi * 32 + (u32) __ffs((int) m) - 1;
                            }
                            ncand = hi - lo;
                        }
                    }
                    // (2) the warp's (query, candidate) pairs are spread evenly over the lanes: pair p belongs to the lane
                    // whose exclusive prefix of ncand is the largest one <= p
                    u32 incl = ncand;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                        if (lane >= (u32) d) incl += t;
                    }
                    const u32 total = __shfl_sync(0xFFFFFFFFu, incl, 31);
                    if (total == 0) continue;
                    const u32 excl = incl - ncand;
                    const u32 q24 = (u32) qvalue & 0xFFFFFFu;
                    ws.qmin[lane] = 255u;
                    __syncwarp();
                    const u32 rounds = (total + 31) / 32;
                    u32 pk0 = 255, pk1 = 255;           // pairs of rounds 0 and 1 are kept for the emit pass: dd | owner << 8 | j << 16
                    for (u32 r = 0; r < rounds; ++r) {
                        const u32 p = r * 32 + lane;
                        u32 owner = 0;
#pragma unroll
                        for (int st = 16; st > 0; st >>= 1) {
                            const u32 e = __shfl_sync(0xFFFFFFFFu, excl, (owner + st) & 31);
                            if (owner + st < 32 && e <= p) owner += st;
                        }
                        // lanes with ncand == 0 share their prefix with the next lane: the search lands on the LAST lane
                        // with excl <= p, which is the one that owns p (its range is non-empty) when p < total
                        const u32 o_excl = __shfl_sync(0xFFFFFFFFu, excl, owner);
                        const u32 o_lo = __shfl_sync(0xFFFFFFFFu, lo, owner);
                        const u32 o_q = __shfl_sync(0xFFFFFFFFu, q24, owner);
                        u32 dd = 255, j = 0;
                        if (p < total) {
                            j = o_lo + (p - o_excl);
                            // the candidate's info entry (make_match reads it in the emit pass) is asked for now
                            if (MB_MT_INFO_PF) asm volatile("prefetch.global.L1 [%0];" ::"l"(db.info + c.kidx + j));
                            const u32 t24 = ws.dna[j];
                            dd = t24 == o_q ? 0u : hamming_sum_lut(tb.lut2, (u64) o_q, (u64) t24);
                            atomicMin(&ws.qmin[owner], dd);
                        }
                        const u32 pk = dd | (owner << 8) | (j << 16);
                        if (r == 0) pk0 = pk;
                        else if (r == 1) pk1 = pk;
                    }
                    __syncwarp();
                    // (3) emit: distance <= min(2 * min, 7) (KmerMatcher.cpp:1126); one staging slot per surviving pair
                    for (u32 r = 0; r < rounds; ++r) {
                        const u32 p = r * 32 + lane;
                        u32 owner, dd, j;
                        if (r < 2) { const u32 pk = r == 0 ? pk0 : pk1; dd = pk & 255u; owner = (pk >> 8) & 31u; j = pk >> 16; }
                        else {
                            owner = 0;
#pragma unroll
                            for (int st = 16; st > 0; st >>= 1) {
                                const u32 e = __shfl_sync(0xFFFFFFFFu, excl, (owner + st) & 31);
                                if (owner + st < 32 && e <= p) owner += st;
                            }
                            const u32 o_excl = __shfl_sync(0xFFFFFFFFu, excl, owner);
                            const u32 o_lo = __shfl_sync(0xFFFFFFFFu, lo, owner);
                            const u32 o_q = __shfl_sync(0xFFFFFFFFu, q24, owner);
                            dd = 255; j = 0;
                            if (p < total) {
                                j = o_lo + (p - o_excl);
                                const u32 t24 = ws.dna[j];
                                dd = t24 == o_q ? 0u : hamming_sum_lut(tb.lut2, (u64) o_q, (u64) t24);
                            }
                        }
                        const u32 mind = ws.qmin[owner];
                        const u32 maxh = mind * 2 < 7 ? mind * 2 : 7;
                        const bool ok = p < total && dd <= maxh;
                        const u32 okmask = __ballot_sync(0xFFFFFFFFu, ok);
                        // the owner's query record for the match (only the DNA part of its value is needed)
                        mb_kmer oq;
                        oq.value = (u64) __shfl_sync(0xFFFFFFFFu, q24, owner);
                        oq.qinfo = __shfl_sync(0xFFFFFFFFu, qinfo, owner);
                        if (okmask == 0) continue;
                        const u32 nok = (u32) __popc(okmask);
                        if (fill + nok > MT_STAGE) warp_flush<ROUTED>(mo, ws.stage, fill);
                        if (ok) {
                            u64 rec[3];
                            make_match(db, mo, tb.lutp, oq, (u64) ws.dna[j], c.kidx + j, dd, format, rec);
                            const u32 o = fill + (u32) __popc(okmask & ((1u << lane) - 1u));
                            ws.stage[o * 3] = rec[0]; ws.stage[o * 3 + 1] = rec[1]; ws.stage[o * 3 + 2] = rec[2];
                        }
                        fill += nok;
                        __syncwarp();
                    }
                }
                // behind the entries joined: the next piece of this unit, if any
                if (c.kidx + proc >= kend) break;
                const DbCursor nx = cursor_after(ws, c, wl, proc);
                __syncwarp();
                c = nx;
            }
        }
        if (!asked) prefetch_unit(nxt);
        cur = nxt;
        nxt = __shfl_sync(0xFFFFFFFFu, drawn, 0);
    }
    warp_flush<ROUTED>(mo, ws.stage, fill);
}

// ---- load-time unit index ----------------------------------------------------------------------------------------
// pass A, one warp per raw block of UN_W words: how many k-mers END in the block and the sum of their deltas.
__global__ void __launch_bounds__(256)
k_db_raw_summary(DbView db, u64 n_raw, u32 *__restrict__ raw_cnt, u64 *__restrict__ raw_sum) {
    const u64 b = (u64) blockIdx.x * 8 + (threadIdx.x >> 5);
    const u32 lane = threadIdx.x & 31;
    if (b >= n_raw) return;
    const u64 w0 = b * UN_W;
    constexpr int R = UN_W / 32;
    const u64 l0 = w0 + (u64) lane * R;
    u64 carry = 0;
    {
        u64 st = l0;
        while (st > 0 && l0 - st < 4 && st - 1 < db.n_words && !(db.words[st - 1] & 0x8000u)) --st;
        for (u64 k = st; k < l0; ++k) carry = (carry | db.words[k]) << 15;
    }
    u32 cnt = 0;
    u64 sum = 0, diff = carry;
    for (int i = 0; i < R; ++i) {
        const u64 idx = l0 + i;
        if (idx >= db.n_words) break;
        const u32 frag = db.words[idx];
        if (frag & 0x8000u) { diff |= (frag & 0x7FFFu); sum += diff; ++cnt; diff = 0; }
        else diff = (diff | frag) << 15;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, d);
        sum += __shfl_xor_sync(0xFFFFFFFFu, sum, d);
    }
    if (lane == 0) { raw_cnt[b] = cnt; raw_sum[b] = sum; }
}

// pass B (after the host scanned counts and sums into kidx / vbefore per raw block), one warp per raw block: the first
// k-mer that STARTS in the block and opens a new T bucket (or the stream).  out[b].t_lo = ~0 when there is none.
__global__ void __launch_bounds__(256)
k_db_raw_own(DbView db, u64 n_raw, const u64 *__restrict__ raw_kidx, const u64 *__restrict__ raw_vbefore, SortT st, Unit *__restrict__ out) {
    const u64 b = (u64) blockIdx.x * 8 + (threadIdx.x >> 5);
    const u32 lane = threadIdx.x & 31;
    if (b >= n_raw) return;
    const u64 w0 = b * UN_W, w1 = w0 + UN_W < db.n_words ? w0 + UN_W : db.n_words;
    constexpr int R = UN_W / 32;
    const u64 l0 = w0 + (u64) lane * R;
    // deltas ENDING in my run (as in pass A), then absolute values; a k-mer's start word is the word behind the previous terminator
    u64 carry = 0;
    u64 first_start = l0;                    // start word of the delta that is unfinished at l0 (if any)
    {
        u64 s0 = l0;
        while (s0 > 0 && l0 - s0 < 4 && s0 - 1 < db.n_words && !(db.words[s0 - 1] & 0x8000u)) --s0;
        for (u64 k = s0; k < l0; ++k) carry = (carry | db.words[k]) << 15;
        first_start = s0;
    }
    u32 cnt = 0;
    u64 sum = 0, diff = carry;
    for (int i = 0; i < R; ++i) {
        const u64 idx = l0 + i;
        if (idx >= db.n_words) break;
        const u32 frag = db.words[idx];
        if (frag & 0x8000u) { diff |= (frag & 0x7FFFu); sum += diff; ++cnt; diff = 0; }
        else diff = (diff | frag) << 15;
    }
    u32 icnt = cnt;
    u64 isum = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 tc = __shfl_up_sync(0xFFFFFFFFu, icnt, d);
        const u64 ts = __shfl_up_sync(0xFFFFFFFFu, isum, d);
        if (lane >= (u32) d) { icnt += tc; isum += ts; }
    }
    u64 run = raw_vbefore[b] + (isum - sum);
    u64 ord = raw_kidx[b] + (icnt - cnt);
    // candidates: k-mers ENDING in my run whose start word lies in [w0, w1)
    u64 best_word = ~0ull, best_vb = 0, best_k = 0, best_t = ~0ull;
    diff = carry;
    u64 start = first_start;
    for (int i = 0; i < R; ++i) {
        const u64 idx = l0 + i;
        if (idx >= db.n_words) break;
        const u32 frag = db.words[idx];
        if (frag & 0x8000u) {
            diff |= (frag & 0x7FFFu);
            const u64 nv = run + diff;
            if (start >= w0 && start < w1 && best_word == ~0ull) {
                const bool opens = ord == 0 || sort_t(st, nv) != sort_t(st, run);
                if (opens) { best_word = start; best_vb = run; best_k = ord; best_t = sort_t(st, nv); }
            }
            run = nv;
            ++ord;
            diff = 0;
            start = idx + 1;
        } else diff = (diff | frag) << 15;
    }
    // a k-mer that starts in this block but ends in the next one: the last lane sees its start (unfinished delta at w1)
    // and finishes it from the following words
    if (lane == 31 && best_word == ~0ull && start < w1 && start >= w0 && w1 < db.n_words) {
        u64 d2 = diff;
        u64 idx = w1;
        bool done = false;
        for (int k = 0; k < 5 && idx < db.n_words; ++k, ++idx) {
            const u32 frag = db.words[idx];
            if (frag & 0x8000u) { d2 |= (frag & 0x7FFFu); done = true; break; }
            d2 = (d2 | frag) << 15;
        }
        if (done) {
            const u64 nv = run + d2;
            if (ord == 0 || sort_t(st, nv) != sort_t(st, run)) { best_word = start; best_vb = run; best_k = ord; best_t = sort_t(st, nv); }
        }
    }
    // lowest start word wins
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        const u64 ow = __shfl_xor_sync(0xFFFFFFFFu, best_word, d);
        const u64 ov = __shfl_xor_sync(0xFFFFFFFFu, best_vb, d);
        const u64 ok = __shfl_xor_sync(0xFFFFFFFFu, best_k, d);
        const u64 ot = __shfl_xor_sync(0xFFFFFFFFu, best_t, d);
        if (ow < best_word) { best_word = ow; best_vb = ov; best_k = ok; best_t = ot; }
    }
    if (lane == 0) {
        Unit u;
        u.word = best_word; u.vbefore = best_vb; u.kidx = best_k; u.t_lo = best_t;
        out[b] = u;
    }
}

// plain decode of the resident stream (parity tap): one warp per unit writes the values of its k-mers
__global__ void __launch_bounds__(MT_THREADS)
k_db_decode(const __grid_constant__ DbView db, const Unit *__restrict__ units, u32 n_units, u64 *__restrict__ values) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpSmem &ws = *reinterpret_cast<WarpSmem *>(smem_raw + (size_t) warp * sizeof(WarpSmem));
    if (lane == 0) mbar_init(&ws.mbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    u32 parity = 0;
    for (u32 u = blockIdx.x * MT_WARPS + warp; u < n_units; u += gridDim.x * MT_WARPS) {
        DbCursor c{units[u].word, units[u].vbefore, units[u].kidx};
        const u64 wend = units[u + 1].word, kend = units[u + 1].kidx;
        while (c.kidx < kend) {
            u32 wl;
            bool complete;
            const u32 n = open_piece(db, ws, parity, c, wend, kend, false, &wl, &complete);
            if (n == 0) break;
            for (u32 j = lane; j < n; j += 32) values[c.kidx + j] = piece_value(ws, j);
            if (c.kidx + n >= kend) break;
            const DbCursor nx = cursor_after(ws, c, wl, n);
            __syncwarp();
            c = nx;
        }
    }
}

// ---- per-batch partition: where each unit's T range starts in the sorted query array -----------------------------
// The queries are ordered by T(aa) (SortT, mb_common.cuh); unit u owns the T values [t_lo(u), t_lo(u+1)), so its query
// slice is [qstart[u], qstart[u+1]) — disjoint between units and exact.
__device__ __forceinline__ u64 query_lower_bound_t(const mb_kmer *__restrict__ q, u64 nq, const SortT &st, u64 t) {
    u64 lo = 0, hi = nq;
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        if (sort_t(st, __ldg(&q[mid].value)) < t) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__global__ void k_match_partition(const Unit *__restrict__ units, u32 n_units, const mb_kmer *__restrict__ q, u64 nq, SortT st,
                                  u64 *__restrict__ qstart) {
    const u32 u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u > n_units) return;
    qstart[u] = u == n_units ? nq : query_lower_bound_t(q, nq, st, units[u].t_lo);
}

} // namespace mb
