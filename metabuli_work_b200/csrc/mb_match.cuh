// mb_match.cuh — K3+K4: block-parallel diffIdx decode feeding the amino-acid merge join with the DNA
// Hamming filter and match compaction.
//
// Replaces KmerMatcher::matchKmers (src/commons/KmerMatcher.cpp:123-481) with getNextTargetKmer
// (KmerMatcher.h:282-297) and compareDna (KmerMatcher.cpp:1117-1146).
//
// What the reference computes, stated without its thread ranges: a query k-mer q is compared with
// ALL DB k-mers whose amino-acid part (value >> 24) equals q's — except the very last k-mer of the
// DB, which the `diffIdxPos != numOfDiffIdx` guards never offer (KmerMatcher.cpp:363-406) — and
// every candidate whose codon-id Hamming sum is <= min(2*min_sum, 7) yields a Match.  The result
// per query depends on nothing else, so the join parallelises over DB blocks.
//
// B200 design.  The compressed stream stays in HBM (about 9.7 B per k-mer with `info`).  At load
// time a block index is built: for every 2048-word block the decode cursor (word, value before,
// k-mer ordinal) of the first amino-acid group that STARTS in the block, and that group's AA.  A
// group is owned by the block it starts in.  Per batch, a partition kernel binary-searches the
// sorted query list once per block (merge-path style: block b joins queries [qstart[b],
// qstart[b+1])); blocks without queries never touch their words.  A CTA then
//   1. loads its block plus a 512-word overhang with 16-byte loads into shared memory,
//   2. decodes it cooperatively: each thread assembles the deltas ending in its 10 words, a block
//      scan of (count, 64-bit delta sum) turns them into absolute values in shared memory,
//   3. joins its query slice, one query per thread, warps running independently (32-query slices dealt
//      round robin to the warps, no block barrier): binary search of the AA part in the decoded block, ONE Hamming evaluation per candidate
//      through a two-codon table in shared memory,
//   4. compacts matches per warp: records are staged in a per-warp shared-memory buffer and flushed with
//      one atomic reservation and coalesced 8-byte stores when the buffer fills.
// The kernel is persistent (one CTA per resident slot, DB blocks handed out by an atomic ticket), so the
// tables are loaded into shared memory once per CTA and staged matches survive from block to block.  Tickets
// are drawn one block ahead: while a CTA joins block b, one of its warps pulls the words, the info entries and
// the query slice of its NEXT block into L2 (prefetch.global.L2), and the query loads of the ticket loop are
// software-pipelined, so the join itself waits on L2, not on HBM.
// Groups longer than the overhang are re-decoded from their own start; groups longer than a whole
// chunk take a streaming two-pass path.  Capacity misses are counted, never written out of bounds;
// the host grows the buffer and reruns (the reference's retry, Classifier.cpp:127-131).
#pragma once
#include "mb_common.cuh"

namespace mb {

constexpr int DB_W = 2048;                 // words per block
constexpr int DB_E = 256;                  // overhang words
constexpr int DB_C = DB_W + DB_E;          // words decoded per chunk
constexpr int MT_THREADS = 256;
constexpr int DB_R = DB_C / MT_THREADS;    // words per thread (10: odd word stride, conflict-free)
constexpr int DB_PAD = DB_C + 64;          // zero words the device stream is padded with

static_assert(DB_C % MT_THREADS == 0, "chunk must divide among the threads");

struct BlockIndex {      // per 2048-word block: cursor of the first AA group starting in it
    u64 word;            // first word of that group's first k-mer
    u64 vbefore;         // value of the preceding k-mer (0 at the stream start)
    u64 kidx;            // ordinal of that k-mer
    u64 own_aa;          // its AA part; for a block that owns nothing: the next owning block's AA
    u32 owns;            // 1 when a group starts in this block
    u32 pad;
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

constexpr int MT_STAGE = 96;               // matches staged per warp before a flush

struct MatchSmem {
    __align__(16) u16 w[DB_C + 16];
    u64 vals[DB_C];
    u16 wend[DB_C];
    u64 scan_sum[MT_THREADS / 32];
    u32 scan_cnt[MT_THREADS / 32];
    u32 n;
    int last_start;
    int own_end;
    u32 group_end;
    u64 reserve_base;
    u32 warp_tot[MT_THREADS / 32];
    u32 qnext;                             // next 32-query ticket of the current chunk
    u32 tick[2];                           // persistent loop: DB block tickets, fetched one block ahead
    u64 qadv;
};

// the part of the shared memory that lives for the whole (persistent) kernel
struct MatchTables {
    u8 lut2[4096];
    u8 lutp[4096];
    u64 stage[MT_THREADS / 32][MT_STAGE * 3];
    u32 fill[MT_THREADS / 32];
    u32 qmin[MT_THREADS / 32][32];           // per query of the warp's ticket: smallest distance in its group
};

// Decode up to DB_C words starting at cursor c (which must sit on the first word of a k-mer).
// Leaves values / terminator word offsets in shared memory, returns the number of k-mers decoded.
__device__ __forceinline__ u32 decode_chunk(const DbView &db, const DbCursor &c, MatchSmem &s) {
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u64 wbase = c.word & ~7ull;
    const u32 pad = (u32) (c.word - wbase);
    // 16-byte loads; the stream is zero padded so reads past n_words are harmless (no terminators)
    const uint4 *src = reinterpret_cast<const uint4 *>(db.words + wbase);
    uint4 *dst = reinterpret_cast<uint4 *>(s.w);
    for (u32 i = tid; i < (DB_C + 16) / 8; i += MT_THREADS) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (wbase + (u64) i * 8 < db.n_words) v = __ldg(src + i);
        dst[i] = v;
    }
    __syncthreads();
    const u16 *w = s.w + pad;
    const u32 l0 = tid * DB_R;
    // words of an unfinished delta in front of this thread's run (at most 4 for a valid stream)
    u32 st = l0;
    while (st > 0 && l0 - st < 4 && !(w[st - 1] & 0x8000u)) --st;
    u64 carry = 0;
    for (u32 k = st; k < l0; ++k) carry = (carry | w[k]) << 15;
    // pass 1: count and sum the deltas terminating in my run (getNextTargetKmer, KmerMatcher.h:282-297)
    u32 cnt = 0;
    u64 sum = 0, diff = carry;
#pragma unroll
    for (int i = 0; i < DB_R; ++i) {
        const u32 frag = w[l0 + i];
        const bool in_range = c.word + l0 + i < db.n_words;
        if ((frag & 0x8000u) && in_range) { diff |= (frag & 0x7FFFu); sum += diff; ++cnt; diff = 0; }
        else diff = (diff | frag) << 15;
    }
    // block exclusive scan of (cnt, sum)
    u32 icnt = cnt;
    u64 isum = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 tc = __shfl_up_sync(0xFFFFFFFFu, icnt, d);
        const u64 ts = __shfl_up_sync(0xFFFFFFFFu, isum, d);
        if (lane >= (u32) d) { icnt += tc; isum += ts; }
    }
    if (lane == 31) { s.scan_cnt[warp] = icnt; s.scan_sum[warp] = isum; }
    __syncthreads();
    u32 cbase = 0, ctot = 0;
    u64 sbase = 0;
#pragma unroll
    for (int k = 0; k < MT_THREADS / 32; ++k) {
        const u32 tc = s.scan_cnt[k];
        const u64 ts = s.scan_sum[k];
        if ((u32) k < warp) { cbase += tc; sbase += ts; }
        ctot += tc;
    }
    u32 pos = cbase + icnt - cnt;
    u64 run = c.vbefore + sbase + isum - sum;
    // pass 2: absolute values
    diff = carry;
#pragma unroll
    for (int i = 0; i < DB_R; ++i) {
        const u32 frag = w[l0 + i];
        const bool in_range = c.word + l0 + i < db.n_words;
        if ((frag & 0x8000u) && in_range) {
            diff |= (frag & 0x7FFFu);
            run += diff;
            s.vals[pos] = run;
            s.wend[pos] = (u16) (l0 + i);
            ++pos;
            diff = 0;
        } else diff = (diff | frag) << 15;
    }
    // candidates stop at n_kmers_eff
    u64 room = db.n_kmers_eff > c.kidx ? db.n_kmers_eff - c.kidx : 0;
    if ((u64) ctot > room) ctot = (u32) room;
    __syncthreads();
    return ctot;
}

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// One warp warms L2 with what DB block nb will touch: its words, its info entries and its query slice.  Also asks for
// the index entries that some CTA will read `ahead` tickets from now.  Every address stays inside its allocation.
__device__ __forceinline__ void prefetch_block(const DbView &db, const BlockIndex *__restrict__ blocks, u32 n_blocks,
                                               const mb_kmer *__restrict__ q, const u64 *__restrict__ qstart,
                                               const u64 *__restrict__ qend_of, u32 nb, u32 ahead) {
    const u32 lane = threadIdx.x & 31;
    if (nb >= n_blocks) return;
    if (lane == 0 && nb + ahead < n_blocks) { prefetch_l2(blocks + nb + ahead); prefetch_l2(qstart + nb + ahead); prefetch_l2(qend_of + nb + ahead); }
    if (!blocks[nb].owns) return;
    const u64 qs = qstart[nb], qe = qend_of[nb];
    if (qs >= qe) return;
    const u64 word = blocks[nb].word & ~7ull, kidx = blocks[nb].kidx;
    u64 nk = DB_W / 2;
    if (nb + 1 < n_blocks && blocks[nb + 1].owns && blocks[nb + 1].kidx > kidx) nk = blocks[nb + 1].kidx - kidx;
    if (nk > (u64) DB_C) nk = DB_C;
    if (kidx + nk > db.n_kmers_eff) nk = db.n_kmers_eff > kidx ? db.n_kmers_eff - kidx : 0;
    const char *w = reinterpret_cast<const char *>(db.words + word);
    u64 wbytes = (db.n_words + DB_PAD - word) * 2;
    if (wbytes > (u64) (DB_C + 16) * 2) wbytes = (u64) (DB_C + 16) * 2;
    for (u64 o = (u64) lane * 128; o < wbytes; o += 32 * 128) prefetch_l2(w + o);
    const char *inf = reinterpret_cast<const char *>(db.info + kidx);
    for (u64 o = (u64) lane * 128; o < nk * 4; o += 32 * 128) prefetch_l2(inf + o);
    u64 qn = qe - qs;
    if (qn > 8192) qn = 8192;
    const char *qp = reinterpret_cast<const char *>(q + qs);
    for (u64 o = (u64) lane * 128; o < qn * 16; o += 32 * 128) prefetch_l2(qp + o);
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

// Warp-level emission: lane l adds `cnt` matches (entries [lo,hi) of s.vals whose distance is <= maxh; the
// distances of the first 16 entries are cached 4 bits each in dpack).
template <bool ROUTED>
__device__ __forceinline__ void warp_emit(const DbView &db, const DbCursor &c, const MatchSmem &s, MatchTables &tb, const MatchOut &mo,
                                          bool active, const mb_kmer &q, u32 lo, u32 hi, u32 maxh, u32 cnt, u64 dpack, int format,
                                          u32 &fill) {
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u32 incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (u32) d) incl += t;
    }
    const u32 total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (total == 0) return;
    u64 *stage = tb.stage[warp];
    if (total > MT_STAGE) {
        // more than the staging buffer holds (a few very long groups): reserve and write straight to HBM
        warp_flush<ROUTED>(mo, stage, fill);
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(mo.counter, (unsigned long long) total);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (active && cnt) {
            u64 o = base + incl - cnt;
            for (u32 j = lo; j < hi; ++j) {
                const u64 t = s.vals[j];
                const u32 d = (j - lo < 16) ? (u32) (dpack >> (4 * (j - lo))) & 15u : hamming_sum_lut(tb.lut2, q.value, t);
                if (d > maxh) continue;
                if (ROUTED) {                                     // rare path: one reservation per record at its home
                    u64 rec[3];
                    make_match(db, mo, tb.lutp, q, t, c.kidx + j, d, format, rec);
                    const u32 h = route_home(*mo.route, qi_seq(rec[0]));
                    const unsigned long long slot = atomicAdd(mo.route->sent_m + h, 1ull);
                    if (slot < mo.route->match_cap[h]) {
                        u64 *p = reinterpret_cast<u64 *>(mo.route->matches[h]) + slot * 3;
                        p[0] = rec[0]; p[1] = rec[1]; p[2] = rec[2];
                    }
                } else if (o < mo.cap) {
                    u64 rec[3];
                    make_match(db, mo, tb.lutp, q, t, c.kidx + j, d, format, rec);
                    u64 *p = reinterpret_cast<u64 *>(mo.out + o);
                    p[0] = rec[0]; p[1] = rec[1]; p[2] = rec[2];
                }
                ++o;
            }
        }
        return;
    }
    if (fill + total > MT_STAGE) warp_flush<ROUTED>(mo, stage, fill);
    if (active && cnt) {
        u32 o = fill + incl - cnt;
        for (u32 j = lo; j < hi; ++j) {
            const u64 t = s.vals[j];
            const u32 d = (j - lo < 16) ? (u32) (dpack >> (4 * (j - lo))) & 15u : hamming_sum_lut(tb.lut2, q.value, t);
            if (d > maxh) continue;
            u64 rec[3];
            make_match(db, mo, tb.lutp, q, t, c.kidx + j, d, format, rec);
            stage[o * 3] = rec[0]; stage[o * 3 + 1] = rec[1]; stage[o * 3 + 2] = rec[2];
            ++o;
        }
    }
    fill += total;
    __syncwarp();
}

// Streaming two-pass join for an amino-acid group that does not fit one chunk: for every batch of
// 256 queries of that group the group is decoded twice from its start cursor (minimum, then emit).
template <bool ROUTED>
__device__ __noinline__ void giant_group(const DbView &db, const DbCursor &c, MatchSmem &s, MatchTables &tb, const MatchOut &mo,
                                         const mb_kmer *__restrict__ q, u64 qcur, u64 qend, u64 gaa, int format, u32 &fill) {
    const u32 tid = threadIdx.x;
    for (; qcur < qend; qcur += MT_THREADS) {
        const u64 qi = qcur + tid;
        mb_kmer qk;
        qk.value = 0; qk.qinfo = 0;
        bool mine = false;
        if (qi < qend) { qk = q[qi]; mine = (qk.value >> 24) == gaa; }
        if (__syncthreads_count(mine) == 0) continue;
        u32 mind = 255, maxh = 0;
        for (int phase = 0; phase < 2; ++phase) {
            DbCursor g = c;
            while (true) {
                const u32 n = decode_chunk(db, g, s);
                if (n == 0) break;
                if (tid == 0) s.group_end = n;
                __syncthreads();
                for (u32 j = tid; j < n; j += MT_THREADS)
                    if ((s.vals[j] >> 24) != gaa) atomicMin(&s.group_end, j);
                __syncthreads();
                const u32 ge = s.group_end;
                const bool finished = ge < n || (g.kidx + n == db.n_kmers_eff);
                if (phase == 0) {
                    if (mine)
                        for (u32 j = 0; j < ge; ++j) {
                            const u32 d = hamming_sum_lut(tb.lut2, qk.value, s.vals[j]);
                            mind = d < mind ? d : mind;
                        }
                } else {
                    u32 cnt = 0;
                    u64 dpack = 0;
                    if (mine)
                        for (u32 j = 0; j < ge; ++j) {
                            const u32 d = hamming_sum_lut(tb.lut2, qk.value, s.vals[j]);
                            if (j < 16) dpack |= (u64) (d < 15 ? d : 15) << (4 * j);
                            cnt += d <= maxh ? 1u : 0u;
                        }
                    warp_emit<ROUTED>(db, g, s, tb, mo, mine, qk, 0, ge, maxh, cnt, dpack, format, fill);
                }
                __syncthreads();
                if (finished) break;
                const u64 nw = g.word + (u64) s.wend[n - 1] + 1;
                const u64 nv = s.vals[n - 1];
                __syncthreads();
                g.word = nw; g.vbefore = nv; g.kidx += n;
            }
            if (phase == 0) maxh = mind * 2 < 7 ? mind * 2 : 7;
        }
    }
}

// first index in q[lo, hi) whose AA part is >= aa
__device__ __forceinline__ u64 query_lower_bound(const mb_kmer *__restrict__ q, u64 lo, u64 hi, u64 aa) {
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        if ((__ldg(&q[mid].value) >> 24) < aa) lo = mid + 1; else hi = mid;
    }
    return lo;
}

template <bool ROUTED>
__global__ void __launch_bounds__(MT_THREADS, 4)
k_match(DbView db, const BlockIndex *__restrict__ blocks, u32 n_blocks, const mb_kmer *__restrict__ q, const u64 *__restrict__ qstart,
        const u64 *__restrict__ qend_of, MatchOut mo_in, const Tables *__restrict__ tables, int format, u32 *block_ticket,
        const Route *__restrict__ route_g) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ Route s_route;
    MatchOut mo = mo_in;
    mo.route = nullptr;
    if (ROUTED) {                                   // multi-GPU: matches leave for their home rank (warp_flush_routed)
        for (u32 i = threadIdx.x; i < sizeof(Route) / 8; i += MT_THREADS) reinterpret_cast<u64 *>(&s_route)[i] = reinterpret_cast<const u64 *>(route_g)[i];
        mo.route = &s_route;
    }
    MatchSmem &s = *reinterpret_cast<MatchSmem *>(smem_raw);
    MatchTables &tb = *reinterpret_cast<MatchTables *>(smem_raw + ((sizeof(MatchSmem) + 15) & ~(size_t) 15));
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const uint4 *src2 = reinterpret_cast<const uint4 *>(tables->lut2);
        const uint4 *srcp = reinterpret_cast<const uint4 *>(tables->lutp);
        uint4 *d2 = reinterpret_cast<uint4 *>(tb.lut2);
        uint4 *dp = reinterpret_cast<uint4 *>(tb.lutp);
        for (u32 i = tid; i < 4096 / 16; i += MT_THREADS) { d2[i] = __ldg(src2 + i); dp[i] = __ldg(srcp + i); }
    }
    u32 fill = 0;                                   // matches staged by this warp (warp-uniform)

    // Block tickets are drawn one block ahead: tick[it & 1] is the block of iteration `it`, tick[(it + 1) & 1] the next one
    // (known from the start, so that it can be prefetched).  Thread 0 draws the ticket after that at the start of the
    // iteration, keeps it in a register and publishes it in the slot of the current block just before the closing barrier.
    if (tid == 0) { s.tick[0] = atomicAdd(block_ticket, 1u); s.tick[1] = atomicAdd(block_ticket, 1u); }
    __syncthreads();
    u32 b = s.tick[0];
    for (u32 it = 0; b < n_blocks; ++it) {
        const u32 nb = s.tick[(it + 1) & 1];
        u32 drawn = 0;
        if (tid == 0) drawn = atomicAdd(block_ticket, 1u);
        if (warp == MT_THREADS / 32 - 1) prefetch_block(db, blocks, n_blocks, q, qstart, qend_of, nb, gridDim.x);
        do {
        const BlockIndex bi = blocks[b];
        if (!bi.owns) break;
        const u64 qcur = qstart[b];
        const u64 qend = qend_of[b];
        if (qcur >= qend) break;                    // no query falls into this block's AA range

        const u64 tile_end = (u64) (b + 1) * DB_W;  // groups starting at or after this word belong to later blocks
        DbCursor c{bi.word, bi.vbefore, bi.kidx};

        while (true) {
            const u32 n = decode_chunk(db, c, s);
            if (n == 0) break;
            if (tid == 0) { s.last_start = 0; s.own_end = (int) n; s.qnext = 0; }
            __syncthreads();
            const u64 own_limit = tile_end > c.word ? tile_end - c.word : 0;   // local word offset of the block end
            {
                // one sweep over the group boundaries: last group start and first group that starts in a later block
                int lmax = 0, lmin = (int) n;
                for (u32 j = tid + 1; j < n; j += MT_THREADS) {
                    const u64 a = s.vals[j] >> 24, pa = s.vals[j - 1] >> 24;
                    if (a != pa) {
                        lmax = (int) j;
                        // first word of k-mer j is the word after the previous terminator
                        if ((u64) s.wend[j - 1] + 1 >= own_limit && (int) j < lmin) lmin = (int) j;
                    }
                }
                lmax = __reduce_max_sync(0xFFFFFFFFu, lmax);
                lmin = __reduce_min_sync(0xFFFFFFFFu, lmin);
                if (lane == 0) {
                    if (lmax) atomicMax(&s.last_start, lmax);
                    if (lmin < (int) n) atomicMin(&s.own_end, lmin);
                }
            }
            __syncthreads();
            const bool db_end = (c.kidx + n == db.n_kmers_eff);
            u32 proc_end = db_end ? n : (u32) s.last_start;     // the last group may continue past the chunk
            if ((u32) s.own_end < proc_end) proc_end = (u32) s.own_end;
            const u32 own_end = (u32) s.own_end;

            if (proc_end == 0) {
                // One group fills the whole chunk and is not known to be complete: streaming path.  The
                // group is owned (the cursor always sits on a group starting inside this block) and, being
                // longer than a block plus overhang, it is the last one this block owns.
                const u64 gaa = s.vals[0] >> 24;
                __syncthreads();
                giant_group<ROUTED>(db, c, s, tb, mo, q, qcur, qend, gaa, format, fill);
                break;
            }

            // ---- fast path: all groups in [0, proc_end) are complete and owned.  Warps run on their own; every query
            // of the block's range is looked up (the range is not ordered below the sorted bits, and it also holds
            // queries of the neighbouring blocks: those find no group of their amino-acid part here).  The query
            // record of the NEXT ticket is requested before the current one is processed. ----
            u32 base = warp * 32u;
            ulonglong2 v_next = make_ulonglong2(0, 0);
            if (qcur + base + lane < qend) v_next = __ldg(reinterpret_cast<const ulonglong2 *>(q + qcur + base + lane));
            while (qcur + base < qend) {
                mb_kmer qk;
                qk.value = v_next.x; qk.qinfo = v_next.y;
                const bool mine = qcur + base + lane < qend;
                {
                    base += MT_THREADS;                  // tickets are dealt round robin: warp w takes 32-query slices w, w + 8, ...
                    if (qcur + base + lane < qend) v_next = __ldg(reinterpret_cast<const ulonglong2 *>(q + qcur + base + lane));
                }
                // (1) every lane: candidate range [lo, lo + ncand) of its query = its amino-acid group in the block
                u32 lo = 0, ncand = 0;
                bool exact = false;                 // the candidates are copies of the query's value: distance 0, no table lookups
                if (mine) {
                    // binary search of the decoded block (measured on B200 against a 1024-bucket index and against a hash
                    // directory of the groups built per block: both cost more to build than the search they save)
                    // The search is on the FULL value.  When the block holds the query's value itself, the minimum distance
                    // of its group is 0, the filter min(2 * min, 7) = 0 keeps only distance 0, and distance 0 means the same
                    // codon ids, i.e. the same value (the Hamming table is 0 on its diagonal only): the candidates are just
                    // the run of equal values (35 % of the queries and 68 % of the matches of configs[1] end here, with one
                    // or two pairs instead of the whole group).  Otherwise the candidates are the amino-acid group around
                    // the insertion point.
                    const u64 v = qk.value, aa = v >> 24;
                    u32 l = 0, h = proc_end;
                    while (l < h) {
                        const u32 mid = (l + h) >> 1;
                        if (s.vals[mid] < v) l = mid + 1; else h = mid;
                    }
                    lo = l;
                    u32 hi = l;
                    if (l < proc_end && s.vals[l] == v) {
                        exact = true;
                        do ++hi; while (hi < proc_end && s.vals[hi] == v);
                    } else {
                        while (lo > 0 && (s.vals[lo - 1] >> 24) == aa) --lo;
                        while (hi < proc_end && (s.vals[hi] >> 24) == aa) ++hi;
                    }
                    ncand = hi - lo;
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
                const u32 q24 = (u32) qk.value & 0xFFFFFFu;
                u32 *qmin = tb.qmin[warp];
                qmin[lane] = exact ? 0u : 255u;
                const u32 lo_x = lo | (exact ? 0x80000000u : 0u);
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
                    const u32 o_lo = __shfl_sync(0xFFFFFFFFu, lo_x, owner);
                    const u32 o_q = __shfl_sync(0xFFFFFFFFu, q24, owner);
                    u32 dd = 255, j = 0;
                    if (p < total) {
                        j = (o_lo & 0x7FFFFFFFu) + (p - o_excl);
                        dd = 0;
                        // the candidate's info entry (make_match reads it in the emit pass) is asked for now
                        asm volatile("prefetch.global.L1 [%0];" ::"l"(db.info + c.kidx + j));
                        if (!(o_lo >> 31)) {
                            dd = hamming_sum_lut(tb.lut2, (u64) o_q, s.vals[j]);
                            atomicMin(&qmin[owner], dd);
                        }
                    }
                    const u32 pk = dd | (owner << 8) | (j << 16);
                    if (r == 0) pk0 = pk;
                    else if (r == 1) pk1 = pk;
                }
                __syncwarp();
                // (3) emit: distance <= min(2 * min, 7) (KmerMatcher.cpp:1126); one staging slot per surviving pair
                u64 *stage = tb.stage[warp];
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
                        const u32 o_lo = __shfl_sync(0xFFFFFFFFu, lo_x, owner);
                        const u32 o_q = __shfl_sync(0xFFFFFFFFu, q24, owner);
                        dd = 255; j = 0;
                        if (p < total) {
                            j = (o_lo & 0x7FFFFFFFu) + (p - o_excl);
                            dd = (o_lo >> 31) ? 0u : hamming_sum_lut(tb.lut2, (u64) o_q, s.vals[j]);
                        }
                    }
                    const u32 mind = qmin[owner];
                    const u32 maxh = mind * 2 < 7 ? mind * 2 : 7;
                    const bool ok = p < total && dd <= maxh;
                    const u32 okmask = __ballot_sync(0xFFFFFFFFu, ok);
                    // the owner's full query record (value and qinfo) for the match
                    mb_kmer oq;
                    oq.value = __shfl_sync(0xFFFFFFFFu, qk.value, owner);
                    oq.qinfo = __shfl_sync(0xFFFFFFFFu, qk.qinfo, owner);
                    if (okmask == 0) continue;
                    const u32 nok = (u32) __popc(okmask);
                    if (fill + nok > MT_STAGE) warp_flush<ROUTED>(mo, stage, fill);
                    if (ok) {
                        u64 rec[3];
                        make_match(db, mo, tb.lutp, oq, s.vals[j], c.kidx + j, dd, format, rec);
                        const u32 o = fill + (u32) __popc(okmask & ((1u << lane) - 1u));
                        stage[o * 3] = rec[0]; stage[o * 3 + 1] = rec[1]; stage[o * 3 + 2] = rec[2];
                    }
                    fill += nok;
                    __syncwarp();
                }
            }
            __syncthreads();
            if (proc_end >= n && db_end) break;                        // end of the stream
            if (proc_end >= own_end) break;                            // the next group starts in a later block
            // continue from the first group that was not processed (it starts inside this block)
            const u64 nw = c.word + (u64) s.wend[proc_end - 1] + 1;
            const u64 nv = s.vals[proc_end - 1];
            __syncthreads();
            c.word = nw; c.vbefore = nv; c.kidx += proc_end;
        }
        } while (false);
        if (tid == 0) s.tick[it & 1] = drawn;
        __syncthreads();                            // this block is done with the shared arrays; the new ticket is visible
        b = nb;
    }
    warp_flush<ROUTED>(mo, tb.stage[warp], fill);
}

constexpr size_t match_smem_bytes() { return ((sizeof(MatchSmem) + 15) & ~(size_t) 15) + sizeof(MatchTables); }

// ---- load-time block index ---------------------------------------------------------------------------
// pass A: per block, how many words in front of the block belong to the first k-mer ending in it, how
// many k-mers end in it and the sum of their deltas.
__global__ void __launch_bounds__(MT_THREADS)
k_db_block_summary(DbView db, u32 n_blocks, u32 *__restrict__ blk_back, u32 *__restrict__ blk_cnt, u64 *__restrict__ blk_sum) {
    __shared__ MatchSmem s;
    __shared__ u32 s_back;
    __shared__ u32 s_cnt;
    const u32 b = blockIdx.x, tid = threadIdx.x;
    if (b >= n_blocks) return;
    const u64 w0 = (u64) b * DB_W;
    if (tid == 0) {
        u32 back = 0;
        while (back < 4 && w0 > back && !(db.words[w0 - back - 1] & 0x8000u)) ++back;
        s_back = back;
        s_cnt = 0;
    }
    __syncthreads();
    const u32 back = s_back;
    DbCursor c{w0 - back, 0, 0};
    const u32 n = decode_chunk(db, c, s);
    const u32 limit = DB_W + back;                 // local offset of the block end
    // entries are ordered: count those whose terminator lies inside the block
    for (u32 j = tid; j < n; j += MT_THREADS)
        if (s.wend[j] < limit && (j + 1 == n || s.wend[j + 1] >= limit)) s_cnt = j + 1;
    __syncthreads();
    if (tid == 0) {
        const u32 cnt = s_cnt;
        blk_back[b] = back;
        blk_cnt[b] = cnt;
        blk_sum[b] = cnt ? s.vals[cnt - 1] : 0;
    }
}

// pass B (after the host scanned counts and sums into kidx/vbefore per block): first group starting in the block
__global__ void __launch_bounds__(MT_THREADS)
k_db_block_own(DbView db, u32 n_blocks, const u32 *__restrict__ blk_back, const u64 *__restrict__ blk_kidx,
               const u64 *__restrict__ blk_vbefore, BlockIndex *__restrict__ out) {
    __shared__ MatchSmem s;
    __shared__ u32 s_first;
    const u32 b = blockIdx.x, tid = threadIdx.x;
    if (b >= n_blocks) return;
    const u32 back = blk_back[b];
    DbCursor c{(u64) b * DB_W - back, blk_vbefore[b], blk_kidx[b]};
    if (tid == 0) s_first = 0xFFFFFFFFu;
    const u32 n = decode_chunk(db, c, s);
    const u32 lo = back, hi = DB_W + back;         // local word range of the block
    for (u32 j = tid; j < n; j += MT_THREADS) {
        const u32 first_word = j == 0 ? 0u : (u32) s.wend[j - 1] + 1u;
        if (first_word < lo || first_word >= hi) continue;
        const bool starts = (j == 0) ? (c.kidx == 0 || (s.vals[0] >> 24) != (c.vbefore >> 24))
                                     : ((s.vals[j] >> 24) != (s.vals[j - 1] >> 24));
        if (starts) atomicMin(&s_first, j);
    }
    __syncthreads();
    if (tid == 0) {
        BlockIndex bi;
        const u32 j = s_first;
        if (j == 0xFFFFFFFFu) { bi.word = 0; bi.vbefore = 0; bi.kidx = 0; bi.own_aa = ~0ull; bi.owns = 0; }
        else {
            bi.word = c.word + (j == 0 ? 0u : (u32) s.wend[j - 1] + 1u);
            bi.vbefore = j == 0 ? c.vbefore : s.vals[j - 1];
            bi.kidx = c.kidx + j;
            bi.own_aa = s.vals[j] >> 24;
            bi.owns = 1;
        }
        bi.pad = 0;
        out[b] = bi;
    }
}

// plain decode of the resident stream (parity tap): values of the k-mers ending in each block
__global__ void __launch_bounds__(MT_THREADS)
k_db_decode(DbView db, u32 n_blocks, const u32 *__restrict__ blk_back, const u32 *__restrict__ blk_cnt,
            const u64 *__restrict__ blk_kidx, const u64 *__restrict__ blk_vbefore, u64 *__restrict__ values) {
    __shared__ MatchSmem s;
    const u32 b = blockIdx.x, tid = threadIdx.x;
    if (b >= n_blocks) return;
    DbCursor c{(u64) b * DB_W - blk_back[b], blk_vbefore[b], blk_kidx[b]};
    decode_chunk(db, c, s);
    const u32 cnt = blk_cnt[b];
    for (u32 j = tid; j < cnt; j += MT_THREADS) values[c.kidx + j] = s.vals[j];
}

// ---- per-batch partition: the query range a block has to look at ------------------------------------------
// The queries are ordered by T(aa) only (SortT, mb_common.cuh: leading bits or the re-coded 24-bit key).  Block b owns the amino-acid groups [own_aa(b), own_aa(b+1)),
// so it scans the queries with T(own_aa(b)) <= T(aa) <= T(own_aa(b+1)): the buckets at both ends are shared with the
// neighbouring blocks, which is harmless because a query only finds its group in the block that owns it.
__device__ __forceinline__ u64 query_lower_bound_t(const mb_kmer *__restrict__ q, u64 nq, const SortT &st, u64 t) {
    u64 lo = 0, hi = nq;
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        if (sort_t(st, __ldg(&q[mid].value)) < t) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__global__ void k_match_partition(const BlockIndex *__restrict__ blocks, u32 n_blocks, const mb_kmer *__restrict__ q, u64 nq, SortT st,
                                  u64 *__restrict__ qstart, u64 *__restrict__ qend) {
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    if (!blocks[b].owns) { qstart[b] = 0; qend[b] = 0; return; }
    qstart[b] = query_lower_bound_t(q, nq, st, sort_t(st, blocks[b].own_aa << 24));
    const u64 nxt = b + 1 < n_blocks ? blocks[b + 1].own_aa : ~0ull;
    qend[b] = nxt == ~0ull ? nq : query_lower_bound_t(q, nq, st, sort_t(st, nxt << 24) + (sort_t_lossy(st) ? 1 : 0));
}

} // namespace mb
