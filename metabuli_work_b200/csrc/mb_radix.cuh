// mb_radix.cuh — K2/K5: hand-written LSD radix sort (8-bit digits, one read + one write per pass).
//
// Replaces SORT_PARALLEL(..., Kmer::compareQueryKmer) (src/commons/KmerExtractor.cpp:79, Kmer.h:89-94)
// and KmerMatcher::sortMatches / compareMatches (src/commons/KmerMatcher.cpp:1071-1078,1149-1166), and
// serves as the stable multi-way partition of the multi-GPU exchange.
//
// B200 design ("one-sweep"): the digit histograms of ALL passes are taken in one read of the input;
// each pass is then a single kernel in which a CTA ranks a tile of records (warp-level match_any
// multisplit, stable), publishes its per-digit counts, resolves its global offsets by decoupled
// look-back over the preceding tiles, stages the tile in digit order in shared memory and writes
// runs to HBM.  Traffic per pass is exactly one read and one write of the records.  Tiles are
// handed out by an atomic ticket so a tile only ever waits on tiles that are already running.
#pragma once
#include "mb_common.cuh"

namespace mb {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_BINS = 256;
constexpr int RS_MAX_PASSES = 16;

constexpr u64 RS_FLAG_AGG = 1ull << 62;   // tile aggregate published
constexpr u64 RS_FLAG_INC = 2ull << 62;   // inclusive prefix published
constexpr u64 RS_VALUE_MASK = (1ull << 62) - 1;

// ---- key functors: digit(record, pass) in [0,256) -------------------------------------------------
struct KmerKey {            // sort by the amino-acid part and then the DNA part of Kmer::value
    int first_bit;          // 24 = order the 40 AA bits only (all the join needs); 0 = full value
    __device__ __forceinline__ u32 digit(const mb_kmer &r, int pass) const { return (u32) (r.value >> (first_bit + 8 * pass)) & 255u; }
};
struct KmerKeyPacked {      // digits of the 24-bit re-coded key sort_t_packed (mb_common.cuh), format 2 only
    __device__ __forceinline__ u32 digit(const mb_kmer &r, int pass) const { return (sort_t_packed(r.value) >> (8 * pass)) & 255u; }
};

struct KmerKeyPacked7 {     // digits of the 32-bit re-coded key sort_t_packed7 (seven residues), format 2 only
    __device__ __forceinline__ u32 digit(const mb_kmer &r, int pass) const { return (sort_t_packed7(r.value) >> (8 * pass)) & 255u; }
};

struct KmerOwnerKey {       // multi-GPU bucketing: owner shard = number of bounds <= value
    u64 bounds[8];
    int n_bounds;
    __device__ __forceinline__ u32 digit(const mb_kmer &r, int) const {
        u32 o = 0;
        for (int i = 0; i < n_bounds; ++i) o += (r.value >= bounds[i]) ? 1u : 0u;
        return o;
    }
};

// compareMatches order: (sequenceID, speciesId, frame, pos, hamming, dnaEncoding), packed into one
// integer with per-batch field widths so that only the significant bytes are sorted.
struct MatchKey {
    int pos_bits, sp_bits, seq_bits;
    u32 seq_base;          // ids of this batch are seq_base+1 .. seq_base+n_reads
    __device__ __forceinline__ unsigned __int128 key(const mb_match &m) const {
        unsigned __int128 k = (u32) (qi_seq(m.qinfo) - seq_base);
        k = (k << sp_bits) | (u32) m.speciesId;
        k = (k << 3) | qi_frame(m.qinfo);
        k = (k << pos_bits) | qi_pos(m.qinfo);
        k = (k << 3) | (m.hamming & 7u);
        k = (k << 24) | (m.dnaEncoding & 0xFFFFFFu);
        return k;
    }
    int skip_bits = 0;     // 24: the radix passes leave the DNA bits out; k_fix_dna_ties orders the (short) runs that tie above them
    __device__ __forceinline__ u32 digit(const mb_match &m, int pass) const { return (u32) (key(m) >> (skip_bits + 8 * pass)) & 255u; }
    int total_bits() const { return 24 + 3 + pos_bits + 3 + sp_bits + seq_bits; }
};

// After a full-key sort without the 24 DNA bits: records that agree in everything above them form runs (same read, species,
// frame, position and Hamming sum: a handful of DB k-mers at most); the first record of a run orders it by the DNA bits with
// an insertion sort in place.  Three radix passes over every record become one read of the keys.
__global__ void __launch_bounds__(256) k_fix_dna_ties(mb_match *__restrict__ m, u64 n, MatchKey key) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const unsigned __int128 hi = key.key(m[i]) >> 24;
    if (i && (key.key(m[i - 1]) >> 24) == hi) return;          // not the head of its run
    u64 e = i + 1;
    while (e < n && (key.key(m[e]) >> 24) == hi) ++e;
    if (e - i < 2) return;
    for (u64 a = i + 1; a < e; ++a) {
        const mb_match x = m[a];
        const u32 dx = x.dnaEncoding & 0xFFFFFFu;
        u64 b = a;
        while (b > i && (m[b - 1].dnaEncoding & 0xFFFFFFu) > dx) { m[b] = m[b - 1]; --b; }
        if (b != a) m[b] = x;
    }
}

struct MatchHomeKey {       // multi-GPU: home rank of a match = (sequenceID-1) / reads_per_rank
    u32 reads_per_rank;
    __device__ __forceinline__ u32 digit(const mb_match &m, int) const { return (qi_seq(m.qinfo) - 1u) / reads_per_rank; }
};

// ---- record moves ------------------------------------------------------------------------------------
// 16-byte records travel as one 128-bit access (their arrays are 16-byte aligned: cudaMalloc bases, region sizes in
// whole records); other sizes as 64-bit words.
template <typename Rec>
__device__ __forceinline__ void rec_load(Rec &dst, const Rec *src) {
    if constexpr (sizeof(Rec) == 16) *reinterpret_cast<ulonglong2 *>(&dst) = *reinterpret_cast<const ulonglong2 *>(src);
    else dst = *src;
}
template <typename Rec>
__device__ __forceinline__ void rec_store(Rec *dst, const Rec &src) {
    if constexpr (sizeof(Rec) == 16) *reinterpret_cast<ulonglong2 *>(dst) = *reinterpret_cast<const ulonglong2 *>(&src);
    else *dst = src;
}

// ---- histogram of every pass in one read -----------------------------------------------------------
template <typename Rec, typename Key, bool SEG>
__global__ void __launch_bounds__(RS_THREADS)
k_radix_hist(const Rec *__restrict__ in, u64 n, Key key, int n_passes, u64 *__restrict__ hist /*[n_passes][256]*/,
             const __grid_constant__ SegIn seg) {
    extern __shared__ u32 s_hist[]; // n_passes * 256
    for (int i = threadIdx.x; i < n_passes * RS_BINS; i += RS_THREADS) s_hist[i] = 0;
    __syncthreads();
    const u64 stride = (u64) gridDim.x * RS_THREADS;
    for (u64 i = (u64) blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
        const Rec r = SEG ? *seg_ptr(seg, in, i) : in[i];
        for (int p = 0; p < n_passes; ++p) atomicAdd(&s_hist[p * RS_BINS + key.digit(r, p)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_passes * RS_BINS; i += RS_THREADS) {
        const u32 v = s_hist[i];
        if (v) atomicAdd(reinterpret_cast<unsigned long long *>(hist + i), (unsigned long long) v);
    }
}

// exclusive scan of each pass' histogram (one CTA per pass, 256 threads); also flags trivial passes
__global__ void __launch_bounds__(RS_BINS) k_radix_scan_hist(u64 *__restrict__ hist, u64 n, u32 *__restrict__ trivial) {
    __shared__ u64 s[RS_BINS];
    const int p = blockIdx.x, d = threadIdx.x;
    const u64 v = hist[p * RS_BINS + d];
    s[d] = v;
    __syncthreads();
    // simple Hillis-Steele over 256 entries
    for (int off = 1; off < RS_BINS; off <<= 1) {
        u64 t = d >= off ? s[d - off] : 0;
        __syncthreads();
        s[d] += t;
        __syncthreads();
    }
    hist[p * RS_BINS + d] = s[d] - v;
    if (v == n && n > 0) trivial[p] = 1; // every record shares this digit: the pass is the identity
}

// ---- one pass -----------------------------------------------------------------------------------
// resident CTAs per SM the register allocation aims at: tiles of up to 48 KB fit three times into an SM
template <typename Rec, int ITEMS>
constexpr int radix_min_blocks() { return sizeof(Rec) * ITEMS <= 192 ? 3 : 2; }

template <typename Rec, typename Key, int ITEMS, bool SEG>
__global__ void __launch_bounds__(RS_THREADS, radix_min_blocks<Rec, ITEMS>())
k_radix_pass(const Rec *__restrict__ in, Rec *__restrict__ out, u64 n, Key key, int pass,
             const u64 *__restrict__ digit_base /*[256] exclusive*/, u64 *tile_state /*[tiles][256]*/,
             u32 *ticket, int dbg, const __grid_constant__ SegIn seg) {
    constexpr int TILE = RS_THREADS * ITEMS;
    extern __shared__ __align__(16) unsigned char s_raw[];
    Rec *s_rec = reinterpret_cast<Rec *>(s_raw);                                   // TILE records
    u32 *s_whist = reinterpret_cast<u32 *>(s_raw + sizeof(Rec) * TILE);            // [RS_WARPS][256]
    u32 *s_binstart = s_whist + RS_WARPS * RS_BINS;                                // [256]
    u64 *s_gbase = reinterpret_cast<u64 *>(s_binstart + RS_BINS);                  // [256]
    u8 *s_dig = reinterpret_cast<u8 *>(s_gbase + RS_BINS);                         // [TILE] digit of every staged record
    __shared__ u32 s_tile;
    __shared__ u32 s_scan[RS_WARPS];

    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < RS_WARPS * RS_BINS; i += RS_THREADS) s_whist[i] = 0;
    __syncthreads();
    const u32 tile = s_tile;
    const u64 tile_base = (u64) tile * TILE;

    // warp-striped load: item j of lane l is record tile_base + warp*32*ITEMS + j*32 + l
    Rec rec[ITEMS];
    u32 dig[ITEMS];
    u32 rank[ITEMS];
    const u64 wbase = tile_base + (u64) warp * 32 * ITEMS;
    // SEG: the input lies in a few regions (SegIn); a tile that falls inside one region — all but n-1 tiles — reads
    // through one rebased pointer, so that its loads stay one independent batch
    const Rec *in_eff = in;
    bool split = false;
    if (SEG) {
        u32 r = 0;
        while (r + 1 < seg.n && tile_base >= seg.prefix[r + 1]) ++r;
        in_eff = reinterpret_cast<const Rec *>(seg.base[r]) - seg.prefix[r];
        split = r + 1 < seg.n && tile_base + TILE > seg.prefix[r + 1];
    }
    // tiles are taken in ticket order, so the tile that a CTA starting about one machine-full of CTAs later will load is
    // known now: ask L2 for it (that CTA's load phase then waits on L2 instead of HBM)
    if (!SEG) {
        constexpr u64 AHEAD = 148ull;                // tiles: about one machine-full of CTAs ahead on a B200 (a prefetch distance, not a grid size)
        const u64 pf_base = (tile_base + AHEAD * TILE) * sizeof(Rec);
        const u64 end_bytes = n * sizeof(Rec);
        for (u64 o = pf_base + (u64) tid * 128; o < pf_base + (u64) TILE * sizeof(Rec) && o < end_bytes; o += (u64) RS_THREADS * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(in) + o));
    }
    // all loads first (one independent batch per thread), digits afterwards
    if (SEG && split) {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const u64 idx = wbase + (u64) j * 32 + lane;
            if (idx < n) rec_load(rec[j], seg_ptr(seg, in, idx));
        }
    } else {
#pragma unroll
        for (int j = 0; j < ITEMS; ++j) {
            const u64 idx = wbase + (u64) j * 32 + lane;
            if (idx < n) rec_load(rec[j], in_eff + idx);
        }
    }
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u64 idx = wbase + (u64) j * 32 + lane;
        dig[j] = idx < n ? key.digit(rec[j], pass) : 0xFFFFFFFFu;
    }
    // early counts: the tile's digit histogram is published BEFORE the ranking, so that by the time this tile looks
    // back its predecessors (which started earlier) have long published theirs
    __shared__ u32 s_cnt[RS_BINS];
    s_cnt[tid] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < ITEMS; ++j)
        if (dig[j] != 0xFFFFFFFFu) atomicAdd(&s_cnt[dig[j]], 1u);
    __syncthreads();
    volatile u64 *state = tile_state;
    const u32 my_cnt = s_cnt[tid];
    state[(u64) tile * RS_BINS + tid] = (tile == 0 ? RS_FLAG_INC : RS_FLAG_AGG) | my_cnt;

    // stable ranking inside the warp: items are ordered by (j, lane) == memory order.  The leader of every group of
    // equal digits advances the warp's running count with a shared-memory atomic and hands the old value to its peers.
    u32 *wh = s_whist + warp * RS_BINS;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 d = dig[j];
        // lanes holding the same digit: eight ballots (one per digit bit) — measured faster than __match_any_sync
        // here (3.8 -> 2.6 ms per pass); padding lanes (digit 0xFFFFFFFF) only ever match each other
        u32 peers = __ballot_sync(0xFFFFFFFFu, d == 0xFFFFFFFFu);
        if (d != 0xFFFFFFFFu) peers = ~peers;
        if (dbg == 3) peers = 1u << lane;
        else if (dbg == 5) peers = __match_any_sync(0xFFFFFFFFu, d);
        else {
#pragma unroll
            for (int bit = 0; bit < 8; ++bit) {
                const u32 bal = __ballot_sync(0xFFFFFFFFu, (d >> bit) & 1u);
                peers &= ((d >> bit) & 1u) ? bal : ~bal;
            }
        }
        const int leader = __ffs(peers) - 1;
        u32 pre = 0;
        if (d != 0xFFFFFFFFu && lane == (u32) leader) pre = atomicAdd(&wh[d], (u32) __popc(peers));
        pre = __shfl_sync(0xFFFFFFFFu, pre, leader);
        rank[j] = pre + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();

    // per digit (thread == digit): exclusive scan over the warps, look-back
    {
        const u32 d = tid;
        u32 sum = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            const u32 t = s_whist[w * RS_BINS + d];
            s_whist[w * RS_BINS + d] = sum;
            sum += t;
        }
        u64 excl = 0;
        if (tile > 0 && dbg != 1) {
            // predecessors are read four at a time (independent loads); a flag of 0 means "not published yet"
            long long p = (long long) tile - 1;
            bool done = false;
            while (!done) {
                u64 v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) { v[k] = 2ull << 62; if (p - k >= 0) v[k] = state[(u64) (p - k) * RS_BINS + d]; }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (done) break;
                    if ((v[k] >> 62) == 0) break;        // spin: reload from here
                    excl += v[k] & RS_VALUE_MASK;
                    --p;
                    if (v[k] & RS_FLAG_INC) done = true;
                }
            }
            state[(u64) tile * RS_BINS + d] = RS_FLAG_INC | (excl + sum);
        }
        // block exclusive scan of the tile's digit counts
        u32 incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= (u32) o) incl += t;
        }
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        u32 wb = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) if ((u32) w < warp) wb += s_scan[w];
        const u32 start = wb + incl - sum;
        s_binstart[d] = start;
        s_gbase[d] = digit_base[d] + excl - start;   // out index = s_gbase[digit] + position in the staged tile
    }
    __syncthreads();

    // stage the tile in digit order; the digit of every staged record is kept beside it
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 d = dig[j];
        if (d != 0xFFFFFFFFu) {
            const u32 pos = s_binstart[d] + s_whist[warp * RS_BINS + d] + rank[j];
            rec_store(s_rec + pos, rec[j]);
            s_dig[pos] = (u8) d;
        }
    }
    __syncthreads();
    const u32 valid = (u32) (n - tile_base < (u64) TILE ? n - tile_base : (u64) TILE);
    if constexpr (sizeof(Rec) == 16) {
        // one 128-bit store per record: a warp writes 512 contiguous bytes of a run
        for (u32 i = tid; i < valid; i += RS_THREADS) {
            Rec r;
            rec_load(r, s_rec + i);
            rec_store(dbg == 2 ? out + tile_base + i : out + s_gbase[s_dig[i]] + i, r);
        }
    } else {
        // other record sizes leave word by word: consecutive lanes write consecutive 8-byte words of a run (a record-wise
        // store of a 24-byte struct is three strided stores that each touch every sector of the span)
        constexpr u32 W = sizeof(Rec) / 8;
        static_assert(sizeof(Rec) % 8 == 0, "records are whole 64-bit words");
        const u64 *s_w = reinterpret_cast<const u64 *>(s_rec);
        u64 *out_w = reinterpret_cast<u64 *>(out);
        for (u32 wi = tid; wi < valid * W; wi += RS_THREADS) {
            const u32 r = wi / W, c = wi - r * W;
            const u64 dst = dbg == 2 ? tile_base + r : s_gbase[s_dig[r]] + r;
            out_w[dst * W + c] = s_w[wi];
        }
    }
}

template <typename Rec, int ITEMS>
constexpr size_t radix_pass_smem() {
    return sizeof(Rec) * RS_THREADS * ITEMS + sizeof(u32) * (RS_WARPS * RS_BINS + RS_BINS) + sizeof(u64) * RS_BINS + RS_THREADS * ITEMS;
}


// ---- one pass, second form: the tile stays in shared memory ------------------------------------------------
// k_radix_pass keeps a thread's records in registers (126 of them: 16 warps per SM, and every warp issues on half of its
// cycles while it waits on its own ballots and shared-memory atomics).  Here the tile is brought in by ONE bulk copy
// (cp.async.bulk global -> shared, completion on an mbarrier) and never enters the register file: a thread reads the
// keys of its items from shared memory, ranks them (same stable warp multisplit), and the tile leaves through a
// permutation — slot p of the digit-ordered tile holds (digit << 16 | source index) — so that 512 threads x 2 CTAs
// = 32 warps per SM fit (<= 64 registers), twice the warps to cover the same latencies.  Tiles are as large as before
// (same look-back chain, same tile_state).  Whole tiles only take the bulk copy; the last, partial tile is copied by the
// threads, and so are tiles that straddle two regions of a segmented input (peer receive regions) or start off a 16-byte
// boundary there.
#ifndef MB_R2_THREADS
#define MB_R2_THREADS 512
#endif
#ifndef MB_R2_LB
#define MB_R2_LB 6
#endif

constexpr int R2_THREADS = MB_R2_THREADS;  // 512 x 2 CTAs or 256 x 4 CTAs per SM
constexpr int R2_WARPS = R2_THREADS / 32;
constexpr int R2_LB = MB_R2_LB;            // predecessors read per look-back round (independent loads)
constexpr u32 R2_PAD = 0x100u;              // digit of a padding slot of the partial tile (bit 8 keeps it apart from digit 0)

// peers &= lanes whose digit agrees with mine in bit BIT (whole warp converged)
template <int BIT>
__device__ __forceinline__ void peers_bit(u32 &peers, u32 d) {
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 ".reg .b32 m, t;\n"
                 "and.b32 t, %1, %2;\n"
                 "setp.ne.u32 p, t, 0;\n"
                 "vote.sync.ballot.b32 m, p, 0xffffffff;\n"
                 "@!p not.b32 m, m;\n"
                 "and.b32 %0, %0, m;\n"
                 "}\n" : "+r"(peers) : "r"(d), "n"(1 << BIT));
}

template <typename Rec, int ITEMS>
constexpr size_t radix_pass2_smem() {
    return sizeof(Rec) * R2_THREADS * ITEMS + sizeof(u64) * RS_BINS + sizeof(u32) * R2_THREADS * ITEMS +
           sizeof(u32) * R2_WARPS * RS_BINS;
}

template <typename Rec, typename Key, int ITEMS, bool SEG>
__global__ void __launch_bounds__(R2_THREADS, 1024 / R2_THREADS)
k_radix_pass2(const Rec *__restrict__ in, Rec *__restrict__ out, u64 n, Key key, int pass,
              const u64 *__restrict__ digit_base /*[256] exclusive*/, u64 *tile_state /*[tiles][256]*/, u32 *ticket,
              const __grid_constant__ SegIn seg) {
    constexpr int TILE = R2_THREADS * ITEMS;
    constexpr u32 W = sizeof(Rec) / 8;
    static_assert(sizeof(Rec) % 8 == 0 && (sizeof(Rec) * TILE) % 16 == 0 && TILE <= 65536, "tile: whole 16-byte units, 16-bit indices");
    extern __shared__ __align__(128) unsigned char s_raw[];
    Rec *s_rec = reinterpret_cast<Rec *>(s_raw);                                   // TILE records, in input order
    u64 *s_gbase = reinterpret_cast<u64 *>(s_raw + sizeof(Rec) * TILE);            // [256]
    u32 *s_perm = reinterpret_cast<u32 *>(s_gbase + RS_BINS);                      // [TILE] digit << 16 | source index
    u32 *s_whist = s_perm + TILE;                                                  // [R2_WARPS][256]
    __shared__ u32 s_tile;
    __shared__ u32 s_scan[RS_BINS / 32];
    __shared__ u32 s_cnt[RS_BINS];
    __shared__ __align__(8) u64 s_mbar;

    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_tile = atomicAdd(ticket, 1u);
        mbar_init(&s_mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = tid; i < R2_WARPS * RS_BINS / 4; i += R2_THREADS) reinterpret_cast<uint4 *>(s_whist)[i] = make_uint4(0, 0, 0, 0);
    if (tid < RS_BINS) s_cnt[tid] = 0;
    __syncthreads();
    const u32 tile = s_tile;
    const u64 tile_base = (u64) tile * TILE;
    const u32 valid = (u32) (n - tile_base < (u64) TILE ? n - tile_base : (u64) TILE);
    const bool full = valid == (u32) TILE;
    // SEG: the input lies in a few regions (SegIn); a tile inside one region reads through one rebased pointer
    const Rec *src = in + tile_base;
    bool split = false;
    if (SEG) {
        u32 r = 0;
        while (r + 1 < seg.n && tile_base >= seg.prefix[r + 1]) ++r;
        src = reinterpret_cast<const Rec *>(seg.base[r]) + (tile_base - seg.prefix[r]);
        split = r + 1 < seg.n && tile_base + valid > seg.prefix[r + 1];
    }
    // the bulk copy wants a whole tile at a 16-byte boundary (24-byte records inside a region start at one every other time)
    const bool bulk = full && !split && (reinterpret_cast<size_t>(src) & 15) == 0;
    if (bulk) {
        if (tid == 0) {
            constexpr u32 BYTES = (u32) (sizeof(Rec) * TILE), CHUNK = 16384;
            mbar_expect_tx(&s_mbar, BYTES);
            const char *srcb = reinterpret_cast<const char *>(src);
            for (u32 o = 0; o < BYTES; o += CHUNK) bulk_g2s(s_raw + o, srcb + o, BYTES - o < CHUNK ? BYTES - o : CHUNK, &s_mbar);
        }
    } else if (SEG && split) {
        for (u32 i = tid; i < valid; i += R2_THREADS) s_rec[i] = *seg_ptr(seg, in, tile_base + i);
    } else {
        const u64 *in_w = reinterpret_cast<const u64 *>(src);
        u64 *s_w = reinterpret_cast<u64 *>(s_rec);
        for (u32 wi = tid; wi < valid * W; wi += R2_THREADS) s_w[wi] = in_w[wi];
    }
    if (!SEG) {   // the tile a CTA starting about one machine-full of CTAs later will ask for: into L2 now
        constexpr u64 AHEAD = 148ull * 4096 / TILE;
        constexpr u32 TILE_BYTES = (u32) (sizeof(Rec) * TILE);
        const u64 pf_base = (tile_base + AHEAD * TILE) * sizeof(Rec);
        const u64 end_bytes = n * sizeof(Rec);
        if (pf_base + TILE_BYTES <= end_bytes) {
#pragma unroll
            for (u32 o = 0; o < TILE_BYTES; o += R2_THREADS * 128)
                if (o + tid * 128 < TILE_BYTES) prefetch_l2(reinterpret_cast<const char *>(in) + pf_base + o + tid * 128);
        }
    }
    if (bulk) mbar_wait(&s_mbar, 0);
    else __syncthreads();

    // digits of my items (warp-striped: item j of lane l is tile record warp*32*ITEMS + j*32 + l)
    u32 dig[ITEMS];
    u32 rank[ITEMS];
    const u32 wbase = warp * 32 * ITEMS;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 idx = wbase + (u32) j * 32 + lane;
        dig[j] = (full || idx < valid) ? key.digit(s_rec[idx], pass) : R2_PAD;
    }
    // early counts: the tile's digit histogram is published BEFORE the ranking (successors look back at it), and the
    // positions of the digits in the digit-ordered tile are scanned now, off the critical path
    volatile u64 *state = tile_state;
#pragma unroll
    for (int j = 0; j < ITEMS; ++j)
        if (dig[j] != R2_PAD) atomicAdd(&s_cnt[dig[j]], 1u);
    __syncthreads();
    u32 cnt_d = 0, incl_c = 0;
    if (tid < RS_BINS) {
        cnt_d = s_cnt[tid];
        state[(u64) tile * RS_BINS + tid] = (tile == 0 ? RS_FLAG_INC : RS_FLAG_AGG) | cnt_d;
        incl_c = cnt_d;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl_c, o);
            if (lane >= (u32) o) incl_c += t;
        }
        if (lane == 31) s_scan[warp] = incl_c;
    }

    // stable ranking inside the warp, items ordered by (j, lane) == memory order
    u32 *wh = s_whist + warp * RS_BINS;
    const u32 lt = (1u << lane) - 1u;
    // Three sweeps, so that nothing waits on a shared-memory atomic it has just issued: (1) per item the lanes holding
    // the same digit — one ballot per digit bit (test, vote, conditional complement, and); (2) the first lane of every
    // group advances the warp's running count of the digit (one predicated atomic per item, issued back to back: they
    // reach shared memory in program order); (3) the old count travels to the group, plus the lane's rank in it.
    u32 peers[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 d = dig[j];
        u32 p = 0xFFFFFFFFu;
        peers_bit<0>(p, d); peers_bit<1>(p, d); peers_bit<2>(p, d); peers_bit<3>(p, d);
        peers_bit<4>(p, d); peers_bit<5>(p, d); peers_bit<6>(p, d); peers_bit<7>(p, d);
        if (!full) peers_bit<8>(p, d);               // (block-uniform) padding slots only ever match each other
        peers[j] = p;
    }
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 d = dig[j];
        const u32 first = (d != R2_PAD && (peers[j] & lt) == 0u) ? 1u : 0u;
        u32 pre = 0;
        asm volatile("{\n"
                     ".reg .pred p;\n"
                     "setp.ne.u32 p, %3, 0;\n"
                     "@p atom.shared.add.u32 %0, [%1], %2;\n"
                     "}\n" : "+r"(pre) : "r"(smem_u32(wh + (d & 255u))), "r"((u32) __popc(peers[j])), "r"(first) : "memory");
        rank[j] = pre;
    }
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const int leader = __ffs(peers[j]) - 1;
        rank[j] = __shfl_sync(0xFFFFFFFFu, rank[j], leader) + __popc(peers[j] & lt);
    }
    __syncthreads();

    // per digit (thread == digit): where each warp's run of the digit starts in the digit-ordered tile, look-back
    if (tid < RS_BINS) {
        const u32 d = tid;
        u32 start = incl_c - cnt_d;
#pragma unroll
        for (int w = 0; w < RS_BINS / 32; ++w) if ((u32) w < warp) start += s_scan[w];
        u32 run = start;
#pragma unroll
        for (int w = 0; w < R2_WARPS; ++w) {
            const u32 t = s_whist[w * RS_BINS + d];
            s_whist[w * RS_BINS + d] = run;
            run += t;
        }
        u64 excl = 0;
        if (tile > 0) {
            long long p = (long long) tile - 1;
            bool done = false;
            while (!done) {
                u64 v[R2_LB];
#pragma unroll
                for (int k = 0; k < R2_LB; ++k) { v[k] = 2ull << 62; if (p - k >= 0) v[k] = state[(u64) (p - k) * RS_BINS + d]; }
#pragma unroll
                for (int k = 0; k < R2_LB; ++k) {
                    if (done) break;
                    if ((v[k] >> 62) == 0) break;        // not published yet: reload from here
                    excl += v[k] & RS_VALUE_MASK;
                    --p;
                    if (v[k] & RS_FLAG_INC) done = true;
                }
            }
            state[(u64) tile * RS_BINS + d] = RS_FLAG_INC | (excl + cnt_d);
        }
        s_gbase[d] = digit_base[d] + excl - start;       // out index = s_gbase[digit] + slot in the digit-ordered tile
    }
    __syncthreads();

    // the permutation: slot of every item in the digit-ordered tile
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const u32 d = dig[j];
        if (d != R2_PAD) {
            const u32 pos = s_whist[warp * RS_BINS + d] + rank[j];
            s_perm[pos] = (d << 16) | (wbase + (u32) j * 32 + lane);
        }
    }
    __syncthreads();
    if constexpr (sizeof(Rec) == 16) {
        for (u32 i = tid; i < valid; i += R2_THREADS) {
            const u32 pk = s_perm[i];
            Rec r;
            rec_load(r, s_rec + (pk & 0xFFFFu));
            rec_store(out + s_gbase[pk >> 16] + i, r);
        }
    } else {
        // other record sizes leave word by word: consecutive lanes write consecutive 8-byte words of a run
        const u64 *s_w = reinterpret_cast<const u64 *>(s_rec);
        u64 *out_w = reinterpret_cast<u64 *>(out);
        for (u32 wi = tid; wi < valid * W; wi += R2_THREADS) {
            const u32 r = wi / W, c = wi - r * W;
            const u32 pk = s_perm[r];
            out_w[(s_gbase[pk >> 16] + r) * W + c] = s_w[(pk & 0xFFFFu) * W + c];
        }
    }
}

} // namespace mb
