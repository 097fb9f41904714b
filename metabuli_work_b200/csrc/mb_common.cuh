// mb_common.cuh — shared types, lookup tables and small device helpers of libmetabuli_b200.
//
// Every table here restates a table of the reference (cited per item; paths relative to the
// reference tree).  Functions marked MB_HD are pure per-element logic: they compile for the device
// (product) and for the host (tests/emu only — a checker build that lets the CPU test-suite exercise
// the exact device logic; the product never calls the host instantiation).
#pragma once

#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define MB_HD __host__ __device__ __forceinline__
#else
#define MB_HD inline
#endif

#include "../../include/metabuli_b200.h"

namespace mb {

typedef unsigned long long u64;
typedef unsigned int u32;
typedef unsigned short u16;
typedef unsigned char u8;

// ---- qinfo bit layout: src/commons/Kmer.h:11-16 -------------------------------------------------
MB_HD u64 make_qinfo(u32 seqId, u32 pos, u32 frame) {
    return (u64) pos | ((u64) (seqId & 0x1FFFFFFFu) << 32) | ((u64) (frame & 7u) << 61);
}
MB_HD u32 qi_pos(u64 q) { return (u32) q; }
MB_HD u32 qi_seq(u64 q) { return (u32) ((q >> 32) & 0x1FFFFFFFu); }
MB_HD u32 qi_frame(u64 q) { return (u32) (q >> 61); }

// ---- multi-GPU routing (mb_peer_attach): where the producers send their records ---------------------
// k-mers go to the shard that owns their value (owner = number of bounds <= value; bounds = first value of shards
// 1..n-1, i.e. the `split` cut points of KmerMatcher.cpp:127-271); matches go to the home rank of their read.
struct Route {
    u32 n, rank, reads_per_rank;
    float inv_reads_per_rank;
    u64 bounds[MB_MAX_RANKS];
    void *kmers[MB_MAX_RANKS];                 // start of THIS rank's region inside rank d's k-mer receive buffer
    void *matches[MB_MAX_RANKS];               // same for matches
    u64 kmer_cap[MB_MAX_RANKS], match_cap[MB_MAX_RANKS];     // region capacities (records)
    unsigned long long *sent_k, *sent_m;       // LOCAL per-destination counters [MB_MAX_RANKS]: slots reserved so far
};
// A record array that lives in up to MB_MAX_RANKS pieces (the per-source regions of a receive buffer): element idx of the
// concatenation.  n == 0: not segmented.
struct SegIn {
    u32 n, pad;
    u64 prefix[MB_MAX_RANKS + 1];
    const void *base[MB_MAX_RANKS];
};
#ifdef __CUDACC__
template <typename Rec>
__device__ __forceinline__ const Rec *seg_ptr(const SegIn &s, const Rec *in, u64 idx) {
    if (s.n == 0) return in + idx;
    u32 r = 0;
    while (r + 1 < s.n && idx >= s.prefix[r + 1]) ++r;
    return reinterpret_cast<const Rec *>(s.base[r]) + (idx - s.prefix[r]);
}
#endif
MB_HD u32 route_owner(const Route &r, u64 value) {
    u32 d = 0;
    for (u32 i = 0; i + 1 < r.n; ++i) d += r.bounds[i] <= value ? 1u : 0u;
    return d;
}
MB_HD u32 route_home(const Route &r, u32 seq) {
#ifdef __CUDA_ARCH__
    // (seq-1) / reads_per_rank without the integer-division sequence: float estimate, exact correction (x < 2^29)
    const u32 x = seq - 1u;
    u32 q = (u32) (__uint2float_rz(x) * r.inv_reads_per_rank);
    if (q * r.reads_per_rank > x) --q;
    else if ((q + 1u) * r.reads_per_rank <= x) ++q;
    return q < r.n ? q : r.n - 1u;
#else
    const u32 h = (seq - 1u) / r.reads_per_rank;
    return h < r.n ? h : r.n - 1u;
#endif
}

// ---- order key of the query k-mer sort ---------------------------------------------------------------
// The join only needs the queries of one DB block contiguous (mb_match.cuh), so the sort orders a monotone function T of the
// 40-bit amino-acid part instead of all of it.  mode 0: T = aa >> shift (the leading 40-shift bits).  mode 1 (format 2 only):
// the 5-bit codes of a format-2 metamer take 21 values (GeneticCode.h:34-193), so the five leading residues are re-coded
// as one base-21 number (21^5 < 2^22) and the sixth residue adds two more bits: a 24-bit key — three 8-bit radix passes —
// that resolves as much as the leading 27 raw bits would.  Monotone for codes <= 20, which is all the translation tables emit.
struct SortT { int mode, shift; };
MB_HD u32 sort_t_packed(u64 value) {
    const u32 hi = (u32) (value >> 32);          // amino-acid bits 8..39
    u32 t = hi >> 27;
    t = t * 21u + ((hi >> 22) & 31u);
    t = t * 21u + ((hi >> 17) & 31u);
    t = t * 21u + ((hi >> 12) & 31u);
    t = t * 21u + ((hi >> 7) & 31u);
    return (t << 2) | ((((hi >> 2) & 31u) * 49u) >> 8);
}
// mode 2: all seven leading residues as a base-21 number (21^7 < 2^31) plus one bit of the eighth: a 32-bit key — four
// passes — that resolves 7.3 residues (the raw leading 32 bits resolve 6.4).  Chosen for large databases, where a coarser
// key leaves T buckets that span several units of the DB stream and the join walks a unit's queries once per piece.
MB_HD u32 sort_t_packed7(u64 value) {
    const u64 aa = value >> 24;                  // 8 x 5 bits, first residue most significant
    u32 t = (u32) (aa >> 35) & 31u;
    t = t * 21u + ((u32) (aa >> 30) & 31u);
    t = t * 21u + ((u32) (aa >> 25) & 31u);
    t = t * 21u + ((u32) (aa >> 20) & 31u);
    t = t * 21u + ((u32) (aa >> 15) & 31u);
    t = t * 21u + ((u32) (aa >> 10) & 31u);
    t = t * 21u + ((u32) (aa >> 5) & 31u);
    return (t << 1) | (u32) (((u32) aa & 31u) >= 11u);
}
MB_HD u64 sort_t(const SortT &k, u64 value) {
    return k.mode == 1 ? (u64) sort_t_packed(value) : k.mode == 2 ? (u64) sort_t_packed7(value) : (value >> 24) >> k.shift;
}
MB_HD bool sort_t_lossy(const SortT &k) { return k.mode != 0 || k.shift != 0; }

// first index in vals[lo, hi) whose amino-acid part is >= aa (k_match: lookup of a query in a decoded DB block)
MB_HD u32 lower_bound_aa_range(const u64 *vals, u32 lo, u32 hi, u64 aa) {
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if ((vals[mid] >> 24) < aa) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ---- read-length arithmetic: src/commons/LocalUtil.h:45-59 ----------------------------------------
MB_HD int max_covered_length(int len) {
    int r = len % 3;
    return r == 2 ? len - 2 : (r == 1 ? len - 4 : len - 3);
}
// slot bound of one mate; < 1 means the mate is "empty" (KmerExtractor.cpp:457-460)
MB_HD int query_kmer_number(int len) { return (max_covered_length(len) / 3 - 8 + 1) * 6; }

// ---- translation tables ---------------------------------------------------------------------------
// Built on the host by build_tables() and copied to __constant__ memory.
//   nuc[c]    : ASCII -> base code A0 C1 T2 G3, 4 = not a base.  Restates the atcg map
//               (src/commons/common.cpp:13-23) followed by nuc2int (src/commons/GeneticCode.h:6).
//   codonF[i] : i = b0 | b1<<2 | b2<<4 (three consecutive stored bases, little-endian) ->
//               aa | cid<<5 of the forward codon b0 b1 b2 (GeneticCode.h:34-193: 21-letter alphabet
//               ARNDCQEGHILKMFPSTWYVX, 3-bit codon ids)
//   codonR[i] : same index -> reverse-strand codon comp(b2) comp(b1) comp(b0) (KmerScanner.h:94-98)
//   ham4      : 8x8 codon-id Hamming table, 4 bits per cell (KmerMatcher.h:66-70 hammingLookup)
//   lut2      : the same table for two codons at a time: index = qpair << 6 | tpair, where a pair is six bits
//               of the DNA part (codon 2j in the low three bits); cell = sum of the two Hamming cells (<= 8)
//   lutp      : per pair, the two 2-bit fields of getHammings (low nibble: v0 | v1 << 2) and of
//               getHammings_reverse (high nibble: v1 | v0 << 2), v = cell & 3 (KmerMatcher.h:72-158, 386-416)
struct Tables {
    u8 nuc[256];
    u8 codonF[64];
    u8 codonR[64];
    u64 ham4[4];
    u8 lut2[4096];
    u8 lutp[4096];
};

void build_tables(Tables &t);

// hammingLookup[a][b], a/b 3-bit codon ids (KmerMatcher.h:66-70); cell (a,b) sits at nibble a*8+b
MB_HD u32 ham_cell(const u64 (&h)[4], u32 a, u32 b) {
    u32 idx = a * 8 + b;
    u64 w = (idx & 32) ? ((idx & 16) ? h[3] : h[2]) : ((idx & 16) ? h[1] : h[0]);
    return (u32) (w >> ((idx & 15) * 4)) & 15u;
}

// getHammingDistanceSum (KmerMatcher.h:348-360): sum over the 8 codon ids of the low 24 bits
MB_HD u32 hamming_sum(const u64 (&h)[4], u64 q, u64 t) {
    u32 s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += ham_cell(h, (u32) (q >> (3 * k)) & 7u, (u32) (t >> (3 * k)) & 7u);
    return s;
}

// getHammings / getHammings_reverse (KmerMatcher.h:386-416) with the HAMMING_LUT0..7 cells
// (KmerMatcher.h:72-158): field = hamming & 3 (so the cells holding 4 pack to 0), except that
// HAMMING_LUT7 holds 1<<14 for query ids 4,5 against target ids 6,7.
MB_HD u32 hamming_pack(const u64 (&h)[4], u64 q, u64 t, bool reverse) {
    u32 out = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        u32 a = (u32) (q >> (3 * k)) & 7u, b = (u32) (t >> (3 * k)) & 7u;
        int field = reverse ? 7 - k : k;
        u32 v = ham_cell(h, a, b) & 3u;
        if (field == 7 && (a == 4 || a == 5) && (b == 6 || b == 7)) v = 1;
        out |= v << (2 * field);
    }
    return out;
}

// The same two functions through the two-codon tables (4 lookups instead of 8 cell extractions).
MB_HD u32 hamming_sum_lut(const u8 *lut2, u64 q, u64 t) {
    u32 s = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) s += lut2[(((u32) (q >> (6 * j)) & 63u) << 6) | ((u32) (t >> (6 * j)) & 63u)];
    return s;
}
MB_HD u32 hamming_pack_lut(const u8 *lutp, u64 q, u64 t, bool reverse) {
    u32 out = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const u32 p = lutp[(((u32) (q >> (6 * j)) & 63u) << 6) | ((u32) (t >> (6 * j)) & 63u)];
        out |= reverse ? ((p >> 4) << (4 * (3 - j))) : ((p & 15u) << (4 * j));
    }
    // HAMMING_LUT7 quirk: field 7 (codon 7, or codon 0 when reversed) holds 1 for query ids 4,5 vs target ids 6,7
    const int kq = reverse ? 0 : 7;
    const u32 a = (u32) (q >> (3 * kq)) & 7u, b = (u32) (t >> (3 * kq)) & 7u;
    if ((a == 4 || a == 5) && (b == 6 || b == 7)) out |= 1u << 14;
    return out;
}

// ---- metamer assembly from 24 stored bases --------------------------------------------------------
// win48: 24 bases, base p at bits 2p..2p+1 (A0 C1 T2 G3).  Returns the metamer of the window on the
// forward strand (reverse=false) or on the reverse strand (reverse=true).
//   format 2 (MetamerScanner::next, KmerScanner.h:82-117): first translated codon most significant,
//            AA 5 bits each << 24 | 3-bit codon ids.
//   format 1 (OldMetamerScanner::next, KmerScanner.h:137-181): codons consumed from the far end, so
//            the first translated codon is LEAST significant; AA part is a base-21 number.
MB_HD u64 metamer_from_window(const u8 *codonF, const u8 *codonR, u64 win48, bool reverse, int format) {
    u64 aa = 0, dna = 0;
    if (format == 2) {
        if (!reverse) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {           // j = codon index along the stored bases = translation order
                u32 c = codonF[(win48 >> (6 * j)) & 63];
                aa = (aa << 5) | (c & 31u);
                dna = (dna << 3) | (c >> 5);
            }
        } else {
#pragma unroll
            for (int j = 7; j >= 0; --j) {          // translation order runs from the high end down
                u32 c = codonR[(win48 >> (6 * j)) & 63];
                aa = (aa << 5) | (c & 31u);
                dna = (dna << 3) | (c >> 5);
            }
        }
    } else {
        if (!reverse) {
#pragma unroll
            for (int j = 7; j >= 0; --j) {          // highest stored codon most significant
                u32 c = codonF[(win48 >> (6 * j)) & 63];
                aa = aa * 21 + (c & 31u);
                dna = (dna << 3) | (c >> 5);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {           // lowest stored codon most significant
                u32 c = codonR[(win48 >> (6 * j)) & 63];
                aa = aa * 21 + (c & 31u);
                dna = (dna << 3) | (c >> 5);
            }
        }
    }
    return (aa << 24) | dna;
}

// Closed-syncmer test of SyncmerScanner::next (SyncmerScanner.h:36-102) on a format-2 metamer:
// the leftmost minimum of the 8-s+1 amino-acid s-mers must sit at offset 0 or 8-s.
template <int S>
MB_HD bool is_syncmer_fixed(u64 metamer) {         // s known at compile time: constant shifts, 32-bit s-mers (5*S <= 30)
    const u64 aa = metamer >> 24;
    constexpr u32 mask = (1u << (5 * S)) - 1u;
    constexpr int last = 8 - S;
    const u32 v0 = (u32) (aa >> (5 * last)) & mask, vl = (u32) aa & mask;
    u32 mid = 0xFFFFFFFFu;                          // smallest s-mer strictly between the two ends
#pragma unroll
    for (int i = 1; i < last; ++i) {
        const u32 v = (u32) (aa >> (5 * (last - i))) & mask;
        mid = v < mid ? v : mid;
    }
    // leftmost minimum at offset 0: v0 <= every other one; at offset `last`: vl < every other one
    const u32 rest0 = mid < vl ? mid : vl, restl = mid < v0 ? mid : v0;
    return v0 <= rest0 || vl < restl;
}
MB_HD bool is_syncmer(u64 metamer, int s) {
    if (s == 5) return is_syncmer_fixed<5>(metamer);        // the default --smer-len (LocalParameters.cpp)
    if (s == 6) return is_syncmer_fixed<6>(metamer);
    if (s == 4) return is_syncmer_fixed<4>(metamer);
    const u64 aa = metamer >> 24;
    const u64 mask = (1ull << (5 * s)) - 1;
    const int last = 8 - s;
    u64 best = (aa >> (5 * last)) & mask;           // s-mer at offset 0
    int arg = 0;
    for (int i = 1; i <= last; ++i) {
        u64 v = (aa >> (5 * (last - i))) & mask;
        if (v < best) { best = v; arg = i; }
    }
    return arg == 0 || arg == last;
}

// Device-resident taxonomy view (see mb_load_taxonomy in metabuli_b200.h)
struct TaxView {
    const int32_t *parent;         // -1 = taxon does not exist
    const int32_t *depth;          // root 0
    const int32_t *species_of;
    const int32_t *species_rank_of;
    int32_t max_taxid;
    int32_t node0_taxid;
    int32_t eukaryota;
    const u64 *node;               // optional: (u32) parent | (u64) depth << 32 per taxon, so that one load answers "exists",
                                   // "depth" and "parent" (tax_lca: one round trip instead of three for the common cases)
};

MB_HD bool tax_exists(const TaxView &t, int32_t x) { return x >= 0 && x <= t.max_taxid && t.parent[x] >= 0; }

// lowest common ancestor by depth-equalising parent walk == lcaHelper's Euler-tour RMQ on a tree
// (NcbiTaxonomy.cpp:261-280), including its node-index-0 short circuit.
MB_HD int32_t tax_lca_existing(const TaxView &t, int32_t a, int32_t b) {
    if (a == t.node0_taxid || b == t.node0_taxid) return t.node0_taxid;
    if (a == b) return a;
    int da = t.depth[a], db = t.depth[b];
    while (da > db) { a = t.parent[a]; --da; }
    while (db > da) { b = t.parent[b]; --db; }
    while (a != b) { a = t.parent[a]; b = t.parent[b]; }
    return a;
}
// NcbiTaxonomy::LCA(TaxID, TaxID) (NcbiTaxonomy.cpp:303-310)
MB_HD int32_t tax_lca(const TaxView &t, int32_t a, int32_t b) {
    if (a == b) return a;                      // every branch below returns a (== b) in this case; saves the table loads
    if (t.node) {
        // the two node records are requested together; the parents they carry serve the first step of each walk
        const bool ia = a >= 0 && a <= t.max_taxid, ib = b >= 0 && b <= t.max_taxid;
        const u64 na = ia ? t.node[a] : 0xFFFFFFFFull, nb = ib ? t.node[b] : 0xFFFFFFFFull;
        int32_t pa = (int32_t) (u32) na, pb = (int32_t) (u32) nb;
        if (pa < 0) return b;                  // !tax_exists(a)
        if (pb < 0) return a;                  // !tax_exists(b)
        if (a == t.node0_taxid || b == t.node0_taxid) return t.node0_taxid;
        int da = (int) (na >> 32), db = (int) (nb >> 32);
        bool have_pa = true, have_pb = true;
        while (da > db) { a = have_pa ? pa : (int32_t) (u32) t.node[a]; have_pa = false; --da; }
        while (db > da) { b = have_pb ? pb : (int32_t) (u32) t.node[b]; have_pb = false; --db; }
        while (a != b) {
            a = have_pa ? pa : (int32_t) (u32) t.node[a];
            b = have_pb ? pb : (int32_t) (u32) t.node[b];
            have_pa = have_pb = false;
        }
        return a;
    }
    if (!tax_exists(t, a)) return b;
    if (!tax_exists(t, b)) return a;
    return tax_lca_existing(t, a, b);
}
// NcbiTaxonomy::IsAncestor (NcbiTaxonomy.cpp:282-300)
MB_HD bool tax_is_ancestor(const TaxView &t, int32_t anc, int32_t child) {
    if (anc == child) return true;
    if (anc == 0 || child == 0) return false;
    if (!tax_exists(t, child) || !tax_exists(t, anc)) return false;
    return tax_lca_existing(t, child, anc) == anc;
}


#ifdef __CUDACC__
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// ---- mbarrier + bulk copy (sm_90+/sm_100: UBLKCP, SYNCS in SASS) --------------------------------------------------
__device__ __forceinline__ u32 smem_u32(const void *p) { return (u32) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, u32 parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy of `bytes` (multiple of 16; both addresses 16-byte aligned), completion on `bar`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, u32 bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
#endif

} // namespace mb
