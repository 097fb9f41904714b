// mb_segsort.cuh — K5 (short-read form): matches are first grouped by read with a few radix passes on
// the sequence ID alone; each read's segment (a few hundred 24-byte records) is then put into the
// compareMatches order (KmerMatcher.cpp:1149-1166: speciesId, frame, pos, hamming, dnaEncoding) by ONE
// warp in shared memory.  This replaces ~7 full passes over the match list by one read and one write.
#pragma once
#include "mb_common.cuh"

namespace mb {

constexpr int SS_THREADS = 256;
constexpr int SS_WARPS = SS_THREADS / 32;
constexpr int SS_CAP = 1024;                    // longest segment the shared-memory path takes
constexpr u32 SS_HUGE_MAX = 1u << 16;           // segments beyond SS_CAP that are sorted one CTA each (more: full radix sort)
struct HugeSeg { u32 read, n_pad; unsigned long long off; };

struct MatchSeqKey {                            // radix digit = byte `pass` of the batch-local read index
    u32 seq_base;
    __device__ __forceinline__ u32 digit(const mb_match &m, int pass) const { return ((qi_seq(m.qinfo) - seq_base) >> (8 * pass)) & 255u; }
};

struct SegKey {                                 // the compareMatches key below the sequence ID, packed
    int pos_bits, sp_bits;
    __device__ __forceinline__ u64 key(u64 qinfo, u32 species, u32 dna, u32 ham) const {
        u64 k = species;
        k = (k << 3) | qi_frame(qinfo);
        k = (k << pos_bits) | qi_pos(qinfo);
        k = (k << 3) | (ham & 7u);
        k = (k << 24) | (dna & 0xFFFFFFu);
        return k;
    }
    __host__ __device__ int total_bits() const { return sp_bits + 3 + pos_bits + 3 + 24; }
};

// branch-free selects (the compiler otherwise turns the conditional exchanges into divergent branches)
__device__ __forceinline__ u64 sel64(bool p, u64 a, u64 b) {
    u64 r;
    asm("{.reg .pred q; setp.ne.u32 q, %3, 0; selp.b64 %0, %1, %2, q;}" : "=l"(r) : "l"(a), "l"(b), "r"((u32) p));
    return r;
}
__device__ __forceinline__ u32 sel32(bool p, u32 a, u32 b) {
    u32 r;
    asm("{.reg .pred q; setp.ne.u32 q, %3, 0; selp.b32 %0, %1, %2, q;}" : "=r"(r) : "r"(a), "r"(b), "r"((u32) p));
    return r;
}

// Bitonic sort of the first n_pad (a power of two <= ITEMS*32) elements held in registers by one warp; element
// e = i*32 + lane (striped), so partners less than 32 apart are exchanged with shuffles and the others are
// register-to-register.  PACKED keys carry the source index in their low 9 bits (possible when the compareMatches
// key needs at most 55 bits): one 64-bit compare per exchange; otherwise ties on the key are broken by src[].
// Either way the order is total and equals a stable sort of the input order.  The stage loops stay rolled (a fully
// unrolled network is larger than the instruction cache); only the per-register loop is unrolled.
template <int ITEMS, int DJ, bool PACKED>
__device__ __forceinline__ void bitonic_in_lane(u64 (&k)[ITEMS], u32 (&src)[ITEMS], int kk) {
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        if ((i & DJ) == 0 && (i | DJ) < ITEMS) {
            const int p = i | DJ;
            {
                const bool up = ((i * 32) & kk) == 0;
                const u64 a = k[i], c = k[p];
                const bool a_gt = PACKED ? a > c : (a > c || (a == c && src[i] > src[p]));
                const bool sw = a_gt == up;
                k[i] = sel64(sw, c, a);
                k[p] = sel64(sw, a, c);
                if (!PACKED) { const u32 sa = src[i], sc = src[p]; src[i] = sel32(sw, sc, sa); src[p] = sel32(sw, sa, sc); }
            }
        }
    }
}

template <int ITEMS, bool PACKED>
__device__ __forceinline__ void warp_bitonic(u64 (&k)[ITEMS], u32 (&src)[ITEMS], int n_pad) {
    const u32 lane = threadIdx.x & 31;
#pragma unroll 1
    for (int kk = 2; kk <= n_pad; kk <<= 1) {
#pragma unroll 1
        for (int j = kk >> 1; j > 0; j >>= 1) {
            if (j < 32) {
                const bool lower = (lane & j) == 0;
#pragma unroll
                for (int i = 0; i < ITEMS; ++i) {
                    {
                        const u64 ok = __shfl_xor_sync(0xFFFFFFFFu, k[i], j);
                        const bool up = kk < 32 ? (lane & kk) == 0 : ((i * 32) & kk) == 0;   // kk == 32: bit 5 of e = bit 0 of i
                        bool other_less = ok < k[i];
                        if (!PACKED) {
                            const u32 os = __shfl_xor_sync(0xFFFFFFFFu, src[i], j);
                            other_less = other_less || (ok == k[i] && os < src[i]);
                            const bool take = (lower == up) == other_less;
                            src[i] = sel32(take, os, src[i]);
                            k[i] = sel64(take, ok, k[i]);
                        } else {
                            const bool take = (lower == up) == other_less;      // packed keys never tie
                            k[i] = sel64(take, ok, k[i]);
                        }
                    }
                }
            } else {
                switch (j >> 5) {
                    case 1: bitonic_in_lane<ITEMS, 1, PACKED>(k, src, kk); break;
                    case 2: bitonic_in_lane<ITEMS, 2, PACKED>(k, src, kk); break;
                    case 4: bitonic_in_lane<ITEMS, 4, PACKED>(k, src, kk); break;
                    default: bitonic_in_lane<ITEMS, 8, PACKED>(k, src, kk); break;
                }
            }
        }
    }
}

// Sort one segment of n <= ITEMS*32 records with the register network and write it in order.
template <int ITEMS, bool PACKED>
__device__ __forceinline__ void sort_segment_regs(const u64 *__restrict__ in64, u64 *__restrict__ out64, u64 b, u32 n, const SegKey &sk) {
    const u32 lane = threadIdx.x & 31;
    u64 k[ITEMS];
    u32 src[ITEMS];
    int n_pad = 32;
    while ((u32) n_pad < n) n_pad <<= 1;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const u32 e = i * 32 + lane;
        k[i] = ~0ull;
        src[i] = e;
        if (e < n) {
            const u64 w0 = __ldg(in64 + (b + e) * 3), w1 = __ldg(in64 + (b + e) * 3 + 1), w2 = __ldg(in64 + (b + e) * 3 + 2);
            const u64 key = sk.key(w0, (u32) (w1 >> 32), (u32) w2, (u32) (w2 >> 48));
            k[i] = PACKED ? ((key << 9) | e) : key;
        }
    }
    warp_bitonic<ITEMS, PACKED>(k, src, n_pad);
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const u32 e = i * 32 + lane;
        if (e < n) {
            const u64 s = b + (PACKED ? (u32) (k[i] & 511u) : src[i]);
            const u64 w0 = __ldg(in64 + s * 3), w1 = __ldg(in64 + s * 3 + 1), w2 = __ldg(in64 + s * 3 + 2);
            out64[(b + e) * 3] = w0; out64[(b + e) * 3 + 1] = w1; out64[(b + e) * 3 + 2] = w2;
        }
    }
}

constexpr size_t segsort_smem_bytes() { return (size_t) SS_WARPS * SS_CAP * (sizeof(u64) + sizeof(u16)); }

// One warp per read segment [seg_begin[r], seg_end[r]) of `in` (grouped by read); sorted records go to the same
// range of `out`.  Segments of up to 512 records are sorted in registers here; those of 513..SS_CAP are appended to
// `long_list` for k_sort_segments_long (keeping this kernel free of shared memory: 3 CTAs per SM instead of 2);
// longer ones are listed in `huge` with a slice of key scratch each (k_sort_segments_huge sorts them one CTA each); when
// there are more than SS_HUGE_MAX of them they are only counted and the host takes the full radix path.
template <bool PACKED>
__global__ void __launch_bounds__(SS_THREADS, 3)
k_sort_segments(const mb_match *__restrict__ in, mb_match *__restrict__ out, const u64 *__restrict__ seg_begin,
                const u64 *__restrict__ seg_end, u32 n_reads, SegKey sk, u32 *ticket, u32 *too_long, u32 *__restrict__ long_list,
                u32 *long_count, HugeSeg *__restrict__ huge, unsigned long long *huge_words) {
    const u32 lane = threadIdx.x & 31;
    const u64 *in64 = reinterpret_cast<const u64 *>(in);
    u64 *out64 = reinterpret_cast<u64 *>(out);
    while (true) {
        u32 r0 = 0;
        if (lane == 0) r0 = atomicAdd(ticket, 8u);
        r0 = __shfl_sync(0xFFFFFFFFu, r0, 0);
        if (r0 >= n_reads) break;
        const u32 r1 = r0 + 8 < n_reads ? r0 + 8 : n_reads;
        for (u32 r = r0; r < r1; ++r) {
            const u64 b = seg_begin[r], e = seg_end[r];
            // the records of the NEXT read are pulled into L2 while this one is sorted (a warp otherwise waits on HBM
            // at the start of every segment)
            if (r + 1 < n_reads) {
                const u64 nb = seg_begin[r + 1], ne = seg_end[r + 1];
                const char *pb = reinterpret_cast<const char *>(in64 + nb * 3);
                const u64 bytes = (ne - nb) * sizeof(mb_match);
                for (u64 o = (u64) lane * 128; o < bytes && o < 16384; o += 32 * 128)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + o));
            }
            const u32 n = (u32) (e - b);
            if (n == 0) continue;
            if (n > SS_CAP) {
                if (lane == 0) {
                    const u32 slot = atomicAdd(too_long, 1u);
                    if (slot < SS_HUGE_MAX) {
                        u32 n_pad = 2048;
                        while (n_pad < n) n_pad <<= 1;
                        huge[slot].read = r;
                        huge[slot].n_pad = n_pad;
                        huge[slot].off = atomicAdd(huge_words, (unsigned long long) n_pad);
                    }
                }
                continue;
            }
            if (n <= 32) { sort_segment_regs<1, PACKED>(in64, out64, b, n, sk); continue; }
            if (n <= 64) { sort_segment_regs<2, PACKED>(in64, out64, b, n, sk); continue; }
            if (n <= 128) { sort_segment_regs<4, PACKED>(in64, out64, b, n, sk); continue; }
            if (n <= 256) { sort_segment_regs<8, PACKED>(in64, out64, b, n, sk); continue; }
            if (n <= 512) { sort_segment_regs<16, PACKED>(in64, out64, b, n, sk); continue; }
            if (lane == 0) long_list[atomicAdd(long_count, 1u)] = r;
        }
    }
}

// The segments of 513..SS_CAP records k_sort_segments left over: one warp each, bitonic network in shared memory.
__global__ void __launch_bounds__(SS_THREADS)
k_sort_segments_long(const mb_match *__restrict__ in, mb_match *__restrict__ out, const u64 *__restrict__ seg_begin,
                     const u64 *__restrict__ seg_end, SegKey sk, u32 *ticket, const u32 *__restrict__ long_list,
                     const u32 *__restrict__ long_count) {
    extern __shared__ __align__(16) unsigned char ss_raw[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64 *keys = reinterpret_cast<u64 *>(ss_raw) + (size_t) warp * SS_CAP;
    u16 *idx = reinterpret_cast<u16 *>(ss_raw + (size_t) SS_WARPS * SS_CAP * sizeof(u64)) + (size_t) warp * SS_CAP;
    const u64 *in64 = reinterpret_cast<const u64 *>(in);
    u64 *out64 = reinterpret_cast<u64 *>(out);
    const u32 total = *long_count;
    while (true) {
        u32 t = 0;
        if (lane == 0) t = atomicAdd(ticket, 1u);
        t = __shfl_sync(0xFFFFFFFFu, t, 0);
        if (t >= total) break;
        const u32 r = long_list[t];
        const u64 b = seg_begin[r];
        const u32 n = (u32) (seg_end[r] - b);
        const u32 n_pad = 1024;
        for (u32 i = lane; i < n_pad; i += 32) {
            u64 k = ~0ull;
            if (i < n) {
                const u64 w0 = __ldg(in64 + (b + i) * 3), w1 = __ldg(in64 + (b + i) * 3 + 1), w2 = __ldg(in64 + (b + i) * 3 + 2);
                k = sk.key(w0, (u32) (w1 >> 32), (u32) w2, (u32) (w2 >> 48));
            }
            keys[i] = k;
            idx[i] = (u16) i;
        }
        __syncwarp();
        for (u32 kk = 2; kk <= n_pad; kk <<= 1) {
            for (u32 j = kk >> 1; j > 0; j >>= 1) {
                for (u32 tt = lane; tt < n_pad / 2; tt += 32) {
                    const u32 i = ((tt & ~(j - 1)) << 1) | (tt & (j - 1));
                    const u32 p = i | j;
                    const u64 a = keys[i], c = keys[p];
                    const u16 ia = idx[i], ic = idx[p];
                    const bool up = (i & kk) == 0;
                    const bool a_gt = a > c || (a == c && ia > ic);
                    if (a_gt == up) { keys[i] = c; keys[p] = a; idx[i] = ic; idx[p] = ia; }
                }
                __syncwarp();
            }
        }
        for (u32 i = lane; i < n; i += 32) {
            const u64 s = b + idx[i];
            const u64 w0 = __ldg(in64 + s * 3), w1 = __ldg(in64 + s * 3 + 1), w2 = __ldg(in64 + s * 3 + 2);
            out64[(b + i) * 3] = w0; out64[(b + i) * 3 + 1] = w1; out64[(b + i) * 3 + 2] = w2;
        }
        __syncwarp();
    }
}

// Segments of more than SS_CAP records (a short read that hits very many species): one CTA each, bitonic network over
// (key, index) pairs in a global scratch slice (L2 resident), then the gather.  Rare; before this kernel existed ONE such
// read sent the whole batch through the full-key radix sort (8+ passes: 53 ms at configs[2] on the ranks that had one).
__global__ void __launch_bounds__(SS_THREADS)
k_sort_segments_huge(const mb_match *__restrict__ in, mb_match *__restrict__ out, const u64 *__restrict__ seg_begin,
                     const u64 *__restrict__ seg_end, SegKey sk, const HugeSeg *__restrict__ huge, u32 n_huge, u64 *__restrict__ keys_all,
                     u32 *__restrict__ idx_all) {
    const u64 *in64 = reinterpret_cast<const u64 *>(in);
    u64 *out64 = reinterpret_cast<u64 *>(out);
    for (u32 h = blockIdx.x; h < n_huge; h += gridDim.x) {
        const HugeSeg hs = huge[h];
        const u64 b = seg_begin[hs.read];
        const u32 n = (u32) (seg_end[hs.read] - b), n_pad = hs.n_pad;
        u64 *keys = keys_all + hs.off;
        u32 *idx = idx_all + hs.off;
        for (u32 i = threadIdx.x; i < n_pad; i += SS_THREADS) {
            u64 k = ~0ull;
            if (i < n) {
                const u64 w0 = __ldg(in64 + (b + i) * 3), w1 = __ldg(in64 + (b + i) * 3 + 1), w2 = __ldg(in64 + (b + i) * 3 + 2);
                k = sk.key(w0, (u32) (w1 >> 32), (u32) w2, (u32) (w2 >> 48));
            }
            keys[i] = k;
            idx[i] = i;
        }
        __syncthreads();
        for (u32 kk = 2; kk <= n_pad; kk <<= 1) {
            for (u32 j = kk >> 1; j > 0; j >>= 1) {
                for (u32 tt = threadIdx.x; tt < n_pad / 2; tt += SS_THREADS) {
                    const u32 i = ((tt & ~(j - 1)) << 1) | (tt & (j - 1));
                    const u32 p = i | j;
                    const u64 a = keys[i], c = keys[p];
                    const u32 ia = idx[i], ic = idx[p];
                    const bool up = (i & kk) == 0;
                    const bool a_gt = a > c || (a == c && ia > ic);
                    if (a_gt == up) { keys[i] = c; keys[p] = a; idx[i] = ic; idx[p] = ia; }
                }
                __syncthreads();
            }
        }
        for (u32 i = threadIdx.x; i < n; i += SS_THREADS) {
            const u64 s = b + idx[i];
            const u64 w0 = __ldg(in64 + s * 3), w1 = __ldg(in64 + s * 3 + 1), w2 = __ldg(in64 + s * 3 + 2);
            out64[(b + i) * 3] = w0; out64[(b + i) * 3 + 1] = w1; out64[(b + i) * 3 + 2] = w2;
        }
        __syncthreads();
    }
}

} // namespace mb
