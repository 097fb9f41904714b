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
    int total_bits() const { return sp_bits + 3 + pos_bits + 3 + 24; }
};

constexpr size_t segsort_smem_bytes() { return (size_t) SS_WARPS * SS_CAP * (sizeof(u64) + sizeof(u16)); }

// One warp per read segment [seg_begin[r], seg_end[r]) of `in` (grouped by read); sorted records go to the same
// range of `out`.  Segments longer than SS_CAP are copied unsorted and counted in *too_long (the host then takes
// the full radix path).
__global__ void __launch_bounds__(SS_THREADS)
k_sort_segments(const mb_match *__restrict__ in, mb_match *__restrict__ out, const u64 *__restrict__ seg_begin,
                const u64 *__restrict__ seg_end, u32 n_reads, SegKey sk, u32 *ticket, u32 *too_long) {
    extern __shared__ __align__(16) unsigned char ss_raw[];
    const u32 lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    u64 *keys = reinterpret_cast<u64 *>(ss_raw) + (size_t) warp * SS_CAP;
    u16 *idx = reinterpret_cast<u16 *>(ss_raw + (size_t) SS_WARPS * SS_CAP * sizeof(u64)) + (size_t) warp * SS_CAP;
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
            const u32 n = (u32) (e - b);
            if (n == 0) continue;
            if (n > SS_CAP) {
                if (lane == 0) atomicAdd(too_long, 1u);
                for (u64 i = lane; i < (u64) n * 3; i += 32) out64[b * 3 + i] = in64[b * 3 + i];
                continue;
            }
            if (n <= 32) {
                // register path: one record per lane, bitonic network over the warp with shuffles
                u64 k = ~0ull;
                u64 w0 = 0, w1 = 0, w2 = 0;
                if (lane < n) {
                    w0 = __ldg(in64 + (b + lane) * 3); w1 = __ldg(in64 + (b + lane) * 3 + 1); w2 = __ldg(in64 + (b + lane) * 3 + 2);
                    k = sk.key(w0, (u32) (w1 >> 32), (u32) w2, (u32) (w2 >> 48));
                }
                u32 src = lane;
#pragma unroll
                for (int kk = 2; kk <= 32; kk <<= 1) {
#pragma unroll
                    for (int j = kk >> 1; j > 0; j >>= 1) {
                        const u64 ok = __shfl_xor_sync(0xFFFFFFFFu, k, j);
                        const u32 os = __shfl_xor_sync(0xFFFFFFFFu, src, j);
                        const bool up = (lane & kk) == 0;
                        const bool lower = (lane & j) == 0;
                        const bool other_less = ok < k || (ok == k && os < src);
                        // the lower lane keeps the smaller element when the run ascends, the larger when it descends
                        const bool take = (lower == up) ? other_less : !other_less;
                        if (take) { k = ok; src = os; }
                    }
                }
                // lane now holds the element of rank `lane`; fetch its record from the lane that loaded it
                const u64 s0 = __shfl_sync(0xFFFFFFFFu, w0, src), s1 = __shfl_sync(0xFFFFFFFFu, w1, src), s2 = __shfl_sync(0xFFFFFFFFu, w2, src);
                if (lane < n) { out64[(b + lane) * 3] = s0; out64[(b + lane) * 3 + 1] = s1; out64[(b + lane) * 3 + 2] = s2; }
                continue;
            }
            u32 n_pad = 64;
            while (n_pad < n) n_pad <<= 1;
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
                    for (u32 t = lane; t < n_pad / 2; t += 32) {
                        const u32 i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
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
}

} // namespace mb
