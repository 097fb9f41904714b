// mb_extract.cuh — K1: six-frame metamer / syncmer extraction.
//
// Replaces KmerExtractor::fillQueryKmerBufferParallel[_paired] / processSequence / fillQueryKmerBuffer
// (src/commons/KmerExtractor.cpp:83-386) and the scanners MetamerScanner / OldMetamerScanner /
// SyncmerScanner (src/commons/KmerScanner.h:49-182, src/commons/SyncmerScanner.h:36-102).
//
// B200 design.  The scanners are sequential state machines, but what they emit is a pure function of
// a 24-base window: window x of a mate is emitted on a strand iff x lies inside that strand's frame
// segment, none of its 24 bases is ambiguous (the scanners restart right after an ambiguous codon,
// KmerScanner.h:104-109) and, for syncmer DBs, the closed-syncmer test holds.  So the kernel is a
// flat map over the base stream of the whole batch: a CTA stages 2048 (+24 halo) consecutive bases
// of the packed batch in shared memory as 2-bit codes plus an ambiguity bit (one coalesced 8-byte
// load per thread), each thread then slides over 8 window starts with the window held in a 64-bit
// register, and both strands come out of the same register through two 64-entry codon tables.
// Output is compacted exactly (count pass -> exclusive scan of tile counts -> write pass), so the
// sort never sees the zero-filled holes the reference reserves (KmerMatcher.cpp:143-151 skips them).
#pragma once
#include "mb_common.cuh"

namespace mb {

constexpr int EX_THREADS = 256;
constexpr int EX_PER_THREAD = 8;
constexpr int EX_TILE = EX_THREADS * EX_PER_THREAD; // 2048 window starts per CTA

struct ExtractArgs {
    const u8 *bases;       // packed batch, padded with >= 64 readable bytes
    u64 n_bases;
    const u64 *offsets;    // [2*n_reads]
    const u32 *lens;       // [2*n_reads]
    u32 n_mates;           // 2*n_reads
    int seq_mode, format, syncmer, smer_len;
    u32 seq_base;          // added to the 1-based read index (multi-GPU: ids unique across home ranks)
};

// largest mate index m with offsets[m] <= g, searching [lo, hi] (offsets are non-decreasing)
__device__ __forceinline__ u32 mate_of(const u64 *__restrict__ offsets, u32 lo, u32 hi, u64 g) {
    while (lo < hi) {
        u32 mid = (lo + hi + 1) >> 1;
        if (__ldg(offsets + mid) <= g) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// one thread per tile: first mate that can own a base of the tile
__global__ void k_tile_first_mate(ExtractArgs a, u32 n_tiles, u32 *__restrict__ tile_mate) {
    u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tiles) return;
    u64 g = (u64) t * EX_TILE;
    tile_mate[t] = (t == n_tiles || a.n_mates == 0) ? (a.n_mates ? a.n_mates - 1 : 0) : mate_of(a.offsets, 0, a.n_mates - 1, g);
}

template <bool WRITE>
__global__ void __launch_bounds__(EX_THREADS)
k_extract(ExtractArgs a, const u32 *__restrict__ tile_mate, const Tables *__restrict__ tables,
          u32 *__restrict__ tile_counts, const u64 *__restrict__ tile_offsets, mb_kmer *__restrict__ out) {
    __shared__ u16 s_code[EX_THREADS + 4];
    __shared__ u8 s_amb[EX_THREADS + 4];
    __shared__ u8 s_nuc[256];
    __shared__ u8 s_codonF[64];
    __shared__ u8 s_codonR[64];
    __shared__ u32 s_warp[EX_THREADS / 32];

    const u32 tid = threadIdx.x;
    const u32 tile = blockIdx.x;
    const u64 g0 = (u64) tile * EX_TILE;

    s_nuc[tid] = tables->nuc[tid];
    if (tid < 64) { s_codonF[tid] = tables->codonF[tid]; s_codonR[tid] = tables->codonR[tid]; }
    __syncthreads();

    // stage 8 bases per thread (+ 4 halo groups by the first threads)
    for (u32 grp = tid; grp < EX_THREADS + 4; grp += EX_THREADS) {
        const u64 gb = g0 + (u64) grp * 8;
        u64 raw = 0;
        if (gb < a.n_bases) raw = *reinterpret_cast<const u64 *>(a.bases + gb); // buffer is padded
        u32 code = 0, amb = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            u32 c = s_nuc[(raw >> (8 * i)) & 255];
            if (gb + i >= a.n_bases) c = 4;
            code |= (c & 3u) << (2 * i);
            amb |= (c >> 2) << i;
        }
        s_code[grp] = (u16) code;
        s_amb[grp] = (u8) amb;
    }
    __syncthreads();

    // window register: bases [8*tid, 8*tid+32)
    const u64 w = (u64) s_code[tid] | ((u64) s_code[tid + 1] << 16) | ((u64) s_code[tid + 2] << 32) | ((u64) s_code[tid + 3] << 48);
    const u32 ambw = (u32) s_amb[tid] | ((u32) s_amb[tid + 1] << 8) | ((u32) s_amb[tid + 2] << 16) | ((u32) s_amb[tid + 3] << 24);

    // mate of the first window start of this thread
    const u64 gfirst = g0 + (u64) tid * EX_PER_THREAD;
    u32 m = 0;
    u64 moff = 0;
    u32 mlen = 0;
    const bool any = gfirst < a.n_bases && a.n_mates > 0;
    if (any) {
        m = mate_of(a.offsets, tile_mate[tile], tile_mate[tile + 1], gfirst);
        moff = __ldg(a.offsets + m);
        mlen = __ldg(a.lens + m);
    }

    mb_kmer local[2 * EX_PER_THREAD];
    int cnt = 0;
    const bool paired = a.seq_mode == 2;
    bool have_meta = false, usable = false;
    u32 pos_offset = 0;
    int used = 0;

    if (any) {
#pragma unroll
        for (int i = 0; i < EX_PER_THREAD; ++i) {
            const u64 g = gfirst + i;
            // advance to the mate that owns base g (mates are back to back, zero-length ones share an offset)
            while (m + 1 < a.n_mates && __ldg(a.offsets + m + 1) <= g) {
                ++m;
                moff = __ldg(a.offsets + m);
                mlen = __ldg(a.lens + m);
                have_meta = false;
            }
            if (g < moff || g >= moff + mlen) continue;
            if (!have_meta) {
                have_meta = true;
                // KmerExtractor.cpp:261-279, 457-460: a mate whose slot bound is < 1 is empty; in paired
                // mode one empty mate drops the pair; other modes use mate 0 only.
                const bool self_ok = query_kmer_number((int) mlen) >= 1;
                if (paired) {
                    const u32 sib = __ldg(a.lens + (m ^ 1u));
                    usable = self_ok && query_kmer_number((int) sib) >= 1;
                    pos_offset = (m & 1u) ? (u32) max_covered_length((int) sib) + 3u : 0u;
                } else {
                    usable = self_ok && (m & 1u) == 0;
                    pos_offset = 0;
                }
                used = max_covered_length((int) mlen);
            }
            if (!usable) continue;
            const u32 x = (u32) (g - moff);
            // window [x, x+23] must end inside the frame segment [x%3, x%3 + used - 1]
            if ((int) (x + 24) > (int) (x % 3) + used) continue;
            if ((ambw >> i) & 0xFFFFFFu) continue;          // an ambiguous base: no window may cover it
            const u64 win = (w >> (2 * i)) & 0xFFFFFFFFFFFFull;
            const u32 seqId = a.seq_base + (m >> 1) + 1;
            // forward strand: frame = begin = x % 3 (KmerExtractor.cpp:369-373)
            {
                const u64 v = metamer_from_window(s_codonF, s_codonR, win, false, a.format);
                if (a.format == 1 || !a.syncmer || is_syncmer(v, a.smer_len)) {
                    local[cnt].value = v;
                    local[cnt].qinfo = make_qinfo(seqId, x + pos_offset, x % 3);
                    ++cnt;
                }
            }
            // reverse strand: begin = (L%3 - frame%3) mod 3 and begin == x%3  =>  frame = 3 + (L - x) mod 3
            {
                const u64 v = metamer_from_window(s_codonF, s_codonR, win, true, a.format);
                if (a.format == 1 || !a.syncmer || is_syncmer(v, a.smer_len)) {
                    const u32 fr = 3u + (u32) (((int) (mlen % 3) - (int) (x % 3) + 3) % 3);
                    local[cnt].value = v;
                    local[cnt].qinfo = make_qinfo(seqId, x + pos_offset, fr);
                    ++cnt;
                }
            }
        }
    }

    // block-wide exclusive scan of cnt
    const u32 lane = tid & 31, warp = tid >> 5;
    u32 incl = (u32) cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        u32 v = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (u32) d) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    u32 wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < EX_THREADS / 32; ++k) {
        const u32 v = s_warp[k];
        if ((u32) k < warp) wbase += v;
        total += v;
    }
    if (!WRITE) {
        if (tid == 0) tile_counts[tile] = total;
        return;
    }
    const u64 base = tile_offsets[tile] + wbase + (incl - (u32) cnt);
    for (int k = 0; k < cnt; ++k) {
        // 16-byte vector store
        *reinterpret_cast<ulonglong2 *>(out + base + k) = make_ulonglong2(local[k].value, local[k].qinfo);
    }
}

// single-CTA exclusive scan of u32 counts into u64 offsets; offsets[n] = total
__global__ void __launch_bounds__(1024) k_scan_counts(const u32 *__restrict__ counts, u64 *__restrict__ offsets, u32 n) {
    __shared__ u64 s_warp[32];
    __shared__ u64 s_carry;
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (u32 base = 0; base < n; base += 1024) {
        const u32 i = base + tid;
        const u64 v = i < n ? counts[i] : 0;
        u64 incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            u64 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if (lane >= (u32) d) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            u64 x = s_warp[lane];
            u64 xi = x;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                u64 t = __shfl_up_sync(0xFFFFFFFFu, xi, d);
                if (lane >= (u32) d) xi += t;
            }
            s_warp[lane] = xi - x; // exclusive warp bases
        }
        __syncthreads();
        const u64 carry = s_carry;
        if (i < n) offsets[i] = carry + s_warp[warp] + incl - v;
        __syncthreads();
        if (tid == 1023) s_carry = carry + s_warp[warp] + incl;
        __syncthreads();
    }
    if (tid == 0) offsets[n] = s_carry;
}

} // namespace mb
