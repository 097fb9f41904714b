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

// Layout check and sizing of a staged batch, on the device (one thread per read): mates back to back in increasing
// offset order inside the base buffer; k-mer slot bound (QueryIndexer.cpp:62-67); longest query.
struct StageInfo { unsigned long long bound; u32 max_q; u32 bad; };
__global__ void __launch_bounds__(256) k_validate_reads(const u64 *__restrict__ offsets, const u32 *__restrict__ lens, u32 n_reads,
                                                        u64 n_bases, int paired, StageInfo *__restrict__ info) {
    const u32 i = blockIdx.x * 256 + threadIdx.x;
    unsigned long long bound = 0;
    u32 ql = 0, bad = 0;
    if (i < n_reads) {
        u64 prev_end = i ? offsets[2 * i - 1] + lens[2 * i - 1] : 0;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
            const u64 off = offsets[2 * i + mt];
            const u32 len = lens[2 * i + mt];
            if (off < prev_end || off + len > n_bases) bad = 1;
            prev_end = off + len;
            if (mt == 1 && !paired) continue;
            const int kn = query_kmer_number((int) len);
            if (kn > 0) { bound += (unsigned long long) kn; ql += (u32) max_covered_length((int) len); }
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        bound += __shfl_xor_sync(0xFFFFFFFFu, bound, d);
        ql = max(ql, __shfl_xor_sync(0xFFFFFFFFu, ql, d));
        bad |= __shfl_xor_sync(0xFFFFFFFFu, bad, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (bound) atomicAdd(&info->bound, bound);
        if (ql) atomicMax(&info->max_q, ql);
        if (bad) atomicOr(&info->bad, 1u);
    }
}

constexpr size_t extract_smem_bytes() { return (size_t) EX_TILE * 2 * sizeof(mb_kmer); }   // every window, both strands

// ONE pass: a CTA extracts the k-mers of its 2048 window starts into a shared-memory staging buffer (one region per
// warp, slots = running count + ballot rank), every warp reserves its output range with one global atomic and copies
// its region out with coalesced 16-byte stores.  The query buffer is unordered by contract (the reference fills it from
// many threads, KmerExtractor.cpp:83-386, and sorts afterwards), so no prefix sum over tiles is needed.
//
// FMT2 (current metamer format): the three codon phases of a thread's 32 staged bases are translated ONCE into packed
// streams (5-bit amino acids, 3-bit codon ids; 10 codons per phase); the metamer of every window is then a bit-field
// of its phase's stream — forward strand with the first codon most significant, reverse strand with the stream built
// in the opposite order (MetamerScanner::next, KmerScanner.h:82-117).  The legacy base-21 format keeps the
// per-window translation.
//
// ROUTED (multi-GPU, mb_extract_kmers_routed): the staged k-mers of half a tile are grouped by owner shard in shared
// memory (second half of the staging area), one reservation per owner on this rank's LOCAL per-owner counter (every
// source owns a fixed region of every owner's receive buffer, so no remote atomic is needed), then every owner's run is
// written with coalesced 16-byte peer stores over NVLink.
template <bool FMT2, bool ROUTED>
__global__ void __launch_bounds__(EX_THREADS)
k_extract(ExtractArgs a, const u32 *__restrict__ tile_mate, const Tables *__restrict__ tables, mb_kmer *__restrict__ out,
          unsigned long long *__restrict__ out_count, u64 out_cap, const Route *__restrict__ route_g) {
    extern __shared__ __align__(16) unsigned char ex_raw[];
    ulonglong2 *s_out = reinterpret_cast<ulonglong2 *>(ex_raw);          // EX_TILE * 2 records
    __shared__ u16 s_code[EX_THREADS + 4];
    __shared__ u8 s_amb[EX_THREADS + 4];
    __shared__ u8 s_nuc[256];
    __shared__ u8 s_codonF[64];
    __shared__ u8 s_codonR[64];
    __shared__ u32 s_wfill[EX_THREADS / 32];
    __shared__ Route s_route;
    __shared__ u32 s_dcnt[MB_MAX_RANKS], s_dloc[MB_MAX_RANKS + 1];
    __shared__ u32 s_wcnt[EX_THREADS / 32][MB_MAX_RANKS];
    __shared__ unsigned long long s_dbase[MB_MAX_RANKS];

    const u32 tid = threadIdx.x, lane = tid & 31;
    if (ROUTED) {
        for (u32 i = tid; i < sizeof(Route) / 8; i += EX_THREADS) reinterpret_cast<u64 *>(&s_route)[i] = reinterpret_cast<const u64 *>(route_g)[i];
        if (tid < MB_MAX_RANKS * (EX_THREADS / 32)) (&s_wcnt[0][0])[tid] = 0;
    }
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

    // packed codon streams of the three phases (FMT2)
    u64 aaF[3], aaR[3];
    u32 dnaF[3], dnaR[3];
    if (FMT2) {
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            u64 af = 0, ar = 0;
            u32 df = 0, dr = 0;
#pragma unroll
            for (int k = 0; k < 10; ++k) {
                const int p = r + 3 * k;                                  // <= 29: bases p..p+2 lie inside the 32 staged bases
                const u32 six = (u32) (w >> (2 * p)) & 63u;
                const u32 cf = s_codonF[six], cr = s_codonR[six];
                af = (af << 5) | (cf & 31u);
                df = (df << 3) | (cf >> 5);
                ar |= (u64) (cr & 31u) << (5 * k);
                dr |= (cr >> 5) << (3 * k);
            }
            aaF[r] = af; dnaF[r] = df; aaR[r] = ar; dnaR[r] = dr;
        }
    }

    // mate of the first window start of this thread
    const u64 gfirst = g0 + (u64) tid * EX_PER_THREAD;
    u32 m = 0;
    u64 moff = 0;
    u32 mlen = 0;
    u64 next_off = ~0ull;                        // offset of mate m + 1: the mate changes when a window start reaches it
    const bool any = gfirst < a.n_bases && a.n_mates > 0;
    if (any) {
        m = mate_of(a.offsets, tile_mate[tile], tile_mate[tile + 1], gfirst);
        moff = __ldg(a.offsets + m);
        mlen = __ldg(a.lens + m);
        if (m + 1 < a.n_mates) next_off = __ldg(a.offsets + m + 1);
    }
    // staging: every warp owns a fixed region of the buffer (as many records as its lanes can emit), so a slot is the
    // warp's running count plus the lane's rank in the ballot: no shared-memory atomics
    constexpr u32 RW = (ROUTED ? EX_PER_THREAD / 2 : EX_PER_THREAD) * 2 * 32;
    const u32 wbase = (tid >> 5) * RW;
    u32 wfill = 0;
    const bool paired = a.seq_mode == 2;
    bool have_meta = false, usable = false;
    u32 pos_offset = 0;
    int used = 0;

#pragma unroll
    for (int half = 0; half < (ROUTED ? 2 : 1); ++half) {
#pragma unroll
    for (int i = half * (EX_PER_THREAD / 2); i < (ROUTED ? (half + 1) * (EX_PER_THREAD / 2) : EX_PER_THREAD); ++i) {
        bool ok = false;
        u32 x = 0;
        if (any) {
            const u64 g = gfirst + i;
            // advance to the mate that owns base g (mates are back to back, zero-length ones share an offset)
            while (g >= next_off) {
                ++m;
                moff = next_off;
                mlen = __ldg(a.lens + m);
                next_off = m + 1 < a.n_mates ? __ldg(a.offsets + m + 1) : ~0ull;
                have_meta = false;
            }
            if (g >= moff && g < moff + mlen) {
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
                x = (u32) (g - moff);
                // window [x, x+23] must end inside the frame segment [x%3, x%3 + used - 1]; an ambiguous base
                // inside the window: the scanners restart after it, so no window may cover it
                ok = usable && (int) (x + 24) <= (int) (x % 3) + used && !((ambw >> i) & 0xFFFFFFu);
            }
        }
        u64 vf, vr;
        if (FMT2) {
            const int r = i % 3, k0 = i / 3;
            vf = (((aaF[r] >> (5 * (2 - k0))) & 0xFFFFFFFFFFull) << 24) | ((dnaF[r] >> (3 * (2 - k0))) & 0xFFFFFFu);
            vr = (((aaR[r] >> (5 * k0)) & 0xFFFFFFFFFFull) << 24) | ((dnaR[r] >> (3 * k0)) & 0xFFFFFFu);
        } else {
            const u64 win = (w >> (2 * i)) & 0xFFFFFFFFFFFFull;
            vf = metamer_from_window(s_codonF, s_codonR, win, false, 1);
            vr = metamer_from_window(s_codonF, s_codonR, win, true, 1);
        }
        const u32 seqId = a.seq_base + (m >> 1) + 1;
        const u32 xm = x % 3;
        // forward strand: frame = begin = x % 3 (KmerExtractor.cpp:369-373)
        // reverse strand: begin = (L%3 - frame%3) mod 3 and begin == x%3  =>  frame = 3 + (L - x) mod 3
        const u32 fr_rev = 3u + (u32) (((int) (mlen % 3) - (int) xm + 3) % 3);
#pragma unroll
        for (int strand = 0; strand < 2; ++strand) {
            const u64 v = strand ? vr : vf;
            const bool emit = ok && (!FMT2 || !a.syncmer || is_syncmer(v, a.smer_len));
            const u32 mask = __ballot_sync(0xFFFFFFFFu, emit);
            if (emit) s_out[wbase + wfill + __popc(mask & ((1u << lane) - 1u))] = make_ulonglong2(v, make_qinfo(seqId, x + pos_offset, strand ? fr_rev : xm));
            wfill += __popc(mask);
        }
    }
    if (ROUTED) {
        // ---- group the <= EX_TILE staged records by owner, reserve, write runs to the owners ----
        if (lane == 0) s_wfill[tid >> 5] = wfill;
        wfill = 0;
        __syncthreads();
        const u32 nr = s_route.n;
        constexpr int PER = EX_TILE / EX_THREADS;          // == number of warps: pass j handles warp j's region, RW == EX_THREADS
        constexpr int NW = EX_THREADS / 32;
        static_assert(PER == NW && RW == EX_THREADS || !ROUTED, "routed grouping walks one staging region per pass");
        const u32 warp = tid >> 5;
        u32 total = 0;
#pragma unroll
        for (int j = 0; j < NW; ++j) total += s_wfill[j];
        u32 pos[PER];
        // rank of every record among the records of its warp with the same owner: one shared-memory atomic per record
        // on a per-warp counter (cost independent of the number of ranks)
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const u32 k = tid + j * EX_THREADS;
            pos[j] = 0;
            if (tid < s_wfill[j]) {
                const u32 d = route_owner(s_route, s_out[k].x);
                pos[j] = (d << 16) | atomicAdd(&s_wcnt[warp][d], 1u);
            }
        }
        __syncthreads();
        if (tid < nr) {                                       // per owner: offsets of the warps' runs, total, reservation
            u32 acc = 0;
#pragma unroll
            for (int w = 0; w < NW; ++w) { const u32 c = s_wcnt[w][tid]; s_wcnt[w][tid] = acc; acc += c; }
            s_dcnt[tid] = acc;
            s_dbase[tid] = acc ? atomicAdd(s_route.sent_k + tid, (unsigned long long) acc) : 0ull;
        }
        __syncthreads();
        if (tid == 0) {
            u32 acc = 0;
            for (u32 dd = 0; dd < nr; ++dd) { s_dloc[dd] = acc; acc += s_dcnt[dd]; }
            s_dloc[nr] = acc;
            if (total) atomicAdd(out_count, (unsigned long long) total);          // local statistic: k-mers this rank produced
        }
        __syncthreads();
        ulonglong2 *s_ord = s_out + EX_TILE;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const u32 k = tid + j * EX_THREADS;
            if (tid < s_wfill[j]) { const u32 d = pos[j] >> 16; s_ord[s_dloc[d] + s_wcnt[warp][d] + (pos[j] & 0xFFFFu)] = s_out[k]; }
        }
        __syncthreads();
        for (u32 k = tid; k < total; k += EX_THREADS) {
            u32 d = 0;
            while (d + 1 < nr && s_dloc[d + 1] <= k) ++d;
            const u64 idx = s_dbase[d] + (k - s_dloc[d]);
            if (idx < s_route.kmer_cap[d]) reinterpret_cast<ulonglong2 *>(s_route.kmers[d])[idx] = s_ord[k];
        }
        __syncthreads();
        if (tid < MB_MAX_RANKS * (EX_THREADS / 32)) (&s_wcnt[0][0])[tid] = 0;
        __syncthreads();
    }
    }   // half
    if (!ROUTED) {
        // every warp reserves the output range of its own region (one global atomic per warp) and copies it out with
        // coalesced 16-byte stores; no block barrier is needed
        __syncwarp();
        unsigned long long base = 0;
        if (lane == 0 && wfill) base = atomicAdd(out_count, (unsigned long long) wfill);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(out);
        for (u32 k = lane; k < wfill; k += 32)
            if (base + k < out_cap) dst[base + k] = s_out[wbase + k];
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
