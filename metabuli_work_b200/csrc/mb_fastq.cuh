// mb_fastq.cuh — input pipeline: FASTQ text -> packed read batch, on the device.
//
// Replaces, for plain four-line FASTQ, the host loop that the reference runs twice per input file:
// QueryIndexer::indexQueryFile (src/commons/QueryIndexer.cpp:30-147: a full parse only to count reads and
// k-mer slots) and KmerExtractor::loadChunkOfReads (src/commons/KmerExtractor.cpp:312-386) over KSeqWrapper.
// What the rest of the path needs from the text is (a) the bases of every read packed back to back with
// offsets and lengths — exactly the input of mb_stage_reads — and (b) where each read's name lies, so that
// the host can print it without scanning the text itself.
//
// B200 design.  The raw text goes to HBM as it is (one H2D copy, no host parsing).  Lines are found by a
// count / scan / scatter over 4 KB tiles (every '\n' gets its ordinal, line_start[j] = byte after the j-th
// newline), record i is lines 4i..4i+3 (checked: '@' and '+' markers, equal sequence and quality length,
// optional '\r'), the sequence lengths of both mates are scanned into packed offsets and one warp per mate
// copies its sequence line into place.  Everything is streaming byte work at HBM speed; the host gets back
// the number of reads, how many bytes of each text were consumed (whole records only, so it can carry the
// tail of a chunk over to the next one) and the name spans.
#pragma once
#include "mb_common.cuh"

namespace mb {

constexpr int FQ_THREADS = 256;
constexpr int FQ_PER = 16;                        // bytes per thread
constexpr int FQ_TILE = FQ_THREADS * FQ_PER;      // 4096 bytes per CTA

__device__ __forceinline__ u32 fq_newline_mask(uint4 v) {      // bit i <=> byte i of the 16 is '\n'
    u32 m = 0;
    const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int b = 0; b < 4; ++b) m |= (((w[k] >> (8 * b)) & 255u) == 10u ? 1u : 0u) << (4 * k + b);
    return m;
}
__device__ __forceinline__ uint4 fq_load16(const u8 *text, u64 n, u64 p) {   // text is padded to a multiple of 16 readable bytes
    uint4 v = make_uint4(0, 0, 0, 0);
    if (p < n) v = *reinterpret_cast<const uint4 *>(text + p);
    if (p + 16 > n && p < n) {                                   // bytes past the end never count
        u8 *b = reinterpret_cast<u8 *>(&v);
        for (u64 i = n - p; i < 16; ++i) b[i] = 0;
    }
    return v;
}

// pass 1: newlines per tile
__global__ void __launch_bounds__(FQ_THREADS) k_fq_count(const u8 *__restrict__ text, u64 n, u32 *__restrict__ tile_counts) {
    const u64 p = ((u64) blockIdx.x * FQ_THREADS + threadIdx.x) * FQ_PER;
    u32 c = (u32) __popc(fq_newline_mask(fq_load16(text, n, p)));
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
    __shared__ u32 s[FQ_THREADS / 32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        u32 t = 0;
        for (int i = 0; i < FQ_THREADS / 32; ++i) t += s[i];
        tile_counts[blockIdx.x] = t;
    }
}

// pass 2 (after the exclusive scan of the tile counts): line_start[j + 1] = byte after the j-th newline
__global__ void __launch_bounds__(FQ_THREADS) k_fq_lines(const u8 *__restrict__ text, u64 n, const u64 *__restrict__ tile_offsets,
                                                         u64 *__restrict__ line_start) {
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u64 p = ((u64) blockIdx.x * FQ_THREADS + tid) * FQ_PER;
    u32 m = fq_newline_mask(fq_load16(text, n, p));
    const u32 c = (u32) __popc(m);
    u32 incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (u32) d) incl += t;
    }
    __shared__ u32 s[FQ_THREADS / 32];
    if (lane == 31) s[warp] = incl;
    __syncthreads();
    u32 wb = 0;
    for (u32 i = 0; i < warp; ++i) wb += s[i];
    u64 idx = tile_offsets[blockIdx.x] + wb + incl - c;
    if (blockIdx.x == 0 && tid == 0) line_start[0] = 0;
    while (m) {
        const u32 b = (u32) __ffs((int) m) - 1u;
        m &= m - 1u;
        line_start[++idx] = p + b + 1;
    }
}

struct FqInfo {                 // device-side scalars of one indexing call
    u32 bad;                    // first kind of format error seen (0 = none): 1 '@', 2 '+', 3 length mismatch
    u32 pad;
    unsigned long long bad_record;
};

// pass 3: one thread per record of one file: checks, sequence span, name span
__global__ void __launch_bounds__(256) k_fq_records(const u8 *__restrict__ text, const u64 *__restrict__ line_start, u32 n_records, int mate,
                                                    int n_mates_filled, u32 *__restrict__ lens, u64 *__restrict__ src_off,
                                                    u64 *__restrict__ name_off, u32 *__restrict__ name_len, FqInfo *__restrict__ info) {
    const u32 i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n_records) return;
    const u64 s0 = line_start[4 * (u64) i], s1 = line_start[4 * (u64) i + 1], s2 = line_start[4 * (u64) i + 2],
              s3 = line_start[4 * (u64) i + 3], s4 = line_start[4 * (u64) i + 4];
    u32 bad = 0;
    if (text[s0] != '@') bad = 1;
    else if (text[s2] != '+') bad = 2;
    u64 seq_len = s2 - s1 - 1, qual_len = s4 - s3 - 1;
    if (seq_len && text[s1 + seq_len - 1] == '\r') --seq_len;
    if (qual_len && text[s3 + qual_len - 1] == '\r') --qual_len;
    if (!bad && seq_len != qual_len) bad = 3;
    if (seq_len > 0xFFFFFFFFull) { seq_len = 0; bad = 3; }
    if (bad && atomicCAS(&info->bad, 0u, bad) == 0u) info->bad_record = i;
    lens[2 * (u64) i + mate] = (u32) seq_len;
    src_off[2 * (u64) i + mate] = s1;
    if (n_mates_filled == 1) { lens[2 * (u64) i + 1] = 0; src_off[2 * (u64) i + 1] = s1; }
    if (name_off) {
        // the name runs from after '@' to the first white space (KSeqWrapper / kseq.h: name = first token of the header)
        u64 e = s0 + 1;
        const u64 line_end = s1 - 1;
        while (e < line_end && text[e] != ' ' && text[e] != '\t' && text[e] != '\r') ++e;
        name_off[i] = s0 + 1;
        name_len[i] = (u32) (e - (s0 + 1));
    }
}

// ---- exclusive scan of u32 -> u64 over any length: tile sums, scan of the sums (k_scan_counts), apply ----
constexpr int SCAN_PER = 8;
constexpr int SCAN_TILE = 256 * SCAN_PER;
__global__ void __launch_bounds__(256) k_scan_tile_sums(const u32 *__restrict__ v, u64 n, u32 *__restrict__ sums) {
    const u64 base = (u64) blockIdx.x * SCAN_TILE + (u64) threadIdx.x * SCAN_PER;
    u32 c = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER; ++k) if (base + k < n) c += v[base + k];
#pragma unroll
    for (int d = 16; d; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
    __shared__ u32 s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { u32 t = 0; for (int i = 0; i < 8; ++i) t += s[i]; sums[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(256) k_scan_tile_apply(const u32 *__restrict__ v, u64 n, const u64 *__restrict__ tile_base, u64 *__restrict__ out) {
    const u32 tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const u64 base = (u64) blockIdx.x * SCAN_TILE + (u64) tid * SCAN_PER;
    u32 x[SCAN_PER];
    u32 c = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER; ++k) { x[k] = base + k < n ? v[base + k] : 0u; c += x[k]; }
    u32 incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= (u32) d) incl += t;
    }
    __shared__ u32 s[8];
    if (lane == 31) s[warp] = incl;
    __syncthreads();
    u32 wb = 0;
    for (u32 i = 0; i < warp; ++i) wb += s[i];
    u64 run = tile_base[blockIdx.x] + wb + incl - c;
#pragma unroll
    for (int k = 0; k < SCAN_PER; ++k) { if (base + k < n) out[base + k] = run; run += x[k]; }
}

// pass 4: one warp per mate copies its sequence line into the packed batch
__global__ void __launch_bounds__(256) k_fq_pack(const u8 *__restrict__ text1, const u8 *__restrict__ text2, const u64 *__restrict__ src_off,
                                                 const u32 *__restrict__ lens, const u64 *__restrict__ offsets, u32 n_mates,
                                                 u8 *__restrict__ bases) {
    const u32 lane = threadIdx.x & 31;
    const u64 w = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (w >= n_mates) return;
    const u8 *src = ((w & 1) && text2 ? text2 : text1) + src_off[w];
    u8 *dst = bases + offsets[w];
    const u32 len = lens[w];
    for (u32 k = lane; k < len; k += 32) dst[k] = src[k];
}

} // namespace mb
