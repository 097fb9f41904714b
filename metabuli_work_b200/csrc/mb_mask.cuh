// mb_mask.cuh — `--mask 1`: low-complexity (tandem repeat) masking of the staged reads before k-mer extraction.
//
// Replaces SeqIterator::maskLowComplexityRegions (src/commons/SeqIterator.cpp:154-177) as KmerExtractor::processSequence
// calls it per mate (src/commons/KmerExtractor.cpp:328-335): tantan::maskSequences (lib/mmseqs/lib/tantan/tantan.cpp:482-527)
// with maxRepeatOffset 50, repeatProb 0.005, repeatEndProb 0.05, repeatOffsetProbDecay 0.9 and no gap states, i.e. the
// forward / backward recursions of tantan.cpp:312-450 over 50 repeat-offset states + background, rescaled every 16
// letters; a base whose repeat probability (a float) reaches --mask-prob, or whose letter is not a nucleotide code,
// becomes 'N'.
//
// One thread per mate: the recursion is sequential in the position, the 50 states of a position are independent except
// for one running sum, and the reference adds that sum in index order in double precision — so does this kernel, with
// every product and sum an explicit round-to-nearest operation (__dmul_rn / __dadd_rn: the compiler must not contract
// them into FMAs), which makes the probabilities bit-identical to tantan's scalar build (mcf_simd.h:507-513, simdDblLen 1;
// a build with -mavx2 adds four partial sums instead and can differ in the last bit of a probability).  The 50 state values
// of a thread live in shared memory, state-major (conflict-free); keeping them in registers instead (140 registers, fully
// unrolled state loops) measured slower: 35.9 vs 23.4 ms for mask + extraction of a million read pairs.  The letters of the
// window are four 64-bit occupancy masks (bit i of occ[c]: "the letter i+1 positions back is c"), so the likelihood ratio
// of state i is a bit test.  Algorithmic bytes: 2 reads + 1 write of the bases, 2 x 4 bytes of probability per base
// (global scratch), 8 bytes of scale factor per 16 bases; the kernel is bound by FP64 issue (about 1000 double
// operations per base), not by memory.
#pragma once

#include "mb_common.cuh"

namespace mb {

constexpr int MK_OFF = 50;            // maxRepeatOffset
constexpr int MK_THREADS = 96;        // 96 x 50 x 8 B = 37.5 KB of state per CTA

struct MaskModel {                    // Tantan::Tantan (tantan.cpp:81-135), computed on the host with the host's pow()
    double b2b, f2b, f2f0;
    double lr_match, lr_mismatch;     // ProbabilityMatrix over NucleotideMatrix("nucleotide.out") (BaseMatrix.h:83-96)
    double min_prob;                  // (double) (float) --mask-prob
    double b2f[MK_OFF];
};

// NucleotideMatrix::setupLetterMapping (lib/mmseqs/src/commons/NucleotideMatrix.cpp:17-58): A0 C1 T2 G3, anything else 4
__host__ __device__ inline u32 mask_letter_code(u32 ch) {
    ch &= ~0x20u;                     // toupper for letters; the other characters stay non-letters
    switch (ch) {
        case 'A': return 0;
        case 'C': case 'M': case 'Y': case 'H': return 1;
        case 'T': case 'U': case 'W': return 2;
        case 'G': case 'K': case 'B': case 'D': case 'V': case 'R': case 'S': return 3;
        default: return 4;
    }
}

__global__ void __launch_bounds__(MK_THREADS)
k_mask_mates(u8 *bases, const u64 *offsets, const u32 *lens, u32 n_mates, const __grid_constant__ MaskModel M, float *probs, double *scales) {
    __shared__ double s_fg[MK_OFF * MK_THREADS];
    __shared__ u8 s_code[256];
    const u32 tid = threadIdx.x;
    for (u32 c = tid; c < 256; c += MK_THREADS) {
        // only ASCII letters fold their case; '@' | 0x20 etc. must not turn into letters
        const bool letter = (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z');
        // byte 255 keeps the table's initial 0: the reference's mapping loops stop before UCHAR_MAX (NucleotideMatrix.cpp:18)
        s_code[c] = (u8) (c == 255 ? 0u : letter ? mask_letter_code(c) : 4u);
    }
    __syncthreads();
    const u32 m = blockIdx.x * MK_THREADS + tid;
    if (m >= n_mates) return;
    const u32 L = lens[m];
    if (L == 0) return;
    const u64 off = offsets[m];
    u8 *seq = bases + off;
    float *pr = probs + off;
    double *sc = scales + (off >> 4) + m;
    double *fg = s_fg + tid;                                   // fg[i * MK_THREADS]
#pragma unroll 10
    for (int i = 0; i < MK_OFF; ++i) fg[i * MK_THREADS] = 0.0;
    // ---- forward (tantan.cpp:312-354, 421-431) ----
    u64 occ[4] = {0, 0, 0, 0};
    double bg = 1.0;
    for (u32 pos = 0; pos < L; ++pos) {
        const u32 cur = s_code[seq[pos]];
        const u64 same = cur < 4 ? occ[cur] : 0ull;
        const int maxOff = pos < (u32) MK_OFF ? (int) pos : MK_OFF;
        const double b = bg;
        double from = 0.0;
#pragma unroll 5
        for (int i = 0; i < maxOff; ++i) {
            const double f = fg[i * MK_THREADS];
            from = __dadd_rn(from, f);
            const double lr = ((same >> i) & 1ull) ? M.lr_match : M.lr_mismatch;
            fg[i * MK_THREADS] = __dmul_rn(__dadd_rn(__dmul_rn(b, M.b2f[i]), __dmul_rn(f, M.f2f0)), lr);
        }
        bg = __dadd_rn(__dmul_rn(b, M.b2b), __dmul_rn(from, M.f2b));
        if ((pos & 15u) == 15u) {
            const double s = __ddiv_rn(1.0, bg);
            sc[pos >> 4] = s;
            bg = __dmul_rn(bg, s);
#pragma unroll 10
            for (int i = 0; i < MK_OFF; ++i) fg[i * MK_THREADS] = __dmul_rn(fg[i * MK_THREADS], s);
        }
        pr[pos] = __double2float_rn(bg);
#pragma unroll
        for (int c = 0; c < 4; ++c) occ[c] = (occ[c] << 1) | (u64) (cur == (u32) c);
    }
    // forwardTotal (tantan.cpp:136-143): all 50 states, in order
    double all = 0.0;
#pragma unroll 10
    for (int i = 0; i < MK_OFF; ++i) all = __dadd_rn(all, fg[i * MK_THREADS]);
    const double z = __dadd_rn(__dmul_rn(bg, M.b2b), __dmul_rn(all, M.f2b));
    // ---- backward (tantan.cpp:356-397, 433-446) ----
    bg = M.b2b;
#pragma unroll 10
    for (int i = 0; i < MK_OFF; ++i) fg[i * MK_THREADS] = M.f2b;
    // occ now describes the window of position L (bit i: letter L-1-i); position L-1 needs bit i: letter L-2-i
    for (u32 pos = L; pos-- > 0;) {
#pragma unroll
        for (int c = 0; c < 4; ++c) occ[c] >>= 1;
        // the masks hold 64 letters but lose one at the top per step: the letter that enters state 49's view is re-read
        if (pos >= (u32) MK_OFF) {
            const u32 far = s_code[seq[pos - MK_OFF]];
            if (far < 4) occ[far] |= 1ull << (MK_OFF - 1);
        }
        const u32 cur = s_code[seq[pos]];
        const double nonRepeat = __ddiv_rn(__dmul_rn((double) pr[pos], bg), z);
        const float p = __fsub_rn(1.0f, __double2float_rn(nonRepeat));
        if ((double) p >= M.min_prob || cur == 4u) seq[pos] = 'N';       // tantan.cpp:503-517, SeqIterator.cpp:172-175
        if ((pos & 15u) == 15u) {
            const double s = sc[pos >> 4];
            bg = __dmul_rn(bg, s);
#pragma unroll 10
            for (int i = 0; i < MK_OFF; ++i) fg[i * MK_THREADS] = __dmul_rn(fg[i * MK_THREADS], s);
        }
        const u64 same = cur < 4 ? occ[cur] : 0ull;
        const int maxOff = pos < (u32) MK_OFF ? (int) pos : MK_OFF;
        const double toBackground = __dmul_rn(M.f2b, bg);
        double toFg = 0.0;
#pragma unroll 5
        for (int i = 0; i < maxOff; ++i) {
            const double lr = ((same >> i) & 1ull) ? M.lr_match : M.lr_mismatch;
            const double f = __dmul_rn(fg[i * MK_THREADS], lr);
            toFg = __dadd_rn(toFg, __dmul_rn(M.b2f[i], f));
            fg[i * MK_THREADS] = __dadd_rn(toBackground, __dmul_rn(M.f2f0, f));
        }
        bg = __dadd_rn(__dmul_rn(M.b2b, bg), toFg);
    }
}

} // namespace mb
