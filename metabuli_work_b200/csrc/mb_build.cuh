// mb_build.cuh — the DB-build inner loop on the device.
//
// Replaces, for one buffer of target k-mers, what IndexCreator::createIndex does between extraction and the files
// (src/commons/IndexCreator.cpp:262-280):
//   SORT_PARALLEL(buffer, Kmer::compareTargetKmer)            Kmer.h:77-87                 -> the hand-written radix sort
//   IndexCreator::filterKmers<FilterMode::DB_CREATION>        IndexCreator.h:475-615       -> k_build_heads + scan + k_build_reduce
//   IndexCreator::writeTargetFilesAndSplits / getDiffIdx      IndexCreator.cpp:811-886     -> k_build_wordcount + scan +
//                                                                                             k_build_encode, k_build_splits
// A target record is {value, taxId, speciesId}: 16 bytes, the layout of mb_kmer with (taxId | speciesId << 32) in the
// second word.  All kernels are streaming passes over 16-byte records or 8-byte values (HBM-bound); the sort dominates:
// 8 value passes + 1-3 species passes of 32 bytes per record each.
#pragma once

#include "mb_common.cuh"

namespace mb {

struct TargetKey {          // LSD order: species digits first, then the eight value bytes -> sorted by (value, species)
    int sp_passes;
    __device__ __forceinline__ u32 digit(const mb_kmer &r, int pass) const {
        return pass < sp_passes ? (u32) (r.qinfo >> (32 + 8 * pass)) & 255u : (u32) (r.value >> (8 * (pass - sp_passes))) & 255u;
    }
};

__global__ void __launch_bounds__(256) k_build_pack(const u64 *__restrict__ values, const int32_t *__restrict__ taxids,
                                                    const int32_t *__restrict__ species, u64 n, mb_kmer *__restrict__ out) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    mb_kmer r;
    r.value = values[i];
    r.qinfo = (u64) (u32) taxids[i] | ((u64) (u32) species[i] << 32);
    out[i] = r;
}

// areKmersDuplicate<DB_CREATION> (IndexCreator.h:617-625): same value and same speciesId; a record that is no duplicate of
// its predecessor heads a run
__global__ void __launch_bounds__(256) k_build_heads(const mb_kmer *__restrict__ rec, u64 n, u32 *__restrict__ head) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    bool h = true;
    if (i) {
        const mb_kmer a = rec[i - 1], b = rec[i];
        h = a.value != b.value || (a.qinfo >> 32) != (b.qinfo >> 32);
    }
    head[i] = h ? 1u : 0u;
}

// the run's first record is kept, its taxId becomes taxonomy->LCA(all taxIds of the run) (IndexCreator.h:535-577,
// NcbiTaxonomy::LCA(vector) NcbiTaxonomy.cpp:313-333: taxa without a node are skipped)
__global__ void __launch_bounds__(256) k_build_reduce(const mb_kmer *__restrict__ rec, u64 n, const u32 *__restrict__ head,
                                                      const u64 *__restrict__ pos, const TaxView tv, u64 *__restrict__ out_v,
                                                      u32 *__restrict__ out_t) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n || !head[i]) return;
    const mb_kmer r = rec[i];
    int32_t acc = (int32_t) (u32) r.qinfo;
    for (u64 j = i + 1; j < n && !head[j]; ++j) acc = tax_lca(tv, acc, (int32_t) (u32) rec[j].qinfo);
    out_v[pos[i]] = r.value;
    out_t[pos[i]] = (u32) acc;
}

// getDiffIdx (IndexCreator.cpp:868-886): the difference to the previous k-mer in 15-bit groups, most significant first, the
// last group flagged; the first k-mer is a difference from 0
__device__ __forceinline__ u32 build_words_of(u64 d) {
    return 1u + (u32) ((d >> 15) != 0) + (u32) ((d >> 30) != 0) + (u32) ((d >> 45) != 0) + (u32) ((d >> 60) != 0);
}
__global__ void __launch_bounds__(256) k_build_wordcount(const u64 *__restrict__ v, u64 n, u32 *__restrict__ cnt) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    cnt[i] = build_words_of(v[i] - (i ? v[i - 1] : 0ull));
}
__global__ void __launch_bounds__(256) k_build_encode(const u64 *__restrict__ v, u64 n, const u64 *__restrict__ woff, u16 *__restrict__ words) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const u64 d = v[i] - (i ? v[i - 1] : 0ull);
    const u32 nw = build_words_of(d);
    u16 *w = words + woff[i];
    for (u32 k = 0; k < nw; ++k) {
        const u32 sh = 15u * (nw - 1 - k);
        const u16 g = (u16) ((d >> sh) & 0x7FFFu);
        w[k] = k + 1 == nw ? (u16) (g | 0x8000u) : g;
    }
}

// writeTargetFilesAndSplits (IndexCreator.cpp:811-866): when the os-th multiple of uniq / (splitNum - 1) k-mers has been
// written (k-mer j0), the next k-mer whose amino-acid part differs from k-mer j0's opens a split.  One thread per
// threshold finds that k-mer; the host replays the reference's one-flag state machine over the answers (two thresholds
// inside one amino-acid group share their split).
struct SplitProbe { u64 j; u64 adkmer; u64 words_through; };
__global__ void __launch_bounds__(256) k_build_splits(const u64 *__restrict__ v, u64 n, const u64 *__restrict__ woff, u64 size_of_split,
                                                      int split_num, SplitProbe *__restrict__ out) {
    const int os = blockIdx.x * 256 + threadIdx.x;
    if (os >= split_num) return;
    SplitProbe p = {~0ull, 0, 0};
    if (os >= 1 && size_of_split > 0 && (u64) os * size_of_split <= n) {
        const u64 j0 = (u64) os * size_of_split - 1;
        const u64 aa = v[j0] & ~0xFFFFFFull;
        u64 j = j0 + 1;
        while (j < n && (v[j] & ~0xFFFFFFull) == aa) ++j;
        if (j < n) { p.j = j; p.adkmer = v[j]; p.words_through = woff[j + 1]; }
    }
    out[os] = p;
}

} // namespace mb
