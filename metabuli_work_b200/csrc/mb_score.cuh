// mb_score.cuh — K6: per-read scoring, best-species selection, LCA and sub-species assignment.
//
// Replaces Classifier::assignTaxonomy (src/commons/Classifier.cpp:166-208) and
// Taxonomer::chooseBestTaxon with everything below it (src/commons/Taxonomer.cpp:130-699):
// getBestSpeciesMatches, getMatchPaths, combineMatchPaths / trimMatchPath, filterRedundantMatches,
// lowerRankClassification / getSpeciesCladeCounts / BFS, and the Match score accessors
// (src/commons/Match.h:32-86).
//
// The per-read logic is written once as score_read(), a pure function over the read's slice of the
// sorted match list and a scratch arena.  On the GPU one warp owns one read at a time (reads are
// handed out by an atomic ticket, so long and short reads balance); the taxonomy (parent / depth /
// species tables, a few MB) stays L2-resident and LCA is a depth-equalising parent walk, which on a
// tree equals the reference's Euler-tour RMQ.  All partial scores are multiples of 0.5 and the only
// rounding operations (one division by the read length, the tie average) are IEEE fp32 operations
// performed in the reference's order, so scores come out bit-identical.
#pragma once
#include "mb_common.cuh"
#include <string.h>

namespace mb {

struct ScoreCfg {
    int format;            // Kmer_format
    int dna_shift;         // Taxonomer.cpp:34-43: 3, or 3*(8-s) with syncmers
    int max_codon_shift;   // 1, or 8-s with syncmers
    int denominator;       // Taxonomer.cpp:44-48: 100 short reads, 1000 long reads
    int min_cons_cnt, min_cons_cnt_euk;
    float min_score, min_sp_score, tie_ratio;
};

struct PathRec {           // Taxonomer.h:35-58 MatchPath
    int start, end;
    float score;
    int ham, depth;
    int sm, em;            // indices of the first / last match of the path
};

struct ReadScratch {
    PathRec *local;        // [cap_n] one per match of the current (species, frame) group
    PathRec *paths;        // [cap_n]
    PathRec *combined;     // [cap_n]
    u8 *linked;            // [cap_n]
    int32_t *sp_id;        // [cap_n] species with a score, in match order
    float *sp_score;       // [cap_n]
    int32_t *bucket_tax;   // [cap_b] filterRedundantMatches buckets
    u8 *bucket_ham;        // [cap_b] 255 = empty
    int32_t *tc_id;        // [cap_b] taxCnt in ascending taxID order
    int32_t *tc_cnt;       // [cap_b]
    int32_t *cl_id;        // [cap_c] clade table of lowerRankClassification
    int32_t *cl_cnt;       // [cap_c]
    int32_t *cl_parent;    // [cap_c] index into cl_* of the parent, -1 for the root
    int cap_n, cap_b, cap_c;
};

// Match::getScore pieces (Match.h:32-44): a codon with Hamming 0 scores 3, else 2 - 0.5*h
MB_HD float codon_score(u32 h) { return h == 0 ? 3.0f : 2.0f - 0.5f * (float) h; }
MB_HD float match_score(const mb_match &m) {
    float s = 0.f;
    for (int k = 0; k < 8; ++k) s += codon_score((m.rightEndHamming >> (2 * k)) & 3u);
    return s;
}
// Match::getRightPartScore / getLeftPartScore / get*PartHammingDist (Match.h:46-86)
MB_HD float right_part_score(const mb_match &m, int range) {
    float s = 0.f;
    for (int k = 0; k < range; ++k) s += codon_score((m.rightEndHamming >> (2 * k)) & 3u);
    return s;
}
MB_HD float left_part_score(const mb_match &m, int range) {
    float s = 0.f;
    for (int k = 0; k < range; ++k) s += codon_score((m.rightEndHamming >> (14 - 2 * k)) & 3u);
    return s;
}
MB_HD int right_part_ham(const mb_match &m, int range) {
    int s = 0;
    for (int k = 0; k < range; ++k) s += (m.rightEndHamming >> (2 * k)) & 3;
    return s;
}
MB_HD int left_part_ham(const mb_match &m, int range) {
    int s = 0;
    for (int k = 0; k < range; ++k) s += (m.rightEndHamming >> (14 - 2 * k)) & 3;
    return s;
}

// Taxonomer::isConsecutive / isConsecutive2 (Taxonomer.cpp:671-699): DNA overlap of two matches
// `shift` codons apart; which end overlaps depends on the metamer format.
MB_HD bool dna_consecutive(const mb_match &a, const mb_match &b, int shift, int format) {
    const u32 keep = (1u << (24 - 3 * shift)) - 1u;
    if (format == 2) return (a.dnaEncoding & keep) == (b.dnaEncoding >> (3 * shift));
    return (a.dnaEncoding >> (3 * shift)) == (b.dnaEncoding & keep);
}

MB_HD void path_init(PathRec &p, const mb_match *m, int i) {
    p.start = (int) qi_pos(m[i].qinfo);
    p.end = p.start + 23;
    p.score = match_score(m[i]);
    p.ham = m[i].hamming;
    p.depth = 1;
    p.sm = i;
    p.em = i;
}

// Taxonomer::getMatchPaths (Taxonomer.cpp:487-648): matches [b,e) share species and frame and are
// ordered by (pos, hamming, dnaEncoding).  Appends to scr.paths[n_paths...].
MB_HD void match_paths(const mb_match *m, int b, int e, int min_depth, const ScoreCfg &cfg, ReadScratch &scr, int &n_paths) {
    const bool forward = qi_frame(m[b].qinfo) < 3;
    PathRec *local = scr.local;
    u8 *linked = scr.linked;
    for (int i = 0; i < e - b; ++i) linked[i] = 0;
    int i = b;
    u32 cur_pos = qi_pos(m[b].qinfo);
    const int cur_b0 = i;
    while (i < e && qi_pos(m[i].qinfo) == cur_pos) { path_init(local[i - b], m, i); ++i; }
    int cur_b = cur_b0, cur_e = i;
    while (i < e) {
        const u32 next_pos = qi_pos(m[i].qinfo);
        const int next_b = i;
        while (i < e && qi_pos(m[i].qinfo) == next_pos) { path_init(local[i - b], m, i); ++i; }
        const int next_e = i;
        const int shift = (int) ((next_pos - cur_pos) / 3);
        if (shift > 0 && shift <= cfg.max_codon_shift) {
            for (int nx = next_b; nx < next_e; ++nx) {
                // calScoreIncrement / calHammingDistIncrement (Taxonomer.cpp:650-669)
                float inc = 0.f;
                int hinc = 0;
                for (int k = 0; k < shift; ++k) {
                    const u32 h = (m[nx].rightEndHamming >> (2 * k)) & 3u;
                    inc += codon_score(h);
                    hinc += (int) h;
                }
                int best = -1;
                float best_score = 0.f;
                for (int c = cur_b; c < cur_e; ++c) {
                    const bool ok = forward ? dna_consecutive(m[c], m[nx], shift, cfg.format)
                                            : dna_consecutive(m[nx], m[c], shift, cfg.format);
                    if (ok) {
                        linked[c - b] = 1;
                        if (local[c - b].score > best_score) { best = c; best_score = local[c - b].score; }
                    }
                }
                if (best >= 0) {
                    const PathRec &bp = local[best - b];
                    PathRec &t = local[nx - b];
                    t.start = bp.start;
                    t.score = bp.score + inc;
                    t.ham = bp.ham + hinc;
                    t.depth = bp.depth + shift;
                    t.sm = bp.sm;
                }
            }
        }
        for (int c = cur_b; c < cur_e; ++c)
            if (!linked[c - b] && local[c - b].depth >= min_depth) scr.paths[n_paths++] = local[c - b];
        if (i == e)
            for (int nx = next_b; nx < next_e; ++nx)
                if (local[nx - b].depth >= min_depth) scr.paths[n_paths++] = local[nx - b];
        cur_b = next_b; cur_e = next_e; cur_pos = next_pos;
    }
}

// comparator of combineMatchPaths (Taxonomer.cpp:417-426)
MB_HD bool path_before(const PathRec &a, const PathRec &b) {
    if (a.score != b.score) return a.score > b.score;
    if (a.ham != b.ham) return a.ham < b.ham;
    return a.start > b.start;
}

// The reference sorts the paths with std::sort.  Its comparator can tie, so the order of tied
// paths is whatever libstdc++'s introsort leaves; to stay bit-exact this is the same algorithm
// (bits/stl_algo.h: __introsort_loop with median-of-three + __final_insertion_sort, threshold 16).
MB_HD void path_swap(PathRec &a, PathRec &b) { PathRec t = a; a = b; b = t; }
MB_HD void path_unguarded_linear_insert(PathRec *last) {
    PathRec val = *last;
    PathRec *next = last - 1;
    while (path_before(val, *next)) { *last = *next; last = next; --next; }
    *last = val;
}
MB_HD void path_insertion_sort(PathRec *first, PathRec *last) {
    if (first == last) return;
    for (PathRec *i = first + 1; i != last; ++i) {
        if (path_before(*i, *first)) {
            PathRec val = *i;
            for (PathRec *p = i; p != first; --p) *p = *(p - 1);
            *first = val;
        } else path_unguarded_linear_insert(i);
    }
}
MB_HD void path_adjust_heap(PathRec *first, long hole, long len, PathRec value) {
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (path_before(first[child], first[child - 1])) --child;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    long parent = (hole - 1) / 2;                       // __push_heap
    while (hole > top && path_before(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}
MB_HD void path_heap_sort(PathRec *first, PathRec *last) { // __partial_sort(first, last, last)
    const long len = last - first;
    if (len >= 2)
        for (long parent = (len - 2) / 2;; --parent) {   // __make_heap
            path_adjust_heap(first, parent, len, first[parent]);
            if (parent == 0) break;
        }
    while (last - first > 1) {                           // __sort_heap
        --last;
        PathRec value = *last;
        *last = *first;
        path_adjust_heap(first, 0, last - first, value);
    }
}
MB_HD void path_sort(PathRec *first, PathRec *last) {
    const long n = last - first;
    if (n <= 1) return;
    if (n > 16) {
        // explicit stack for the recursion on the right part
        PathRec *stk_first[64];
        PathRec *stk_last[64];
        int stk_depth[64];
        int sp = 0;
        int lg = 0;
        for (long t = n; t > 1; t >>= 1) ++lg;
        stk_first[sp] = first; stk_last[sp] = last; stk_depth[sp] = 2 * lg; ++sp;
        while (sp > 0) {
            --sp;
            PathRec *f = stk_first[sp], *l = stk_last[sp];
            int depth = stk_depth[sp];
            while (l - f > 16) {
                if (depth == 0) { path_heap_sort(f, l); break; }
                --depth;
                // __move_median_to_first(f, f+1, mid, l-1)
                PathRec *a = f + 1, *bm = f + (l - f) / 2, *c = l - 1;
                if (path_before(*a, *bm)) {
                    if (path_before(*bm, *c)) path_swap(*f, *bm);
                    else if (path_before(*a, *c)) path_swap(*f, *c);
                    else path_swap(*f, *a);
                } else if (path_before(*a, *c)) path_swap(*f, *a);
                else if (path_before(*bm, *c)) path_swap(*f, *c);
                else path_swap(*f, *bm);
                // __unguarded_partition(f+1, l, pivot = f)
                PathRec *lo = f + 1, *hi = l;
                while (true) {
                    while (path_before(*lo, *f)) ++lo;
                    --hi;
                    while (path_before(*f, *hi)) --hi;
                    if (!(lo < hi)) break;
                    path_swap(*lo, *hi);
                    ++lo;
                }
                // recurse on [lo, l), continue with [f, lo)
                if (sp < 64) { stk_first[sp] = lo; stk_last[sp] = l; stk_depth[sp] = depth; ++sp; }
                l = lo;
            }
        }
        path_insertion_sort(first, first + 16);
        for (PathRec *i = first + 16; i != last; ++i) path_unguarded_linear_insert(i);
    } else path_insertion_sort(first, last);
}

// Taxonomer::combineMatchPaths + trimMatchPath (Taxonomer.cpp:410-485) over scr.paths[0, n_paths)
MB_HD float combine_paths(const mb_match *m, ReadScratch &scr, int n_paths, int read_length) {
    PathRec *paths = scr.paths, *comb = scr.combined;
    path_sort(paths, paths + n_paths);
    int n_comb = 0;
    float score = 0.f;
    for (int i = 0; i < n_paths; ++i) {
        PathRec cand = paths[i];
        if (n_comb == 0) { comb[n_comb++] = cand; score += cand.score; continue; }
        bool rejected = false;
        for (int j = 0; j < n_comb; ++j) {
            const PathRec &kept = comb[j];
            if (cand.end < kept.start || kept.end < cand.start) continue;
            const int lo = cand.start > kept.start ? cand.start : kept.start;
            const int hi = cand.end < kept.end ? cand.end : kept.end;
            const int overlap = hi - lo + 1;
            if (overlap == cand.end - cand.start + 1 || overlap >= 24) { rejected = true; break; }
            if (cand.start < kept.start) {
                cand.end = kept.start - 1;
                int h = cand.ham - right_part_ham(m[cand.em], overlap / 3);
                cand.ham = h > 0 ? h : 0;
                cand.score = cand.score - right_part_score(m[cand.em], overlap / 3) - (float) (overlap % 3);
            } else {
                cand.start = kept.end + 1;
                int h = cand.ham - left_part_ham(m[cand.sm], overlap / 3);
                cand.ham = h > 0 ? h : 0;
                cand.score = cand.score - left_part_score(m[cand.sm], overlap / 3) - (float) (overlap % 3);
            }
        }
        if (!rejected) { comb[n_comb++] = cand; score += cand.score; }
    }
    return score / (float) read_length;
}

struct BestSpecies { int32_t tax; float score; bool lca; int rb, re; };

// Taxonomer::getBestSpeciesMatches (Taxonomer.cpp:316-408); matches [b,e) belong to one read
MB_HD BestSpecies best_species(const mb_match *m, int b, int e, int query_length, const ScoreCfg &cfg,
                               const TaxView &tax, ReadScratch &scr) {
    BestSpecies out;
    out.tax = 0; out.score = 0.f; out.lca = false; out.rb = 0; out.re = 0;
    int n_sp = 0, meaningful = 0;
    float best_sp = 0.f;
    int i = b;
    while (i < e) {
        const int32_t species = m[i].speciesId;
        const int sp_begin = i;
        int n_paths = 0;
        int min_depth = cfg.min_cons_cnt;
        if (tax_is_ancestor(tax, tax.eukaryota, species)) min_depth = cfg.min_cons_cnt_euk;
        while (i < e && m[i].speciesId == species) {
            const u32 frame = qi_frame(m[i].qinfo);
            const int fb = i;
            while (i < e && m[i].speciesId == species && qi_frame(m[i].qinfo) == frame) ++i;
            if (i - fb > 1) match_paths(m, fb, i, min_depth, cfg, scr, n_paths);
        }
        if (n_paths > 0) {
            float score = combine_paths(m, scr, n_paths, query_length);
            score = score < 1.0f ? score : 1.0f;
            if (score < cfg.min_score) continue;
            scr.sp_id[n_sp] = species;
            scr.sp_score[n_sp] = score;
            ++n_sp;
            if (score > 0.f) ++meaningful;
            if (score > best_sp) { best_sp = score; out.rb = sp_begin; out.re = i; }
        }
    }
    if (meaningful == 0) { out.score = 0.f; return out; }
    // species within tie_ratio of the best (Taxonomer.cpp:388-402)
    int n_top = 0;
    int32_t lca = 0;
    bool have = false;
    int32_t single = 0;
    const float cut = best_sp * cfg.tie_ratio;
    for (int k = 0; k < n_sp; ++k) {
        if (scr.sp_score[k] >= cut) {
            const int32_t t = scr.sp_id[k];
            if (n_top == 0) single = t;
            ++n_top;
            out.score += scr.sp_score[k];
            // NcbiTaxonomy::LCA(vector) (NcbiTaxonomy.cpp:313-333): fold over the existing taxa
            if (tax_exists(tax, t)) {
                if (!have) { lca = t; have = true; }
                else lca = tax_lca_existing(tax, lca, t);
            }
        }
    }
    if (n_top > 1) {
        out.lca = true;
        out.tax = have ? lca : 0;
        out.score /= (float) n_top;
        return out;
    }
    out.tax = single;
    return out;
}

// Everything of Taxonomer::chooseBestTaxon after getBestSpeciesMatches.  Returns the number of taxCnt entries left
// in scr.tc_id / scr.tc_cnt.
// buckets_ready: scr.bucket_ham / scr.bucket_tax were already filled (fill_buckets_warp); n_tc_ready >= 0: so was the taxCnt
// list scr.tc_id / scr.tc_cnt (count_taxa_warp), with that many entries.
MB_HD int finish_read(const mb_match *m, const BestSpecies &best, int query_length, const ScoreCfg &cfg, const TaxView &tax,
                      ReadScratch &scr, mb_read_result &r, u32 *error_flag, bool buckets_ready = false, int n_tc_ready = -1) {
    // Taxonomer::chooseBestTaxon (Taxonomer.cpp:130-202), em == false
    if (best.score == 0.f || best.score < cfg.min_score) { r.score = best.score; return 0; }
    if (best.lca) { r.is_classified = 1; r.classification = best.tax; r.score = best.score; return 0; }

    // filterRedundantMatches (Taxonomer.cpp:205-241)
    const int max_q = (query_length + 3) / cfg.dna_shift;
    if (max_q + 1 > scr.cap_b) { *error_flag = 2; return 0; }
    if (!buckets_ready) {
    for (int k = 0; k <= max_q; ++k) { scr.bucket_ham[k] = 255; scr.bucket_tax[k] = 0; }
    {
        // the record of the next match is requested while the current one is bucketed (one thread per read: every load
        // is a round trip to L1/L2 that nothing else hides)
        const u64 *mw = reinterpret_cast<const u64 *>(m);
        u64 n0 = 0, n1 = 0, n2 = 0;
        if (best.rb < best.re) { n0 = mw[best.rb * 3]; n1 = mw[best.rb * 3 + 1]; n2 = mw[best.rb * 3 + 2]; }
        for (int i = best.rb; i < best.re; ++i) {
            const u64 w0 = n0, w1 = n1, w2 = n2;
            if (i + 1 < best.re) { n0 = mw[(i + 1) * 3]; n1 = mw[(i + 1) * 3 + 1]; n2 = mw[(i + 1) * 3 + 2]; }
            // dna_shift is 3 (plain) or 9 (syncmers, s = 5) in practice: constant divisions instead of the general sequence
            const u32 pos_i = qi_pos(w0);
            const int qd = (int) (cfg.dna_shift == 3 ? pos_i / 3u : cfg.dna_shift == 9 ? pos_i / 9u : pos_i / (u32) cfg.dna_shift);
            if (qd > max_q) continue;
            const u8 h = (u8) (w2 >> 48);
            const int32_t target = (int32_t) (u32) w1;
            const u8 bh = scr.bucket_ham[qd];
            if (bh == 255 || h < bh) { scr.bucket_tax[qd] = target; scr.bucket_ham[qd] = h; }
            else if (h == bh) {
                const int32_t cur_tax = scr.bucket_tax[qd];
                if (cur_tax != target) scr.bucket_tax[qd] = tax_lca(tax, cur_tax, target);
            }
        }
    }
    }
    int n_tc = 0;                                   // taxCnt as a sorted array (std::map order)
    if (n_tc_ready >= 0) n_tc = n_tc_ready;
    else {
        // neighbouring buckets mostly carry the same taxon: runs are counted in a register and entered once (the arrays
        // live in the per-thread scratch slice, every access is a round trip)
        int32_t run_t = 0;
        int run = 0;
        for (int k = 0; k <= max_q + 1; ++k) {
            const bool last = k > max_q;
            int32_t t = 0;
            if (!last) {
                if (scr.bucket_ham[k] == 255) continue;
                t = scr.bucket_tax[k];
                if (run && t == run_t) { ++run; continue; }
            }
            if (run) {
                int p = 0;
                while (p < n_tc && scr.tc_id[p] < run_t) ++p;
                if (p < n_tc && scr.tc_id[p] == run_t) scr.tc_cnt[p] += run;
                else {
                    for (int z = n_tc; z > p; --z) { scr.tc_id[z] = scr.tc_id[z - 1]; scr.tc_cnt[z] = scr.tc_cnt[z - 1]; }
                    scr.tc_id[p] = run_t; scr.tc_cnt[p] = run;
                    ++n_tc;
                }
            }
            run_t = t; run = last ? 0 : 1;
        }
    }
    r.is_classified = 1;
    r.score = best.score;
    if (best.score < cfg.min_sp_score) {
        // parent of the species-rank ancestor (Taxonomer.cpp:186-190)
        int32_t sp = 0;
        if (best.tax > 1 && tax_exists(tax, best.tax)) sp = tax.species_rank_of[best.tax];
        r.classification = (sp != 0 && tax_exists(tax, sp)) ? tax.parent[sp] : 0;
        return n_tc;
    }
    // lowerRankClassification / getSpeciesCladeCounts / BFS (Taxonomer.cpp:252-314), accessionLevel != 2
    const int32_t species = best.tax;
    int n_cl = 0;
    bool overflow = false;
    for (int k = 0; k < n_tc && !overflow; ++k) {
        int32_t t = scr.tc_id[k];
        const int cnt = scr.tc_cnt[k];
        if (t == 0 || !tax_exists(tax, t)) continue;
        int child_slot = -1;
        int guard = 0;
        while (true) {
            int slot = 0;
            while (slot < n_cl && scr.cl_id[slot] != t) ++slot;
            if (slot == n_cl) {
                if (n_cl >= scr.cap_c) { overflow = true; break; }
                scr.cl_id[slot] = t; scr.cl_cnt[slot] = 0; scr.cl_parent[slot] = -1;
                ++n_cl;
            }
            scr.cl_cnt[slot] += cnt;
            if (child_slot >= 0) scr.cl_parent[child_slot] = slot;
            child_slot = slot;
            if (t == species || guard++ >= 1000) break;
            const int32_t p = tax.parent[t];
            if (p == t || !tax_exists(tax, p)) break;
            t = p;
        }
    }
    if (overflow) { *error_flag = 3; r.classification = species; return n_tc; }
    int root = 0;
    while (root < n_cl && scr.cl_id[root] != species) ++root;
    if (root == n_cl) { r.classification = species; return n_tc; }
    const int floor_cnt = (query_length - 1) / cfg.denominator;
    int cur = root;
    while (true) {
        int max_cnt = floor_cnt, nb = 0, pick = -1, n_children = 0;
        for (int k = 0; k < n_cl; ++k) {
            if (scr.cl_parent[k] != cur) continue;
            ++n_children;
            const int cnt = scr.cl_cnt[k];
            if (cnt > max_cnt) { nb = 1; pick = k; max_cnt = cnt; }
            else if (cnt == max_cnt) { ++nb; pick = k; }
        }
        if (n_children == 0 || nb != 1) break;
        cur = pick;
    }
    r.classification = scr.cl_id[cur];
    return n_tc;
}

MB_HD void result_init(mb_read_result &r, int query_length, int n_matches) {
    r.classification = 0; r.score = 0.f; r.is_classified = 0; r.query_length = query_length;
    r.hamming_dist = 0; r.taxcnt_begin = 0; r.taxcnt_end = 0; r.n_matches = n_matches;
}

// Scores one read (general form: any number of matches, paths and species; scratch sized by the match count).
MB_HD int score_read(const mb_match *m, int b, int e, int query_length, const ScoreCfg &cfg, const TaxView &tax,
                     ReadScratch &scr, mb_read_result &r, u32 *error_flag) {
    result_init(r, query_length, e - b);
    if (e <= b) return 0;
    const BestSpecies best = best_species(m, b, e, query_length, cfg, tax, scr);
    return finish_read(m, best, query_length, cfg, tax, scr, r, error_flag);
}

constexpr int SF_P = 16;     // paths per species: the size up to which std::sort is a plain insertion sort

// ---- compact streaming form ---------------------------------------------------------------------------------
// The streaming state packed into 32-bit words so that it lives in SHARED memory on the device (one odd-strided
// slice per thread: bank-conflict free, no local-memory traffic).  Scores are kept in half units as integers: every
// partial score of the reference is a multiple of 0.5 far below 2^23, so its float sums are exact and equal to
// these integer sums; the only rounding operations (division by the read length, tie average) stay in fp32.
// Two phases per read so that the lanes of a warp stay together: (1) one loop over the matches runs the path DP and
// collects the paths of every species, (2) one loop over the species that have paths sorts and combines them.
constexpr int SC_G = 4;                          // matches per (species, frame, position) group
constexpr int SC_NP = 40;                        // paths per read
#ifndef MB_SC_NS
#define MB_SC_NS 32
#endif
constexpr int SC_NS = MB_SC_NS;                  // species with paths per read
constexpr int SC_GWORDS = 2 * SC_G * 3;          // DP state `w`: 2 groups x SC_G entries x 3 words (shared memory on the device)
constexpr int SC_GSTRIDE = SC_GWORDS + 1;        // odd: lanes touching the same word hit different banks
constexpr int SC_W_SPEC = SC_NP * 3;             // path list `wp`: SC_NP paths x 3 words, then SC_NS species x 2 words:
constexpr int SC_PWORDS = SC_W_SPEC + SC_NS * 2; // id | paths << 24, then begin | end << 16 (later the score).  Touched a few
                                                 // times per read only, so it lives in per-thread local memory (L1) and the
                                                 // kernel keeps 25 words of shared memory per thread instead of 177.

// half-unit score and Hamming sum of the lowest `cnt` 2-bit fields of r (codon score 3 / 1.5 / 1 / 0.5 for h = 0..3)
MB_HD void fields_score(u32 r, int cnt, int &score2, int &ham) {
    r &= (cnt >= 16) ? 0xFFFFFFFFu : ((1u << (2 * cnt)) - 1u);
    const u32 lo = r & 0x55555555u, hi = (r >> 1) & 0x55555555u;
#if defined(__CUDA_ARCH__)
    const int n1 = __popc(lo & ~hi), n2 = __popc(hi & ~lo), n3 = __popc(hi & lo);
#else
    const int n1 = __builtin_popcount(lo & ~hi), n2 = __builtin_popcount(hi & ~lo), n3 = __builtin_popcount(hi & lo);
#endif
    const int n0 = cnt - n1 - n2 - n3;
    score2 = 6 * n0 + 3 * n1 + 2 * n2 + n3;
    ham = n1 + 2 * n2 + 3 * n3;
}

// packed path (3 words): w0 = start | end << 16, w1 = (score2 & 0xFFFF) | ham << 16, w2 = sm | em << 16
MB_HD int pk_score2(u32 w1) { return (int) (short) (w1 & 0xFFFFu); }
MB_HD int pk_ham(u32 w1) { return (int) (w1 >> 16); }
// comparator of combineMatchPaths (Taxonomer.cpp:417-426) on packed paths: score desc, hamming asc, start desc
MB_HD bool pk_before(u32 a0, u32 a1, u32 b0, u32 b1) {
    const int sa = pk_score2(a1), sb = pk_score2(b1);
    if (sa != sb) return sa > sb;
    const int ha = pk_ham(a1), hb = pk_ham(b1);
    if (ha != hb) return ha < hb;
    return (a0 & 0xFFFFu) > (b0 & 0xFFFFu);
}


MB_HD u32 f2u_hd(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    u32 u; memcpy(&u, &f, 4); return u;
#endif
}
MB_HD float u2f_hd(u32 u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
// one group entry that is a finished path -> packed path list
MB_HD void emit_path(u32 *wp, int &n_paths, int &overflow, u32 e0, u32 e1, u32 end_pos, u32 em) {
    if (n_paths < SC_NP) {
        u32 *pp = wp + n_paths * 3;
        pp[0] = (e0 & 0xFFFFu) | (end_pos << 16);
        pp[1] = (e1 & 0xFFFu) | (((e1 >> 12) & 0xFFFu) << 16);
        pp[2] = (e0 >> 16) | (em << 16);
    } else overflow = 2;
    ++n_paths;
}

// Returns false when a bound of the compact state is exceeded (the caller then uses the general form).
// (reason codes: 1 position group, 2 paths per read, 3 species per read, 4 paths per species, 5 field width, 6 read size)
MB_HD int best_species_compact(const mb_match *m, int n, int query_length, const ScoreCfg &cfg, const TaxView &tax, u32 *w, u32 *wp,
                               BestSpecies &out) {
    out.tax = 0; out.score = 0.f; out.lca = false; out.rb = 0; out.re = 0;
    if (n > 60000 || query_length > 60000) return 6;
    int n_paths = 0, n_spp = 0, sp_paths0 = 0;
    int32_t species = 0;
    u32 frame = 0;
    int sp_begin = 0, min_depth = cfg.min_cons_cnt;
    // group `cur` = previous position group (na entries from match ia0 at pos_a), cur ^ 1 = the group being filled
    int cur = 0, na = 0, nb = 0, ia0 = 0, ib0 = 0;
    u32 pos_a = 0, pos_b = 0;
    bool have_cur = false, started = false, forward = true;
    int overflow = 0;

    // ---- phase 1: path DP over the matches (Taxonomer.cpp:487-648); the next record is loaded one step ahead ----
    const u64 *mw = reinterpret_cast<const u64 *>(m);
    u64 nx0 = mw[0], nx1 = mw[1], nx2 = mw[2];
    for (int i = 0; i <= n; ++i) {
        if (overflow) break;                     // nothing below may run with a group count past SC_G
        const bool at_end = i == n;
        const u64 w0 = nx0, w1 = nx1, w2 = nx2;
        if (i + 1 < n) { nx0 = mw[(i + 1) * 3]; nx1 = mw[(i + 1) * 3 + 1]; nx2 = mw[(i + 1) * 3 + 2]; }
        u64 qinfo = 0;
        int32_t sp = 0;
        u32 dna = 0, reh = 0, ham = 0;
        if (!at_end) {
            qinfo = w0; sp = (int32_t) (w1 >> 32); dna = (u32) w2 & 0xFFFFFFu; reh = (u32) (w2 >> 32) & 0xFFFFu; ham = (u32) (w2 >> 48) & 0xFFu;
        }
        const u32 fr = qi_frame(qinfo), pos = qi_pos(qinfo);
        const bool new_species = at_end || i == 0 || sp != species;
        const bool new_frame = new_species || fr != frame;
        const bool rotate = !new_frame && started && pos != pos_b;
        if (have_cur && (rotate || (new_frame && i > 0))) {
            // paths of the previous group that were not extended (Taxonomer.cpp:631-640)
            const u32 *ga = w + cur * SC_G * 3;
            for (int c = 0; c < na; ++c) {
                const u32 e1 = ga[c * 3 + 1], e2 = ga[c * 3 + 2];
                if (!(e2 >> 24) && (int) (e1 >> 24) >= min_depth) emit_path(wp, n_paths, overflow, ga[c * 3], e1, pos_a + 23u, (u32) (ia0 + c));
            }
            if (new_frame) {
                // the last group of the (species, frame) group: every path that is long enough (Taxonomer.cpp:641-646)
                const u32 *gb = w + (cur ^ 1) * SC_G * 3;
                for (int c = 0; c < nb; ++c) {
                    const u32 e1 = gb[c * 3 + 1];
                    if ((int) (e1 >> 24) >= min_depth) emit_path(wp, n_paths, overflow, gb[c * 3], e1, pos_b + 23u, (u32) (ib0 + c));
                }
            }
        }
        if (new_species && i > 0) {
            // species with paths are scored in phase 2
            if (n_paths > sp_paths0) {
                if (n_spp >= SC_NS || (u32) species >= (1u << 24)) overflow = 3;
                else if (n_paths - sp_paths0 > SF_P) overflow = 4;
                else {
                    u32 *sw = wp + SC_W_SPEC + n_spp * 2;
                    sw[0] = (u32) species | ((u32) (n_paths - sp_paths0) << 24);
                    sw[1] = (u32) sp_begin | ((u32) i << 16);
                }
                ++n_spp;
            }
            sp_paths0 = n_paths;
        }
        if (at_end || overflow) break;
        if (new_species) {
            species = sp; sp_begin = i;
            min_depth = tax_is_ancestor(tax, tax.eukaryota, species) ? cfg.min_cons_cnt_euk : cfg.min_cons_cnt;
        }
        if (new_frame) {
            frame = fr; forward = fr < 3;
            have_cur = false; started = false;
            na = 0; nb = 0; cur = 0;
        }
        if (!started) { started = true; pos_b = pos; ib0 = i; }
        if (pos != pos_b) {                      // == rotate
            cur ^= 1;
            have_cur = true;
            na = nb; pos_a = pos_b; ia0 = ib0;
            nb = 0; pos_b = pos; ib0 = i;
        }
        // path_init (Taxonomer.h:39-46): start = pos, score of the 8 codons, depth 1, sm = em = i
        int t_score2, t_ham8;
        fields_score(reh, 8, t_score2, t_ham8);
        u32 t_start = pos, t_sm = (u32) i;
        int t_ham = (int) ham, t_depth = 1;
        if (have_cur) {
            const int shift = (int) ((pos - pos_a) / 3);
            if (shift > 0 && shift <= cfg.max_codon_shift) {
                int inc2, hinc;
                fields_score(reh, shift, inc2, hinc);     // calScoreIncrement / calHammingDistIncrement (Taxonomer.cpp:650-669)
                const u32 keep = (1u << (24 - 3 * shift)) - 1u;
                u32 *ga = w + cur * SC_G * 3;
                int best = -1, best_score2 = 0;
                for (int c = 0; c < na; ++c) {
                    const u32 e2 = ga[c * 3 + 2];
                    const u32 adna = e2 & 0xFFFFFFu;
                    // isConsecutive(cur, next) forward, isConsecutive(next, cur) reverse (Taxonomer.cpp:671-699)
                    const u32 x = forward ? adna : dna, y = forward ? dna : adna;
                    const bool ok = cfg.format == 2 ? (x & keep) == (y >> (3 * shift)) : (x >> (3 * shift)) == (y & keep);
                    if (ok) {
                        ga[c * 3 + 2] = e2 | (1u << 24);
                        const int s2 = (int) (ga[c * 3 + 1] & 0xFFFu);
                        if (s2 > best_score2) { best = c; best_score2 = s2; }
                    }
                }
                if (best >= 0) {
                    const u32 e0 = ga[best * 3], e1 = ga[best * 3 + 1];
                    t_start = e0 & 0xFFFFu; t_sm = e0 >> 16;
                    t_score2 = (int) (e1 & 0xFFFu) + inc2;
                    t_ham = (int) ((e1 >> 12) & 0xFFFu) + hinc;
                    t_depth = (int) (e1 >> 24) + shift;
                }
            }
        }
        if (nb >= SC_G) overflow = 1;
        else if (t_score2 >= 4096 || t_ham >= 4096 || t_depth >= 256) overflow = 5;
        else {
            u32 *gb = w + (cur ^ 1) * SC_G * 3 + nb * 3;
            gb[0] = t_start | (t_sm << 16);
            gb[1] = (u32) t_score2 | ((u32) t_ham << 12) | ((u32) t_depth << 24);
            gb[2] = dna;
        }
        ++nb;
    }
    if (overflow) return overflow;

    // ---- phase 2: per species with paths, combineMatchPaths (Taxonomer.cpp:410-485) ----
    int meaningful = 0;
    float best_sp = 0.f;
    u32 *comb = w;                 // kept paths (start | end << 16): the DP groups are dead now
    int p0 = 0;
    for (int s = 0; s < n_spp; ++s) {
        u32 *sw = wp + SC_W_SPEC + s * 2;
        const int pc = (int) (sw[0] >> 24);
        u32 *pp = wp + p0 * 3;
        p0 += pc;
        // std::sort on <= 16 elements is libstdc++'s __insertion_sort (bits/stl_algo.h): a stable sort by the comparator.
        // Up to four paths (the usual case) are not moved at all: their comparator keys (score desc, hamming asc, start
        // desc, then the index) are ranked in registers and the list is read through the resulting order.
        u64 ord = 0xFEDCBA9876543210ull;         // ord nibble k = index of the k-th path in sorted order
        if (pc <= 4) {
            u64 key[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                key[i] = ~0ull;
                if (i < pc) {
                    const u32 v0 = pp[i * 3], v1 = pp[i * 3 + 1];
                    key[i] = ((((u64) (u32) (32767 - pk_score2(v1)) << 32) | ((u64) (u32) pk_ham(v1) << 16) | (u64) (0xFFFFu - (v0 & 0xFFFFu))) << 2) | (u64) i;
                }
            }
            const u32 c01 = key[0] < key[1], c02 = key[0] < key[2], c03 = key[0] < key[3];
            const u32 c12 = key[1] < key[2], c13 = key[1] < key[3], c23 = key[2] < key[3];
            const u32 r0 = 3u - c01 - c02 - c03, r1 = c01 + 2u - c12 - c13, r2 = c02 + c12 + 1u - c23, r3 = c03 + c13 + c23;
            ord = (1ull << (4 * r1)) | (2ull << (4 * r2)) | (3ull << (4 * r3));
            (void) r0;
        } else
        for (int k = 1; k < pc; ++k) {
            const u32 v0 = pp[k * 3], v1 = pp[k * 3 + 1], v2 = pp[k * 3 + 2];
            int j = k;
            if (pk_before(v0, v1, pp[0], pp[1])) {
                for (; j > 0; --j) { pp[j * 3] = pp[(j - 1) * 3]; pp[j * 3 + 1] = pp[(j - 1) * 3 + 1]; pp[j * 3 + 2] = pp[(j - 1) * 3 + 2]; }
            } else {
                while (pk_before(v0, v1, pp[(j - 1) * 3], pp[(j - 1) * 3 + 1])) {
                    pp[j * 3] = pp[(j - 1) * 3]; pp[j * 3 + 1] = pp[(j - 1) * 3 + 1]; pp[j * 3 + 2] = pp[(j - 1) * 3 + 2];
                    --j;
                }
            }
            pp[j * 3] = v0; pp[j * 3 + 1] = v1; pp[j * 3 + 2] = v2;
        }
        int n_comb = 0, sum2 = 0;
        for (int k = 0; k < pc; ++k) {
            const int o = (int) ((ord >> (4 * k)) & 15u);
            const u32 q0 = pp[o * 3], q1 = pp[o * 3 + 1], q2 = pp[o * 3 + 2];
            int c_start = (int) (q0 & 0xFFFFu), c_end = (int) (q0 >> 16);
            int c_score2 = pk_score2(q1), c_ham = pk_ham(q1);
            const int c_sm = (int) (q2 & 0xFFFFu), c_em = (int) (q2 >> 16);
            bool rejected = false;
            for (int j = 0; j < n_comb; ++j) {
                const int k_start = (int) (comb[j] & 0xFFFFu), k_end = (int) (comb[j] >> 16);
                if (c_end < k_start || k_end < c_start) continue;
                const int lo = c_start > k_start ? c_start : k_start;
                const int hi = c_end < k_end ? c_end : k_end;
                const int overlap = hi - lo + 1;
                if (overlap == c_end - c_start + 1 || overlap >= 24) { rejected = true; break; }
                int part2, part_ham;
                if (c_start < k_start) {         // trim the right end: Match::getRightPartScore / getRightPartHammingDist
                    c_end = k_start - 1;
                    fields_score((u32) m[c_em].rightEndHamming, overlap / 3, part2, part_ham);
                } else {                         // trim the left end: the top overlap/3 fields (Match.h:46-86)
                    c_start = k_end + 1;
                    const int range = overlap / 3;
                    fields_score(range ? ((u32) m[c_sm].rightEndHamming >> (16 - 2 * range)) : 0u, range, part2, part_ham);
                }
                const int h = c_ham - part_ham;
                c_ham = h > 0 ? h : 0;
                c_score2 = c_score2 - part2 - 2 * (overlap % 3);
            }
            if (!rejected) { comb[n_comb++] = (u32) c_start | ((u32) c_end << 16); sum2 += c_score2; }
        }
        float score = ((float) sum2 * 0.5f) / (float) query_length;
        score = score < 1.0f ? score : 1.0f;
        const u32 range = sw[1];
        if (score < cfg.min_score) { sw[1] = 0xFFFFFFFFu; continue; }     // not recorded (Taxonomer.cpp:361)
        sw[1] = f2u_hd(score);
        if (score > 0.f) ++meaningful;
        if (score > best_sp) { best_sp = score; out.rb = (int) (range & 0xFFFFu); out.re = (int) (range >> 16); }
    }
    if (meaningful == 0) { out.score = 0.f; return 0; }
    // species within tie_ratio of the best (Taxonomer.cpp:388-402)
    int n_top = 0;
    int32_t lca = 0;
    bool have = false;
    int32_t single = 0;
    const float cut = best_sp * cfg.tie_ratio;
    for (int s = 0; s < n_spp; ++s) {
        const u32 *sw = wp + SC_W_SPEC + s * 2;
        if (sw[1] == 0xFFFFFFFFu) continue;
        const float sc = u2f_hd(sw[1]);
        if (sc >= cut) {
            const int32_t t = (int32_t) (sw[0] & 0xFFFFFFu);
            if (n_top == 0) single = t;
            ++n_top;
            out.score += sc;
            if (tax_exists(tax, t)) {
                if (!have) { lca = t; have = true; }
                else lca = tax_lca_existing(tax, lca, t);
            }
        }
    }
    if (n_top > 1) {
        out.lca = true;
        out.tax = have ? lca : 0;
        out.score /= (float) n_top;
        return 0;
    }
    out.tax = single;
    return 0;
}

// Compact streaming scorer.  Returns -(reason) when a bound of the state was exceeded (the caller then uses score_read).
MB_HD int score_read_compact(const mb_match *m, int n, int query_length, const ScoreCfg &cfg, const TaxView &tax, u32 *w, u32 *wp,
                             ReadScratch &scr, mb_read_result &r, u32 *error_flag) {
    result_init(r, query_length, n);
    if (n <= 0) return 0;
    BestSpecies best;
    const int why = best_species_compact(m, n, query_length, cfg, tax, w, wp, best);
    if (why) return -why;
    return finish_read(m, best, query_length, cfg, tax, scr, r, error_flag);
}

#if defined(__CUDACC__)
// ---- segmentation of the sorted match list by read (Classifier.cpp:174-185) ----------------------------
__global__ void k_segment_matches(const mb_match *__restrict__ m, u64 n, u32 n_reads, u32 seq_base, u64 *__restrict__ seg_begin,
                                  u64 *__restrict__ seg_end) {
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 sg = qi_seq(m[i].qinfo);
    if (sg <= seq_base || sg - seq_base > n_reads) return;       // not a read of this batch
    const u32 s = sg - seq_base;
    const bool first = i == 0 || qi_seq(m[i - 1].qinfo) != sg;
    const bool last = i + 1 == n || qi_seq(m[i + 1].qinfo) != sg;
    if (first) seg_begin[s - 1] = i;
    if (last) seg_end[s - 1] = i + 1;
}
__global__ void k_segment_max(const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, u32 n_reads, u32 *__restrict__ max_len) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    u32 len = 0;
    if (i < n_reads) len = (u32) (seg_end[i] - seg_begin[i]);
    for (int d = 16; d > 0; d >>= 1) { const u32 o = __shfl_xor_sync(0xFFFFFFFFu, len, d); len = o > len ? o : len; }
    if ((threadIdx.x & 31) == 0 && len) atomicMax(max_len, len);
}

struct ScoreArena {
    // per-thread slices (filterRedundantMatches buckets, taxCnt, clade table): slot-indexed
    unsigned char *base;
    u64 slice_bytes;
    int cap_b, cap_c;
    // per-match scratch, indexed from the read's first match (a read never needs more entries than it has matches)
    PathRec *local, *paths, *combined;
    u8 *linked;
    int32_t *sp_id;
    float *sp_score;
};

__host__ __device__ inline u64 arena_slice_bytes(int cap_b, int cap_c) {
    u64 b = 0;
    b += (u64) cap_b * sizeof(int32_t) * 3;
    b += (u64) cap_c * sizeof(int32_t) * 3;
    b += (u64) cap_b;
    return (b + 15) & ~15ull;
}

__device__ __forceinline__ void arena_carve(const ScoreArena &a, u64 slot, u64 first_match, int n_matches, ReadScratch &s) {
    unsigned char *p = a.base + slot * a.slice_bytes;
    s.bucket_tax = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_b * sizeof(int32_t);
    s.tc_id = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_b * sizeof(int32_t);
    s.tc_cnt = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_b * sizeof(int32_t);
    s.cl_id = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_c * sizeof(int32_t);
    s.cl_cnt = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_c * sizeof(int32_t);
    s.cl_parent = reinterpret_cast<int32_t *>(p); p += (u64) a.cap_c * sizeof(int32_t);
    s.bucket_ham = p;
    s.local = a.local + first_match;
    s.paths = a.paths + first_match;
    s.combined = a.combined + first_match;
    s.linked = a.linked + first_match;
    s.sp_id = a.sp_id + first_match;
    s.sp_score = a.sp_score + first_match;
    s.cap_n = n_matches; s.cap_b = a.cap_b; s.cap_c = a.cap_c;
}

constexpr int SC_THREADS = 128;

// Shared tail of both scoring kernels: copy the read's taxCnt list out and store the result.
__device__ __forceinline__ void store_result(mb_read_result &res, int n_tc, const ReadScratch &scr, u32 r, mb_read_result *__restrict__ results,
                                             int32_t *__restrict__ tc_id_out, int32_t *__restrict__ tc_cnt_out, u64 tc_cap,
                                             unsigned long long *tc_counter, u32 *error_flag) {
    if (n_tc > 0) {
        const u64 o = atomicAdd(tc_counter, (unsigned long long) n_tc);
        if (o + (u64) n_tc <= tc_cap) {
            for (int k = 0; k < n_tc; ++k) { tc_id_out[o + k] = scr.tc_id[k]; tc_cnt_out[o + k] = scr.tc_cnt[k]; }
            res.taxcnt_begin = (int32_t) o;
            res.taxcnt_end = (int32_t) (o + n_tc);
        } else *error_flag = 5;
    }
    results[r] = res;
}

// General form, one THREAD per read.  With `list` == nullptr a warp takes 32 consecutive reads by an atomic ticket;
// otherwise the reads are list[0 .. *list_count) (the reads the streaming kernel left over).
__global__ void __launch_bounds__(SC_THREADS)
k_score(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end,
        const u32 *__restrict__ lens, u32 n_reads, int paired, ScoreCfg cfg, TaxView tax, ScoreArena arena,
        mb_read_result *__restrict__ results, int32_t *__restrict__ tc_id_out, int32_t *__restrict__ tc_cnt_out,
        u64 tc_cap, unsigned long long *tc_counter, u32 *ticket, u32 *error_flag, const u32 *__restrict__ list,
        const u32 *__restrict__ list_count, const u64 *__restrict__ list_off, u64 scratch_cap) {
    const u32 lane = threadIdx.x & 31;
    const u64 slot = (u64) blockIdx.x * SC_THREADS + threadIdx.x;
    const u32 total = list ? *list_count : n_reads;
    // Left-over reads are few and heavy, and the general form diverges completely between lanes: one read per WARP
    // (lane 0 works) finishes them sooner than 32 per warp.  The full-batch form takes 32 consecutive reads per warp.
    // ... unless there are so many that every warp gets 32 anyway (a database with dozens of species per genus leaves a
    // large share of the reads over): then lanes that idle cost more than divergence does
    const u32 per_ticket = (list && total < 32u * gridDim.x * (SC_THREADS / 32)) ? 1u : 32u;
    while (true) {
        u32 r0 = 0;
        if (lane == 0) r0 = atomicAdd(ticket, per_ticket);
        r0 = __shfl_sync(0xFFFFFFFFu, r0, 0);
        if (r0 >= total) break;
        if (lane < per_ticket && r0 + lane < total) {
            const u32 r = list ? list[r0 + lane] : r0 + lane;
            // queryLength = max covered length of each mate (KmerExtractor.cpp:452,476)
            int ql = max_covered_length((int) lens[2 * r]);
            if (paired) ql += max_covered_length((int) lens[2 * r + 1]);
            const u64 b = seg_begin[r], e = seg_end[r];
            mb_read_result res;
            int n_tc = 0;
            ReadScratch scr;
            // per-match scratch: a listed read got a compact offset when it was listed (k_score_fast), so that the scratch
            // is sized by the matches of the listed reads and not by the whole batch
            const u64 first = list ? list_off[r0 + lane] : b;
            if (e > b && (e - b) < 0x7FFFFFFFull && first + (e - b) <= scratch_cap) {
                arena_carve(arena, slot, first, (int) (e - b), scr);
                n_tc = score_read(m + b, 0, (int) (e - b), ql, cfg, tax, scr, res, error_flag);
            } else {
                if (e > b) *error_flag = first + (e - b) > scratch_cap ? 6 : 4;
                result_init(res, ql, (int32_t) (e - b));
            }
            store_result(res, n_tc, scr, r, results, tc_id_out, tc_cnt_out, tc_cap, tc_counter, error_flag);
        }
        __syncwarp();
    }
}

// What k_score_fast hands to k_score_finish: the outcome of getBestSpeciesMatches for one read.
struct BestRec {
    int32_t tax;
    float score;
    u32 range;             // rb | re << 16 : the best species' matches inside the read's segment
    u32 state;             // 0 = left to the general kernel, 1 = no match, 2 = best species, 3 = LCA of tied species
};

// Reads ordered by match count (classes of 4 matches, longest first): the 32 reads of a warp of the per-thread scorers then
// run match loops of about the same length — in batch order a warp waits for its longest read (11.9 of 32 lanes were active
// on average).  A counting sort over 256 classes; the order inside a class is arbitrary.
__device__ __forceinline__ u32 order_class(u64 b, u64 e) {
    const u64 n = (e - b) >> 2;
    return 255u - (u32) (n < 255 ? n : 255);
}
__global__ void __launch_bounds__(256) k_order_count(const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, u32 n_reads,
                                                     u32 *__restrict__ hist) {
    __shared__ u32 s[256];
    s[threadIdx.x] = 0;
    __syncthreads();
    for (u32 r = blockIdx.x * 256 + threadIdx.x; r < n_reads; r += gridDim.x * 256) atomicAdd(&s[order_class(seg_begin[r], seg_end[r])], 1u);
    __syncthreads();
    if (s[threadIdx.x]) atomicAdd(&hist[threadIdx.x], s[threadIdx.x]);
}
__global__ void __launch_bounds__(256) k_order_scan(u32 *hist) {      // one block: counts -> exclusive offsets, in place
    __shared__ u32 s[256];
    const u32 t = threadIdx.x;
    s[t] = hist[t];
    __syncthreads();
    u32 acc = 0;
    for (u32 i = 0; i < t; ++i) acc += s[i];
    hist[t] = acc;
}
__global__ void __launch_bounds__(256) k_order_scatter(const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, u32 n_reads,
                                                       u32 *__restrict__ cursor, u32 *__restrict__ order) {
    __shared__ u32 s_cnt[256], s_base[256];
    const u32 r = blockIdx.x * 256 + threadIdx.x;
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    u32 k = 0, local = 0;
    if (r < n_reads) { k = order_class(seg_begin[r], seg_end[r]); local = atomicAdd(&s_cnt[k], 1u); }
    __syncthreads();
    if (s_cnt[threadIdx.x]) s_base[threadIdx.x] = atomicAdd(&cursor[threadIdx.x], s_cnt[threadIdx.x]);
    __syncthreads();
    if (r < n_reads) order[s_base[k] + local] = r;
}

// Compact streaming form, one THREAD per read (best_species_compact): the 32 reads of a warp advance through their match
// lists in lock step.  Reads that exceed a bound of the streaming state are appended to `list` for k_score.
__global__ void __launch_bounds__(SC_THREADS)
k_score_fast(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end,
             const u32 *__restrict__ lens, u32 n_reads, int paired, ScoreCfg cfg, TaxView tax, BestRec *__restrict__ best_out,
             u32 *ticket, u32 *__restrict__ list, u32 *list_count, u32 *overflow_reasons, u64 *__restrict__ list_off,
             unsigned long long *slow_matches, const u32 *__restrict__ order) {
    extern __shared__ u32 sc_state[];            // SC_THREADS slices of SC_GSTRIDE words
    u32 wp[SC_PWORDS];                           // path / species lists of this thread's read (local memory)
    const u32 lane = threadIdx.x & 31;
    while (true) {
        u32 r0 = 0;
        if (lane == 0) r0 = atomicAdd(ticket, 32u);
        r0 = __shfl_sync(0xFFFFFFFFu, r0, 0);
        if (r0 >= n_reads) break;
        if (r0 + lane < n_reads) {
            const u32 r = order ? order[r0 + lane] : r0 + lane;
            int ql = max_covered_length((int) lens[2 * r]);
            if (paired) ql += max_covered_length((int) lens[2 * r + 1]);
            const u64 b = seg_begin[r], e = seg_end[r];
            BestRec rec;
            rec.tax = 0; rec.score = 0.f; rec.range = 0; rec.state = 1;
            if (e > b) {
                BestSpecies best;
                const int why = (e - b) < 0x7FFFFFFFull ? best_species_compact(m + b, (int) (e - b), ql, cfg, tax, sc_state + threadIdx.x * SC_GSTRIDE, wp, best) : 6;
                if (why) {
                    rec.state = 0;
                    const u32 li = atomicAdd(list_count, 1u);
                    list[li] = r;
                    list_off[li] = atomicAdd(slow_matches, (unsigned long long) (e - b));
                    atomicAdd(overflow_reasons + (why > 7 ? 7 : why), 1u);
                } else {
                    rec.tax = best.tax; rec.score = best.score; rec.range = (u32) best.rb | ((u32) best.re << 16);
                    rec.state = best.lca ? 3u : 2u;
                }
            }
            best_out[r] = rec;
        }
        __syncwarp();
    }
}

// Second half for the reads k_score_fast scored: chooseBestTaxon after getBestSpeciesMatches (finish_read).  The
// redundancy buckets live in shared memory (cap_b entries per thread), so this kernel runs at high occupancy.
__global__ void __launch_bounds__(SC_THREADS)
k_score_finish(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end,
               const u32 *__restrict__ lens, u32 n_reads, int paired, ScoreCfg cfg, TaxView tax, ScoreArena arena,
               const BestRec *__restrict__ best_in, int bucket_stride, mb_read_result *__restrict__ results,
               int32_t *__restrict__ tc_id_out, int32_t *__restrict__ tc_cnt_out, u64 tc_cap, unsigned long long *tc_counter,
               u32 *ticket, u32 *error_flag, const u32 *__restrict__ order) {
    extern __shared__ u32 fin_buckets[];         // SC_THREADS slices of bucket_stride words: cap_b taxa, then cap_b bytes
    const u32 lane = threadIdx.x & 31;
    const u64 slot = (u64) blockIdx.x * SC_THREADS + threadIdx.x;
    while (true) {
        u32 r0 = 0;
        if (lane == 0) r0 = atomicAdd(ticket, 32u);
        r0 = __shfl_sync(0xFFFFFFFFu, r0, 0);
        if (r0 >= n_reads) break;
        if (r0 + lane < n_reads) {
            const u32 r = order ? order[r0 + lane] : r0 + lane;
            const BestRec rec = best_in[r];
            if (rec.state != 0) {
                int ql = max_covered_length((int) lens[2 * r]);
                if (paired) ql += max_covered_length((int) lens[2 * r + 1]);
                const u64 b = seg_begin[r], e = seg_end[r];
                mb_read_result res;
                ReadScratch scr;
                arena_carve(arena, slot, 0, 0, scr);
                u32 *mine = fin_buckets + (size_t) threadIdx.x * bucket_stride;
                scr.bucket_tax = reinterpret_cast<int32_t *>(mine);
                scr.bucket_ham = reinterpret_cast<u8 *>(mine + arena.cap_b);
                result_init(res, ql, (int) (e - b));
                int n_tc = 0;
                if (rec.state != 1) {
                    BestSpecies best;
                    best.tax = rec.tax; best.score = rec.score; best.lca = rec.state == 3;
                    best.rb = (int) (rec.range & 0xFFFFu); best.re = (int) (rec.range >> 16);
                    n_tc = finish_read(m + b, best, ql, cfg, tax, scr, res, error_flag);
                }
                store_result(res, n_tc, scr, r, results, tc_id_out, tc_cnt_out, tc_cap, tc_counter, error_flag);
            }
        }
        __syncwarp();
    }
}

// ---- long reads and contigs (--seq-mode 3): the general per-read functions spread over four kernels ----------------
// A long read has thousands of matches and a contig hundreds of thousands, so one thread per read (k_score) serialises
// what is independent: getMatchPaths works on one (species, frame) run of the read's matches at a time
// (Taxonomer.cpp:316-408 calls it per frame), combineMatchPaths on one species.  The four kernels keep every function and
// every order of the general scorer (scores stay bit-identical) and only change who runs them:
//   k_lr_heads    one thread per match: lists the first match of every piece of a (read, species, frame) run (runs are cut
//                 where no link is possible) and of every (read, species) run
//   k_lr_paths    one thread per piece: match_paths -> the piece's paths, left beside its matches
//   k_lr_species  one thread per (read, species) run: gathers the paths of its frames in frame order and runs
//                 combine_paths -> the species' score
//   k_lr_best     one thread per read: walks its species in match order exactly as best_species does (first maximum wins,
//                 ties within tie_ratio go to the LCA), then finish_read
// filterRedundantMatches' bucket phase (Taxonomer.cpp:205-241) by a whole thread block: per position bucket the smallest
// Hamming distance (atomicMin), then the LCA of the targets at that distance (compare-and-swap fold).  LCA is associative
// and commutative, so the fold order does not matter and the buckets equal the sequential ones.  `ham32` is a cap_b-word
// scratch array.  Every thread of the block calls it with the same arguments.
__device__ __forceinline__ void fill_buckets_block(const mb_match *m, int rb, int re, int max_q, const ScoreCfg &cfg, const TaxView &tax,
                                                   ReadScratch &scr, u32 *ham32) {
    const int tid = (int) threadIdx.x, nt = (int) blockDim.x;
    for (int k = tid; k <= max_q; k += nt) { ham32[k] = 255u; scr.bucket_tax[k] = -1; }
    __syncthreads();
    const u64 *mw = reinterpret_cast<const u64 *>(m);
    for (int i = rb + tid; i < re; i += nt) {
        const u32 pos_i = qi_pos(mw[(u64) i * 3]);
        const int qd = (int) (cfg.dna_shift == 3 ? pos_i / 3u : cfg.dna_shift == 9 ? pos_i / 9u : pos_i / (u32) cfg.dna_shift);
        if (qd <= max_q) atomicMin(&ham32[qd], (u32) (mw[(u64) i * 3 + 2] >> 48) & 255u);
    }
    __syncthreads();
    for (int i = rb + tid; i < re; i += nt) {
        const u32 pos_i = qi_pos(mw[(u64) i * 3]);
        const int qd = (int) (cfg.dna_shift == 3 ? pos_i / 3u : cfg.dna_shift == 9 ? pos_i / 9u : pos_i / (u32) cfg.dna_shift);
        if (qd > max_q) continue;
        const u32 h = (u32) (mw[(u64) i * 3 + 2] >> 48) & 255u;
        if (h != ham32[qd]) continue;
        const int32_t target = (int32_t) (u32) mw[(u64) i * 3 + 1];
        int32_t old = scr.bucket_tax[qd];
        while (true) {
            const int32_t nw = old < 0 ? target : (old == target ? old : tax_lca(tax, old, target));
            if (nw == old) break;
            const int32_t seen = atomicCAS(&scr.bucket_tax[qd], old, nw);
            if (seen == old) break;
            old = seen;
        }
    }
    __syncthreads();
    for (int k = tid; k <= max_q; k += nt) {
        const u32 h = ham32[k];
        scr.bucket_ham[k] = (u8) h;
        if (h == 255u) scr.bucket_tax[k] = 0;
    }
    __syncthreads();
}

// The taxCnt list of filterRedundantMatches (one count per non-empty bucket, ascending taxID = std::map order) by a whole
// warp: 32 buckets at a time, equal taxa of the window grouped with match_any, one sorted insert per distinct taxon of the
// window (neighbouring buckets mostly share their taxon).  Every lane returns the same n_tc.  Warp-synchronous.
__device__ __forceinline__ int count_taxa_warp(int max_q, ReadScratch &scr) {
    const int lane = (int) (threadIdx.x & 31);
    int n_tc = 0;
    for (int k0 = 0; k0 <= max_q; k0 += 32) {
        const int k = k0 + lane;
        const bool valid = k <= max_q && scr.bucket_ham[k] != 255;
        const int32_t t = valid ? scr.bucket_tax[k] : 0;
        u32 todo = __ballot_sync(0xFFFFFFFFu, valid);
        while (todo) {
            const int leader = __ffs((int) todo) - 1;
            const int32_t lt = __shfl_sync(0xFFFFFFFFu, t, leader);
            const u32 same = __ballot_sync(0xFFFFFFFFu, valid && t == lt) & todo;
            const int cnt = __popc(same);
            todo &= ~same;
            // sorted insert of (lt, cnt): all lanes follow the same control flow, lane 0 writes
            int p = 0;
            while (p < n_tc && scr.tc_id[p] < lt) ++p;
            if (p < n_tc && scr.tc_id[p] == lt) { if (lane == 0) scr.tc_cnt[p] += cnt; }
            else {
                if (lane == 0) {
                    for (int z = n_tc; z > p; --z) { scr.tc_id[z] = scr.tc_id[z - 1]; scr.tc_cnt[z] = scr.tc_cnt[z - 1]; }
                    scr.tc_id[p] = lt; scr.tc_cnt[p] = cnt;
                }
                ++n_tc;
            }
            __syncwarp();
        }
    }
    return n_tc;
}

// combine_paths by a whole warp, for species with many paths (contigs): the paths are sorted by lane 0 exactly as
// combine_paths does; the greedy pass keeps its order (candidates one after another, each checked against the accepted
// paths in acceptance order) but the 32 lanes look at 32 accepted paths at a time and only the first one that overlaps the
// candidate is handled, after which the scan resumes behind it with the trimmed candidate — the same sequence of trims and
// rejections as the sequential loop, since accepted paths that do not overlap leave the candidate unchanged.
// Warp-synchronous; every lane calls it with the same arguments and gets the same result.
// The sort of combineMatchPaths for a species with many paths, by the whole warp.  std::sort's result is unique when no
// two paths tie under the comparator; then ANY correct sort reproduces it.  So: a copy of the paths is sorted with a
// bitonic network in global memory (all exchanges ascending, the virtual padding beyond n stays in place), ties are looked
// for among neighbours, and only if there is one the untouched original goes through the literal introsort (path_sort)
// on one lane as before.  (Profile at configs[4]: the one-lane sort was 55 % of k_lr_species.)
__device__ __forceinline__ void warp_sort_paths(PathRec *paths, PathRec *tmp, int n) {
    const int lane = (int) (threadIdx.x & 31);
    for (int i = lane; i < n; i += 32) tmp[i] = paths[i];
    __syncwarp();
    int n_pad = 1;
    while (n_pad < n) n_pad <<= 1;
    for (int k = 2; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < n_pad / 2; t += 32) {
                const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                const int p = j == (k >> 1) ? (i ^ (k - 1)) : (i | j);        // first step of a merge: flip partner
                if (p < n && p > i) {
                    const PathRec a = tmp[i], b = tmp[p];
                    if (path_before(b, a)) { tmp[i] = b; tmp[p] = a; }
                }
            }
            __syncwarp();
        }
    }
    bool tie = false;
    for (int i = lane; i + 1 < n; i += 32) tie = tie || !path_before(tmp[i], tmp[i + 1]);
    if (__any_sync(0xFFFFFFFFu, tie)) {
        if (lane == 0) path_sort(paths, paths + n);
    } else {
        for (int i = lane; i < n; i += 32) paths[i] = tmp[i];
    }
    __syncwarp();
}

__device__ __noinline__ float combine_paths_warp(const mb_match *m, PathRec *paths, PathRec *comb, int n_paths, int read_length) {
    const int lane = (int) (threadIdx.x & 31);
    warp_sort_paths(paths, comb, n_paths);
    int n_comb = 0;
    float score = 0.f;
    for (int i = 0; i < n_paths; ++i) {
        PathRec cand = paths[i];
        bool rejected = false;
        int j = 0;
        while (j < n_comb) {
            const int jj = j + lane;
            bool ov = false;
            if (jj < n_comb) {
                const int ks = comb[jj].start, ke = comb[jj].end;
                ov = !(cand.end < ks || ke < cand.start);
            }
            const u32 mask = __ballot_sync(0xFFFFFFFFu, ov);
            if (!mask) { j += 32; continue; }
            const int hit = j + (__ffs((int) mask) - 1);
            const PathRec kept = comb[hit];
            const int lo = cand.start > kept.start ? cand.start : kept.start;
            const int hi = cand.end < kept.end ? cand.end : kept.end;
            const int overlap = hi - lo + 1;
            if (overlap == cand.end - cand.start + 1 || overlap >= 24) { rejected = true; break; }
            if (cand.start < kept.start) {
                cand.end = kept.start - 1;
                int h = cand.ham - right_part_ham(m[cand.em], overlap / 3);
                cand.ham = h > 0 ? h : 0;
                cand.score = cand.score - right_part_score(m[cand.em], overlap / 3) - (float) (overlap % 3);
            } else {
                cand.start = kept.end + 1;
                int h = cand.ham - left_part_ham(m[cand.sm], overlap / 3);
                cand.ham = h > 0 ? h : 0;
                cand.score = cand.score - left_part_score(m[cand.sm], overlap / 3) - (float) (overlap % 3);
            }
            j = hit + 1;
        }
        if (!rejected) {
            if (lane == 0) comb[n_comb] = cand;
            ++n_comb;
            score += cand.score;
        }
        __syncwarp();
    }
    return score / (float) read_length;
}

struct LongScratch {
    PathRec *local, *paths, *combined;     // per match, as in ScoreArena
    u8 *linked;
    int32_t *grp_np, *grp_end;             // at a run's first match: number of paths, end of the run (index inside the read)
    float *sp_score;                       // at a species' first match: its score, < 0 = no entry (no path, or below min_score)
    int32_t *sp_end;                       // at a species' first match: end of its matches (index inside the read)
    u32 *grp_heads, *sp_heads;             // lists of first matches (global match indices)
    u32 *n_grp, *n_sp;                     // their lengths
};

__global__ void __launch_bounds__(256)
k_lr_heads(const mb_match *__restrict__ m, u64 n, LongScratch ls, int max_gap) {
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    const u32 lane = threadIdx.x & 31;
    bool gh = false, sh = false;
    if (i < n) {
        const u64 q = m[i].qinfo;
        const int32_t sp = m[i].speciesId;
        if (i == 0) gh = sh = true;
        else {
            const u64 qp = m[i - 1].qinfo;
            const int32_t spp = m[i - 1].speciesId;
            sh = qi_seq(q) != qi_seq(qp) || sp != spp;
            // a (read, species, frame) run is cut wherever two neighbouring positions are too far apart to be linked
            // (getMatchPaths only links position groups up to maxCodonShift codons apart, Taxonomer.cpp:530-560): the
            // pieces are independent and emit their paths in the same order as the whole run would
            gh = sh || qi_frame(q) != qi_frame(qp) || (int) (qi_pos(q) - qi_pos(qp)) > max_gap;
        }
    }
    // one reservation per warp and list
    const u32 gm = __ballot_sync(0xFFFFFFFFu, gh), sm = __ballot_sync(0xFFFFFFFFu, sh);
    u32 gbase = 0, sbase = 0;
    if (lane == 0) {
        if (gm) gbase = atomicAdd(ls.n_grp, (u32) __popc(gm));
        if (sm) sbase = atomicAdd(ls.n_sp, (u32) __popc(sm));
    }
    gbase = __shfl_sync(0xFFFFFFFFu, gbase, 0);
    sbase = __shfl_sync(0xFFFFFFFFu, sbase, 0);
    if (gh) ls.grp_heads[gbase + __popc(gm & ((1u << lane) - 1u))] = (u32) i;
    if (sh) ls.sp_heads[sbase + __popc(sm & ((1u << lane) - 1u))] = (u32) i;
}

__global__ void __launch_bounds__(SC_THREADS)
k_lr_paths(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, u32 seq_base,
           ScoreCfg cfg, TaxView tax, LongScratch ls, u32 *ticket) {
    const u32 lane = threadIdx.x & 31;
    const u32 total = *ls.n_grp;
    while (true) {
        u32 g0 = 0;
        if (lane == 0) g0 = atomicAdd(ticket, 32u);
        g0 = __shfl_sync(0xFFFFFFFFu, g0, 0);
        if (g0 >= total) break;
        if (g0 + lane < total) {
            const u64 h = ls.grp_heads[g0 + lane];
            const u32 r = qi_seq(m[h].qinfo) - seq_base - 1u;
            const u64 rb = seg_begin[r], re = seg_end[r];
            const mb_match *mr = m + rb;
            const int fb = (int) (h - rb), n = (int) (re - rb);
            const int32_t species = mr[fb].speciesId;
            const u32 frame = qi_frame(mr[fb].qinfo);
            const int max_gap = 3 * cfg.max_codon_shift;
            int fe = fb + 1;
            while (fe < n && mr[fe].speciesId == species && qi_frame(mr[fe].qinfo) == frame &&
                   (int) (qi_pos(mr[fe].qinfo) - qi_pos(mr[fe - 1].qinfo)) <= max_gap) ++fe;
            // getBestSpeciesMatches skips a (species, frame) run of a single match (Taxonomer.cpp:340-350): a piece of one
            // match is scored only when its run has more
            const bool run_has_more = fe - fb > 1 ||
                                      (fb > 0 && mr[fb - 1].speciesId == species && qi_frame(mr[fb - 1].qinfo) == frame) ||
                                      (fe < n && mr[fe].speciesId == species && qi_frame(mr[fe].qinfo) == frame);
            int n_paths = 0;
            if (run_has_more) {
                int min_depth = cfg.min_cons_cnt;
                if (tax_is_ancestor(tax, tax.eukaryota, species)) min_depth = cfg.min_cons_cnt_euk;
                ReadScratch scr;
                scr.local = ls.local + h; scr.linked = ls.linked + h; scr.paths = ls.paths + h;
                match_paths(mr, fb, fe, min_depth, cfg, scr, n_paths);
            }
            ls.grp_np[h] = n_paths;
            ls.grp_end[h] = fe;
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(SC_THREADS)
k_lr_species(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, const u32 *__restrict__ lens,
             u32 seq_base, ScoreCfg cfg, LongScratch ls, u32 *ticket) {
    const u32 lane = threadIdx.x & 31;
    const u32 total = *ls.n_sp;
    while (true) {
        u32 s0 = 0;
        if (lane == 0) s0 = atomicAdd(ticket, 32u);
        s0 = __shfl_sync(0xFFFFFFFFu, s0, 0);
        if (s0 >= total) break;
        u64 sp_start = 0, rb_mine = 0;
        int sp_stop = 0, np_mine = 0, ql_mine = 1;
        bool active = false;
        if (s0 + lane < total) {
            const u64 s = ls.sp_heads[s0 + lane];
            const u32 r = qi_seq(m[s].qinfo) - seq_base - 1u;
            const u64 rb = seg_begin[r], re = seg_end[r];
            const mb_match *mr = m + rb;
            const int n = (int) (re - rb);
            const int32_t species = mr[s - rb].speciesId;
            // the paths of the species' frames, in frame order, moved to the front of the species' slice
            int g = (int) (s - rb), n_paths = 0;
            PathRec *dst = ls.paths + s;
            while (g < n && mr[g].speciesId == species) {
                const int np = ls.grp_np[rb + g];
                const PathRec *src = ls.paths + rb + g;
                if (dst + n_paths != src)
                    for (int k = 0; k < np; ++k) dst[n_paths + k] = src[k];
                n_paths += np;
                g = ls.grp_end[rb + g];
            }
            sp_start = s; sp_stop = g; np_mine = n_paths; rb_mine = rb;
            ql_mine = max_covered_length((int) lens[2 * r]);
            active = true;
        }
        // species with few paths: every lane combines its own; species with many: the whole warp, one after another
        constexpr int LR_WARP_PATHS = 48;
        float score = -1.f;
        if (active && np_mine > 0 && np_mine <= LR_WARP_PATHS) {
            ReadScratch scr;
            scr.paths = ls.paths + sp_start; scr.combined = ls.combined + sp_start;
            score = combine_paths(m + rb_mine, scr, np_mine, ql_mine);
        }
        __syncwarp();                                    // the gathered paths of every lane are visible to the warp
        u32 big = __ballot_sync(0xFFFFFFFFu, active && np_mine > LR_WARP_PATHS);
        while (big) {
            const int src = __ffs((int) big) - 1;
            big &= big - 1;
            const u64 s_b = __shfl_sync(0xFFFFFFFFu, sp_start, src), rb_b = __shfl_sync(0xFFFFFFFFu, rb_mine, src);
            const int np_b = __shfl_sync(0xFFFFFFFFu, np_mine, src), ql_b = __shfl_sync(0xFFFFFFFFu, ql_mine, src);
            const float sc = combine_paths_warp(m + rb_b, ls.paths + s_b, ls.combined + s_b, np_b, ql_b);
            if ((int) lane == src) score = sc;
        }
        if (active) {
            if (np_mine > 0) {
                score = score < 1.0f ? score : 1.0f;
                if (score < cfg.min_score) score = -1.f;
            }
            ls.sp_score[sp_start] = score;
            ls.sp_end[sp_start] = sp_stop;
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(SC_THREADS)
k_lr_best(const mb_match *__restrict__ m, const u64 *__restrict__ seg_begin, const u64 *__restrict__ seg_end, const u32 *__restrict__ lens,
          u32 n_reads, ScoreCfg cfg, TaxView tax, ScoreArena arena, LongScratch ls, mb_read_result *__restrict__ results,
          int32_t *__restrict__ tc_id_out, int32_t *__restrict__ tc_cnt_out, u64 tc_cap, unsigned long long *tc_counter, u32 *ticket,
          u32 *error_flag) {
    // one thread BLOCK per read at a time (reads are few and heavy); thread 0 does what is sequential, the block fills the
    // redundancy buckets, warp 0 counts the taxa; one arena slice per block
    __shared__ u32 s_r;
    __shared__ BestSpecies s_best;
    __shared__ int s_need, s_ntc;
    const u32 tid = threadIdx.x;
    const u64 slot = blockIdx.x;
    while (true) {
        __syncthreads();
        if (tid == 0) s_r = atomicAdd(ticket, 1u);
        __syncthreads();
        const u32 r = s_r;
        if (r >= n_reads) break;
        const int ql = max_covered_length((int) lens[2 * r]);
        const u64 b = seg_begin[r], e = seg_end[r];
        const bool usable = e > b && (e - b) < 0x7FFFFFFFull;
        const int n = usable ? (int) (e - b) : 0;
        const int max_q = (ql + 3) / cfg.dna_shift;
        ReadScratch scr;
        arena_carve(arena, slot, b, n, scr);
        if (tid == 0) {
            BestSpecies best;
            best.tax = 0; best.score = 0.f; best.lca = false; best.rb = 0; best.re = 0;
            if (usable) {
                // Taxonomer::getBestSpeciesMatches (Taxonomer.cpp:316-408) over the species scores, in match order
                int meaningful = 0;
                float best_sp = 0.f;
                for (int i = 0; i < n;) {
                    const float sc = ls.sp_score[b + i];
                    const int ie = ls.sp_end[b + i];
                    if (sc >= 0.f) {
                        if (sc > 0.f) ++meaningful;
                        if (sc > best_sp) { best_sp = sc; best.rb = i; best.re = ie; }
                    }
                    i = ie;
                }
                if (meaningful) {
                    int n_top = 0;
                    int32_t lca = 0, single = 0;
                    bool have = false;
                    const float cut = best_sp * cfg.tie_ratio;
                    for (int i = 0; i < n;) {
                        const float sc = ls.sp_score[b + i];
                        const int ie = ls.sp_end[b + i];
                        if (sc >= 0.f && sc >= cut) {
                            const int32_t t = m[b + i].speciesId;
                            if (n_top == 0) single = t;
                            ++n_top;
                            best.score += sc;
                            if (tax_exists(tax, t)) {
                                if (!have) { lca = t; have = true; }
                                else lca = tax_lca_existing(tax, lca, t);
                            }
                        }
                        i = ie;
                    }
                    if (n_top > 1) { best.lca = true; best.tax = have ? lca : 0; best.score /= (float) n_top; }
                    else best.tax = single;
                }
            }
            s_best = best;
            s_need = usable && !(best.score == 0.f || best.score < cfg.min_score) && !best.lca && max_q + 1 <= scr.cap_b;
            s_ntc = -1;
        }
        __syncthreads();
        if (s_need) {
            fill_buckets_block(m + b, s_best.rb, s_best.re, max_q, cfg, tax, scr, reinterpret_cast<u32 *>(scr.cl_id));
            if (tid < 32) {
                const int n_tc_w = count_taxa_warp(max_q, scr);
                if (tid == 0) s_ntc = n_tc_w;
            }
            __syncthreads();
        }
        if (tid == 0) {
            mb_read_result res;
            int n_tc = 0;
            if (usable) {
                result_init(res, ql, n);
                const BestSpecies best = s_best;
                n_tc = finish_read(m + b, best, ql, cfg, tax, scr, res, error_flag, s_need != 0, s_ntc);
            } else {
                if (e > b) *error_flag = 4;
                result_init(res, ql, (int32_t) (e - b));
            }
            store_result(res, n_tc, scr, r, results, tc_id_out, tc_cnt_out, tc_cap, tc_counter, error_flag);
        }
    }
}
#endif // __CUDACC__

} // namespace mb
