// tests/emu/emu.cpp — TEST-ONLY host instantiation of the per-element device logic (MB_HD functions of
// metabuli_work_b200/csrc/*.cuh).  It lets the CPU test-suite run the exact code the kernels inline
// (window -> metamer, syncmer test, Hamming tables, LCA walk, score_read) against the oracle before a GPU
// is available.  The product never builds or loads this file.
#include "../../metabuli_work_b200/csrc/mb_common.cuh"
#include "../../metabuli_work_b200/csrc/mb_score.cuh"

#include <cstring>
#include <vector>

using namespace mb;

static Tables g_t;
static bool g_init = false;
static void init() { if (!g_init) { build_tables(g_t); g_init = true; } }

extern "C" {

// mirrors the per-position body of k_extract (mb_extract.cuh) with a plain loop over positions
size_t emu_extract(int format, int syncmer, int smer_len, int seq_mode, const unsigned char *bases, const uint64_t *offsets,
                   const uint32_t *lens, uint32_t n_reads, mb_kmer *out, size_t cap) {
    init();
    size_t n = 0;
    const bool paired = seq_mode == 2;
    for (uint32_t m = 0; m < 2 * n_reads; ++m) {
        const uint32_t mlen = lens[m];
        const bool self_ok = query_kmer_number((int) mlen) >= 1;
        bool usable;
        uint32_t pos_offset = 0;
        if (paired) {
            const uint32_t sib = lens[m ^ 1u];
            usable = self_ok && query_kmer_number((int) sib) >= 1;
            pos_offset = (m & 1u) ? (uint32_t) max_covered_length((int) sib) + 3u : 0u;
        } else usable = self_ok && (m & 1u) == 0;
        if (!usable) continue;
        const int used = max_covered_length((int) mlen);
        const unsigned char *s = bases + offsets[m];
        for (uint32_t x = 0; x + 24 <= mlen; ++x) {
            if ((int) (x + 24) > (int) (x % 3) + used) continue;
            u64 win = 0;
            bool amb = false;
            for (int i = 0; i < 24; ++i) {
                const u32 c = g_t.nuc[s[x + i]];
                if (c > 3) { amb = true; break; }
                win |= (u64) c << (2 * i);
            }
            if (amb) continue;
            const uint32_t seqId = (m >> 1) + 1;
            for (int rev = 0; rev < 2; ++rev) {
                const u64 v = metamer_from_window(g_t.codonF, g_t.codonR, win, rev != 0, format);
                if (!(format == 1 || !syncmer || is_syncmer(v, smer_len))) continue;
                const uint32_t fr = rev ? 3u + (uint32_t) (((int) (mlen % 3) - (int) (x % 3) + 3) % 3) : x % 3;
                if (n < cap) { out[n].value = v; out[n].qinfo = make_qinfo(seqId, x + pos_offset, fr); }
                ++n;
            }
        }
    }
    return n;
}

// Hamming sum / packed per-codon Hammings for pairs (q[i], t[i]); reverse[i] selects getHammings_reverse
void emu_hamming(const uint64_t *q, const uint64_t *t, const uint8_t *reverse, size_t n, uint8_t *sum, uint16_t *packed) {
    init();
    for (size_t i = 0; i < n; ++i) {
        sum[i] = (uint8_t) hamming_sum(g_t.ham4, q[i], t[i]);
        packed[i] = (uint16_t) hamming_pack(g_t.ham4, q[i], t[i], reverse[i] != 0);
    }
}

// the two-codon table forms the match kernel uses
void emu_hamming_lut(const uint64_t *q, const uint64_t *t, const uint8_t *reverse, size_t n, uint8_t *sum, uint16_t *packed) {
    init();
    for (size_t i = 0; i < n; ++i) {
        sum[i] = (uint8_t) hamming_sum_lut(g_t.lut2, q[i], t[i]);
        packed[i] = (uint16_t) hamming_pack_lut(g_t.lutp, q[i], t[i], reverse[i] != 0);
    }
}

int32_t emu_lca(const int32_t *parent, const int32_t *depth, int32_t max_taxid, int32_t node0, int32_t a, int32_t b) {
    TaxView tv;
    tv.parent = parent; tv.depth = depth; tv.species_of = nullptr; tv.species_rank_of = nullptr;
    tv.max_taxid = max_taxid; tv.node0_taxid = node0; tv.eukaryota = 0; tv.node = nullptr;
    const int32_t plain = tax_lca(tv, a, b);
    // the packed node table (what the device uses) must give the same answer
    std::vector<u64> node((size_t) max_taxid + 1);
    for (int32_t i = 0; i <= max_taxid; ++i) node[i] = (u64) (u32) parent[i] | ((u64) (u32) depth[i] << 32);
    tv.node = node.data();
    const int32_t packed = tax_lca(tv, a, b);
    return plain == packed ? plain : -12345;
}

static int g_fast = 0;
static long g_fallbacks = 0;
void emu_set_fast(int on) { g_fast = on; g_fallbacks = 0; }
long emu_fast_fallbacks() { return g_fallbacks; }

// score all reads of a sorted match list with score_read (mb_score.cuh), like k_score does per warp
int64_t emu_score(const mb_match *m, size_t n, const uint32_t *lens, uint32_t n_reads, int seq_mode, int format, int syncmer,
                  int smer_len, int min_cons_cnt, int min_cons_cnt_euk, float min_score, float min_sp_score, float tie_ratio,
                  const int32_t *parent, const int32_t *depth, const int32_t *species_of, const int32_t *species_rank_of,
                  int32_t max_taxid, int32_t node0, int32_t euk, mb_read_result *res, int32_t *tc_id, int32_t *tc_cnt, size_t tc_cap) {
    ScoreCfg cfg;
    cfg.format = format;
    if (syncmer) { cfg.dna_shift = (8 - smer_len) * 3; cfg.max_codon_shift = 8 - smer_len; } else { cfg.dna_shift = 3; cfg.max_codon_shift = 1; }
    cfg.denominator = (seq_mode == 1 || seq_mode == 2) ? 100 : 1000;
    cfg.min_cons_cnt = min_cons_cnt; cfg.min_cons_cnt_euk = min_cons_cnt_euk;
    cfg.min_score = min_score; cfg.min_sp_score = min_sp_score; cfg.tie_ratio = tie_ratio;
    TaxView tv;
    tv.parent = parent; tv.depth = depth; tv.species_of = species_of; tv.species_rank_of = species_rank_of;
    tv.max_taxid = max_taxid; tv.node0_taxid = node0; tv.eukaryota = euk;
    std::vector<u64> node((size_t) max_taxid + 1);
    for (int32_t i = 0; i <= max_taxid; ++i) node[i] = (u64) (u32) parent[i] | ((u64) (u32) depth[i] << 32);
    tv.node = node.data();
    std::vector<size_t> sb(n_reads, 0), se(n_reads, 0);
    size_t max_seg = 16;
    for (size_t i = 0; i < n;) {
        const u32 s = qi_seq(m[i].qinfo);
        size_t j = i;
        while (j < n && qi_seq(m[j].qinfo) == s) ++j;
        if (s >= 1 && s <= n_reads) { sb[s - 1] = i; se[s - 1] = j; if (j - i > max_seg) max_seg = j - i; }
        i = j;
    }
    uint32_t max_q = 0;
    for (uint32_t r = 0; r < n_reads; ++r) {
        int ql = max_covered_length((int) lens[2 * r]);
        if (seq_mode == 2) ql += max_covered_length((int) lens[2 * r + 1]);
        if (ql > 0 && (uint32_t) ql > max_q) max_q = (uint32_t) ql;
    }
    const int cap_n = (int) max_seg, cap_b = (int) ((max_q + 11) / (u32) cfg.dna_shift + 2), cap_c = 4 * cap_b + 64;
    std::vector<PathRec> local(cap_n), paths(cap_n), comb(cap_n);
    std::vector<u8> linked(cap_n), bham(cap_b);
    std::vector<int32_t> spid(cap_n), btax(cap_b), tcid(cap_b), tccnt(cap_b), clid(cap_c), clcnt(cap_c), clpar(cap_c);
    std::vector<float> spsc(cap_n);
    ReadScratch scr;
    scr.local = local.data(); scr.paths = paths.data(); scr.combined = comb.data(); scr.linked = linked.data();
    scr.sp_id = spid.data(); scr.sp_score = spsc.data(); scr.bucket_tax = btax.data(); scr.bucket_ham = bham.data();
    scr.tc_id = tcid.data(); scr.tc_cnt = tccnt.data(); scr.cl_id = clid.data(); scr.cl_cnt = clcnt.data(); scr.cl_parent = clpar.data();
    scr.cap_n = cap_n; scr.cap_b = cap_b; scr.cap_c = cap_c;
    size_t used = 0;
    u32 err = 0;
    for (uint32_t r = 0; r < n_reads; ++r) {
        int ql = max_covered_length((int) lens[2 * r]);
        if (seq_mode == 2) ql += max_covered_length((int) lens[2 * r + 1]);
        int n_tc = -1;
        if (g_fast) {
            // the streaming form k_score_fast runs; reads it cannot hold fall back to the general form, as on the device
            u32 state[SC_GSTRIDE], paths[SC_PWORDS];
            n_tc = score_read_compact(m + sb[r], (int) (se[r] - sb[r]), ql, cfg, tv, state, paths, scr, res[r], &err);
            if (n_tc < 0) ++g_fallbacks;
        }
        if (n_tc < 0) n_tc = score_read(m + sb[r], 0, (int) (se[r] - sb[r]), ql, cfg, tv, scr, res[r], &err);
        if (n_tc > 0) {
            if (used + (size_t) n_tc > tc_cap) return -1;
            for (int k = 0; k < n_tc; ++k) { tc_id[used + k] = scr.tc_id[k]; tc_cnt[used + k] = scr.tc_cnt[k]; }
            res[r].taxcnt_begin = (int32_t) used;
            res[r].taxcnt_end = (int32_t) (used + n_tc);
            used += (size_t) n_tc;
        }
    }
    return err ? -(int64_t) (100 + err) : (int64_t) used;
}

void emu_is_syncmer(const uint64_t *values, size_t n, int s, unsigned char *out) {
    for (size_t i = 0; i < n; ++i) out[i] = is_syncmer(values[i], s) ? 1 : 0;
}

// the order key of the query k-mer sort (SortT, mb_common.cuh)
void emu_sort_t(int mode, int shift, const uint64_t *values, size_t n, uint64_t *out) {
    SortT st{mode, shift};
    for (size_t i = 0; i < n; ++i) out[i] = sort_t(st, values[i]);
}

} // extern "C"
