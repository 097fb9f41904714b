/*
 * mb_oracle.h — CPU restatement of Metabuli's classify hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (metabuli_work_b200/,
 * include/) may include, link or call this.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it, as the checker.
 *
 * Parity pinning: the reference ships no golden vectors for this path
 * (SURVEY.md §4).  This oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF:
 *   - tests/golden/: k-mer, match and per-read dumps written by oracle/_ref/ref_harness
 *     (the reference's own KmerScanner/SyncmerScanner/KmerMatcher/Taxonomer classes
 *     compiled from /root/reference by oracle/Makefile) and classification TSVs
 *     written by the reference `metabuli classify` binary (tests/golden/make_golden.sh).
 *   - tests/test_oracle_golden.py compares every stage of this file with them.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference).
 */
#ifndef MB_ORACLE_H
#define MB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Query k-mer record: src/commons/Kmer.h:24-47 (16 bytes).
 * qinfo bit layout (src/commons/Kmer.h:11-16): pos 0-31, sequenceID 32-60, frame 61-63. */
typedef struct {
    uint64_t value;
    uint64_t qinfo;
} orc_kmer;

/* Match payload: src/commons/Match.h:9-25 without the vptr (24 bytes). */
typedef struct {
    uint64_t qinfo;
    int32_t  targetId;
    int32_t  speciesId;
    uint32_t dnaEncoding;
    uint16_t rightEndHamming;
    uint8_t  hamming;
    uint8_t  pad;
} orc_match;

/* src/commons/Kmer.h:111-119 */
typedef struct {
    uint64_t ADkmer;
    uint64_t diffIdxOffset;
    uint64_t infoIdxOffset;
} orc_split;

/* Classification parameters (defaults: src/workflow/classify.cpp:10-37). */
typedef struct {
    int   kmer_format;      /* 1 legacy base-21, 2 current 5-bit */
    int   syncmer;          /* 0/1 */
    int   smer_len;         /* --smer-len, default 5 */
    int   seq_mode;         /* 1 single, 2 paired, 3 long */
    int   min_cons_cnt;     /* 4 */
    int   min_cons_cnt_euk; /* 9 */
    float min_score;        /* 0 */
    float min_sp_score;     /* 0 */
    float tie_ratio;        /* 0.95 */
    int   skip_redundancy;  /* from db.parameters */
    int   accession_level;  /* 0 */
    int   threads;          /* OpenMP threads for the baseline port */
} orc_params;

/* Per-read result (the fields of src/commons/common.h:95-128 that reach the TSV). */
typedef struct {
    int32_t classification; /* internal taxID, 0 = unclassified */
    float   score;
    int32_t is_classified;
    int32_t query_length;   /* queryLength + queryLength2 */
    int32_t taxcnt_begin;   /* range into the taxcnt arrays */
    int32_t taxcnt_end;
} orc_read_result;

typedef struct orc_db orc_db;

void orc_default_params(orc_params *p);

/* ---- database / taxonomy (F1-F5) ---- */
orc_db *orc_db_open(const char *dbdir);            /* reads diffIdx, info, split, taxonomyDB, taxID_list, db.parameters */
void    orc_db_close(orc_db *db);
const char *orc_last_error(void);
/* db.parameters overrides (src/commons/common.cpp:88-133) applied onto p */
void    orc_db_apply_parameters(const orc_db *db, orc_params *p);
size_t  orc_db_num_kmers(const orc_db *db);
size_t  orc_db_num_words(const orc_db *db);
/* decode the whole diffIdx stream: values[n], infos[n] (KmerMatcher.h:282-297) */
size_t  orc_db_decode(const orc_db *db, uint64_t *values, uint32_t *infos, size_t cap);
/* taxonomy queries used by tests */
int32_t orc_tax_max_taxid(const orc_db *db);
int32_t orc_tax_lca(const orc_db *db, int32_t a, int32_t b);
int32_t orc_tax_parent(const orc_db *db, int32_t t);
int32_t orc_tax_species_of(const orc_db *db, int32_t t);      /* taxId2speciesId, 0 if absent */
int32_t orc_tax_original(const orc_db *db, int32_t t);
const char *orc_tax_rank(const orc_db *db, int32_t t);
const char *orc_tax_name(const orc_db *db, int32_t t);
int32_t orc_tax_is_ancestor(const orc_db *db, int32_t anc, int32_t child);
int32_t orc_tax_eukaryota(const orc_db *db);
int32_t orc_tax_at_rank(const orc_db *db, int32_t t, const char *rank);

/* ---- stage A2-A5: extraction ---- */
/* Upper bound of k-mer slots for a read: LocalUtil.h:45-59 */
int64_t orc_query_kmer_number(int64_t len);
int64_t orc_max_covered_length(int64_t len);
/* Extract all query k-mers of a batch. seqs: concatenated bases; for read i,
 * mate m: offset off[2*i+m], length len[2*i+m] (len==0 -> mate absent).
 * Returns number of k-mers written (unsorted, reference order: read, mate, frame, pos). */
size_t orc_extract(const orc_params *p, const char *seqs, const uint64_t *off,
                   const uint32_t *len, uint32_t n_reads, orc_kmer *out, size_t cap);
/* ---- stage A6: sort by (value, qinfo) — a refinement of compareQueryKmer ---- */
void   orc_sort_kmers(orc_kmer *k, size_t n, int threads);
/* ---- stage A8/A9: match ---- */
size_t orc_match_kmers(const orc_db *db, const orc_params *p, const orc_kmer *sorted,
                       size_t n, orc_match *out, size_t cap);
/* ---- stage A10 ---- */
void   orc_sort_matches(orc_match *m, size_t n, int threads);
/* ---- stage A11-A17: per-read scoring ---- */
/* qlen1/qlen2: RAW mate lengths (0 = absent). taxcnt arrays receive per-read
 * (internal taxID, count) pairs in std::map order. Returns entries used, or -1 on overflow. */
int64_t orc_score(const orc_db *db, const orc_params *p, const orc_match *sorted, size_t n,
                  const uint32_t *raw_len, uint32_t n_reads, orc_read_result *res,
                  int32_t *taxcnt_id, int32_t *taxcnt_n, size_t taxcnt_cap);

/* ---- TSV writers (Reporter.cpp:38-83, 175-253) ---- */
int orc_write_classifications_lineage(const orc_db *db, const char *path, const char *const *names, const orc_read_result *res,
                                      uint32_t n_reads, const int32_t *taxcnt_id, const int32_t *taxcnt_n);
int orc_write_classifications(const orc_db *db, const char *path, const char *const *names,
                              const orc_read_result *res, uint32_t n_reads,
                              const int32_t *taxcnt_id, const int32_t *taxcnt_n, int append);
int orc_write_report(const orc_db *db, const char *path, const orc_read_result *res, uint32_t n_reads);

/* ---- synthetic DB writer (format spec: IndexCreator.cpp:811-886, 1245-1266) ---- */
/* values/taxids must already be sorted by (value, species, taxid) and unique per (value, species). */
int orc_write_db_files(const char *dbdir, const uint64_t *values, const uint32_t *taxids, size_t n,
                       int split_num, int kmer_format);

/* ---- DB-build inner loop (orc_build.inc): sort, (value, species) duplicates folded with LCA, diffIdx / info / split written.
 * db supplies the taxonomy (any directory with a taxonomyDB opened by orc_db_open).  Returns the number of DB k-mers. */
int64_t orc_build_index(const orc_db *db, const uint64_t *values, const int32_t *taxids, const int32_t *species, size_t n,
                        int split_num, int kmer_format, const char *dbdir);

/* ---- --mask 1: tantan low-complexity masking of every mate in place (orc_mask.inc) ----
 * seqs: packed bases; mate m at off[m], len[m] characters; probs_out (may be NULL): repeat probability per base. */
void orc_mask_sequences(char *seqs, const uint64_t *off, const uint32_t *len, uint32_t n_mates, float mask_prob, float *probs_out);
double orc_mask_lr(int match);    /* the two likelihood-ratio values (match / mismatch) */

#ifdef __cplusplus
}
#endif
#endif
