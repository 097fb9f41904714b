/*
 * metabuli_b200.h — C ABI of libmetabuli_b200.so: the B200-native replacement of Metabuli's
 * classification hot path (query k-mer extraction -> k-mer sort -> diffIdx decode + AA merge join ->
 * match sort -> per-read scoring / LCA).
 *
 * The reference has no FFI for this path; the path sits behind four C++ member calls made by
 * Classifier::startClassify (src/commons/Classifier.cpp:105-119).  Each entry point below names
 * the reference interface it replaces (paths relative to the reference tree).  INTEGRATION.md shows
 * the binding a maintainer of the reference would add.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative
 * MB_E* code on failure, with a message available from mb_last_error(); nothing throws or exits
 * across the boundary.  The host owns host memory, the library owns device memory.  One context
 * drives one GPU; a context is not re-entrant.  There is NO CPU fallback: mb_create fails when no
 * CUDA device is usable.
 */
#ifndef METABULI_B200_H
#define METABULI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MB_OK          0
#define MB_E_CUDA     -1   /* CUDA runtime error (message has the call site)                */
#define MB_E_ARG      -2   /* bad argument / call order                                     */
#define MB_E_CAPACITY -3   /* caller buffer too small (required size reported by the call)  */
#define MB_E_DATA     -4   /* malformed database / taxonomy (e.g. DB tax ID without species) */

/* Query k-mer record: src/commons/Kmer.h:24-47 (16 bytes).
 * qinfo (src/commons/Kmer.h:11-16): pos bits 0-31, sequenceID bits 32-60 (1-based read index within
 * the batch), frame bits 61-63. */
typedef struct {
    uint64_t value;
    uint64_t qinfo;
} mb_kmer;

/* Match record: the payload of src/commons/Match.h:9-25 without the vptr (24 bytes). */
typedef struct {
    uint64_t qinfo;
    int32_t  targetId;        /* internal taxID of the DB k-mer (info & mask)            */
    int32_t  speciesId;       /* taxId2speciesId[targetId] (KmerMatcher.cpp:56-120)      */
    uint32_t dnaEncoding;     /* low 24 bits of the DB k-mer                             */
    uint16_t rightEndHamming; /* 8 x 2-bit per-codon Hamming (KmerMatcher.h:386-416)     */
    uint8_t  hamming;         /* Hamming sum (KmerMatcher.h:348-360)                     */
    uint8_t  pad;
} mb_match;

/* src/commons/Kmer.h:111-119 — one entry of the DB `split` file. */
typedef struct {
    uint64_t ADkmer;
    uint64_t diffIdxOffset;
    uint64_t infoIdxOffset;
} mb_split;

/* The parameters of LocalParameters the path reads (defaults: src/workflow/classify.cpp:10-37;
 * DB-side overrides: src/commons/common.cpp:88-133). */
typedef struct {
    int32_t device;           /* CUDA device ordinal                                           */
    int32_t kmer_format;      /* 1 legacy base-21 metamer (OldMetamerScanner), 2 current       */
    int32_t syncmer;          /* 0/1 (SyncmerScanner)                                          */
    int32_t smer_len;         /* --smer-len (5)                                                */
    int32_t seq_mode;         /* 1 single-end, 2 paired-end, 3 long read                       */
    int32_t min_cons_cnt;     /* --min-cons-cnt (4)                                            */
    int32_t min_cons_cnt_euk; /* --min-cons-cnt-euk (9)                                        */
    float   min_score;        /* --min-score (0)                                               */
    float   min_sp_score;     /* --min-sp-score (0)                                            */
    float   tie_ratio;        /* --tie-ratio (0.95)                                            */
    int32_t skip_redundancy;  /* db.parameters Skip_redundancy; 0 masks bit 31 of info         */
    int32_t accession_level;  /* only 0 is supported on this path                              */
    int32_t last_shard;       /* 1 when this context holds the END of the DB (the reference never
                                 offers the final DB k-mer as a candidate, KmerMatcher.cpp:363-406) */
    int32_t kmer_sort_bits;   /* leading amino-acid bits the query k-mer sort orders: 24, 32 or 40; 21 (format 2 only) =
                                 the re-coded key: five leading residues as a base-21 number plus two bits of the
                                 sixth, 24 bits in all (three passes); 31 (format 2 only) = seven residues as a base-21
                                 number plus one bit of the eighth, 32 bits (four passes).  0 = default (format 2: 21,
                                 or 31 when the resident database stream has more than four words per query k-mer of the
                                 first large batch; format 1: 32).
                                 The join only needs the queries of a DB block to be contiguous, not ordered, so
                                 fewer bits = fewer radix passes; blocks re-check the full amino-acid part.        */
    int32_t mask_mode;        /* --mask (0): 1 = tantan low-complexity masking of every mate before extraction
                                 (SeqIterator::maskLowComplexityRegions, SeqIterator.cpp:154-177)              */
    float   mask_prob;        /* --mask-prob (0.9): bases whose repeat probability reaches it become 'N'       */
} mb_params;

/* Per-read result: the fields of Query (src/commons/common.h:95-128) that reach the outputs. */
typedef struct {
    int32_t classification;   /* internal taxID, 0 = unclassified */
    float   score;
    int32_t is_classified;
    int32_t query_length;     /* queryLength + queryLength2       */
    int32_t hamming_dist;     /* reserved (always 0 without --em) */
    int32_t taxcnt_begin;     /* range in the taxcnt arrays (Query::taxCnt in std::map order) */
    int32_t taxcnt_end;
    int32_t n_matches;        /* matches of this read (diagnostic) */
} mb_read_result;

/* CUDA-event timings (ms) and counters of the last batch, per phase. */
typedef struct {
    float    h2d_ms, extract_ms, sort_kmers_ms, match_ms, sort_matches_ms, score_ms, d2h_ms, total_ms;
    uint64_t n_bases, n_kmers, n_matches, n_reads;
    uint64_t db_words, db_kmers;
    uint32_t kernel_launches;     /* kernels of this library launched for the last batch */
    uint32_t match_retries;       /* reruns of the match kernel after a capacity miss     */
} mb_phase_times;

typedef struct mb_ctx mb_ctx;

void        mb_default_params(mb_params *p);                 /* classify.cpp:10-37 */
mb_ctx     *mb_create(const mb_params *p);                   /* NULL on failure; see mb_global_error() */
void        mb_destroy(mb_ctx *ctx);
const char *mb_last_error(const mb_ctx *ctx);
const char *mb_global_error(void);                           /* failure of mb_create */
int         mb_set_params(mb_ctx *ctx, const mb_params *p);  /* device / last_shard cannot change */

/* ---- database (replaces the per-call fopen/fread/mmap of KmerMatcher::matchKmers, KmerMatcher.cpp:127-271) ----
 * diffIdx: uint16 delta stream (IndexCreator.cpp:868-886); info: one uint32 per k-mer; splits: the
 * `split` file.  The library keeps the compressed stream resident in HBM and builds a unit index
 * (the stream cut every 448 words where a new value of the query sort key starts; per unit: first word,
 * value before, k-mer ordinal, key) so that every unit decodes independently and owns an exact slice of
 * the sorted queries.  A SHARD (multi-GPU) is passed as a self-contained stream: the host
 * cuts the DB at `split` entries and re-encodes the shard's first k-mer as a delta from 0
 * (INTEGRATION.md), so the library never needs to know where a shard came from. */
int mb_load_db(mb_ctx *ctx, const uint16_t *diffIdx, size_t n_words, const uint32_t *info, size_t n_kmers);
/* decode the resident shard on the GPU (parity tap for getNextTargetKmer, KmerMatcher.h:282-297) */
int mb_decode_db(mb_ctx *ctx, uint64_t *values_out, size_t cap, size_t *n_out);

/* ---- taxonomy (replaces TaxonomyWrapper lookups: NcbiTaxonomy.cpp:261-333, TaxonomyWrapper.cpp:479-498,
 *      KmerMatcher::loadTaxIdList KmerMatcher.cpp:56-120) ----
 * Dense arrays indexed by internal taxID 0..max_taxid, derived by the host from taxonomyDB/taxID_list:
 *   parent[t]      internal parent taxID (root is its own parent); -1 when t does not exist
 *   species_of[t]  taxId2speciesId (0 when t is not in the DB's taxon list)
 *   species_rank_of[t]  getTaxIdAtRank(t, "species")
 * node0_taxid: taxID stored in taxonomy node index 0 (lcaHelper short-circuits on node index 0);
 * eukaryota_taxid: TaxonomyWrapper::getEukaryotaTaxID (0 if absent). */
int mb_load_taxonomy(mb_ctx *ctx, const int32_t *parent, const int32_t *species_of, const int32_t *species_rank_of,
                     int32_t max_taxid, int32_t node0_taxid, int32_t eukaryota_taxid);

/* ---- the hot path ----
 * Reads are passed packed: `bases` holds the raw sequence characters of all mates back to back;
 * mate m (0/1) of read i starts at offsets[2*i+m] and has lens[2*i+m] characters (0 = absent).
 * Replaces, for one batch (QuerySplit):
 *   KmerExtractor::extractQueryKmers   src/commons/KmerExtractor.h:77-83
 *   KmerMatcher::matchKmers            src/commons/KmerMatcher.h:228-230
 *   KmerMatcher::sortMatches           src/commons/KmerMatcher.h:244
 *   Classifier::assignTaxonomy         src/commons/Classifier.h:77-80
 * results[n_reads]; taxcnt_id/taxcnt_n receive the per-read (internal taxID, count) lists; when
 * taxcnt_cap is too small the call returns MB_E_CAPACITY and *taxcnt_used holds the required size. */
int mb_classify_batch(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                      uint32_t n_reads, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n,
                      size_t taxcnt_cap, size_t *taxcnt_used);

/* The same path in stages (each leaves its output resident on the device): */
int mb_stage_reads(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                   uint32_t n_reads);                                   /* H2D only                      */
/* DB-build inner loop.  Replaces, for one buffer of n target k-mers
 * {value, taxId, speciesId} (internal ids, any order), the three steps of IndexCreator::createIndex between extraction
 * and the files (src/commons/IndexCreator.cpp:262-280):
 *   SORT_PARALLEL(buffer, Kmer::compareTargetKmer)         src/commons/Kmer.h:77-87
 *   IndexCreator::filterKmers<FilterMode::DB_CREATION>     src/commons/IndexCreator.h:475-615 (equal (value, speciesId):
 *                                                          one k-mer, taxId = LCA of the duplicates' taxa)
 *   IndexCreator::writeTargetFilesAndSplits / getDiffIdx   src/commons/IndexCreator.cpp:811-886
 * Outputs are the contents of the diffIdx, info and split files: diffIdx_out[cap_words], info_out[cap_kmers],
 * splits_out[split_num] (may be NULL).  *n_words / *n_kmers always receive the sizes; MB_E_CAPACITY when a buffer is too
 * small (n_kmers <= n, n_words <= 5 * n_kmers).  Needs mb_load_taxonomy (for the LCA); touches no staged batch. */
int mb_build_index(mb_ctx *ctx, const uint64_t *values, const int32_t *taxids, const int32_t *species, size_t n, int split_num,
                   uint16_t *diffIdx_out, size_t cap_words, size_t *n_words, uint32_t *info_out, size_t cap_kmers, size_t *n_kmers,
                   mb_split *splits_out);
/* --mask 1 on the staged batch, in place.  Replaces SeqIterator::maskLowComplexityRegions
 * (src/commons/SeqIterator.cpp:154-177: tantan::maskSequences with maxRepeatOffset 50, repeatProb 0.005, repeatEndProb
 * 0.05, decay 0.9, no gaps) as KmerExtractor::processSequence applies it to every mate (KmerExtractor.cpp:328-335).
 * mb_extract_kmers / mb_classify_* call it themselves when mb_params.mask_mode is 1; the explicit call is the parity tap
 * (mb_fetch_staged_reads then returns the masked bases).  Idempotent per staged batch. */
int mb_mask_reads(mb_ctx *ctx);
/* Input pipeline on the device.  Replaces, for plain four-line FASTQ text, the two host
 * parses of the reference: QueryIndexer::indexQueryFile (src/commons/QueryIndexer.cpp:30-147) and
 * KmerExtractor::loadChunkOfReads over KSeqWrapper (src/commons/KmerExtractor.cpp:312-386).  text1 / text2 are raw
 * chunks of the two FASTQ files (text2 = NULL: single-end; the second mate of every read is then empty), each starting
 * at a record boundary.  The library uploads the text, finds the lines, checks the records ('@' / '+' markers, equal
 * sequence and quality length, optional CR), packs the sequence lines of up to max_reads whole records (the same number
 * from both files) exactly as mb_stage_reads expects them and leaves the batch staged.  final_chunk != 0: the text ends
 * the file, so a last line without a newline is complete.  out->consumed1/2: bytes of each text that belong to the
 * staged reads (the caller carries the rest over to its next chunk).  MB_E_DATA on a malformed record. */
typedef struct {
    uint32_t n_reads;         /* records staged                                            */
    uint32_t reserved;
    uint64_t consumed1, consumed2;
    uint64_t n_bases;         /* bases of the staged batch (both mates)                    */
} mb_fastq_index;
int mb_stage_fastq(mb_ctx *ctx, const char *text1, size_t n1, const char *text2, size_t n2, uint32_t max_reads,
                   int final_chunk, mb_fastq_index *out);
/* Overlap the host->device copy of the NEXT chunk's text with the device work of the current batch: the copy is enqueued
 * on the copy stream into a second pair of text buffers; a following mb_stage_fastq with the same pointers and sizes adopts
 * them instead of copying.  The texts must stay untouched until that call; pinned memory (mb_alloc_host) makes the copy
 * truly asynchronous. */
int mb_prefetch_fastq(mb_ctx *ctx, const char *text1, size_t n1, const char *text2, size_t n2);
/* Device memory of the context's GPU right now (cudaMemGetInfo), for sizing batches: the reference sizes its query
 * splits from --max-ram and bytesPerKmer = 16 + matchPerKmer * 32 (QueryIndexer::indexQueryFile,
 * src/commons/QueryIndexer.cpp:62-67); a host on top of this library does the same against free HBM. */
int mb_mem_info(mb_ctx *ctx, size_t *free_bytes, size_t *total_bytes);
/* Page-locked host memory for read / text buffers (cudaMallocHost); NULL on failure. */
void *mb_alloc_host(size_t bytes);
void mb_free_host(void *p);
/* Name span of every staged read inside text1 (offset of the first character after '@', length up to the first
 * white space: kseq's `name`), n_reads entries each; valid after mb_stage_fastq. */
int mb_fetch_fastq_names(mb_ctx *ctx, uint64_t *name_off, uint32_t *name_len);
/* The staged batch as mb_stage_reads would have received it (parity tap; any pointer may be NULL). */
int mb_fetch_staged_reads(mb_ctx *ctx, char *bases, uint64_t *offsets, uint32_t *lens);

/* Overlap the host->device copy of the NEXT batch with the device work of the current one: the copy is
 * enqueued on a second stream into a second set of buffers; a following mb_stage_reads (or
 * mb_classify_batch) with the SAME arguments adopts the copied batch instead of copying again.  The
 * host buffers must stay unchanged until then (pinned memory for a truly asynchronous copy).
 * Typical loop: stage(cur) -> prefetch(next) -> classify_staged -> fetch_results.                    */
int mb_prefetch_reads(mb_ctx *ctx, const char *bases, size_t n_bases, const uint64_t *offsets, const uint32_t *lens,
                      uint32_t n_reads);
int mb_extract_kmers(mb_ctx *ctx, size_t *n_kmers);                     /* KmerExtractor.cpp:83-386      */
int mb_sort_kmers(mb_ctx *ctx);                                         /* KmerExtractor.cpp:79          */
int mb_match_kmers(mb_ctx *ctx, size_t *n_matches);                     /* KmerMatcher.cpp:123-481       */
int mb_sort_matches(mb_ctx *ctx);                                       /* KmerMatcher.cpp:1071-1078     */
int mb_score_reads(mb_ctx *ctx, size_t *taxcnt_used);                   /* Classifier.cpp:166-208        */
int mb_fetch_results(mb_ctx *ctx, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n, size_t taxcnt_cap);
/* The same copy without waiting for it: enqueued on the copy stream behind the scoring, so the caller can stage and run the
 * next batch while the results travel (page-locked host arrays for a truly asynchronous copy).  The host arrays are valid
 * after mb_fetch_wait; one transfer is in flight at a time, and the next mb_score_reads waits for it before it writes the
 * result arrays again.  The results stay fetchable after the NEXT batch has been staged and extracted (only its scoring
 * overwrites them): starting the transfer after the next batch's mb_extract_kmers keeps it off the host link while the
 * staging and extraction read their counts.  (The reference writes its results on the thread that computed them, Classifier.cpp:133-150; this
 * is the overlap its OpenMP batch loop cannot have.) */
int mb_fetch_results_async(mb_ctx *ctx, mb_read_result *results, int32_t *taxcnt_id, int32_t *taxcnt_n, size_t taxcnt_cap);
int mb_fetch_wait(mb_ctx *ctx);
/* all device stages back to back on staged reads (no host copies) */
int mb_classify_staged(mb_ctx *ctx, size_t *taxcnt_used);

/* parity taps / multi-GPU exchange: copy the resident intermediate out, or replace it. */
int mb_copy_kmers(mb_ctx *ctx, mb_kmer *out, size_t cap, size_t *n);
int mb_copy_matches(mb_ctx *ctx, mb_match *out, size_t cap, size_t *n);
int mb_set_kmers(mb_ctx *ctx, const mb_kmer *host_kmers, size_t n);          /* host -> device   */
int mb_set_matches(mb_ctx *ctx, const mb_match *host_matches, size_t n);
/* device-pointer forms used by the multi-GPU driver (pointers are CUDA device addresses on ctx's GPU) */
int mb_kmers_device(mb_ctx *ctx, void **dptr, size_t *n);
int mb_matches_device(mb_ctx *ctx, void **dptr, size_t *n);
int mb_set_kmers_device(mb_ctx *ctx, const void *dptr, size_t n);            /* device -> device */
int mb_set_matches_device(mb_ctx *ctx, const void *dptr, size_t n);
/* device -> caller-owned device buffer (e.g. a torch tensor that feeds an NCCL all-to-all) */
int mb_export_kmers_device(mb_ctx *ctx, void *dst_dptr, size_t cap, size_t *n);
int mb_export_matches_device(mb_ctx *ctx, void *dst_dptr, size_t cap, size_t *n);
/* multi-GPU: sequenceIDs written by extraction become base + 1-based local read index, so k-mers and
 * matches of different home ranks can share one owner GPU; scoring maps them back to local reads. */
int mb_set_read_id_base(mb_ctx *ctx, uint32_t base);
/* multi-GPU bucketing: stable partition of the resident k-mers by the owner shard of their amino-acid
 * part (owner = number of bounds[] <= value; bounds are the first values of shards 1..n_shards-1) and
 * of the resident matches by home rank ((sequenceID-1) / reads_per_rank).  counts[n] receives sizes. */
int mb_partition_kmers(mb_ctx *ctx, const uint64_t *bounds, int n_shards, uint64_t *counts);
int mb_partition_matches(mb_ctx *ctx, uint32_t reads_per_rank, int n_ranks, uint64_t *counts);

/* multi-GPU, fused form: the producing kernels route their records straight into the destination GPU's
 * receive buffer through peer memory (NVLink P2P stores + one remote atomic reservation per run), so the
 * bucketing passes and the all-to-all collectives above disappear.  Extraction sends every k-mer to the
 * shard that owns its value; the join sends every match to its read's home rank.
 *   mb_peer_alloc    allocate this context's fixed-capacity receive side and describe it
 *   mb_ipc_export/open/close  CUDA IPC plumbing for one-process-per-GPU (handles travel by any host channel)
 *   mb_peer_attach   give the context the receive sides of ALL ranks (device addresses valid in this process;
 *                    entry `rank` is its own) plus the routing rules
 *   mb_extract_kmers_routed / mb_match_kmers_routed   the producers; *n_emitted = records this rank produced
 *   mb_take_routed_kmers / mb_take_routed_matches     after a cross-rank barrier: adopt what arrived as the
 *                    resident k-mers / matches (MB_E_CAPACITY when more arrived than the buffer holds; the
 *                    count is still returned so that the caller can re-allocate on every rank and redo)
 * Between producer and take the CALLER must synchronise all ranks (every producer call returns only after
 * its kernel has finished, so a host barrier is enough). */
#define MB_MAX_RANKS 16
typedef struct {
    void    *kmers;           /* mb_kmer[kmer_cap]   receive buffer                        */
    void    *matches;         /* mb_match[match_cap] receive buffer                        */
    void    *counters;        /* uint64[2]: k-mers arrived, matches arrived                */
    uint64_t kmer_cap, match_cap;
} mb_peer_mem;
int mb_peer_alloc(mb_ctx *ctx, size_t kmer_cap, size_t match_cap, mb_peer_mem *out);
int mb_ipc_export(mb_ctx *ctx, const void *dptr, unsigned char handle[64]);
int mb_ipc_open(mb_ctx *ctx, const unsigned char handle[64], void **dptr);
int mb_ipc_close(mb_ctx *ctx, void *dptr);
/* One process driving several GPUs (the C++ host's --gpus N): lets the context's GPU store into the memory of
 * `peer_device` (cudaDeviceEnablePeerAccess; a no-op for the context's own device or when access is already on).  With it
 * the mb_peer_mem pointers of the other contexts can be passed to mb_peer_attach as they are — no CUDA IPC. */
int mb_enable_peer_access(mb_ctx *ctx, int peer_device);
int mb_peer_attach(mb_ctx *ctx, uint32_t rank, uint32_t n_ranks, const mb_peer_mem *peers, const uint64_t *bounds,
                   uint32_t reads_per_rank);
int mb_extract_kmers_routed(mb_ctx *ctx, size_t *n_emitted);
int mb_take_routed_kmers(mb_ctx *ctx, size_t *n);
int mb_match_kmers_routed(mb_ctx *ctx, size_t *n_emitted);
int mb_take_routed_matches(mb_ctx *ctx, size_t *n);

/* ---- allocation-free query k-mer sort ----
 * No counterpart in the reference: this entry point was added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same radix kernels and the same key as mb_sort_kmers (stable, ordered by the amino-acid bits mb_params.kmer_sort_bits
 * selects), but on caller-owned device memory: d_kmers holds n mb_kmer records and receives the sorted records, d_tmp is a
 * second buffer of n records (both on the context's GPU, 16-byte aligned, distinct).  The sort's scratch (digit histograms,
 * per-tile look-back states, tile ticket: about 2 MB) is a file-scope `static __device__` array that every GPU loading the
 * library holds until exit, so the call performs no cudaMalloc / cudaFree and calls no allocating routine.
 * Limit: n <= MB_NOALLOC_MAX_KMERS, larger n returns MB_E_CAPACITY (use mb_sort_kmers).  stream: a cudaStream_t on the
 * context's GPU (the call is then stream-ordered and does not synchronise), or NULL = the context's stream (the call returns
 * after the sort has finished).  The scratch is shared by every call on a GPU: calls that overlap on the GPU race. */
#define MB_NOALLOC_MAX_KMERS (1u << 22)
int mb_sort_kmers_noalloc(mb_ctx *ctx, void *d_kmers, void *d_tmp, size_t n, void *stream);

int mb_get_timings(mb_ctx *ctx, mb_phase_times *t);
/* Per-kernel timing of the last batch: with profiling on, every kernel launch of the hot path is
 * bracketed by a CUDA event pair on the launching stream; the table sums them per kernel. */
typedef struct {
    char     name[32];
    float    ms;              /* summed over the launches of the last batch */
    uint32_t launches;
} mb_kernel_time;
int mb_set_profile(mb_ctx *ctx, int on);
int mb_get_kernel_times(mb_ctx *ctx, mb_kernel_time *out, size_t cap, size_t *n);   /* *n = table size */
/* CUDA stream the library launches on (cudaStream_t as void*), for event timing by the caller */
void *mb_stream(mb_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* METABULI_B200_H */
