// mb_host.h — C++ host side of the B200 classifier: everything `metabuli classify` does around the hot
// path (src/workflow/classify.cpp, src/commons/Classifier.cpp, Reporter.cpp, common.cpp, TaxonomyWrapper
// read side), written above the C ABI of include/metabuli_b200.h.
//
// The C entry points at the bottom (mbh_*) are what tests/ and bench.py bind through ctypes; the CLI
// (metabuli_main.cpp) uses the C++ classes directly.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/metabuli_b200.h"

namespace mbh {

// ---- taxonomyDB (TaxonomyWrapper::unserialize, src/commons/TaxonomyWrapper.cpp:363-421) --------------
struct TaxNodeRec {                // lib/mmseqs/src/taxonomy/NcbiTaxonomy.h:13-25 (32 bytes on disk)
    int32_t id, taxId, parentTaxId, pad;
    uint64_t rankIdx, nameIdx;
};

class Taxonomy {
public:
    bool load(const std::string &path, std::string &err);
    int32_t maxTaxId() const { return maxTaxID_; }
    size_t numNodes() const { return nodes_.size(); }
    bool exists(int32_t t) const { return t >= 0 && t <= maxTaxID_ && D_[t] != -1; }
    const TaxNodeRec *node(int32_t t) const { return (t == 0 || !exists(t)) ? nullptr : &nodes_[D_[t]]; }
    const TaxNodeRec &nodeAt(size_t i) const { return nodes_[i]; }
    int32_t original(int32_t t) const { return internalIds_ ? (t >= 0 && t <= maxTaxID_ ? internal2org_[t] : 0) : t; }
    const char *rank(int32_t t) const { const TaxNodeRec *n = node(t); return n ? str(n->rankIdx) : ""; }
    const char *name(int32_t t) const { const TaxNodeRec *n = node(t); return n ? str(n->nameIdx) : ""; }
    const char *str(uint64_t idx) const { return strData_.data() + strOffsets_[idx]; }
    int32_t eukaryota() const { return eukaryota_; }
    int32_t node0TaxId() const { return nodes_.empty() ? 0 : nodes_[0].taxId; }
    // TaxonomyWrapper::getTaxIdAtRank (TaxonomyWrapper.cpp:479-498)
    int32_t atRank(int32_t t, const char *rank) const;
    // TaxonomyWrapper::taxLineage2 (TaxonomyWrapper.cpp:431-454): "<short rank>_<name>" from below the root down to t
    std::string lineage(int32_t t) const;
    static int rankIndex(const char *r);       // NcbiTaxonomy::findRankIndex (NcbiTaxonomy.h:52-80)
    // dense arrays for mb_load_taxonomy
    void denseParents(std::vector<int32_t> &parent, std::vector<int32_t> &speciesRankOf) const;

private:
    std::vector<TaxNodeRec> nodes_;
    std::vector<int32_t> D_, internal2org_;
    std::vector<char> strData_;
    std::vector<uint32_t> strOffsets_;
    int32_t maxTaxID_ = 0, eukaryota_ = 0;
    bool internalIds_ = false;
};

// ---- database directory ------------------------------------------------------------------------------
struct DbParameters {              // loadDbParameters, src/commons/common.cpp:88-133
    bool present = false;
    int reducedAA = 0, accessionLevel = -1, skipRedundancy = 0, syncmer = 0, kmerFormat = -1, smerLen = -1;
    std::string dbName, dbDate;
};

struct Options {                   // the classify flag set (src/commons/LocalParameters.cpp:718-740)
    int seqMode = 2;
    int threads = 1;
    float minScore = 0.f, minSpScore = 0.f, tieRatio = 0.95f;
    int minConsCnt = 4, minConsCntEuk = 9;
    int hammingMargin = 0;
    int maskMode = 0;
    float maskProb = 0.9f;
    int ramUsage = 128;
    int matchPerKmer = 4;
    int accessionLevel = 0;
    int printLineage = 0;
    int validateInput = 0, validateDb = 0;
    int syncmer = 0, smerLen = 5;
    int kmerFormat = 1;
    int printLog = 0;
    int reducedAA = 0;
    int em = 0;
    int device = 0;                // --gpu (new): CUDA device ordinal
    std::vector<int> devices;      // --gpus N / --gpu-list a,b,...: (new) more than one entry = one DB shard per listed GPU, reads split over them
    size_t batchReads = 1u << 20;  // --batch-reads (new): reads per GPU batch
    std::string taxonomyPath;
    std::string kronaPrelude;      // --krona-prelude (new): the static head of JOBID_krona.html (lib/mmseqs/data/resources/krona_prelude.html)
};

bool preadParallel(int fd, void *dst, size_t bytes, size_t off, int threads);
int ioThreads(const Options &opt);     // --threads when given, else the host's cores (at most 16): file reads and output formatting

class Database {
public:
    ~Database();
    // Reads diffIdx/info/split/taxonomyDB/taxID_list/db.parameters, applies the DB-side parameter
    // overrides to `opt`, creates the GPU context and makes DB + taxonomy resident.
    // shard/nShards cut the DB at `split` boundaries (multi-GPU).
    // cutFrac (nShards - 1 values in (0, 1), increasing; may be null): where the shard boundaries lie along the usable split
    // entries — null means equal k-mer counts (g / nShards).  Used to re-balance shards by measured join time.
    bool open(const std::string &dir, Options &opt, int shard, int nShards, std::string &err, const double *cutFrac = nullptr);
    mb_ctx *ctx() const { return ctx_; }
    const Taxonomy &taxonomy() const { return tax_; }
    const mb_params &params() const { return par_; }
    const std::vector<uint64_t> &shardBounds() const { return bounds_; }
    size_t numKmers() const { return nKmers_; }
    size_t numWords() const { return nWords_; }

private:
    mb_ctx *ctx_ = nullptr;
    Taxonomy tax_;
    mb_params par_;
    std::vector<uint64_t> bounds_;     // amino-acid part (value with the 24 DNA bits cleared) of the first k-mer of shards 1..n-1
    size_t nKmers_ = 0, nWords_ = 0;
};

// ---- input ------------------------------------------------------------------------------------------
struct ReadBatch {                 // packed layout of mb_classify_batch
    std::vector<char> bases;
    std::vector<uint64_t> offsets;
    std::vector<uint32_t> lens;
    std::vector<std::string> names;
    uint32_t nReads() const { return (uint32_t) names.size(); }
    void clear() { bases.clear(); offsets.clear(); lens.clear(); names.clear(); }
};

class FastxReader {                // KSeqWrapper (lib/mmseqs/src/commons/KSeqWrapper.h) equivalent: FASTA/FASTQ, optional gzip
public:
    FastxReader();
    ~FastxReader();
    bool open(const std::string &path, std::string &err);
    bool next(std::string &name, std::string &seq);   // name up to the first whitespace, like kseq
private:
    bool fill();
    bool getline(std::string &out);
    void *gz_ = nullptr;
    std::vector<char> buf_;
    size_t pos_ = 0, end_ = 0;
    bool eof_ = false;
    std::string pending_;          // header line read ahead (FASTA)
    bool havePending_ = false;
};

// ---- output (Reporter, src/commons/Reporter.cpp:38-83,175-253) ----------------------------------------
struct NameView {                  // read names either as strings or as spans of a FASTQ text (mb_fetch_fastq_names)
    const std::string *names;
    const char *text;
    const uint64_t *off;
    const uint32_t *len;
};

class Reporter {
public:
    Reporter(const Taxonomy &tax, const std::string &outDir, const std::string &jobId, bool lineage, const std::string &kronaPrelude = "");
    bool openClassifications(std::string &err);
    void writeClassifications(const ReadBatch &batch, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt);
    void writeClassifications(const std::string *names, uint32_t n, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt);
    void writeClassifications(const NameView &names, uint32_t n, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt);
    void setThreads(int t) { threads_ = t; }      // lines are formatted by this many threads (the file is written by the caller's)
    void closeClassifications();
    bool writeReport(std::string &err);
    uint64_t totalReads() const { return total_; }
    const std::unordered_map<int32_t, unsigned> &taxCounts() const { return taxCounts_; }
private:
    const Taxonomy &tax_;
    std::string clsPath_, reportPath_, kronaPath_, kronaPrelude_;
    bool lineage_;
    FILE *cls_ = nullptr;
    std::unordered_map<int32_t, unsigned> taxCounts_;
    uint64_t total_ = 0;
    int threads_ = 1;
    std::vector<std::string> parts_;
    void formatSlice(const NameView &nv, uint32_t begin, uint32_t end, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt,
                     std::string &out) const;
};

// whole `classify` run: returns process exit code (opt.devices with more than one entry: runClassifyMulti)
// new shard boundaries (fractions along the split table) that equalise the measured per-shard times, assuming a shard's work
// is spread evenly over its range; frac: the nShards - 1 current boundaries (null = equal counts)
std::vector<double> rebalancedCuts(const std::vector<double> &timePerShard, const double *frac);

int runClassifyMulti(const std::vector<std::string> &inputs, const std::string &dbDir, const std::string &outDir,
                     const std::string &jobId, Options opt, std::string &err);
int runClassify(const std::vector<std::string> &inputs, const std::string &dbDir, const std::string &outDir,
                const std::string &jobId, Options opt, std::string &err);

} // namespace mbh
