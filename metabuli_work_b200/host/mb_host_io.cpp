// mb_host_io.cpp — FASTA/FASTQ input, classification / report output and the batch loop of `classify`.
#include "mb_host.h"

#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <sstream>

namespace mbh {

// ------------------------------------------------------------------------------------------------------
// FastxReader: same record model as kseq (lib/mmseqs/src/commons/KSeqWrapper.h:30): name = header up to
// the first whitespace, sequence = all sequence lines joined; gz files are read through zlib (which
// also reads plain files).
// ------------------------------------------------------------------------------------------------------
FastxReader::FastxReader() : buf_(1 << 20) {}
FastxReader::~FastxReader() { if (gz_) gzclose((gzFile) gz_); }

bool FastxReader::open(const std::string &path, std::string &err) {
    gz_ = gzopen(path.c_str(), "rb");
    if (!gz_) { err = "cannot open " + path; return false; }
    gzbuffer((gzFile) gz_, 1 << 20);
    return true;
}

bool FastxReader::fill() {
    if (eof_) return false;
    const int n = gzread((gzFile) gz_, buf_.data(), (unsigned) buf_.size());
    if (n <= 0) { eof_ = true; return false; }
    pos_ = 0;
    end_ = (size_t) n;
    return true;
}

bool FastxReader::getline(std::string &out) {
    out.clear();
    bool any = false;
    while (true) {
        if (pos_ == end_ && !fill()) return any;
        any = true;
        const char *b = buf_.data() + pos_;
        const char *nl = (const char *) memchr(b, '\n', end_ - pos_);
        if (nl) {
            out.append(b, (size_t) (nl - b));
            pos_ += (size_t) (nl - b) + 1;
            if (!out.empty() && out.back() == '\r') out.pop_back();
            return true;
        }
        out.append(b, end_ - pos_);
        pos_ = end_;
    }
}

bool FastxReader::next(std::string &name, std::string &seq) {
    std::string line;
    if (havePending_) { line = pending_; havePending_ = false; }
    else {
        do { if (!getline(line)) return false; } while (line.empty());
    }
    if (line[0] != '>' && line[0] != '@') return false;
    const bool fastq = line[0] == '@';
    size_t e = 1;
    while (e < line.size() && !isspace((unsigned char) line[e])) ++e;
    name.assign(line, 1, e - 1);
    seq.clear();
    if (fastq) {
        // sequence lines until '+', then as many quality characters as bases
        while (getline(line)) {
            if (!line.empty() && line[0] == '+') break;
            seq += line;
        }
        size_t q = 0;
        while (q < seq.size() && getline(line)) q += line.size();
    } else {
        while (getline(line)) {
            if (!line.empty() && (line[0] == '>' || line[0] == '@')) { pending_ = line; havePending_ = true; break; }
            seq += line;
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------------
// Reporter
// ------------------------------------------------------------------------------------------------------
Reporter::Reporter(const Taxonomy &tax, const std::string &outDir, const std::string &jobId, bool lineage, const std::string &kronaPrelude)
    : tax_(tax), clsPath_(outDir + "/" + jobId + "_classifications.tsv"), reportPath_(outDir + "/" + jobId + "_report.tsv"),
      kronaPath_(outDir + "/" + jobId + "_krona.html"), kronaPrelude_(kronaPrelude), lineage_(lineage) {
    if (kronaPrelude_.empty() && getenv("METABULI_KRONA_PRELUDE")) kronaPrelude_ = getenv("METABULI_KRONA_PRELUDE");
}

bool Reporter::openClassifications(std::string &err) {
    cls_ = fopen(clsPath_.c_str(), "w");
    if (!cls_) { err = "cannot write " + clsPath_; return false; }
    static char iobuf[1 << 22];
    setvbuf(cls_, iobuf, _IOFBF, sizeof(iobuf));
    // Reporter::writeReadClassification header (Reporter.cpp:39-46)
    fputs("#is_classified\tname\ttaxID\tquery_length\tscore\trank", cls_);
    if (lineage_) fputs("\tlineage", cls_);
    fputs("\ttaxID:match_count\n", cls_);
    return true;
}

// std::ostream << float with default flags prints like printf("%g") (precision 6)
static inline int formatScore(char *buf, size_t n, float score) { return snprintf(buf, n, "%g", (double) score); }

void Reporter::writeClassifications(const ReadBatch &batch, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt) {
    writeClassifications(batch.names.data(), batch.nReads(), res, tcId, tcCnt);
}

void Reporter::writeClassifications(const std::string *names, uint32_t nReads, const mb_read_result *res, const int32_t *tcId,
                                    const int32_t *tcCnt) {
    char num[64];
    for (uint32_t i = 0; i < nReads; ++i) {
        const mb_read_result &r = res[i];
        ++taxCounts_[r.classification];          // Classifier.cpp:201-203
        ++total_;
        formatScore(num, sizeof(num), r.score);
        if (r.is_classified) {
            fprintf(cls_, "1\t%s\t%d\t%d\t%s\t%s\t", names[i].c_str(), tax_.original(r.classification), r.query_length, num,
                    tax_.rank(r.classification));
            if (lineage_) { fputs(tax_.lineage(r.classification).c_str(), cls_); fputc('\t', cls_); }   // Reporter.cpp:60-62
            for (int32_t k = r.taxcnt_begin; k < r.taxcnt_end; ++k) fprintf(cls_, "%d:%d ", tax_.original(tcId[k]), tcCnt[k]);
            fputc('\n', cls_);
        } else {
            fprintf(cls_, "0\t%s\t%d\t%d\t%s\t-\t", names[i].c_str(), tax_.original(r.classification), r.query_length, num);
            if (lineage_) fputs("-\t", cls_);
            fputs("-\t\n", cls_);
        }
    }
}

void Reporter::closeClassifications() { if (cls_) { fclose(cls_); cls_ = nullptr; } }

namespace {
struct Clade { unsigned taxCount = 0, cladeCount = 0; std::vector<int32_t> children; };

void reportRec(FILE *fp, const Taxonomy &tax, const std::unordered_map<int32_t, Clade> &cl, unsigned long total, int32_t taxId, int depth) {
    auto it = cl.find(taxId);
    const unsigned cladeCount = it == cl.end() ? 0 : it->second.cladeCount;
    const unsigned taxCount = it == cl.end() ? 0 : it->second.taxCount;
    if (taxId == 0) {
        if (cladeCount > 0) fprintf(fp, "%.4f\t%i\t%i\tno rank\t0\tunclassified\n", 100 * cladeCount / double(total), cladeCount, taxCount);
        reportRec(fp, tax, cl, total, 1, 0);
        return;
    }
    if (cladeCount == 0) return;
    fprintf(fp, "%.4f\t%i\t%i\t%s\t%i\t%s%s\n", 100 * cladeCount / double(total), cladeCount, taxCount, tax.rank(taxId),
            tax.original(taxId), std::string(2 * (size_t) depth, ' ').c_str(), tax.name(taxId));
    std::vector<int32_t> children = it->second.children;
    auto cnt = [&](int32_t t) { auto f = cl.find(t); return f == cl.end() ? 0u : f->second.cladeCount; };
    std::sort(children.begin(), children.end(), [&](int a, int b) { return cnt(a) > cnt(b); });   // Reporter.cpp:243
    for (int32_t c : children) {
        if (!cl.count(c)) break;
        reportRec(fp, tax, cl, total, c, depth + 1);
    }
}
// escapeAttribute (lib/mmseqs/src/taxonomy/taxonomyreport.cpp:76-100)
std::string xmlEscape(const char *s) {
    std::string out;
    for (; *s; ++s) {
        switch (*s) {
            case '&': out += "&amp;"; break;
            case '"': out += "&quot;"; break;
            case '\'': out += "&apos;"; break;
            case '<': out += "&lt;"; break;
            case '>': out += "&gt;"; break;
            default: out += *s;
        }
    }
    return out;
}

// Reporter::kronaReport (Reporter.cpp:146-173): nested <node> elements, children by clade count like the report
void kronaRec(FILE *fp, const Taxonomy &tax, const std::unordered_map<int32_t, Clade> &cl, int32_t taxId) {
    auto it = cl.find(taxId);
    const unsigned cladeCount = it == cl.end() ? 0 : it->second.cladeCount;
    if (taxId == 0) {
        if (cladeCount > 0) fprintf(fp, "<node name=\"unclassified\"><magnitude><val>%d</val></magnitude></node>", cladeCount);
        kronaRec(fp, tax, cl, 1);
        return;
    }
    if (cladeCount == 0) return;
    fprintf(fp, "<node name=\"%s\"><magnitude><val>%d</val></magnitude>", xmlEscape(tax.name(taxId)).c_str(), cladeCount);
    std::vector<int32_t> children = it->second.children;
    auto cnt = [&](int32_t t) { auto f = cl.find(t); return f == cl.end() ? 0u : f->second.cladeCount; };
    std::sort(children.begin(), children.end(), [&](int a, int b) { return cnt(a) > cnt(b); });
    for (int32_t c : children) {
        if (!cl.count(c)) break;
        kronaRec(fp, tax, cl, c);
    }
    fputs("</node>", fp);
}
} // namespace

// Reporter::writeReportFile (Reporter.cpp:175-220) over NcbiTaxonomy::getCladeCounts (NcbiTaxonomy.cpp:520-545)
bool Reporter::writeReport(std::string &err) {
    std::unordered_map<int32_t, std::vector<int32_t>> parentToChildren;
    for (size_t i = 0; i < tax_.numNodes(); ++i) {
        const TaxNodeRec &n = tax_.nodeAt(i);
        if (n.parentTaxId == n.taxId) continue;
        parentToChildren[n.parentTaxId].push_back(n.taxId);
    }
    std::unordered_map<int32_t, Clade> cl;
    for (auto &kv : taxCounts_) {
        cl[kv.first].taxCount = kv.second;
        cl[kv.first].cladeCount += kv.second;
        if (tax_.exists(kv.first)) {
            const TaxNodeRec *n = tax_.node(kv.first);
            if (!n && kv.first == 0) continue;
            while (n && n->parentTaxId != n->taxId && tax_.exists(n->parentTaxId)) {
                n = tax_.node(n->parentTaxId);
                if (!n) break;
                cl[n->taxId].cladeCount += kv.second;
            }
        }
    }
    for (auto &kv : cl) {
        auto f = parentToChildren.find(kv.first);
        if (f != parentToChildren.end()) kv.second.children = f->second;
    }
    FILE *fp = fopen(reportPath_.c_str(), "w");
    if (!fp) { err = "cannot write " + reportPath_; return false; }
    fprintf(fp, "#clade_proportion\tclade_count\ttaxon_count\trank\ttaxID\tname\n");
    reportRec(fp, tax_, cl, (unsigned long) total_, 0, 0);
    fclose(fp);
    // JOBID_krona.html (Reporter.cpp:196-219): static prelude + the clade tree.  The prelude is a 224-KB third-party
    // resource that the reference embeds at build time; here it is read from --krona-prelude / METABULI_KRONA_PRELUDE.
    // Without it the chart data is still written (the node tree), and the file says what is missing.
    FILE *kf = fopen(kronaPath_.c_str(), "w");
    if (!kf) { err = "cannot write " + kronaPath_; return false; }
    bool havePrelude = false;
    if (!kronaPrelude_.empty()) {
        if (FILE *pf = fopen(kronaPrelude_.c_str(), "rb")) {
            char buf[1 << 16];
            size_t n;
            while ((n = fread(buf, 1, sizeof(buf), pf)) > 0) fwrite(buf, 1, n, kf);
            fclose(pf);
            havePrelude = true;
        }
    }
    if (!havePrelude)
        fputs("<!-- krona prelude not supplied: pass --krona-prelude <reference>/lib/mmseqs/data/resources/krona_prelude.html -->\n", kf);
    fprintf(kf, "<node name=\"all\"><magnitude><val>%zu</val></magnitude>", (size_t) total_);
    kronaRec(kf, tax_, cl, 0);
    fputs("</node></krona></div></body></html>", kf);
    fclose(kf);
    return true;
}

// ------------------------------------------------------------------------------------------------------
// the batch loop (Classifier::startClassify, Classifier.cpp:44-164)
// ------------------------------------------------------------------------------------------------------
static bool loadBatch(FastxReader &r1, FastxReader *r2, size_t maxReads, size_t maxBases, ReadBatch &b) {
    b.clear();
    std::string name, seq, name2, seq2;
    while (b.names.size() < maxReads && b.bases.size() < maxBases && r1.next(name, seq)) {
        b.offsets.push_back(b.bases.size());
        b.lens.push_back((uint32_t) seq.size());
        b.bases.insert(b.bases.end(), seq.begin(), seq.end());
        if (r2) {
            if (!r2->next(name2, seq2)) seq2.clear();
        } else seq2.clear();
        b.offsets.push_back(b.bases.size());
        b.lens.push_back((uint32_t) seq2.size());
        b.bases.insert(b.bases.end(), seq2.begin(), seq2.end());
        b.names.push_back(name);
    }
    return !b.names.empty();
}

// Raw text of one plain FASTQ file, read in chunks; the tail that mb_stage_fastq did not consume (an incomplete record,
// or records beyond the batch limit) is carried over to the next chunk.
struct RawText {
    FILE *fp = nullptr;
    std::vector<char> buf;
    size_t fill = 0;
    bool eof = false;
    ~RawText() { if (fp) fclose(fp); }
    bool open(const std::string &path) { fp = fopen(path.c_str(), "rb"); return fp != nullptr; }
    void refill(size_t chunk) {
        if (buf.size() < chunk) buf.resize(chunk);
        while (!eof && fill < chunk) {
            const size_t n = fread(buf.data() + fill, 1, chunk - fill, fp);
            if (n == 0) eof = true;
            fill += n;
        }
    }
    void consume(size_t n) { memmove(buf.data(), buf.data() + n, fill - n); fill -= n; }
};

// plain (not gzip) four-line FASTQ?  Decided on the first bytes: '@' starts the file and the third line starts with '+'.
static bool looksLikePlainFastq(const std::string &path) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) return false;
    char head[1 << 16];
    const size_t n = fread(head, 1, sizeof(head), fp);
    fclose(fp);
    if (n < 4 || head[0] != '@') return false;
    size_t line = 0;
    for (size_t i = 0; i + 1 < n; ++i)
        if (head[i] == '\n' && ++line == 2) return head[i + 1] == '+';
    return false;
}

int runClassify(const std::vector<std::string> &inputs, const std::string &dbDir, const std::string &outDir,
                const std::string &jobId, Options opt, std::string &err) {
    if (opt.devices.size() > 1) return runClassifyMulti(inputs, dbDir, outDir, jobId, opt, err);
    if (opt.devices.size() == 1) opt.device = opt.devices[0];
    using clk = std::chrono::steady_clock;
    const auto t0 = clk::now();
    Database db;
    if (!db.open(dbDir, opt, 0, 1, err)) return 1;
    const auto t1 = clk::now();
    std::cout << "DB resident on GPU " << opt.device << ": " << db.numKmers() << " k-mers, " << db.numWords() * 2 + db.numKmers() * 4
              << " bytes (" << std::chrono::duration<double>(t1 - t0).count() << " s)" << std::endl;
    // Batch size from the memory that is left (QueryIndexer::indexQueryFile sizes its splits the same way from --max-ram,
    // QueryIndexer.cpp:62-67: bytes per k-mer slot = k-mer record(s) + matchPerKmer match records).  Per slot here: two
    // 16-byte k-mer buffers, and per expected match two 24-byte match buffers plus the scorer's per-match scratch
    // (85 bytes with long reads, an eighth of 93 bytes with short reads).  --max-ram (GB) caps the budget like it caps the
    // reference's host memory.
    size_t freeB = 0, totalB = 0;
    if (mb_mem_info(db.ctx(), &freeB, &totalB) != MB_OK) { err = mb_last_error(db.ctx()); return 1; }
    const size_t budget = std::min<size_t>((size_t) (0.9 * (double) freeB), (size_t) std::max(1, opt.ramUsage) << 30);
    const size_t slotBytes = 32 + (size_t) std::max(1, opt.matchPerKmer) * (48 + (opt.seqMode == 3 ? 85 : 12));
    const size_t maxSlots = std::max<size_t>(budget / slotBytes, 1u << 20);
    const size_t maxBatchBases = maxSlots / 2;          // about two k-mer slots per base (six frames, one per codon)
    std::cout << "Batch budget: " << (budget >> 20) << " MiB of HBM (" << (freeB >> 20) << " MiB free), at most " << maxBatchBases
              << " bases per batch" << std::endl;
    Reporter rep(db.taxonomy(), outDir, jobId, opt.printLineage != 0, opt.kronaPrelude);
    if (!rep.openClassifications(err)) return 1;
    const bool paired = opt.seqMode == 2;
    ReadBatch batch;
    std::vector<mb_read_result> res;
    std::vector<int32_t> tcId, tcCnt;
    uint64_t totalKmers = 0, totalMatches = 0;
    double gpuMs = 0;
    // Plain FASTQ goes to the GPU as raw text (mb_stage_fastq: lines, records and packing are found on the device; the host
    // only cuts the read names out at the spans it gets back).  FASTA, multi-line or gzip input takes the host reader.
    const bool deviceParse = !getenv("MB_HOST_PARSE") && looksLikePlainFastq(inputs[0]) && (!paired || looksLikePlainFastq(inputs[1]));
    bool hostFallback = false;           // the first chunk was not four-line FASTQ after all: nothing was written yet
    if (deviceParse) {
        RawText f1, f2;
        uint64_t batchesDone = 0;
        if (!f1.open(inputs[0])) { err = "cannot open " + inputs[0]; return 1; }
        if (paired && !f2.open(inputs[1])) { err = "cannot open " + inputs[1]; return 1; }
        // a chunk holds at least one batch of short reads (~330 text bytes per 150-base record) and never less than 64 MB
        // ... and holds no more bases than the batch budget allows (FASTQ text is a little over twice its bases)
        const size_t chunk = std::min<size_t>(std::max<size_t>(64u << 20, std::min<size_t>(opt.batchReads * 400, 2ull << 30)),
                                              std::max<size_t>(maxBatchBases * 2 + (1u << 20), 4u << 20));
        std::vector<uint64_t> nameOff;
        std::vector<uint32_t> nameLen;
        while (true) {
            f1.refill(chunk);
            if (paired) f2.refill(chunk);
            if (f1.fill == 0) break;
            const bool last = f1.eof && (!paired || f2.eof);
            mb_fastq_index idx;
            int rc = mb_stage_fastq(db.ctx(), f1.buf.data(), f1.fill, paired ? f2.buf.data() : nullptr, paired ? f2.fill : 0,
                                    (uint32_t) std::min<size_t>(opt.batchReads, 0xFFFFFFFFu), last ? 1 : 0, &idx);
            if (rc == MB_E_DATA && batchesDone == 0) {
                std::cout << "Input is not plain four-line FASTQ (" << mb_last_error(db.ctx()) << "): using the host reader" << std::endl;
                hostFallback = true;
                break;
            }
            if (rc != MB_OK) { err = mb_last_error(db.ctx()); return 1; }
            ++batchesDone;
            if (idx.n_reads == 0) {
                if (last) break;
                err = "FASTQ record longer than the read chunk (use a smaller --batch-reads only with short reads)";
                return 1;
            }
            size_t used = 0;
            rc = mb_classify_staged(db.ctx(), &used);
            res.resize(idx.n_reads);
            if (rc == MB_OK) {
                tcId.resize(used + 1);
                tcCnt.resize(used + 1);
                rc = mb_fetch_results(db.ctx(), res.data(), tcId.data(), tcCnt.data(), used + 1);
            }
            nameOff.resize(idx.n_reads);
            nameLen.resize(idx.n_reads);
            if (rc == MB_OK) rc = mb_fetch_fastq_names(db.ctx(), nameOff.data(), nameLen.data());
            if (rc != MB_OK) { err = mb_last_error(db.ctx()); return 1; }
            batch.clear();
            batch.names.resize(idx.n_reads);
            for (uint32_t i = 0; i < idx.n_reads; ++i) batch.names[i].assign(f1.buf.data() + nameOff[i], nameLen[i]);
            mb_phase_times t;
            mb_get_timings(db.ctx(), &t);
            totalKmers += t.n_kmers;
            totalMatches += t.n_matches;
            gpuMs += t.total_ms;
            std::cout << "Batch of " << idx.n_reads << " reads (FASTQ parsed on the GPU): " << t.n_kmers << " query k-mers, " << t.n_matches
                      << " matches; GPU ms: parse+copy " << t.h2d_ms << " extract " << t.extract_ms << " sort " << t.sort_kmers_ms << " match "
                      << t.match_ms << " sort " << t.sort_matches_ms << " score " << t.score_ms << std::endl;
            rep.writeClassifications(batch, res.data(), tcId.data(), tcCnt.data());
            f1.consume(idx.consumed1);
            if (paired) f2.consume(idx.consumed2);
        }
        if (!hostFallback) {
            rep.closeClassifications();
            if (!rep.writeReport(err)) return 1;
            const auto t2 = clk::now();
            std::cout << "Number of query k-mers: " << totalKmers << "\nThe number of matches: " << totalMatches << "\nNumber of sequences: "
                      << rep.totalReads() << "\nGPU time: " << gpuMs / 1000.0 << " s\nTime for processing: "
                      << std::chrono::duration<double>(t2 - t0).count() << " s" << std::endl;
            return 0;
        }
    }
    FastxReader r1, r2;
    if (!r1.open(inputs[0], err)) return 1;
    if (paired && !r2.open(inputs[1], err)) return 1;
    // long reads: bound the batch by bases as well (QueryIndexer sizes batches by k-mer count, QueryIndexer.cpp:62-67)
    const size_t maxBases = std::min<size_t>(4ull << 30, maxBatchBases);
    while (loadBatch(r1, paired ? &r2 : nullptr, opt.batchReads, maxBases, batch)) {
        res.resize(batch.nReads());
        size_t used = 0;
        int rc = mb_stage_reads(db.ctx(), batch.bases.data(), batch.bases.size(), batch.offsets.data(), batch.lens.data(), batch.nReads());
        if (rc == MB_OK) rc = mb_classify_staged(db.ctx(), &used);
        if (rc == MB_OK) {
            tcId.resize(used + 1);
            tcCnt.resize(used + 1);
            rc = mb_fetch_results(db.ctx(), res.data(), tcId.data(), tcCnt.data(), used + 1);
        }
        if (rc != MB_OK) { err = mb_last_error(db.ctx()); return 1; }
        mb_phase_times t;
        mb_get_timings(db.ctx(), &t);
        totalKmers += t.n_kmers;
        totalMatches += t.n_matches;
        gpuMs += t.total_ms;
        std::cout << "Batch of " << batch.nReads() << " reads: " << t.n_kmers << " query k-mers, " << t.n_matches << " matches; GPU ms: extract "
                  << t.extract_ms << " sort " << t.sort_kmers_ms << " match " << t.match_ms << " sort " << t.sort_matches_ms << " score "
                  << t.score_ms << std::endl;
        rep.writeClassifications(batch, res.data(), tcId.data(), tcCnt.data());
    }
    rep.closeClassifications();
    if (!rep.writeReport(err)) return 1;
    const auto t2 = clk::now();
    std::cout << "Number of query k-mers: " << totalKmers << "\nThe number of matches: " << totalMatches << "\nNumber of sequences: "
              << rep.totalReads() << "\nGPU time: " << gpuMs / 1000.0 << " s\nTime for processing: "
              << std::chrono::duration<double>(t2 - t0).count() << " s" << std::endl;
    return 0;
}

} // namespace mbh

// ------------------------------------------------------------------------------------------------------
// C entry points for ctypes (tests/, bench.py)
// ------------------------------------------------------------------------------------------------------
extern "C" {

struct mbh_db {
    mbh::Database db;
    mbh::Options opt;
    std::string err;
};

static thread_local std::string g_err;

const char *mbh_last_error(void) { return g_err.c_str(); }

// seq_mode etc. are the command-line values; DB-side overrides are applied as `classify` does.
mbh_db *mbh_open(const char *dbdir, int device, int seq_mode, float min_score, float min_sp_score, int shard, int n_shards) {
    mbh_db *h = new mbh_db();
    h->opt.device = device;
    h->opt.seqMode = seq_mode;
    h->opt.minScore = min_score;
    h->opt.minSpScore = min_sp_score;
    if (!h->db.open(dbdir, h->opt, shard, n_shards, g_err)) { delete h; return nullptr; }
    return h;
}
void mbh_close(mbh_db *h) { delete h; }
mb_ctx *mbh_ctx(mbh_db *h) { return h->db.ctx(); }
void mbh_params(mbh_db *h, mb_params *out) { *out = h->db.params(); }
int mbh_num_bounds(mbh_db *h) { return (int) h->db.shardBounds().size(); }
void mbh_bounds(mbh_db *h, uint64_t *out) { for (size_t i = 0; i < h->db.shardBounds().size(); ++i) out[i] = h->db.shardBounds()[i]; }
int32_t mbh_original_taxid(mbh_db *h, int32_t t) { return h->db.taxonomy().original(t); }
const char *mbh_rank(mbh_db *h, int32_t t) { return h->db.taxonomy().rank(t); }
uint64_t mbh_db_kmers(mbh_db *h) { return h->db.numKmers(); }
uint64_t mbh_db_words(mbh_db *h) { return h->db.numWords(); }

// the whole CLI path on files: `metabuli classify` (returns exit code)
int mbh_classify_files(const char *reads1, const char *reads2, const char *dbdir, const char *outdir, const char *jobid,
                       int seq_mode, int device, float min_score, float min_sp_score, uint64_t batch_reads) {
    mbh::Options opt;
    opt.seqMode = seq_mode;
    opt.device = device;
    opt.minScore = min_score;
    opt.minSpScore = min_sp_score;
    if (batch_reads) opt.batchReads = (size_t) batch_reads;
    std::vector<std::string> in;
    in.push_back(reads1);
    if (reads2 && *reads2) in.push_back(reads2);
    return mbh::runClassify(in, dbdir, outdir, jobid, opt, g_err);
}

// The output writers alone (no GPU): per-read results in, JOBID_classifications.tsv / JOBID_report.tsv / JOBID_krona.html out,
// exactly as runClassify writes them.  Lets the CPU test-suite compare the writers with the reference Reporter's files.
int mbh_write_outputs(const char *dbdir, const char *outdir, const char *jobid, int lineage, const char *krona_prelude, uint32_t n_reads,
                      const char *const *names, const mb_read_result *res, const int32_t *tc_id, const int32_t *tc_cnt) {
    mbh::Taxonomy tax;
    if (!tax.load(std::string(dbdir) + "/taxonomyDB", g_err)) return 1;
    mbh::Reporter rep(tax, outdir, jobid, lineage != 0, krona_prelude ? krona_prelude : "");
    if (!rep.openClassifications(g_err)) return 1;
    mbh::ReadBatch batch;
    batch.names.assign(names, names + n_reads);
    rep.writeClassifications(batch, res, tc_id, tc_cnt);
    rep.closeClassifications();
    return rep.writeReport(g_err) ? 0 : 1;
}

} // extern "C"
