// mb_host_io.cpp — FASTA/FASTQ input, classification / report output and the batch loop of `classify`.
#include "mb_host.h"

#include <zlib.h>

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <iostream>
#include <mutex>
#include <sstream>
#include <thread>

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

static inline void appendInt(std::string &s, long long v) {     // what "%d" prints
    char b[24];
    int i = 24;
    const bool neg = v < 0;
    unsigned long long u = neg ? 0ull - (unsigned long long) v : (unsigned long long) v;
    do { b[--i] = (char) ('0' + u % 10); u /= 10; } while (u);
    if (neg) b[--i] = '-';
    s.append(b + i, (size_t) (24 - i));
}

void Reporter::writeClassifications(const ReadBatch &batch, const mb_read_result *res, const int32_t *tcId, const int32_t *tcCnt) {
    writeClassifications(batch.names.data(), batch.nReads(), res, tcId, tcCnt);
}

void Reporter::writeClassifications(const std::string *names, uint32_t nReads, const mb_read_result *res, const int32_t *tcId,
                                    const int32_t *tcCnt) {
    writeClassifications(NameView{names, nullptr, nullptr, nullptr}, nReads, res, tcId, tcCnt);
}

// One line per read (Reporter::writeReadClassification, Reporter.cpp:38-83).  The lines of a slice of reads are put together
// in memory, slices by several threads; the file and the per-taxon counters are touched by the calling thread only.
void Reporter::formatSlice(const NameView &nv, uint32_t begin, uint32_t end, const mb_read_result *res, const int32_t *tcId,
                           const int32_t *tcCnt, std::string &out) const {
    char num[64];
    out.clear();
    out.reserve((size_t) (end - begin) * 96);
    for (uint32_t i = begin; i < end; ++i) {
        const mb_read_result &r = res[i];
        const int nNum = formatScore(num, sizeof(num), r.score);
        out.append(r.is_classified ? "1\t" : "0\t", 2);
        if (nv.names) out.append(nv.names[i]);
        else out.append(nv.text + nv.off[i], nv.len[i]);
        out.push_back('\t');
        appendInt(out, tax_.original(r.classification));
        out.push_back('\t');
        appendInt(out, r.query_length);
        out.push_back('\t');
        out.append(num, (size_t) nNum);
        out.push_back('\t');
        if (r.is_classified) {
            out.append(tax_.rank(r.classification));
            out.push_back('\t');
            if (lineage_) { out.append(tax_.lineage(r.classification)); out.push_back('\t'); }   // Reporter.cpp:60-62
            for (int32_t k = r.taxcnt_begin; k < r.taxcnt_end; ++k) {
                appendInt(out, tax_.original(tcId[k]));
                out.push_back(':');
                appendInt(out, tcCnt[k]);
                out.push_back(' ');
            }
            out.push_back('\n');
        } else {
            out.append(lineage_ ? "-\t-\t-\t\n" : "-\t-\t\n");
        }
    }
}

void Reporter::writeClassifications(const NameView &nv, uint32_t nReads, const mb_read_result *res, const int32_t *tcId,
                                    const int32_t *tcCnt) {
    const uint32_t perThread = 1u << 14;
    const int T = (int) std::max<uint32_t>(1, std::min<uint32_t>((uint32_t) std::max(1, threads_), (nReads + perThread - 1) / perThread));
    if (parts_.size() < (size_t) T) parts_.resize((size_t) T);
    auto slice = [&](int t) {
        const uint32_t b = (uint32_t) ((uint64_t) nReads * (uint64_t) t / (uint64_t) T), e = (uint32_t) ((uint64_t) nReads * (uint64_t) (t + 1) / (uint64_t) T);
        formatSlice(nv, b, e, res, tcId, tcCnt, parts_[(size_t) t]);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < T; ++t) th.emplace_back(slice, t);
    slice(0);
    for (uint32_t i = 0; i < nReads; ++i) ++taxCounts_[res[i].classification];          // Classifier.cpp:201-203
    total_ += nReads;
    for (auto &x : th) x.join();
    for (int t = 0; t < T; ++t) fwrite(parts_[(size_t) t].data(), 1, parts_[(size_t) t].size(), cls_);
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

// ---- the three stages of the batch loop run on their own threads: a reader fills text / read buffers, the calling thread
// drives the GPU, a writer formats and writes the classification lines.  Buffers circulate through bounded queues. ----------
template <class T> class Channel {
public:
    void push(T v) { { std::lock_guard<std::mutex> lk(mu_); q_.push_back(std::move(v)); } cv_.notify_all(); }
    bool pop(T &out) {                       // false: closed and drained
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !q_.empty() || closed_; });
        if (q_.empty()) return false;
        out = std::move(q_.front());
        q_.pop_front();
        return true;
    }
    void close() { { std::lock_guard<std::mutex> lk(mu_); closed_ = true; } cv_.notify_all(); }
    void closeAndDrop() { { std::lock_guard<std::mutex> lk(mu_); closed_ = true; q_.clear(); } cv_.notify_all(); }
private:
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<T> q_;
    bool closed_ = false;
};

// Text of one plain FASTQ file for one batch.  The reader thread preads a chunk behind `reserve` free bytes; the tail the
// previous batch did not consume (an incomplete record, or records beyond its mate's) is copied in front of it.
struct TextBuf {
    char *pin = nullptr;               // page-locked (mb_alloc_host) when the driver gives it, else malloc
    size_t cap = 0;
    bool pinned = false;
    std::vector<char> spill;           // a chunk or tail that outgrew the fixed buffer (a record longer than a chunk)
    char *base = nullptr;
    size_t start = 0, fill = 0;
    bool eof = false;
    char *data() const { return base + start; }
    void prepend(const char *src, size_t n) {
        if (n <= start) { memcpy(base + start - n, src, n); start -= n; fill += n; return; }
        std::vector<char> nb(n + fill + 64);
        memcpy(nb.data(), src, n);
        memcpy(nb.data() + n, data(), fill);
        spill.swap(nb);
        base = spill.data();
        start = 0;
        fill += n;
    }
};

struct BatchSlot {
    TextBuf f[2];                      // device-parse path
    ReadBatch batch;                   // host-reader path
    uint32_t nReads = 0;
    std::vector<mb_read_result> res;
    std::vector<int32_t> tcId, tcCnt;
    std::vector<uint64_t> nameOff;
    std::vector<uint32_t> nameLen;
};

// Four-line FASTQ in a regular file, plain or gzip?  Decided on the first (decompressed) bytes: '@' starts the file and the
// third line starts with '+'.  bytesPerRecord: average over the records of the first 64 KB of text.
static bool looksLikeFastq(const std::string &path, size_t &bytesPerRecord, size_t &fileSize, bool &gz) {
    struct stat st;
    if (stat(path.c_str(), &st) != 0 || !S_ISREG(st.st_mode)) return false;
    fileSize = (size_t) st.st_size;
    std::vector<char> headBuf(1 << 16);
    char *const head = headBuf.data();
    const size_t headCap = headBuf.size();
    size_t n = 0;
    {
        FILE *fp = fopen(path.c_str(), "rb");
        if (!fp) return false;
        n = fread(head, 1, headCap, fp);
        fclose(fp);
    }
    gz = n >= 2 && (unsigned char) head[0] == 0x1f && (unsigned char) head[1] == 0x8b;
    if (gz) {
        gzFile g = gzopen(path.c_str(), "rb");
        if (!g) return false;
        const int got = gzread(g, head, (unsigned) headCap);
        gzclose(g);
        n = got > 0 ? (size_t) got : 0;
    }
    if (n < 4 || head[0] != '@') return false;
    size_t line = 0, lastRecordEnd = 0;
    bool plus = false;
    for (size_t i = 0; i < n; ++i)
        if (head[i] == '\n') {
            ++line;
            if (line == 2) plus = i + 1 < n && head[i + 1] == '+';
            if (line % 4 == 0) lastRecordEnd = i + 1;
        }
    bytesPerRecord = line >= 4 ? (lastRecordEnd + line / 4 - 1) / (line / 4) : n;
    return plus;
}

int runClassify(const std::vector<std::string> &inputs, const std::string &dbDir, const std::string &outDir,
                const std::string &jobId, Options opt, std::string &err) {
    if (opt.devices.size() > 1) return runClassifyMulti(inputs, dbDir, outDir, jobId, opt, err);
    if (opt.devices.size() == 1) opt.device = opt.devices[0];
    using clk = std::chrono::steady_clock;
    const auto t0 = clk::now();
    const bool paired = opt.seqMode == 2;
    const int nFiles = paired ? 2 : 1;
    const int ioT = ioThreads(opt);
    constexpr int kSlots = 3;
    BatchSlot slots[kSlots];
    Channel<int> freeSlots, filled, toWrite;
    for (int i = 0; i < kSlots; ++i) freeSlots.push(i);

    // Plain FASTQ goes to the GPU as raw text (mb_stage_fastq: lines, records and packing are found on the device; the host
    // only cuts the read names out at the spans it gets back).  Gzip files are inflated by the reader threads (zlib, one
    // thread per file: a gzip stream cannot be inflated in parallel) and their TEXT takes the same road.  FASTA and
    // multi-line FASTQ take the host reader.
    size_t bpr[2] = {0, 0}, fileSize[2] = {0, 0};
    bool isGz[2] = {false, false};
    const bool deviceParse = !getenv("MB_HOST_PARSE") && looksLikeFastq(inputs[0], bpr[0], fileSize[0], isGz[0]) &&
                             (!paired || looksLikeFastq(inputs[1], bpr[1], fileSize[1], isGz[1]));
    // A chunk holds about --batch-reads records; the batch is bounded by the chunk's bytes, not by a record count, so that
    // what a batch leaves over is an incomplete record and not a pile of whole ones.
    std::atomic<size_t> chunkWant[2];
    size_t reserve[2] = {0, 0};
    for (int f = 0; f < nFiles; ++f) {
        // MB_CHUNK_MB (tuning knob) caps the chunk.  Page-locking costs time in proportion to the buffer size and, while the
        // database is being copied to the device, slows that copy down: measured on B200 with 8M pairs, 128 MB chunks (about
        // 400k short read pairs per batch) finish in 2.6 s, 330 MB chunks in 3.2-4.5 s.
        static const size_t chunkCap = (getenv("MB_CHUNK_MB") ? (size_t) atoll(getenv("MB_CHUNK_MB")) : (size_t) 128) << 20;
        chunkWant[f] = std::min<size_t>(std::max<size_t>(opt.batchReads * bpr[f] + 4096, 64u << 10), chunkCap);
        reserve[f] = std::max<size_t>(64u << 10, chunkWant[f] / 8);
    }
    std::atomic<bool> readerFailed{false};
    std::string readerErr;
    std::mutex readerErrMu;
    // first failure wins; the message is complete before the flag shows (two inflating threads may fail at once)
    auto readerFail = [&](const std::string &msg) {
        std::lock_guard<std::mutex> lk(readerErrMu);
        if (!readerFailed) { readerErr = msg; readerFailed = true; }
    };
    const bool timing = getenv("MB_TIMING") != nullptr;      // busy seconds of the three stages, on stderr
    double readBusy = 0, writeBusy = 0, gpuBusy = 0, pinBusy = 0;
    auto secondsSince = [](clk::time_point t) { return std::chrono::duration<double>(clk::now() - t).count(); };
    std::thread reader;
    if (deviceParse) {
        // starts before the database is resident: page-locking the buffers and reading the first chunk overlap the DB load
        reader = std::thread([&] {
            int fd[2] = {-1, -1};
            gzFile gzf[2] = {nullptr, nullptr};
            size_t pos[2] = {0, 0};
            bool srcEof[2] = {false, false};
            for (int f = 0; f < nFiles; ++f) {
                if (isGz[f]) {
                    gzf[f] = gzopen(inputs[(size_t) f].c_str(), "rb");
                    if (gzf[f]) gzbuffer(gzf[f], 1u << 20);
                } else fd[f] = ::open(inputs[(size_t) f].c_str(), O_RDONLY);
                if (fd[f] < 0 && !gzf[f]) readerFail("cannot open " + inputs[(size_t) f]);
            }
            // one file's share of a slot: `want` bytes of text behind the reserve (plain: pread by several threads; gzip: inflate)
            auto fillOne = [&](int si, int f) {
                TextBuf &tb = slots[si].f[f];
                const size_t want = isGz[f] ? (srcEof[f] ? 0 : (size_t) chunkWant[f]) : std::min<size_t>(chunkWant[f], fileSize[f] - pos[f]);
                if (!tb.pin) {       // first use of this slot (a file that fits one chunk never touches the other slots)
                    tb.cap = reserve[f] + (isGz[f] ? (size_t) chunkWant[f] : std::min<size_t>(chunkWant[f], fileSize[f])) + 64;
                    const auto tp = clk::now();
                    tb.pin = static_cast<char *>(mb_alloc_host(tb.cap));
                    if (f == 0) pinBusy += secondsSince(tp);
                    tb.pinned = tb.pin != nullptr;
                    if (!tb.pin) tb.pin = static_cast<char *>(malloc(tb.cap));
                    if (!tb.pin) { readerFail("out of host memory for the read buffers"); return; }
                }
                if (reserve[f] + want + 64 <= tb.cap) tb.base = tb.pin;
                else { tb.spill.resize(reserve[f] + want + 64); tb.base = tb.spill.data(); }
                tb.start = reserve[f];
                size_t got = 0;
                if (isGz[f]) {
                    while (got < want) {
                        const int r = gzread(gzf[f], tb.data() + got, (unsigned) std::min<size_t>(want - got, 1u << 30));
                        if (r < 0) { readerFail("gzip error in " + inputs[(size_t) f]); return; }
                        if (r == 0) { srcEof[f] = true; break; }
                        got += (size_t) r;
                    }
                    if (!srcEof[f] && gzeof(gzf[f])) srcEof[f] = true;
                } else {
                    got = want;
                    if (want && !preadParallel(fd[f], tb.data(), want, pos[f], ioT)) { readerFail("short read on " + inputs[(size_t) f]); return; }
                    srcEof[f] = pos[f] + want >= fileSize[f];
                }
                pos[f] += got;
                tb.fill = got;
                tb.eof = srcEof[f];
            };
            int si;
            while (!readerFailed && freeSlots.pop(si)) {
                const auto tr = clk::now();
                if (nFiles == 2 && (isGz[0] || isGz[1])) {      // the two inflations run side by side
                    std::thread second(fillOne, si, 1);
                    fillOne(si, 0);
                    second.join();
                } else {
                    for (int f = 0; f < nFiles; ++f) fillOne(si, f);
                }
                bool allEof = true;
                for (int f = 0; f < nFiles; ++f) allEof = allEof && slots[si].f[f].eof;
                readBusy += secondsSince(tr);
                filled.push(si);
                if (allEof || readerFailed) break;
            }
            filled.close();
            for (int f = 0; f < nFiles; ++f) { if (fd[f] >= 0) ::close(fd[f]); if (gzf[f]) gzclose(gzf[f]); }
        });
    }
    auto stopReader = [&] {
        freeSlots.closeAndDrop();
        if (reader.joinable()) reader.join();
    };
    auto freeText = [&] {
        for (BatchSlot &s : slots)
            for (TextBuf &tb : s.f) {
                if (tb.pin) { if (tb.pinned) mb_free_host(tb.pin); else free(tb.pin); }
                tb.pin = nullptr;
            }
    };

    Database db;
    if (!db.open(dbDir, opt, 0, 1, err)) { stopReader(); freeText(); return 1; }
    const auto t1 = clk::now();
    std::cout << "DB resident on GPU " << opt.device << ": " << db.numKmers() << " k-mers, " << db.numWords() * 2 + db.numKmers() * 4
              << " bytes (" << std::chrono::duration<double>(t1 - t0).count() << " s)" << std::endl;
    // Batch size from the memory that is left (QueryIndexer::indexQueryFile sizes its splits the same way from --max-ram,
    // QueryIndexer.cpp:62-67: bytes per k-mer slot = k-mer record(s) + matchPerKmer match records).  Per slot here: two
    // 16-byte k-mer buffers, and per expected match two 24-byte match buffers plus the scorer's per-match scratch
    // (85 bytes with long reads, an eighth of 93 bytes with short reads).  --max-ram (GB) caps the budget like it caps the
    // reference's host memory.
    size_t freeB = 0, totalB = 0;
    if (mb_mem_info(db.ctx(), &freeB, &totalB) != MB_OK) { err = mb_last_error(db.ctx()); stopReader(); freeText(); return 1; }
    const size_t budget = std::min<size_t>((size_t) (0.9 * (double) freeB), (size_t) std::max(1, opt.ramUsage) << 30);
    const size_t slotBytes = 32 + (size_t) std::max(1, opt.matchPerKmer) * (48 + (opt.seqMode == 3 ? 85 : 12));
    const size_t maxSlots = std::max<size_t>(budget / slotBytes, 1u << 20);
    const size_t maxBatchBases = maxSlots / 2;          // about two k-mer slots per base (six frames, one per codon)
    std::cout << "Batch budget: " << (budget >> 20) << " MiB of HBM (" << (freeB >> 20) << " MiB free), at most " << maxBatchBases
              << " bases per batch" << std::endl;
    // FASTQ text is a little over twice its bases: later chunks respect the budget (the first ones are already on their way)
    const size_t chunkMax = std::max<size_t>(maxBatchBases * 2 / (size_t) nFiles + (1u << 20), 4u << 20);
    for (int f = 0; f < nFiles; ++f) if (chunkWant[f] > chunkMax) chunkWant[f] = chunkMax;

    Reporter rep(db.taxonomy(), outDir, jobId, opt.printLineage != 0, opt.kronaPrelude);
    rep.setThreads(ioT);
    if (!rep.openClassifications(err)) { stopReader(); freeText(); return 1; }
    uint64_t totalKmers = 0, totalMatches = 0;
    double gpuMs = 0;
    // the writer: formats a finished batch and hands its buffers back to the reader
    bool namesFromText = deviceParse;
    Channel<int> freeHost;               // host-reader path: its own queue of free slots
    Channel<int> *freeCh = &freeSlots;
    std::thread writer([&] {
        int si;
        while (toWrite.pop(si)) {
            BatchSlot &s = slots[si];
            const auto tw = clk::now();
            if (namesFromText) rep.writeClassifications(NameView{nullptr, s.f[0].data(), s.nameOff.data(), s.nameLen.data()}, s.nReads,
                                                        s.res.data(), s.tcId.data(), s.tcCnt.data());
            else rep.writeClassifications(s.batch, s.res.data(), s.tcId.data(), s.tcCnt.data());
            writeBusy += secondsSince(tw);
            freeCh->push(si);
        }
    });
    auto stopAll = [&] {
        toWrite.close();
        if (writer.joinable()) writer.join();
        stopReader();
        freeText();
    };
    auto classifyAndFetch = [&](BatchSlot &s, uint32_t nReads) -> bool {
        size_t used = 0;
        int rc = mb_classify_staged(db.ctx(), &used);
        s.nReads = nReads;
        s.res.resize(nReads);
        if (rc == MB_OK) {
            s.tcId.resize(used + 1);
            s.tcCnt.resize(used + 1);
            rc = mb_fetch_results(db.ctx(), s.res.data(), s.tcId.data(), s.tcCnt.data(), used + 1);
        }
        if (rc != MB_OK) { err = mb_last_error(db.ctx()); return false; }
        mb_phase_times t;
        mb_get_timings(db.ctx(), &t);
        totalKmers += t.n_kmers;
        totalMatches += t.n_matches;
        gpuMs += t.total_ms;
        std::cout << "Batch of " << nReads << " reads" << (namesFromText ? " (FASTQ parsed on the GPU)" : "") << ": " << t.n_kmers
                  << " query k-mers, " << t.n_matches << " matches; GPU ms: " << (namesFromText ? "parse+copy " : "copy ") << t.h2d_ms
                  << " extract " << t.extract_ms << " sort " << t.sort_kmers_ms << " match " << t.match_ms << " sort " << t.sort_matches_ms
                  << " score " << t.score_ms << std::endl;
        return true;
    };

    bool hostFallback = false;           // the first chunk was not four-line FASTQ after all: nothing was written yet
    if (deviceParse) {
        uint64_t batchesDone = 0;
        int cur = -1;
        bool ok = filled.pop(cur);
        while (ok) {
            BatchSlot &s = slots[cur];
            const bool last = s.f[0].eof && (!paired || s.f[1].eof);
            if (s.f[0].fill == 0) break;
            mb_fastq_index idx;
            const auto tg = clk::now();
            int rc = mb_stage_fastq(db.ctx(), s.f[0].data(), s.f[0].fill, paired ? s.f[1].data() : nullptr, paired ? s.f[1].fill : 0,
                                    (1u << 29) - 2, last ? 1 : 0, &idx);
            if (rc == MB_E_DATA && batchesDone == 0) {
                std::cout << "Input is not plain four-line FASTQ (" << mb_last_error(db.ctx()) << "): using the host reader" << std::endl;
                hostFallback = true;
                break;
            }
            if (rc != MB_OK) { err = mb_last_error(db.ctx()); stopAll(); return 1; }
            ++batchesDone;
            if (idx.n_reads == 0 && last) break;
            int nxt = -1;
            if (!last) {
                // the next chunk is (being) read: what this batch left over goes in front of it, and its copy to the device
                // overlaps this batch's device work
                if (!filled.pop(nxt) || readerFailed) { err = readerFailed ? readerErr : "reader stopped early"; stopAll(); return 1; }
                BatchSlot &nx = slots[nxt];
                nx.f[0].prepend(s.f[0].data() + idx.consumed1, s.f[0].fill - idx.consumed1);
                if (paired) nx.f[1].prepend(s.f[1].data() + idx.consumed2, s.f[1].fill - idx.consumed2);
                if (idx.n_reads == 0) {
                    // not one whole record in the chunk: read twice as much from now on
                    for (int f = 0; f < nFiles; ++f) {
                        const size_t w = chunkWant[f];
                        if (w >= chunkMax) { err = "FASTQ record longer than a batch may be (--max-ram / free GPU memory)"; stopAll(); return 1; }
                        chunkWant[f] = std::min(chunkMax, 2 * w);
                    }
                } else if (mb_prefetch_fastq(db.ctx(), nx.f[0].data(), nx.f[0].fill, paired ? nx.f[1].data() : nullptr, paired ? nx.f[1].fill : 0) != MB_OK) {
                    err = mb_last_error(db.ctx()); stopAll(); return 1;
                }
            }
            if (idx.n_reads) {
                s.nameOff.resize(idx.n_reads);
                s.nameLen.resize(idx.n_reads);
                if (!classifyAndFetch(s, idx.n_reads)) { stopAll(); return 1; }
                if (mb_fetch_fastq_names(db.ctx(), s.nameOff.data(), s.nameLen.data()) != MB_OK) { err = mb_last_error(db.ctx()); stopAll(); return 1; }
                toWrite.push(cur);
            } else freeSlots.push(cur);
            gpuBusy += secondsSince(tg);
            if (last) break;
            cur = nxt;
        }
        if (readerFailed) { err = readerErr; stopAll(); return 1; }
    }
    if (!deviceParse || hostFallback) {
        if (hostFallback) stopReader();      // nothing was written yet; the writer stays
        namesFromText = false;
        freeCh = &freeHost;
        for (int i = 0; i < kSlots; ++i) freeHost.push(i);
        // long reads: bound the batch by bases as well (QueryIndexer sizes batches by k-mer count, QueryIndexer.cpp:62-67)
        const size_t maxBases = std::min<size_t>(4ull << 30, maxBatchBases);
        Channel<int> parsed;
        std::atomic<bool> openFailed{false};
        std::thread hostReader([&] {
            FastxReader r1, r2;
            if (!r1.open(inputs[0], readerErr) || (paired && !r2.open(inputs[1], readerErr))) { openFailed = true; parsed.close(); return; }
            int si;
            while (freeHost.pop(si)) {
                if (!loadBatch(r1, paired ? &r2 : nullptr, opt.batchReads, maxBases, slots[si].batch)) break;
                parsed.push(si);
            }
            parsed.close();
        });
        int cur;
        bool failed = false;
        while (!failed && parsed.pop(cur)) {
            BatchSlot &s = slots[cur];
            const ReadBatch &b = s.batch;
            if (mb_stage_reads(db.ctx(), b.bases.data(), b.bases.size(), b.offsets.data(), b.lens.data(), b.nReads()) != MB_OK) {
                err = mb_last_error(db.ctx());
                failed = true;
            } else if (!classifyAndFetch(s, b.nReads())) failed = true;
            else toWrite.push(cur);
        }
        freeHost.closeAndDrop();
        hostReader.join();
        if (openFailed) { err = readerErr; failed = true; }
        if (failed) { stopAll(); return 1; }
    }
    stopAll();
    rep.closeClassifications();
    if (timing) fprintf(stderr, "[classify] busy seconds: reader %.3f (page-locking %.3f), GPU calls (incl. waiting for the reader) %.3f, writer %.3f; batches done at %.3f s\n",
                        readBusy, pinBusy, gpuBusy, writeBusy, secondsSince(t0));
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
// the same with shard boundaries at given fractions of the split table (n_shards - 1 values; re-balancing by join time)
mbh_db *mbh_open_cuts(const char *dbdir, int device, int seq_mode, float min_score, float min_sp_score, int shard, int n_shards,
                      const double *cut_frac) {
    mbh_db *h = new mbh_db();
    h->opt.device = device;
    h->opt.seqMode = seq_mode;
    h->opt.minScore = min_score;
    h->opt.minSpScore = min_sp_score;
    if (!h->db.open(dbdir, h->opt, shard, n_shards, g_err, cut_frac)) { delete h; return nullptr; }
    return h;
}
// new boundaries from the per-shard times (out: n_shards - 1 values); cur may be NULL (equal counts)
void mbh_rebalanced_cuts(const double *time_per_shard, int n_shards, const double *cur, double *out) {
    std::vector<double> t(time_per_shard, time_per_shard + n_shards);
    const std::vector<double> r = mbh::rebalancedCuts(t, cur);
    for (size_t i = 0; i < r.size(); ++i) out[i] = r[i];
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

// The host reader alone (no GPU): the first batch runClassify's host path would stage from these files — at most max_reads
// reads and, checked before every read, max_bases bases — in the packed layout of mb_stage_reads; names joined by '\n'.
// Returns 0, 1 (error, mbh_last_error) or 2 (a caller's array is too small; the needed sizes are in the out arguments).
int mbh_read_fastx(const char *reads1, const char *reads2, uint64_t max_reads, uint64_t max_bases, char *bases, uint64_t bases_cap,
                   uint64_t *offsets, uint32_t *lens, uint64_t reads_cap, char *names, uint64_t names_cap, uint32_t *n_reads,
                   uint64_t *n_bases, uint64_t *names_len) {
    mbh::FastxReader r1, r2;
    const bool paired = reads2 && *reads2;
    if (!r1.open(reads1, g_err) || (paired && !r2.open(reads2, g_err))) return 1;
    mbh::ReadBatch b;
    mbh::loadBatch(r1, paired ? &r2 : nullptr, (size_t) max_reads, (size_t) max_bases, b);
    size_t nl = 0;
    for (const std::string &n : b.names) nl += n.size() + 1;
    *n_reads = b.nReads();
    *n_bases = b.bases.size();
    *names_len = nl;
    if (b.bases.size() > bases_cap || b.nReads() > reads_cap || nl > names_cap) return 2;
    if (!b.bases.empty()) memcpy(bases, b.bases.data(), b.bases.size());
    for (size_t i = 0; i < b.offsets.size(); ++i) { offsets[i] = b.offsets[i]; lens[i] = b.lens[i]; }
    char *w = names;
    for (const std::string &n : b.names) { memcpy(w, n.data(), n.size()); w += n.size(); *w++ = '\n'; }
    return 0;
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
