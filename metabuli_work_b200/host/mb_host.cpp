// mb_host.cpp — host side of the B200 classifier (see mb_host.h).  File:line citations are relative to
// the reference tree.
#include "mb_host.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <sstream>
#include <thread>

namespace mbh {

// pread of one byte range by several threads (page-cache copies and NVMe queues both want more than one reader)
bool preadParallel(int fd, void *dst, size_t bytes, size_t off, int threads) {
    const size_t minPiece = 8u << 20;
    const int n = (int) std::max<size_t>(1, std::min<size_t>((size_t) std::max(1, threads), bytes / minPiece));
    std::vector<char> ok((size_t) n, 1);
    auto piece = [&](int t) {
        size_t b = bytes * (size_t) t / (size_t) n, e = bytes * (size_t) (t + 1) / (size_t) n;
        char *p = static_cast<char *>(dst) + b;
        while (b < e) {
            const ssize_t got = ::pread(fd, p, e - b, (off_t) (off + b));
            if (got <= 0) { ok[(size_t) t] = 0; return; }
            p += got; b += (size_t) got;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < n; ++t) th.emplace_back(piece, t);
    piece(0);
    for (auto &x : th) x.join();
    for (char c : ok) if (!c) return false;
    return true;
}

std::vector<double> rebalancedCuts(const std::vector<double> &t, const double *frac) {
    const size_t n = t.size();
    std::vector<double> edge(n + 1, 0.0), out;
    for (size_t g = 1; g < n; ++g) edge[g] = frac ? frac[g - 1] : (double) g / (double) n;
    edge[n] = 1.0;
    double total = 0;
    for (double x : t) total += x;
    if (n < 2 || total <= 0) return out;
    // piecewise-constant work density t[g] / width[g]; the new boundaries cut its integral into n equal parts
    size_t g = 0;
    double done = 0;                      // work before shard g
    for (size_t k = 1; k < n; ++k) {
        const double want = total * (double) k / (double) n;
        while (g + 1 < n && done + t[g] < want) { done += t[g]; ++g; }
        const double within = t[g] > 0 ? (want - done) / t[g] : 0.0;
        out.push_back(edge[g] + within * (edge[g + 1] - edge[g]));
    }
    return out;
}

int ioThreads(const Options &opt) {
    if (opt.threads > 1) return opt.threads;
    const unsigned hw = std::thread::hardware_concurrency();
    return (int) std::min<unsigned>(std::max<unsigned>(hw, 1), 16);
}

// ------------------------------------------------------------------------------------------------------
// taxonomyDB
// ------------------------------------------------------------------------------------------------------
namespace {

bool readFile(const std::string &path, std::vector<char> &out, std::string &err) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { err = "cannot open " + path; return false; }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize((size_t) sz);
    size_t got = sz ? fread(out.data(), 1, (size_t) sz, f) : 0;
    fclose(f);
    if (got != (size_t) sz) { err = "short read on " + path; return false; }
    return true;
}

// MathUtil::flog2 (lib/mmseqs/src/commons/MathUtil.h:107-119): the width of the sparse table M in the
// serialized taxonomy is floor(flog2(2*maxNodes)) + 1, computed with this approximation.
float flog2Approx(float x) {
    if (x <= 0) return -128;
    int32_t bits;
    memcpy(&bits, &x, 4);
    float e = (float) (((bits & 0x7F800000) >> 23) - 0x7f);
    bits = (bits & 0x007FFFFF) | 0x3f800000;
    memcpy(&x, &bits, 4);
    x -= 1.0f;
    x *= (1.441740 + x * (-0.7077702 + x * (0.4123442 + x * (-0.1903190 + x * 0.0440047))));
    return x + e;
}

} // namespace

bool Taxonomy::load(const std::string &path, std::string &err) {
    std::vector<char> blob;
    if (!readFile(path, blob, err)) return false;
    const char *p = blob.data();
    const char *end = p + blob.size();
    auto need = [&](size_t n) { return (size_t) (end - p) >= n; };
    if (!need(12)) { err = "taxonomyDB truncated"; return false; }
    int32_t version;
    memcpy(&version, p, 4); p += 4;
    if (version != 2) { err = "taxonomyDB: unsupported serialization version"; return false; }
    uint64_t flag;
    memcpy(&flag, p, 8);
    internalIds_ = false;
    if (flag == 1) { internalIds_ = true; p += 8; }          // TaxonomyWrapper.cpp:372-378
    if (!need(12)) { err = "taxonomyDB truncated"; return false; }
    uint64_t maxNodes;
    memcpy(&maxNodes, p, 8); p += 8;
    memcpy(&maxTaxID_, p, 4); p += 4;
    const size_t nT = (size_t) maxTaxID_ + 1;
    if (!need(maxNodes * sizeof(TaxNodeRec) + nT * 4 * (internalIds_ ? 2 : 1))) { err = "taxonomyDB truncated"; return false; }
    nodes_.resize(maxNodes);
    memcpy(nodes_.data(), p, maxNodes * sizeof(TaxNodeRec)); p += maxNodes * sizeof(TaxNodeRec);
    D_.resize(nT);
    memcpy(D_.data(), p, nT * 4); p += nT * 4;
    internal2org_.clear();
    if (internalIds_) { internal2org_.resize(nT); memcpy(internal2org_.data(), p, nT * 4); p += nT * 4; }
    // E, L (2N each), H (N) and the sparse table M are only needed for the reference's RMQ LCA; the GPU
    // uses parent/depth arrays, so they are skipped.
    const size_t dim = maxNodes * 2;
    const size_t k = (size_t) ((int) flog2Approx((float) dim)) + 1;
    const size_t skip = (dim * 2 + maxNodes + dim * k) * 4;
    if (!need(skip + 16)) { err = "taxonomyDB truncated (tables)"; return false; }
    p += skip;
    // StringBlock<unsigned int> (lib/mmseqs/src/commons/StringBlock.h:116-145)
    uint64_t byteCap;
    uint32_t entryCap, entryCnt;
    memcpy(&byteCap, p, 8); p += 8;
    memcpy(&entryCap, p, 4); p += 4;
    memcpy(&entryCnt, p, 4); p += 4;
    if (!need(byteCap + (size_t) entryCnt * 4)) { err = "taxonomyDB truncated (strings)"; return false; }
    strData_.assign(p, p + byteCap); p += byteCap;
    strOffsets_.resize(entryCnt);
    memcpy(strOffsets_.data(), p, (size_t) entryCnt * 4);
    (void) entryCap;
    eukaryota_ = 0;                                           // TaxonomyWrapper.h:90-100
    for (const TaxNodeRec &n : nodes_) {
        if (n.nameIdx == 0) continue;
        if (strcmp(str(n.nameIdx), "Eukaryota") == 0) { eukaryota_ = n.taxId; break; }
    }
    return true;
}

int Taxonomy::rankIndex(const char *r) {
    static const std::map<std::string, int> ranks = {
        {"forma", 1}, {"varietas", 2}, {"subspecies", 3}, {"species", 4}, {"species subgroup", 5},
        {"species group", 6}, {"subgenus", 7}, {"genus", 8}, {"subtribe", 9}, {"tribe", 10},
        {"subfamily", 11}, {"family", 12}, {"superfamily", 13}, {"parvorder", 14}, {"infraorder", 15},
        {"suborder", 16}, {"order", 17}, {"superorder", 18}, {"infraclass", 19}, {"subclass", 20},
        {"class", 21}, {"superclass", 22}, {"subphylum", 23}, {"phylum", 24}, {"superphylum", 25},
        {"subkingdom", 26}, {"kingdom", 27}, {"superkingdom", 28}, {"domain", 28}};
    auto it = ranks.find(r);
    return it == ranks.end() ? -1 : it->second;
}

std::string Taxonomy::lineage(int32_t t) const {
    // ExtendedShortRanks (TaxonomyWrapper.h:9-26) through findShortRank2 (TaxonomyWrapper.cpp:423-429)
    static const std::map<std::string, std::string> shortRanks = {
        {"subspecies", "ss"}, {"species", "s"}, {"subgenus", "sg"}, {"genus", "g"}, {"subfamily", "sf"}, {"family", "f"},
        {"suborder", "so"}, {"order", "o"}, {"subclass", "sc"}, {"class", "c"}, {"subphylum", "sp"}, {"phylum", "p"},
        {"subkingdom", "sk"}, {"kingdom", "k"}, {"superkingdom", "d"}, {"domain", "d"}, {"realm", "r"}};
    const TaxNodeRec *n = node(t);
    if (!n) return "-";
    std::vector<const TaxNodeRec *> chain;
    do {
        chain.push_back(n);
        n = node(n->parentTaxId);
    } while (n && n->parentTaxId != n->taxId && chain.size() < 1000);
    std::string out;
    for (size_t i = chain.size(); i-- > 0;) {
        const auto it = shortRanks.find(str(chain[i]->rankIdx));
        out += it == shortRanks.end() ? "-" : it->second;
        out += '_';
        out += str(chain[i]->nameIdx);
        if (i > 0) out += ';';
    }
    return out;
}

int32_t Taxonomy::atRank(int32_t t, const char *rank) const {
    if (t == 0 || !exists(t) || t == 1) return 0;
    const int want = rankIndex(rank);
    const TaxNodeRec *cur = &nodes_[D_[t]];
    int steps = 0;
    while (steps < 30 && rankIndex(str(cur->rankIdx)) < want) {
        cur = &nodes_[D_[cur->parentTaxId]];
        ++steps;
    }
    return steps == 30 ? t : cur->taxId;
}

void Taxonomy::denseParents(std::vector<int32_t> &parent, std::vector<int32_t> &speciesRankOf) const {
    const size_t n = (size_t) maxTaxID_ + 1;
    parent.assign(n, -1);
    speciesRankOf.assign(n, 0);
    // rank index per node once, then the <= 30-step walk of getTaxIdAtRank per taxon
    std::vector<int> nodeRank(nodes_.size());
    for (size_t i = 0; i < nodes_.size(); ++i) nodeRank[i] = rankIndex(str(nodes_[i].rankIdx));
    const int want = rankIndex("species");
    for (size_t t = 0; t < n; ++t) {
        if (D_[t] == -1) continue;
        const TaxNodeRec &nd = nodes_[D_[t]];
        parent[t] = nd.parentTaxId;
        if (t <= 1) continue;
        int idx = D_[t], steps = 0;
        while (steps < 30 && nodeRank[idx] < want) { idx = D_[nodes_[idx].parentTaxId]; ++steps; }
        speciesRankOf[t] = steps == 30 ? (int32_t) t : nodes_[idx].taxId;
    }
    if (n > 0 && D_[0] != -1 && parent[0] < 0) parent[0] = 0;
}

// ------------------------------------------------------------------------------------------------------
// database directory
// ------------------------------------------------------------------------------------------------------
static DbParameters readDbParameters(const std::string &dir) {
    DbParameters p;
    std::ifstream f(dir + "/db.parameters");
    if (!f.is_open()) return p;
    p.present = true;
    std::string line;
    while (std::getline(f, line)) {
        const size_t tab = line.find('\t');
        if (tab == std::string::npos) continue;
        const std::string k = line.substr(0, tab), v = line.substr(tab + 1);
        if (k == "Reduced_alphabet") p.reducedAA = atoi(v.c_str());
        else if (k == "Accession_level") p.accessionLevel = atoi(v.c_str());
        else if (k == "DB_name") p.dbName = v;
        else if (k == "Creation_date") p.dbDate = v;
        else if (k == "Skip_redundancy") { if (v == "1") p.skipRedundancy = 1; }
        else if (k == "Syncmer") { if (v == "1") p.syncmer = 1; }
        else if (k == "S-mer_len") p.smerLen = atoi(v.c_str());   // never written by the builder (it writes Syncmer_len)
        else if (k == "Kmer_format") p.kmerFormat = atoi(v.c_str());
    }
    return p;
}

Database::~Database() { if (ctx_) mb_destroy(ctx_); }

bool Database::open(const std::string &dir, Options &opt, int shard, int nShards, std::string &err, const double *cutFrac) {
    // loadDbParameters (common.cpp:88-133): DB-side values win where the reference lets them
    const DbParameters dp = readDbParameters(dir);
    int skipRedundancy = 0;
    if (dp.present) {
        if (dp.reducedAA) opt.reducedAA = dp.reducedAA;
        if (dp.skipRedundancy) skipRedundancy = 1;
        if (dp.syncmer && !opt.syncmer) opt.syncmer = 1;
        if (dp.smerLen > 0) opt.smerLen = dp.smerLen;
        if (dp.kmerFormat > 0) opt.kmerFormat = dp.kmerFormat;
        if (dp.accessionLevel == 0 && opt.accessionLevel == 1) opt.accessionLevel = 0;
        if (dp.accessionLevel == 1 && opt.accessionLevel == 0) opt.accessionLevel = 2;
    }
    if (opt.reducedAA) { err = "reduced amino-acid alphabet databases are outside this path (ReducedKmerMatcher)"; return false; }
    if (opt.accessionLevel != 0) { err = "accession-level classification is outside this path"; return false; }
    if (opt.em) { err = "--em is outside this path"; return false; }

    using clk = std::chrono::steady_clock;
    const bool timing = getenv("MB_TIMING") != nullptr;        // where the load time goes, on stderr
    const auto tStart = clk::now();
    auto lap = [&](const char *what) {
        if (timing) fprintf(stderr, "[db open] %-22s at %.3f s\n", what, std::chrono::duration<double>(clk::now() - tStart).count());
    };
    const std::string taxPath = opt.taxonomyPath.empty() ? dir + "/taxonomyDB" : opt.taxonomyPath + "/taxonomyDB";
    if (!tax_.load(taxPath, err)) return false;
    lap("taxonomy file parsed");

    // Only the shard's byte ranges of diffIdx / info are read (a rank of an 8-GPU run touches an eighth of the files); the
    // split table is small and read whole.
    std::vector<char> splitBytes;
    if (!readFile(dir + "/split", splitBytes, err)) return false;
    struct stat stW, stI;
    if (stat((dir + "/diffIdx").c_str(), &stW) != 0 || stat((dir + "/info").c_str(), &stI) != 0) { err = "cannot stat diffIdx / info in " + dir; return false; }
    const size_t nWords = (size_t) stW.st_size / 2, nKmers = (size_t) stI.st_size / 4;
    const int fdW = ::open((dir + "/diffIdx").c_str(), O_RDONLY), fdI = ::open((dir + "/info").c_str(), O_RDONLY);
    if (fdW < 0 || fdI < 0) { if (fdW >= 0) ::close(fdW); if (fdI >= 0) ::close(fdI); err = "cannot open diffIdx / info in " + dir; return false; }
    struct Closer { int a, b; ~Closer() { ::close(a); ::close(b); } } closer{fdW, fdI};
    auto preadAll = [&](int fd, void *dst, size_t bytes, size_t off) -> bool {
        char *p = static_cast<char *>(dst);
        while (bytes) {
            const ssize_t got = ::pread(fd, p, bytes, (off_t) off);
            if (got <= 0) return false;
            p += got; off += (size_t) got; bytes -= (size_t) got;
        }
        return true;
    };
    const mb_split *splits = reinterpret_cast<const mb_split *>(splitBytes.data());
    const size_t nSplits = splitBytes.size() / sizeof(mb_split);

    // usable split entries (KmerMatcher.cpp:157-164): entry 0 plus every later entry with a real ADkmer
    std::vector<mb_split> cuts;
    for (size_t i = 0; i < nSplits; ++i)
        if (i == 0 || (splits[i].ADkmer != 0 && splits[i].ADkmer != UINT64_MAX)) cuts.push_back(splits[i]);

    // Shard boundaries: nShards-1 split entries spread evenly over the usable ones.  Shard g starts at the
    // k-mer ADkmer of its boundary entry; that k-mer's delta occupies the words just before diffIdxOffset.
    struct Cut { uint64_t value; size_t wordStart; size_t wordAfter; size_t kIndex; };
    std::vector<Cut> cutPoints;       // one per boundary
    bounds_.clear();
    if (nShards > 1) {
        if ((size_t) nShards > cuts.size()) { err = "more shards than usable split entries"; return false; }
        size_t prevCut = 0;
        for (int g = 1; g < nShards; ++g) {
            size_t at = (size_t) g * cuts.size() / (size_t) nShards;
            if (cutFrac) {
                at = (size_t) std::llround(cutFrac[g - 1] * (double) cuts.size());
                // strictly increasing, inside the table
                const size_t lo = prevCut + 1, hi = cuts.size() - (size_t) (nShards - g);
                at = std::min(std::max(at, lo), hi);
            }
            prevCut = at;
            const mb_split &s = cuts[at];
            if (s.diffIdxOffset == 0 || s.diffIdxOffset > nWords || s.infoIdxOffset == 0) { err = "split entry out of range"; return false; }
            // the words in front of the boundary: a delta has at most five
            const size_t lookFrom = s.diffIdxOffset >= 6 ? (size_t) s.diffIdxOffset - 6 : 0;
            uint16_t look[6] = {0, 0, 0, 0, 0, 0};
            const size_t nLook = (size_t) s.diffIdxOffset - lookFrom;
            if (!preadAll(fdW, look, nLook * 2, lookFrom * 2)) { err = "short read on " + dir + "/diffIdx"; return false; }
            size_t start = (size_t) s.diffIdxOffset - 1;       // terminator word of ADkmer's delta
            while (start > lookFrom && !(look[start - 1 - lookFrom] & 0x8000u)) --start;
            cutPoints.push_back({s.ADkmer, start, (size_t) s.diffIdxOffset, (size_t) s.infoIdxOffset - 1});
            // A query is compared with EVERY DB k-mer of its amino-acid group, whatever its DNA part, and a split entry is
            // the first DB k-mer of a group (IndexCreator.cpp:843-846): route on the amino-acid part only, so that a query of
            // the boundary group whose DNA part is below ADkmer's still reaches the shard that holds the group.
            bounds_.push_back(s.ADkmer & ~0xFFFFFFull);
        }
    }
    size_t wBegin = 0, wEnd = nWords, kBegin = 0, kEnd = nKmers;
    std::vector<uint16_t> head;        // the shard's first k-mer re-encoded (shards > 0)
    if (nShards > 1) {
        if (shard > 0) { wBegin = cutPoints[shard - 1].wordAfter; kBegin = cutPoints[shard - 1].kIndex; }
        if (shard + 1 < nShards) { wEnd = cutPoints[shard].wordStart; kEnd = cutPoints[shard].kIndex; }
        if (shard > 0) {
            // re-encode the shard's first k-mer as a delta from 0 (IndexCreator.cpp:868-886) so that the
            // shard is a self-contained stream
            uint64_t v = cutPoints[shard - 1].value;
            uint16_t tmp[5];
            int idx = 3;
            tmp[4] = (uint16_t) (0x8000u | (v & 0x7FFFu));
            v >>= 15;
            while (v) { tmp[idx--] = (uint16_t) (v & 0x7FFFu); v >>= 15; }
            for (int k = idx + 1; k <= 4; ++k) head.push_back(tmp[k]);
        }
    }
    // The two big reads go on while this thread creates the CUDA context and prepares the taxonomy: buffers are not
    // value-initialised (a 2 GB memset would cost as much as the read) and each file is read by several threads.
    const size_t prefix = head.size(), nShardWords = prefix + (wEnd - wBegin), nShardKmers = kEnd - kBegin;
    std::unique_ptr<uint16_t[]> shardWords(new uint16_t[nShardWords + 1]);
    std::unique_ptr<uint32_t[]> shardInfo(new uint32_t[nShardKmers + 1]);
    std::copy(head.begin(), head.end(), shardWords.get());
    const int ioT = ioThreads(opt);
    bool readOk = true;
    std::thread fileReader([&] {
        readOk = preadParallel(fdW, shardWords.get() + prefix, (wEnd - wBegin) * 2, wBegin * 2, ioT) &&
                 preadParallel(fdI, shardInfo.get(), nShardKmers * 4, kBegin * 4, ioT);
    });
    struct Joiner { std::thread &t; ~Joiner() { if (t.joinable()) t.join(); } } joiner{fileReader};

    mb_default_params(&par_);
    par_.device = opt.device;
    par_.kmer_format = opt.kmerFormat;
    par_.syncmer = opt.syncmer;
    par_.smer_len = opt.smerLen;
    par_.seq_mode = opt.seqMode;
    par_.min_cons_cnt = opt.minConsCnt;
    par_.min_cons_cnt_euk = opt.minConsCntEuk;
    par_.min_score = opt.minScore;
    par_.min_sp_score = opt.minSpScore;
    par_.tie_ratio = opt.tieRatio;
    par_.skip_redundancy = skipRedundancy;
    par_.accession_level = 0;
    par_.mask_mode = opt.maskMode ? 1 : 0;          // KmerExtractor.cpp:37-38,328-335
    par_.mask_prob = opt.maskProb;
    par_.last_shard = (shard + 1 == nShards) ? 1 : 0;
    ctx_ = mb_create(&par_);
    if (!ctx_) { err = mb_global_error(); return false; }
    lap("GPU context created");

    // taxonomy tables; taxId2speciesId as KmerMatcher::loadTaxIdList builds it (KmerMatcher.cpp:56-120)
    std::vector<int32_t> parent, speciesRankOf, speciesOf((size_t) tax_.maxTaxId() + 1, 0);
    tax_.denseParents(parent, speciesRankOf);
    {
        std::ifstream f(dir + "/taxID_list");
        if (!f.is_open()) { err = "cannot open " + dir + "/taxID_list"; return false; }
        std::string line;
        while (std::getline(f, line)) {
            if (line.empty()) continue;
            const int32_t taxId = (int32_t) strtoul(line.c_str(), nullptr, 10);
            const TaxNodeRec *n = tax_.node(taxId);
            if (!n) continue;
            const int32_t sp = tax_.atRank(taxId, "species");
            if (taxId != n->taxId) speciesOf[taxId] = sp;
            int guard = 0;
            while (n->taxId != sp && guard++ < 1000) {
                speciesOf[n->taxId] = sp;
                const TaxNodeRec *pn = tax_.node(n->parentTaxId);
                if (!pn || pn == n) break;
                n = pn;
            }
            if (sp >= 0 && sp <= tax_.maxTaxId()) speciesOf[sp] = sp;
        }
    }
    if (mb_load_taxonomy(ctx_, parent.data(), speciesOf.data(), speciesRankOf.data(), tax_.maxTaxId(), tax_.node0TaxId(), tax_.eukaryota()) != MB_OK) {
        err = mb_last_error(ctx_);
        return false;
    }
    lap("taxonomy resident");
    fileReader.join();
    lap("diffIdx / info read");
    if (!readOk) { err = "short read on diffIdx / info in " + dir; return false; }
    const int rc = mb_load_db(ctx_, shardWords.get(), nShardWords, shardInfo.get(), nShardKmers);
    if (rc != MB_OK) { err = mb_last_error(ctx_); return false; }
    lap("DB resident, units cut");
    nKmers_ = nShardKmers;
    nWords_ = nShardWords;
    return true;
}

} // namespace mbh
