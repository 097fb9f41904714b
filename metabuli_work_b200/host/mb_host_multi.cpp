// mb_host_multi.cpp — `metabuli classify --gpus N`: one process, one context and one host thread per GPU.
//
// The database is cut by k-mer range at `split` entries into N shards (Database::open), every GPU is the home of an
// N-th of each batch's reads, and the two exchange steps of SURVEY §8(e) are part of the producing kernels: extraction
// stores every query k-mer straight into the receive buffer of the GPU that owns its amino-acid range, the join stores
// every match into the receive buffer of its read's home GPU (mb_extract_kmers_routed / mb_match_kmers_routed over
// peer memory, cudaDeviceEnablePeerAccess — no CUDA IPC, no NCCL inside one process).  Between a producer and its
// consumer the host threads meet at a barrier.  Replaces the batch loop of Classifier::startClassify
// (src/commons/Classifier.cpp:44-164) for several GPUs; the outputs are those of the one-GPU command, byte for byte.
#include "mb_host.h"

#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <iostream>
#include <memory>
#include <mutex>
#include <thread>

namespace mbh {
namespace {

// Barrier of the per-GPU threads; abort() releases everybody for good (a failed thread must not leave the others waiting).
class Barrier {
public:
    explicit Barrier(int n) : n_(n) {}
    bool wait() {
        std::unique_lock<std::mutex> lk(m_);
        if (aborted_) return false;
        const unsigned gen = gen_;
        if (++count_ == n_) { count_ = 0; ++gen_; cv_.notify_all(); return true; }
        cv_.wait(lk, [&] { return gen_ != gen || aborted_; });
        return !aborted_;
    }
    void abort() { std::lock_guard<std::mutex> lk(m_); aborted_ = true; cv_.notify_all(); }
private:
    std::mutex m_;
    std::condition_variable cv_;
    int n_, count_ = 0;
    unsigned gen_ = 0;
    bool aborted_ = false;
};

struct Shared {
    int n = 0;
    uint32_t readsPerRank = 0;
    std::vector<uint64_t> bounds;
    std::vector<mb_peer_mem> peers;
    size_t kmerCap = 0, matchCap = 0;
    std::atomic<unsigned long long> need{0};
    std::atomic<int> regrows{0};
    std::mutex errMu;
    std::string err;
    void fail(const std::string &e) { std::lock_guard<std::mutex> lk(errMu); if (err.empty()) err = e; }
};

struct Worker {
    std::unique_ptr<Database> db;
    int rank = 0;
    // this batch's slice
    std::vector<uint64_t> offsets;
    std::vector<uint32_t> lens;
    std::vector<mb_read_result> res;
    std::vector<int32_t> tcId, tcCnt;
    uint32_t first = 0, count = 0;
    uint64_t nKmers = 0, nMatches = 0;
};

static void atomicMax(std::atomic<unsigned long long> &a, unsigned long long v) {
    unsigned long long cur = a.load();
    while (cur < v && !a.compare_exchange_weak(cur, v)) {}
}

// (re)allocates every rank's receive side and wires all contexts to all receive sides; all threads call it together
static bool setupPeers(Worker &w, Shared &sh, Barrier &bar, size_t kmerCap, size_t matchCap) {
    mb_ctx *ctx = w.db->ctx();
    if (mb_peer_alloc(ctx, kmerCap, matchCap, &sh.peers[w.rank]) != MB_OK) { sh.fail(mb_last_error(ctx)); bar.abort(); return false; }
    if (!bar.wait()) return false;
    if (mb_peer_attach(ctx, (uint32_t) w.rank, (uint32_t) sh.n, sh.peers.data(), sh.bounds.data(), sh.readsPerRank) != MB_OK) {
        sh.fail(mb_last_error(ctx)); bar.abort(); return false;
    }
    if (w.rank == 0) { sh.kmerCap = kmerCap; sh.matchCap = matchCap; }
    return bar.wait();
}

// one batch on one GPU (thread body)
static void runBatch(Worker &w, Shared &sh, Barrier &bar, const ReadBatch &batch) {
    mb_ctx *ctx = w.db->ctx();
    auto die = [&](const char *what) { sh.fail(std::string(what) + ": " + mb_last_error(ctx)); bar.abort(); };
    const uint32_t n = batch.nReads();
    w.first = std::min<uint32_t>((uint32_t) w.rank * sh.readsPerRank, n);
    w.count = std::min<uint32_t>(sh.readsPerRank, n - w.first);
    const uint64_t b0 = w.count ? batch.offsets[2 * (size_t) w.first] : 0;
    const uint64_t b1 = w.count ? batch.offsets[2 * (size_t) (w.first + w.count) - 1] + batch.lens[2 * (size_t) (w.first + w.count) - 1] : 0;
    w.offsets.resize(2 * (size_t) w.count);
    w.lens.assign(batch.lens.begin() + 2 * (size_t) w.first, batch.lens.begin() + 2 * (size_t) (w.first + w.count));
    for (size_t i = 0; i < w.offsets.size(); ++i) w.offsets[i] = batch.offsets[2 * (size_t) w.first + i] - b0;
    if (mb_set_read_id_base(ctx, (uint32_t) w.rank * sh.readsPerRank) != MB_OK ||
        mb_stage_reads(ctx, batch.bases.data() + b0, (size_t) (b1 - b0), w.offsets.data(), w.lens.data(), w.count) != MB_OK) return die("stage");
    // receive capacity: a rank emits at most two k-mers per base; an owner gets about as many unless the queries pile up in
    // one shard (slack 1.5), as in the torch.distributed driver (exchange.py)
    // MB_EXCHANGE_CAP0 (test knob): first receive capacity in records, so that small inputs reach the regrow paths below
    static const unsigned long long cap0 = getenv("MB_EXCHANGE_CAP0") ? strtoull(getenv("MB_EXCHANGE_CAP0"), nullptr, 10) : 0;
    atomicMax(sh.need, cap0 && !sh.kmerCap ? cap0 : std::max<unsigned long long>(3ull * (b1 - b0), 1ull << 16));
    if (!bar.wait()) return;
    const size_t needK = (size_t) sh.need.load();
    if (!bar.wait()) return;                               // everybody has read `need`
    if (w.rank == 0) sh.need.store(0);
    if (needK > sh.kmerCap && !setupPeers(w, sh, bar, needK, std::max(sh.matchCap, needK / 2))) return;
    size_t got = 0;
    while (true) {
        if (mb_extract_kmers_routed(ctx, &got) != MB_OK) return die("extract");
        if (!bar.wait()) return;                           // every producer is done
        const int rc = mb_take_routed_kmers(ctx, &got);
        if (rc != MB_OK && rc != MB_E_CAPACITY) return die("k-mer exchange");
        if (rc == MB_E_CAPACITY) atomicMax(sh.need, got);
        if (!bar.wait()) return;
        const size_t needMore = (size_t) sh.need.load();
        if (!bar.wait()) return;
        if (w.rank == 0) sh.need.store(0);
        if (!needMore) break;
        // the query k-mers pile up in one shard beyond the slack: everybody re-allocates with the worst demand, extraction is redone
        if (w.rank == 0) ++sh.regrows;
        if (!setupPeers(w, sh, bar, needMore + needMore / 4, sh.matchCap)) return;
    }
    w.nKmers = got;
    if (mb_sort_kmers(ctx) != MB_OK) return die("sort");
    while (true) {
        size_t nm = 0;
        if (mb_match_kmers_routed(ctx, &nm) != MB_OK) return die("match");
        if (!bar.wait()) return;
        const int rc = mb_take_routed_matches(ctx, &nm);
        if (rc != MB_OK && rc != MB_E_CAPACITY) return die("match exchange");
        if (rc == MB_E_CAPACITY) atomicMax(sh.need, nm);
        w.nMatches = nm;
        if (!bar.wait()) return;
        const size_t needM = (size_t) sh.need.load();
        if (!bar.wait()) return;
        if (w.rank == 0) sh.need.store(0);
        if (!needM) break;
        // some GPU received more matches than its buffer holds: everybody re-allocates and the join is redone (the sorted
        // k-mers are still resident)
        if (w.rank == 0) ++sh.regrows;
        if (!setupPeers(w, sh, bar, sh.kmerCap, needM + needM / 4)) return;
    }
    size_t used = 0;
    if (mb_sort_matches(ctx) != MB_OK || mb_score_reads(ctx, &used) != MB_OK) return die("score");
    w.res.resize(w.count);
    w.tcId.resize(used + 1);
    w.tcCnt.resize(used + 1);
    if (mb_fetch_results(ctx, w.res.data(), w.tcId.data(), w.tcCnt.data(), used + 1) != MB_OK) return die("fetch");
}

bool loadBatchMulti(FastxReader &r1, FastxReader *r2, size_t maxReads, size_t maxBases, ReadBatch &b) {
    b.clear();
    std::string name, seq, name2, seq2;
    while (b.names.size() < maxReads && b.bases.size() < maxBases && r1.next(name, seq)) {
        b.offsets.push_back(b.bases.size());
        b.lens.push_back((uint32_t) seq.size());
        b.bases.insert(b.bases.end(), seq.begin(), seq.end());
        if (r2) { if (!r2->next(name2, seq2)) seq2.clear(); } else seq2.clear();
        b.offsets.push_back(b.bases.size());
        b.lens.push_back((uint32_t) seq2.size());
        b.bases.insert(b.bases.end(), seq2.begin(), seq2.end());
        b.names.push_back(name);
    }
    return !b.names.empty();
}

} // namespace

int runClassifyMulti(const std::vector<std::string> &inputs, const std::string &dbDir, const std::string &outDir,
                     const std::string &jobId, Options opt, std::string &err) {
    using clk = std::chrono::steady_clock;
    const auto t0 = clk::now();
    const int n = (int) opt.devices.size();
    if (n > MB_MAX_RANKS) { err = "too many GPUs"; return 1; }
    Shared sh;
    sh.n = n;
    sh.peers.resize(n);
    std::vector<Worker> workers(n);
    // (re)opens all shards side by side: each thread reads its own byte ranges of diffIdx / info.  frac: shard boundaries as
    // fractions of the split table (null: equal k-mer counts)
    auto openShards = [&](const double *frac) -> bool {
        std::vector<std::thread> th;
        std::vector<std::string> errs(n);
        std::vector<Options> opts(n, opt);
        for (int g = 0; g < n; ++g) {
            workers[g].rank = g;
            workers[g].db.reset(new Database());
            opts[g].device = opt.devices[g];
            th.emplace_back([&, g] { if (!workers[g].db->open(dbDir, opts[g], g, n, errs[g], frac)) workers[g].db.reset(); });
        }
        for (auto &t : th) t.join();
        for (int g = 0; g < n; ++g) if (!workers[g].db) { err = "GPU " + std::to_string(opt.devices[g]) + ": " + errs[g]; return false; }
        const std::vector<int> devices = opt.devices;
        opt = opts[0];                                     // DB-side parameter overrides (the same on every shard)
        opt.devices = devices;
        for (int g = 0; g < n; ++g)
            for (int h = 0; h < n; ++h)
                if (mb_enable_peer_access(workers[g].db->ctx(), opt.devices[h]) != MB_OK) { err = mb_last_error(workers[g].db->ctx()); return false; }
        sh.bounds = workers[0].db->shardBounds();
        sh.kmerCap = sh.matchCap = 0;                      // the receive sides belong to the old contexts: set up again
        return true;
    };
    if (!openShards(nullptr)) return 1;
    sh.readsPerRank = (uint32_t) ((opt.batchReads + (size_t) n - 1) / (size_t) n);
    size_t dbKmers = 0, dbBytes = 0;
    for (auto &w : workers) { dbKmers += w.db->numKmers(); dbBytes += w.db->numWords() * 2 + w.db->numKmers() * 4; }
    std::cout << "DB resident on " << n << " GPUs (k-mer range shards): " << dbKmers << " k-mers, " << dbBytes << " bytes ("
              << std::chrono::duration<double>(clk::now() - t0).count() << " s)" << std::endl;
    // batch budget: the smallest GPU's (see runClassify)
    size_t freeMin = ~(size_t) 0;
    for (auto &w : workers) {
        size_t f = 0, t = 0;
        if (mb_mem_info(w.db->ctx(), &f, &t) != MB_OK) { err = mb_last_error(w.db->ctx()); return 1; }
        freeMin = std::min(freeMin, f);
    }
    const size_t budget = std::min<size_t>((size_t) (0.9 * (double) freeMin), (size_t) std::max(1, opt.ramUsage) << 30);
    const size_t slotBytes = 32 + (size_t) std::max(1, opt.matchPerKmer) * (48 + (opt.seqMode == 3 ? 85 : 12)) + 16 + 24 * (size_t) std::max(1, opt.matchPerKmer) / 2;
    const size_t maxBatchBases = std::max<size_t>(budget / slotBytes, 1u << 20) / 2 * (size_t) n;
    const Taxonomy tax = workers[0].db->taxonomy();          // a copy: the shards may be reloaded (re-balancing) while the Reporter lives
    Reporter rep(tax, outDir, jobId, opt.printLineage != 0, opt.kronaPrelude);
    rep.setThreads(ioThreads(opt));
    if (!rep.openClassifications(err)) return 1;
    const bool paired = opt.seqMode == 2;
    FastxReader r1, r2;
    if (!r1.open(inputs[0], err)) return 1;
    if (paired && !r2.open(inputs[1], err)) return 1;
    ReadBatch batch;
    uint64_t totalKmers = 0, totalMatches = 0;
    int rebalanceRounds = 0;
    std::vector<double> cutFrac;
    while (loadBatchMulti(r1, paired ? &r2 : nullptr, opt.batchReads, std::min<size_t>(4ull << 30, maxBatchBases), batch)) {
        Barrier bar(n);
        std::vector<std::thread> th;
        for (int g = 0; g < n; ++g) th.emplace_back([&, g] { runBatch(workers[g], sh, bar, batch); });
        for (auto &t : th) t.join();
        if (!sh.err.empty()) { err = sh.err; return 1; }
        uint64_t bk = 0, bm = 0;
        for (auto &w : workers) {
            rep.writeClassifications(batch.names.data() + w.first, w.count, w.res.data(), w.tcId.data(), w.tcCnt.data());
            bk += w.nKmers; bm += w.nMatches;
        }
        totalKmers += bk; totalMatches += bm;
        std::cout << "Batch of " << batch.nReads() << " reads on " << n << " GPUs: " << bk << " query k-mers, " << bm << " matches" << std::endl;
        // Equal k-mer counts per shard are not equal work: after each of the first two large batches the
        // owners' times (k-mer sort + join) decide whether the shard boundaries move and the shards are reloaded.
        static const size_t minReads = getenv("MB_REBALANCE_MIN_READS") ? (size_t) atoll(getenv("MB_REBALANCE_MIN_READS")) : 100000;
        static const double threshold = getenv("MB_REBALANCE_THRESHOLD") ? atof(getenv("MB_REBALANCE_THRESHOLD")) : 1.06;
        if (rebalanceRounds < 2 && batch.nReads() >= minReads) {
            std::vector<double> t(n);
            double mean = 0, worst = 0;
            for (int g = 0; g < n; ++g) {
                mb_phase_times pt;
                mb_get_timings(workers[g].db->ctx(), &pt);
                t[g] = (double) pt.sort_kmers_ms + (double) pt.match_ms;
                mean += t[g] / n;
                worst = std::max(worst, t[g]);
            }
            ++rebalanceRounds;
            if (worst > threshold * mean) {
                const std::vector<double> moved = rebalancedCuts(t, cutFrac.empty() ? nullptr : cutFrac.data());
                bool sane = moved.size() == (size_t) n - 1;
                for (size_t i = 0; sane && i < moved.size(); ++i) sane = moved[i] > (i ? moved[i - 1] : 0.0) + 1e-4 && moved[i] < 1.0 - 1e-4;
                if (sane) {
                    cutFrac = moved;
                    if (!openShards(cutFrac.data())) return 1;
                    std::cout << "DB shards re-balanced by owner time (slowest " << worst << " ms, mean " << mean << " ms)" << std::endl;
                }
            } else rebalanceRounds = 2;
        }
    }
    rep.closeClassifications();
    if (!rep.writeReport(err)) return 1;
    if (sh.regrows.load()) std::cout << "Exchange receive buffers regrown " << sh.regrows.load() << " time(s)" << std::endl;
    std::cout << "Number of query k-mers: " << totalKmers << "\nThe number of matches: " << totalMatches << "\nNumber of sequences: "
              << rep.totalReads() << "\nTime for processing: " << std::chrono::duration<double>(clk::now() - t0).count() << " s" << std::endl;
    return 0;
}

} // namespace mbh
