/*
 * ref_harness.cpp — drives the REFERENCE's own hot-path classes (TEST INFRASTRUCTURE).
 *
 * This file is ours; everything it calls is compiled from /root/reference by oracle/Makefile:
 *   MetamerScanner / OldMetamerScanner / SyncmerScanner + GeneticCode  (src/commons/*.h)
 *   ips4o SORT_PARALLEL + Kmer::compareQueryKmer                       (KmerExtractor.cpp:79)
 *   KmerMatcher::matchKmers / sortMatches                              (KmerMatcher.cpp:123,1071)
 *   Taxonomer::chooseBestTaxon                                         (Taxonomer.cpp:130)
 *   TaxonomyWrapper / loadTaxonomy / loadDbParameters                  (common.cpp:50-133)
 *   KSeqWrapper (FASTA/FASTQ parsing)
 * Extraction has two forms.  --ref-extractor runs the reference's own QueryIndexer::indexQueryFile and
 * KmerExtractor::extractQueryKmers (incl. --mask 1: SeqIterator::maskLowComplexityRegions / tantan) the way
 * Classifier::startClassify drives them; the nucleotide matrix its constructor needs is handed over as
 * "nucleotide.out:DATA" from a header generated out of lib/mmseqs/data/nucleotide.out.  It parses the files inside the
 * timed extraction, so the timing runs (bench.py) use the default form instead: the ~30-line read loop of KmerExtractor
 * (loadChunkOfReads / processSequence / fillQueryKmerBuffer, KmerExtractor.cpp:312-386,442-494) restated below over
 * reads parsed up front.  tests/golden/make_golden2.py checks on every fixture case that both forms give the same
 * k-mer buffer.
 *
 * LocalParameters cannot be constructed without the MMseqs2 Parameters machinery (generated
 * code again), so a zeroed object is fabricated and only the fields the path reads are set.
 *
 * usage: ref_harness DBDIR OUTPREFIX SEQMODE THREADS READS1 [READS2] [--dump] [--reps N] [--skip-reps W]
 *                    [--max-reads N] [--min-score F] [--min-sp-score F] [--smer-len S] [--report OUTDIR JOBID] [--lineage]
 *                    [--ref-extractor] [--mask 0|1] [--mask-prob P] [--dump-masked FILE]
 * --report: the reference's own Reporter (src/commons/Reporter.cpp:38-83,146-253) writes JOBID_classifications.tsv,
 * JOBID_report.tsv and JOBID_krona.html into OUTDIR exactly as Classifier::startClassify does (Classifier.cpp:133-158).
 * (--skip-reps W: the first W of the N repetitions are warm-up and stay out of the reported means)
 * writes OUTPREFIX.results.tsv and, with --dump, OUTPREFIX.kmers.bin / OUTPREFIX.matches.bin;
 * prints one JSON line with per-phase wall times.
 */
#include "Classifier.h"
#include "KmerExtractor.h"
#include "NucleotideMatrix.h"
#include "QueryIndexer.h"
#include "Reporter.h"
#include "KmerMatcher.h"
#include "KmerScanner.h"
#include "SyncmerScanner.h"
#include "Taxonomer.h"
#include "TaxonomyWrapper.h"
#include "LocalUtil.h"
#include "KSeqWrapper.h"
#include "FastSort.h"
#include "common.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <new>
#include <omp.h>

#include "nucleotide_out.h"   /* oracle/_ref/gen: lib/mmseqs/data/nucleotide.out as a C array (the reference's cmake embeds it the same way) */

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

/* Parameters::getInstance() is referenced by headers only through LocalParameters::getLocalInstance,
 * which we never call, so the symbol need not exist. */
static LocalParameters *fabricate_parameters() {
    void *mem = calloc(1, sizeof(LocalParameters));
    LocalParameters *par = reinterpret_cast<LocalParameters *>(mem);
    new (&par->filenames) std::vector<std::string>();
    new (&par->contamList) std::string();
    new (&par->taxonomyPath) std::string();
    new (&par->dbName) std::string();
    new (&par->dbDate) std::string();
    /* BaseMatrix::unserialize accepts "NAME.out:DATA" (SubstitutionMatrix.cpp:13-20): the matrix text travels in the name */
    new (&par->scoringMatrixFile) MultiParam<NuclAA<std::string>>(NuclAA<std::string>(
        "blosum62.out", std::string("nucleotide.out:") + std::string((const char *) nucleotide_out, nucleotide_out_len)));
    return par;
}

struct Mate { std::string seq; };

/* --dump-lr FILE: the 5x5 likelihood-ratio matrix tantan gets (ProbabilityMatrix over NucleotideMatrix, BaseMatrix.h:83-96),
 * the letter codes (aa2num) and the mask code, as the reference computes them at run time */
static int dump_lr(const char *path) {
    LocalParameters &par = *fabricate_parameters();
    NucleotideMatrix subMat(par.scoringMatrixFile.values.nucleotide().c_str(), 1.0, 0.0);
    ProbabilityMatrix pm(subMat);
    FILE *f = fopen(path, "wb");
    if (!f) return 1;
    for (int i = 0; i < subMat.alphabetSize; ++i) fwrite(pm.probMatrixPointers[i], sizeof(double), subMat.alphabetSize, f);
    unsigned char codes[256];
    for (int c = 0; c < 256; ++c) codes[c] = (unsigned char) subMat.aa2num[c];
    fwrite(codes, 1, 256, f);
    fwrite(pm.hardMaskTable, 1, 1, f);
    fclose(f);
    printf("alphabet %d\n", subMat.alphabetSize);
    return 0;
}

int main(int argc, char **argv) {
    if (argc == 3 && !strcmp(argv[1], "--dump-lr")) return dump_lr(argv[2]);
    if (argc < 6) {
        fprintf(stderr, "usage: %s DBDIR OUTPREFIX SEQMODE THREADS READS1 [READS2] [--dump] [--reps N] [--max-reads N]\n", argv[0]);
        return 2;
    }
    std::string dbDir = argv[1], outPrefix = argv[2];
    int seqMode = atoi(argv[3]);
    int threads = atoi(argv[4]);
    std::string reads1 = argv[5], reads2;
    int argi = 6;
    if (seqMode == 2) { reads2 = argv[argi++]; }
    bool dump = false;
    int reps = 1, skipReps = 0;
    size_t maxReads = SIZE_MAX;
    float minScore = 0.f, minSpScore = 0.f;
    int smerLen = 5, lineage = 0;
    std::string reportDir, reportJob;
    bool refExtractor = false;
    std::string dumpMasked;          /* file: SeqIterator::maskLowComplexityRegions of every mate, one line each */
    int maskMode = 0;
    float maskProb = 0.9f;
    for (; argi < argc; ++argi) {
        if (!strcmp(argv[argi], "--ref-extractor")) { refExtractor = true; continue; }
        if (!strcmp(argv[argi], "--mask")) { maskMode = atoi(argv[++argi]); continue; }
        if (!strcmp(argv[argi], "--mask-prob")) { maskProb = (float) atof(argv[++argi]); continue; }
        if (!strcmp(argv[argi], "--dump-masked")) { dumpMasked = argv[++argi]; continue; }
        if (!strcmp(argv[argi], "--dump")) dump = true;
        else if (!strcmp(argv[argi], "--reps")) reps = atoi(argv[++argi]);
        else if (!strcmp(argv[argi], "--skip-reps")) skipReps = atoi(argv[++argi]);
        else if (!strcmp(argv[argi], "--max-reads")) maxReads = strtoull(argv[++argi], nullptr, 10);
        else if (!strcmp(argv[argi], "--min-score")) minScore = (float) atof(argv[++argi]);
        else if (!strcmp(argv[argi], "--min-sp-score")) minSpScore = (float) atof(argv[++argi]);
        else if (!strcmp(argv[argi], "--smer-len")) smerLen = atoi(argv[++argi]);
        else if (!strcmp(argv[argi], "--lineage")) lineage = 1;
        else if (!strcmp(argv[argi], "--report") && argi + 2 < argc) { reportDir = argv[argi + 1]; reportJob = argv[argi + 2]; argi += 2; }
    }
    if (skipReps >= reps) skipReps = reps - 1;
    omp_set_num_threads(threads);

    LocalParameters &par = *fabricate_parameters();
    /* defaults of setClassifyDefaults (src/workflow/classify.cpp:10-37) */
    par.threads = threads;
    par.seqMode = seqMode;
    par.reducedAA = 0;
    par.minScore = minScore;
    par.minSpScore = minSpScore;
    par.hammingMargin = 0;
    par.verbosity = 3;
    par.ramUsage = 128;
    par.minConsCnt = 4;
    par.minConsCntEuk = 9;
    par.maskMode = maskMode;
    par.maskProb = maskProb;
    par.matchPerKmer = 4;
    par.accessionLevel = 0;
    par.tieRatio = 0.95;
    par.printLog = 0;
    par.syncmer = 0;
    par.smerLen = smerLen;
    par.printLineage = lineage;
    par.kmerFormat = 1;
    par.em = false;
    par.skipRedundancy = 0;
    par.maxGap = 0;
    par.minSSMatch = 0;
    par.filenames.push_back(reads1);
    if (seqMode == 2) par.filenames.push_back(reads2);
    par.filenames.push_back(dbDir);
    par.filenames.push_back(reportDir.empty() ? "/tmp" : reportDir);
    par.filenames.push_back(reportJob.empty() ? "harness" : reportJob);

    loadDbParameters(par, dbDir);                       /* common.cpp:88-133 */
    const int kmerFormat = par.kmerFormat;
    TaxonomyWrapper *taxonomy = loadTaxonomy(dbDir, ""); /* common.cpp:50-86 */
    GeneticCode geneticCode(false);
    KmerMatcher matcher(par, taxonomy, kmerFormat);

    /* ---- read input (KSeqWrapper, same parser as the reference) ---- */
    double t0 = now_s();
    std::vector<Query> queryList;
    std::vector<std::string> seq1, seq2;
    {
        KSeqWrapper *k1 = KSeqFactory(reads1.c_str());
        KSeqWrapper *k2 = seqMode == 2 ? KSeqFactory(reads2.c_str()) : nullptr;
        while (queryList.size() < maxReads && k1->ReadEntry()) {
            Query q;
            q.name = std::string(k1->entry.name.s);
            seq1.emplace_back(k1->entry.sequence.s, k1->entry.sequence.l);
            if (k2) {
                if (!k2->ReadEntry()) break;
                seq2.emplace_back(k2->entry.sequence.s, k2->entry.sequence.l);
            }
            queryList.push_back(q);
        }
        delete k1;
        delete k2;
    }
    const size_t nReads = queryList.size();
    double tParse = now_s() - t0;
    if (!dumpMasked.empty()) {
        NucleotideMatrix subMat(par.scoringMatrixFile.values.nucleotide().c_str(), 1.0, 0.0);
        ProbabilityMatrix probMatrix(subMat);
        FILE *fm = fopen(dumpMasked.c_str(), "w");
        std::vector<char> masked;
        for (size_t i = 0; i < nReads; ++i)
            for (int mate = 0; mate < (seqMode == 2 ? 2 : 1); ++mate) {
                const std::string &s = mate == 0 ? seq1[i] : seq2[i];
                masked.assign(s.size() + 1, 0);
                SeqIterator::maskLowComplexityRegions((const unsigned char *) s.c_str(), (unsigned char *) masked.data(), probMatrix, maskProb, &subMat);
                fprintf(fm, "%s\n", masked.data());
            }
        fclose(fm);
    }

    double tExtract = 0, tSortK = 0, tMatch = 0, tSortM = 0, tScore = 0;
    size_t nKmers = 0, nMatches = 0;
    Buffer<Kmer> kmerBuffer(16);
    Buffer<Match> matchBuffer(16);
    for (int rep = 0; rep < reps; ++rep) {
        if (rep == skipReps) tExtract = tSortK = tMatch = tSortM = tScore = 0;   /* warm-up repetitions end here */
        size_t slots = 0;
        if (refExtractor) {
            /* the reference's own QueryIndexer + KmerExtractor, driven as Classifier::startClassify drives them
             * (Classifier.cpp:57-112): index the files, one split, extract (+ mask) and sort inside extractQueryKmers */
            t0 = now_s();
            QueryIndexer indexer(par);
            indexer.setBytesPerKmer(par.matchPerKmer);
            indexer.indexQueryFile(0);
            const std::vector<QuerySplit> &splits = indexer.getQuerySplits();
            if (splits.size() != 1) { fprintf(stderr, "--ref-extractor: %zu query splits, expected 1\n", splits.size()); return 3; }
            static KmerExtractor *extractor = new KmerExtractor(par, geneticCode, kmerFormat);
            std::vector<std::string> names(nReads);
            for (size_t i = 0; i < nReads; ++i) names[i] = queryList[i].name;
            queryList.clear();
            queryList.resize(splits[0].end - splits[0].start);
            if (queryList.size() != nReads) { fprintf(stderr, "--ref-extractor: read count differs\n"); return 3; }
            kmerBuffer.reallocateMemory(splits[0].kmerCnt);
            kmerBuffer.init();
            KSeqWrapper *k1 = KSeqFactory(reads1.c_str());
            KSeqWrapper *k2 = seqMode == 2 ? KSeqFactory(reads2.c_str()) : nullptr;
            extractor->extractQueryKmers(kmerBuffer, queryList, splits[0], par, k1, k2);
            delete k1;
            delete k2;
            slots = kmerBuffer.startIndexOfReserve;
            tExtract += now_s() - t0;
        } else {
        /* ---- extraction: KmerExtractor.cpp:442-494 (lengths, empties), :312-386 (six frames) ---- */
        t0 = now_s();
        std::vector<size_t> slotStart(nReads + 1);
        std::vector<char> skip(nReads, 0);
        for (size_t i = 0; i < nReads; ++i) {
            Query &q = queryList[i];
            q.queryLength = LocalUtil::getMaxCoveredLength((int) seq1[i].size());
            int c1 = LocalUtil::getQueryKmerNumber<int>((int) seq1[i].size(), 0, 8);
            q.kmerCnt = c1 < 1 ? 0 : c1;
            bool empty = c1 < 1;
            if (seqMode == 2) {
                q.queryLength2 = LocalUtil::getMaxCoveredLength((int) seq2[i].size());
                int c2 = LocalUtil::getQueryKmerNumber<int>((int) seq2[i].size(), 0, 8);
                q.kmerCnt2 = (empty || c2 < 1) ? 0 : c2;
                if (c2 < 1) empty = true; /* shared emptyReads flag drops both mates */
            }
            skip[i] = empty;
            slotStart[i] = slots;
            if (!empty) slots += (size_t) q.kmerCnt + (size_t) (seqMode == 2 ? q.kmerCnt2 : 0);
        }
        slotStart[nReads] = slots;
        kmerBuffer.reallocateMemory(slots + 1);
        kmerBuffer.init();
        kmerBuffer.startIndexOfReserve = slots;
#pragma omp parallel
        {
            KmerScanner *scanner;
            if (kmerFormat == 1) scanner = new OldMetamerScanner(geneticCode);
            else if (par.syncmer) scanner = new SyncmerScanner(par.smerLen, geneticCode);
            else scanner = new MetamerScanner(geneticCode);
#pragma omp for schedule(dynamic, 256)
            for (size_t i = 0; i < nReads; ++i) {
                if (skip[i]) continue;
                size_t pos = slotStart[i];
                for (int mate = 0; mate < (seqMode == 2 ? 2 : 1); ++mate) {
                    const std::string &s = mate == 0 ? seq1[i] : seq2[i];
                    int seqLen = (int) s.size();
                    int usedLen = LocalUtil::getMaxCoveredLength(seqLen);
                    uint32_t offset = mate == 0 ? 0 : (uint32_t) queryList[i].queryLength + 3;
                    size_t w = mate == 0 ? pos : pos + (size_t) queryList[i].kmerCnt;
                    for (int frame = 0; frame < 6; ++frame) {
                        bool isForward = frame < 3;
                        int begin = isForward ? frame % 3 : (seqLen % 3) - (frame % 3);
                        if (begin < 0) begin += 3;
                        scanner->initScanner(s.c_str(), begin, begin + usedLen - 1, isForward);
                        Kmer kmer;
                        while ((kmer = scanner->next()).value != UINT64_MAX)
                            kmerBuffer.buffer[w++] = {kmer.value, (uint32_t) i + 1, kmer.pos + offset, (uint8_t) frame};
                    }
                }
            }
            delete scanner;
        }
        tExtract += now_s() - t0;

        t0 = now_s();
        SORT_PARALLEL(kmerBuffer.buffer, kmerBuffer.buffer + kmerBuffer.startIndexOfReserve, Kmer::compareQueryKmer);
        tSortK += now_s() - t0;
        }

        /* ---- match (retry with a larger buffer on overflow, Classifier.cpp:127-131) ---- */
        t0 = now_s();
        size_t matchCap = std::max<size_t>(slots * (size_t) par.matchPerKmer, 1u << 20);
        for (;;) {
            matchBuffer.reallocateMemory(matchCap);
            matchBuffer.startIndexOfReserve = 0;
            fflush(stdout);
            if (matcher.matchKmers(&kmerBuffer, &matchBuffer)) break;
            matchCap *= 2;
        }
        tMatch += now_s() - t0;
        nMatches = matchBuffer.startIndexOfReserve;

        t0 = now_s();
        matcher.sortMatches(&matchBuffer);
        tSortM += now_s() - t0;

        /* ---- scoring: Classifier::assignTaxonomy (Classifier.cpp:166-208) ---- */
        t0 = now_s();
        for (auto &q : queryList) { q.taxCnt.clear(); q.isClassified = false; q.classification = 0; q.score = 0; }
        {
            const Match *matchList = matchBuffer.buffer;
            size_t numOfMatches = nMatches;
            std::vector<MatchBlock> blocks;
            size_t matchIdx = 0;
            while (matchIdx < numOfMatches) {
                uint32_t cur = matchList[matchIdx].qInfo.sequenceID;
                size_t start = matchIdx;
                while (matchIdx < numOfMatches && cur == matchList[matchIdx].qInfo.sequenceID) ++matchIdx;
                blocks.emplace_back(start, matchIdx - 1, cur);
            }
#pragma omp parallel
            {
                Taxonomer taxonomer(par, taxonomy, kmerFormat);
#pragma omp for schedule(dynamic, 1)
                for (size_t i = 0; i < blocks.size(); ++i)
                    taxonomer.chooseBestTaxon(blocks[i].id - 1, blocks[i].start, blocks[i].end, matchList, queryList, par);
            }
        }
        tScore += now_s() - t0;

        size_t blank = 0;
        while (blank < kmerBuffer.startIndexOfReserve && kmerBuffer.buffer[blank].qInfo.sequenceID == 0) ++blank;
        nKmers = kmerBuffer.startIndexOfReserve - blank;

        if (rep == 0) {
            if (dump) {
                FILE *fk = fopen((outPrefix + ".kmers.bin").c_str(), "wb");
                fwrite(kmerBuffer.buffer + blank, sizeof(Kmer), nKmers, fk);
                fclose(fk);
                FILE *fm = fopen((outPrefix + ".matches.bin").c_str(), "wb");
                for (size_t i = 0; i < nMatches; ++i) {
                    const Match &m = matchBuffer.buffer[i];
                    unsigned char rec[24];
                    memset(rec, 0, sizeof(rec));
                    memcpy(rec, &m.qInfo, 8);
                    memcpy(rec + 8, &m.targetId, 4);
                    memcpy(rec + 12, &m.speciesId, 4);
                    memcpy(rec + 16, &m.dnaEncoding, 4);
                    memcpy(rec + 20, &m.rightEndHamming, 2);
                    rec[22] = m.hamming;
                    fwrite(rec, 1, 24, fm);
                }
                fclose(fm);
            }
            FILE *fr = fopen((outPrefix + ".results.tsv").c_str(), "w");
            for (size_t i = 0; i < nReads; ++i) {
                const Query &q = queryList[i];
                uint32_t bits;
                memcpy(&bits, &q.score, 4);
                fprintf(fr, "%d\t%s\t%d\t%d\t%d\t%08x\t", (int) q.isClassified, q.name.c_str(), q.classification,
                        taxonomy->getOriginalTaxID(q.classification), q.queryLength + q.queryLength2, bits);
                for (auto &kv : q.taxCnt) fprintf(fr, "%d:%d ", kv.first, kv.second);
                fprintf(fr, "\n");
            }
            fclose(fr);
            if (!reportDir.empty()) {
                /* Classifier::startClassify's output calls (Classifier.cpp:133-158) with the counts of assignTaxonomy (:201-203) */
                Reporter reporter(par, taxonomy);
                reporter.openReadClassificationFile();
                reporter.writeReadClassification(queryList);
                reporter.closeReadClassificationFile();
                std::unordered_map<TaxID, unsigned int> taxCounts;
                for (size_t i = 0; i < nReads; ++i) ++taxCounts[queryList[i].classification];
                reporter.writeReportFile((int) nReads, taxCounts, ReportType::Default);
            }
        }
    }
    fflush(stdout);
    printf("\n{\"harness\": \"reference\", \"reads\": %zu, \"kmers\": %zu, \"matches\": %zu, \"threads\": %d, \"reps\": %d, "
           "\"parse_s\": %.6f, \"extract_s\": %.6f, \"sort_kmers_s\": %.6f, \"match_s\": %.6f, \"sort_matches_s\": %.6f, "
           "\"score_s\": %.6f}\n",
           nReads, nKmers, nMatches, threads, reps - skipReps, tParse, tExtract / (reps - skipReps), tSortK / (reps - skipReps),
           tMatch / (reps - skipReps), tSortM / (reps - skipReps), tScore / (reps - skipReps));
    return 0;
}
