/*
 * ref_buildidx.cpp — drives the REFERENCE's own DB-build inner loop (TEST INFRASTRUCTURE).
 *
 * This file is ours; what it calls is compiled from /root/reference by oracle/Makefile:
 *   SORT_PARALLEL(..., Kmer::compareTargetKmer)                       (IndexCreator.cpp:266, Kmer.h:77-87)
 *   IndexCreator::filterKmers<FilterMode::DB_CREATION>                (IndexCreator.h:475-615)
 *   IndexCreator::writeTargetFilesAndSplits / getDiffIdx              (IndexCreator.cpp:811-886)
 * exactly in the order IndexCreator::createIndex runs them on one buffer of target k-mers (IndexCreator.cpp:262-280).
 *
 * usage: ref_buildidx TAXDIR INPUT.bin OUTDIR SPLITNUM THREADS
 *   TAXDIR    directory with a taxonomyDB (LCA of the duplicates' taxa)
 *   INPUT.bin records of 16 bytes: uint64 value, int32 taxId, int32 speciesId (internal ids), any order
 *   OUTDIR    receives diffIdx, info, split
 */
#include "IndexCreator.h"
#include "TaxonomyWrapper.h"
#include "FastSort.h"
#include "common.h"

#include <cstdio>
#include <cstring>
#include <new>
#include <omp.h>

#include "nucleotide_out.h"

static LocalParameters *fabricate_parameters() {
    void *mem = calloc(1, sizeof(LocalParameters));
    LocalParameters *par = reinterpret_cast<LocalParameters *>(mem);
    new (&par->filenames) std::vector<std::string>();
    new (&par->contamList) std::string();
    new (&par->taxonomyPath) std::string();
    new (&par->dbName) std::string();
    new (&par->dbDate) std::string();
    new (&par->cdsInfo) std::string();
    new (&par->scoringMatrixFile) MultiParam<NuclAA<std::string>>(NuclAA<std::string>(
        "blosum62.out", std::string("nucleotide.out:") + std::string((const char *) nucleotide_out, nucleotide_out_len)));
    return par;
}

struct Probe : public IndexCreator {
    Probe(const LocalParameters &par, TaxonomyWrapper *t, int fmt) : IndexCreator(par, t, fmt) {}
    using IndexCreator::writeTargetFilesAndSplits;
};

int main(int argc, char **argv) {
    if (argc < 6) { fprintf(stderr, "usage: %s TAXDIR INPUT.bin OUTDIR SPLITNUM THREADS\n", argv[0]); return 2; }
    const std::string taxDir = argv[1], input = argv[2], outDir = argv[3];
    LocalParameters &par = *fabricate_parameters();
    par.splitNum = atoi(argv[4]);
    par.threads = atoi(argv[5]);
    par.reducedAA = 0;
    par.syncmer = 0;
    par.smerLen = 5;
    par.maskMode = 0;
    par.maskProb = 0.9;
    par.filenames.push_back(outDir);
    par.filenames.push_back("x");
    omp_set_num_threads(par.threads);
    TaxonomyWrapper *taxonomy = loadTaxonomy(taxDir, "");
    Probe creator(par, taxonomy, 2);

    FILE *f = fopen(input.c_str(), "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", input.c_str()); return 1; }
    fseek(f, 0, SEEK_END);
    const size_t n = (size_t) ftell(f) / 16;
    fseek(f, 0, SEEK_SET);
    Buffer<Kmer> kmerBuffer(n + 16);
    for (size_t i = 0; i < n; ++i) {
        struct { uint64_t v; int32_t t, s; } r;
        if (fread(&r, 16, 1, f) != 1) return 1;
        kmerBuffer.buffer[i] = Kmer(r.v, (TaxID) r.t, (TaxID) r.s);
    }
    fclose(f);
    kmerBuffer.startIndexOfReserve = n;
    Buffer<size_t> uniqKmerIdx(n + 16);
    std::vector<std::pair<size_t, size_t>> ranges;
    SORT_PARALLEL(kmerBuffer.buffer, kmerBuffer.buffer + kmerBuffer.startIndexOfReserve, Kmer::compareTargetKmer);
    size_t selected = 0;
    creator.filterKmers<FilterMode::DB_CREATION>(kmerBuffer, uniqKmerIdx.buffer, selected, ranges);
    creator.writeTargetFilesAndSplits(kmerBuffer, uniqKmerIdx.buffer, selected, ranges);
    printf("\n{\"harness\": \"reference-build\", \"records\": %zu, \"selected\": %zu, \"splits\": %d, \"threads\": %d}\n", n, selected,
           par.splitNum, par.threads);
    return 0;
}
