// metabuli_main.cpp — `metabuli classify` command line of the B200 build.
//
// Keeps the reference's call shape (src/workflow/classify.cpp:39-200, flag list
// src/commons/LocalParameters.cpp:718-740):
//   metabuli classify [--seq-mode 1|3] <fastx> <DBDIR> <OUTDIR> <JOBID> [flags]
//   metabuli classify  --seq-mode 2    <fastx1> <fastx2> <DBDIR> <OUTDIR> <JOBID> [flags]
// and writes OUTDIR/JOBID_classifications.tsv, OUTDIR/JOBID_report.tsv and OUTDIR/JOBID_krona.html.
#include "mb_host.h"

#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

static bool fileExists(const std::string &p) { struct stat st; return stat(p.c_str(), &st) == 0; }

static int usage() {
    std::cerr << "usage: metabuli classify [--seq-mode 1|2|3] <reads> [<reads2>] <DBDIR> <OUTDIR> <JOBID> [options]\n"
                 "options: --threads N --min-score F --min-sp-score F --min-cons-cnt N --min-cons-cnt-euk N --tie-ratio F\n"
                 "         --hamming-margin N --taxonomy-path DIR --mask 0|1 --mask-prob F --max-ram N --match-per-kmer N\n"
                 "         --accession-level 0|1 --lineage 0|1 --validate-input 0|1 --validate-db 0|1 --syncmer 0|1\n"
                 "         --smer-len N --kmer-format 1|2 --print-log 0|1 --reduced-aa 0|1 --em 0|1\n"
                 "         --gpu N (CUDA device) --batch-reads N (reads per batch)\n"
                 "         --gpus N | --gpu-list a,b,... (DB sharded by k-mer range over the listed GPUs, reads split over them)\n"
                 "         --krona-prelude FILE (static head of JOBID_krona.html: the reference's krona_prelude.html)\n";
    return 1;
}

int main(int argc, char **argv) {
    if (argc < 2 || strcmp(argv[1], "classify") != 0) return usage();
    mbh::Options opt;                       // defaults: setClassifyDefaults (classify.cpp:10-37)
    std::vector<std::string> pos;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        if (a.rfind("--", 0) != 0) { pos.push_back(a); continue; }
        if (i + 1 >= argc) { std::cerr << "missing value for " << a << "\n"; return 1; }
        const char *v = argv[++i];
        if (a == "--seq-mode") opt.seqMode = atoi(v);
        else if (a == "--threads") opt.threads = atoi(v);
        else if (a == "--min-score") opt.minScore = (float) atof(v);
        else if (a == "--min-sp-score") opt.minSpScore = (float) atof(v);
        else if (a == "--min-cons-cnt") opt.minConsCnt = atoi(v);
        else if (a == "--min-cons-cnt-euk") opt.minConsCntEuk = atoi(v);
        else if (a == "--tie-ratio") opt.tieRatio = (float) atof(v);
        else if (a == "--hamming-margin") opt.hammingMargin = atoi(v);   // parsed, unused on this path like in the reference
        else if (a == "--taxonomy-path") opt.taxonomyPath = v;
        else if (a == "--mask") opt.maskMode = atoi(v);
        else if (a == "--mask-prob") opt.maskProb = (float) atof(v);
        else if (a == "--max-ram") opt.ramUsage = atoi(v);
        else if (a == "--match-per-kmer") opt.matchPerKmer = atoi(v);
        else if (a == "--accession-level") opt.accessionLevel = atoi(v);
        else if (a == "--lineage") opt.printLineage = atoi(v);
        else if (a == "--validate-input") opt.validateInput = atoi(v);
        else if (a == "--validate-db") opt.validateDb = atoi(v);
        else if (a == "--syncmer") opt.syncmer = atoi(v);
        else if (a == "--smer-len") opt.smerLen = atoi(v);
        else if (a == "--kmer-format") opt.kmerFormat = atoi(v);
        else if (a == "--print-log") opt.printLog = atoi(v);
        else if (a == "--reduced-aa") opt.reducedAA = atoi(v);
        else if (a == "--em") opt.em = atoi(v);
        else if (a == "--krona-prelude") opt.kronaPrelude = v;
        else if (a == "--gpu") opt.device = atoi(v);
        else if (a == "--gpus") { opt.devices.clear(); for (int g = 0; g < atoi(v); ++g) opt.devices.push_back(g); }
        else if (a == "--gpu-list") { opt.devices.clear(); for (const char *q = v; *q;) { opt.devices.push_back(atoi(q)); while (*q && *q != ',') ++q; if (*q) ++q; } }
        else if (a == "--batch-reads") opt.batchReads = (size_t) strtoull(v, nullptr, 10);
        else { std::cerr << "unknown option " << a << "\n"; return usage(); }
    }
    // positional rules of classify.cpp:45-139
    const size_t want = opt.seqMode == 2 ? 5 : 4;
    if (pos.size() != want) {
        std::cerr << "Error: expected " << want << " positional arguments for --seq-mode " << opt.seqMode << ", got " << pos.size() << "\n";
        return usage();
    }
    std::vector<std::string> inputs(pos.begin(), pos.begin() + (opt.seqMode == 2 ? 2 : 1));
    const std::string dbDir = pos[want - 3], outDir = pos[want - 2], jobId = pos[want - 1];
    for (const std::string &f : inputs)
        if (!fileExists(f)) { std::cerr << "Error: " << f << " does not exist.\n"; return 1; }
    for (const char *f : {"diffIdx", "info", "split"})
        if (!fileExists(dbDir + "/" + f)) { std::cerr << "Error: " << f << " file is missing in " << dbDir << "\n"; return 1; }
    const std::string taxDir = opt.taxonomyPath.empty() ? dbDir : opt.taxonomyPath;
    if (!fileExists(taxDir + "/taxonomyDB")) { std::cerr << "Error: taxonomyDB file is missing in " << taxDir << "\n"; return 1; }
    if (!fileExists(outDir)) mkdir(outDir.c_str(), 0755);
    std::string err;
    const int rc = mbh::runClassify(inputs, dbDir, outDir, jobId, opt, err);
    if (rc != 0) std::cerr << "Error: " << err << "\n";
    return rc;
}
