/*
 * ref_mkdb.cpp — writes DBDIR/taxonomyDB with the REFERENCE's own TaxonomyWrapper (TEST INFRASTRUCTURE).
 *
 * This file is ours; the classes it drives are compiled from /root/reference by oracle/Makefile:
 *   TaxonomyWrapper(names, nodes, merged, useInternalTaxID=true)   src/commons/TaxonomyWrapper.cpp:67-113
 *   TaxonomyWrapper::writeTaxonomyDB                               src/commons/TaxonomyWrapper.cpp:668-679
 * exactly the two calls `metabuli build` makes (src/workflow/build.cpp:88-100).  The k-mer files
 * (diffIdx/info/split) of the synthetic databases are written by oracle/synth.py through
 * orc_write_db_files and are then READ by the reference's KmerMatcher inside ref_harness, so the
 * format restatement is pinned by the reference reader.
 *
 * usage: ref_mkdb TAXDUMP_DIR DBDIR
 * writes DBDIR/taxonomyDB and DBDIR/tax_map.tsv (original taxID \t internal taxID), the map the
 * DB writer needs because `info` holds internal taxIDs.
 */
#include "TaxonomyWrapper.h"

#include <cstdio>
#include <string>
#include <unordered_map>

int main(int argc, char **argv) {
    if (argc < 3) {
        fprintf(stderr, "usage: %s TAXDUMP_DIR DBDIR\n", argv[0]);
        return 2;
    }
    std::string tax = argv[1], db = argv[2];
    TaxonomyWrapper *t = new TaxonomyWrapper(tax + "/names.dmp", tax + "/nodes.dmp", tax + "/merged.dmp", true);
    t->writeTaxonomyDB(db + "/taxonomyDB");
    std::unordered_map<TaxID, TaxID> ext2int;
    t->getExternal2internalTaxID(ext2int);
    FILE *f = fopen((db + "/tax_map.tsv").c_str(), "w");
    if (!f) return 1;
    for (auto &kv : ext2int) fprintf(f, "%d\t%d\n", kv.first, kv.second);
    fclose(f);
    return 0;
}
