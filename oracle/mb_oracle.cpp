/*
 * mb_oracle.cpp — CPU restatement of Metabuli's classify hot path (TEST INFRASTRUCTURE).
 * See mb_oracle.h for the usage contract and how parity is pinned.
 * All file:line citations are relative to /root/reference.
 */
#include "mb_oracle.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#include <parallel/algorithm>
#endif

namespace {

thread_local std::string g_err;

/* ------------------------------------------------------------------------------------------
 * Nucleotide / codon tables.
 * Restates src/commons/common.cpp:13-23 (atcg / iRCT character maps) and
 * src/commons/GeneticCode.h:6,23-29,34-193 (nuc2int, nuc2aa, nuc2num, standard code).
 * Nucleotide index: A=0 C=1 T=2 G=3, 7 = not a base ('N', '.', anything unmapped).
 * ---------------------------------------------------------------------------------------- */
struct Tables {
    uint8_t nucF[256];    /* forward strand base index               */
    uint8_t nucR[256];    /* complement base index (iRCT after atcg)  */
    int8_t  aa[8][8][8];  /* amino-acid index 0..20, -1 with an N     */
    int8_t  cid[8][8][8]; /* 3-bit codon id, -1 with an N             */
    uint8_t ham[8][8];    /* KmerMatcher.h:66-70 hammingLookup        */
    uint16_t lut[8][64];  /* KmerMatcher.h:72-158 HAMMING_LUT0..7     */

    Tables() {
        for (int c = 0; c < 256; ++c) nucF[c] = 7;
        /* IUPAC collapse exactly as the atcg string does: the listed letters map to one base */
        const char *letters = "ABCDGHKMNRSTUWY";
        const char *mapsTo  = "AGCGGTGCNACTGAT";
        for (int i = 0; letters[i]; ++i) {
            uint8_t idx;
            switch (mapsTo[i]) {
                case 'A': idx = 0; break;
                case 'C': idx = 1; break;
                case 'T': idx = 2; break;
                case 'G': idx = 3; break;
                default:  idx = 7; break;
            }
            nucF[(unsigned char) letters[i]] = idx;
            nucF[(unsigned char) (letters[i] | 0x20)] = idx;
        }
        for (int c = 0; c < 256; ++c) {
            uint8_t f = nucF[c];
            nucR[c] = (f == 7) ? 7 : (uint8_t) (f ^ 2); /* A<->T (0<->2), C<->G (1<->3) */
        }
        /* standard genetic code; amino-acid alphabet "ARNDCQEGHILKMFPSTWYVX" (X = stop) */
        const char *alphabet = "ARNDCQEGHILKMFPSTWYVX";
        /* codon table in T,C,A,G order for each of the three positions */
        const char *tcag = "FFLLSSSSYYXXCCXWLLLLPPPPHHQQRRRRIIIMTTTTNNKKSSRRVVVVAAAADDEEGGGG";
        const int tcag2idx[4] = {2, 1, 0, 3}; /* T,C,A,G -> our base index */
        memset(aa, 0, sizeof(aa));
        memset(cid, 0, sizeof(cid));
        for (int a = 0; a < 4; ++a)
            for (int b = 0; b < 4; ++b)
                for (int c = 0; c < 4; ++c) {
                    char letter = tcag[a * 16 + b * 4 + c];
                    int x = tcag2idx[a], y = tcag2idx[b], z = tcag2idx[c];
                    aa[x][y][z] = (int8_t) (strchr(alphabet, letter) - alphabet);
                    cid[x][y][z] = (int8_t) z; /* default codon id = third base */
                }
        /* six-fold degenerate amino acids and the stop codons get ids 4..7 (GeneticCode.h:168-180) */
        cid[0][3][3] = 4; cid[0][3][0] = 5;   /* Arg AGG, AGA */
        cid[2][2][3] = 4; cid[2][2][0] = 5;   /* Leu TTG, TTA */
        cid[0][3][2] = 6; cid[0][3][1] = 7;   /* Ser AGT, AGC */
        cid[2][3][0] = 5;                     /* stop TGA     */
        for (int i = 0; i < 8; ++i)
            for (int j = 0; j < 8; ++j) {
                aa[7][i][j] = aa[i][7][j] = aa[i][j][7] = -1;
                cid[7][i][j] = cid[i][7][j] = cid[i][j][7] = -1;
            }
        /* Hamming distance between codon ids of the same amino acid (KmerMatcher.h:66-70) */
        static const uint8_t H[8][8] = {
            {0, 1, 1, 1, 2, 1, 3, 3}, {1, 0, 1, 1, 2, 2, 3, 2}, {1, 1, 0, 1, 2, 2, 2, 3},
            {1, 1, 1, 0, 1, 2, 3, 3}, {2, 2, 2, 1, 0, 1, 4, 4}, {1, 2, 2, 2, 1, 0, 4, 4},
            {3, 3, 2, 3, 4, 4, 0, 1}, {3, 2, 3, 3, 4, 4, 1, 0}};
        memcpy(ham, H, sizeof(H));
        /* packed 2-bit tables: entry = (ham & 3) << 2k; cells holding 4 pack to 0, except that
         * HAMMING_LUT7 rows 4,5 x cols 6,7 hold 1<<14 (KmerMatcher.h:148-157). */
        for (int k = 0; k < 8; ++k)
            for (int i = 0; i < 8; ++i)
                for (int j = 0; j < 8; ++j) {
                    uint16_t v = (uint16_t) ((H[i][j] & 3u) << (2 * k));
                    if (k == 7 && (i == 4 || i == 5) && (j == 6 || j == 7)) v = 16384;
                    lut[k][i * 8 + j] = v;
                }
    }
};
const Tables T;

inline uint64_t make_qinfo(uint32_t seqId, uint32_t pos, uint32_t frame) {
    return (uint64_t) pos | ((uint64_t) (seqId & 0x1FFFFFFFu) << 32) | ((uint64_t) (frame & 7u) << 61);
}
inline uint32_t qi_pos(uint64_t q) { return (uint32_t) q; }
inline uint32_t qi_seq(uint64_t q) { return (uint32_t) ((q >> 32) & 0x1FFFFFFFu); }
inline uint32_t qi_frame(uint64_t q) { return (uint32_t) (q >> 61); }

/* ------------------------------------------------------------------------------------------
 * Taxonomy (F4): src/commons/TaxonomyWrapper.cpp:363-421 (unserialize),
 * lib/mmseqs/src/taxonomy/NcbiTaxonomy.cpp:261-333,415-432 (LCA, IsAncestor, nodeExists).
 * ---------------------------------------------------------------------------------------- */
struct TaxNode {          /* lib/mmseqs/src/taxonomy/NcbiTaxonomy.h:13-25, 32 bytes */
    int32_t id;
    int32_t taxId;
    int32_t parentTaxId;
    int32_t pad;
    uint64_t rankIdx;
    uint64_t nameIdx;
};
static_assert(sizeof(TaxNode) == 32, "TaxonNode layout");

/* lib/mmseqs/src/commons/MathUtil.h:107-119 — float log2 approximation used for the table width */
float flog2_ref(float x) {
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

struct Taxonomy {
    std::vector<char> blob;
    size_t maxNodes = 0;
    int32_t maxTaxID = 0;
    bool internalIds = false;
    const TaxNode *nodes = nullptr;
    const int32_t *D = nullptr;
    const int32_t *internal2org = nullptr;
    const int32_t *E = nullptr, *L = nullptr, *H = nullptr, *M = nullptr;
    size_t matrixK = 0;
    const char *strData = nullptr;
    const uint32_t *strOffsets = nullptr;
    size_t strCount = 0;
    int32_t eukaryota = 0;

    bool load(const std::string &path) {
        FILE *f = fopen(path.c_str(), "rb");
        if (!f) { g_err = "cannot open " + path; return false; }
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        blob.resize((size_t) sz);
        if (fread(blob.data(), 1, (size_t) sz, f) != (size_t) sz) { fclose(f); g_err = "short read " + path; return false; }
        fclose(f);
        const char *p = blob.data();
        int32_t version; memcpy(&version, p, 4); p += 4;
        if (version != 2) { g_err = "taxonomyDB version != 2"; return false; }
        uint64_t flag; memcpy(&flag, p, 8);
        if (flag == 1) { internalIds = true; p += 8; }
        uint64_t mn; memcpy(&mn, p, 8); p += 8; maxNodes = mn;
        memcpy(&maxTaxID, p, 4); p += 4;
        nodes = (const TaxNode *) p; p += maxNodes * sizeof(TaxNode);
        D = (const int32_t *) p; p += (size_t) (maxTaxID + 1) * 4;
        internal2org = (const int32_t *) p;
        if (internalIds) p += (size_t) (maxTaxID + 1) * 4;
        E = (const int32_t *) p; p += maxNodes * 2 * 4;
        L = (const int32_t *) p; p += maxNodes * 2 * 4;
        H = (const int32_t *) p; p += maxNodes * 4;
        size_t dim = maxNodes * 2;
        matrixK = (size_t) ((int) flog2_ref((float) dim)) + 1;
        M = (const int32_t *) p; p += dim * matrixK * 4;
        /* StringBlock<unsigned int>: lib/mmseqs/src/commons/StringBlock.h:116-145 */
        uint64_t byteCap; memcpy(&byteCap, p, 8); p += 8;
        uint32_t entryCap, entryCnt; memcpy(&entryCap, p, 4); p += 4; memcpy(&entryCnt, p, 4); p += 4;
        strData = p; p += byteCap;
        strOffsets = (const uint32_t *) p;
        strCount = entryCnt;
        (void) entryCap;
        /* TaxonomyWrapper.h:90-100 setEukaryoteTaxID */
        eukaryota = 0;
        for (size_t i = 0; i < maxNodes; ++i) {
            if (nodes[i].nameIdx == 0) continue;
            if (strcmp(str(nodes[i].nameIdx), "Eukaryota") == 0) { eukaryota = nodes[i].taxId; break; }
        }
        return true;
    }
    const char *str(size_t idx) const { return strData + strOffsets[idx]; }
    bool exists(int32_t t) const { return t >= 0 && t <= maxTaxID && D[t] != -1; }
    const TaxNode *node(int32_t t) const { return (t == 0 || !exists(t)) ? nullptr : &nodes[D[t]]; }
    int32_t original(int32_t t) const { return internalIds ? internal2org[t] : t; }

    /* NcbiTaxonomy.cpp:243-259 RangeMinimumQuery over the sparse table, :261-280 lcaHelper */
    int rmq(int i, int j) const {
        int k = (int) flog2_ref((float) (j - i + 1));
        int A = M[(size_t) i * matrixK + k];
        int B = M[(size_t) (j - (1 << k) + 1) * matrixK + k];
        return (L[A] <= L[B]) ? A : B;
    }
    int lcaNode(int i, int j) const {
        if (i == 0 || j == 0) return 0;
        if (i == j) return i;
        int v1 = H[i], v2 = H[j];
        if (v1 > v2) std::swap(v1, v2);
        return E[rmq(v1, v2)];
    }
    int32_t lca(int32_t a, int32_t b) const {      /* NcbiTaxonomy.cpp:303-310 */
        if (!exists(a)) return b;
        if (!exists(b)) return a;
        return nodes[lcaNode(D[a], D[b])].taxId;
    }
    int32_t lcaMany(const std::vector<int32_t> &v) const { /* NcbiTaxonomy.cpp:313-333 */
        size_t i = 0;
        while (i < v.size() && !exists(v[i])) ++i;
        if (i == v.size()) return 0;
        int red = D[v[i++]];
        for (; i < v.size(); ++i)
            if (exists(v[i])) red = lcaNode(red, D[v[i]]);
        return nodes[red].taxId;
    }
    bool isAncestor(int32_t anc, int32_t child) const { /* NcbiTaxonomy.cpp:282-300 */
        if (anc == child) return true;
        if (anc == 0 || child == 0) return false;
        if (!exists(child) || !exists(anc)) return false;
        return lcaNode(D[child], D[anc]) == D[anc];
    }
    static int rankIndex(const char *r) {           /* NcbiTaxonomy.h:52-80 */
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
    int32_t atRank(int32_t t, const char *rank) const { /* TaxonomyWrapper.cpp:479-498 */
        if (t == 0 || !exists(t) || t == 1) return 0;
        int want = rankIndex(rank);
        const TaxNode *cur = &nodes[D[t]];
        int cnt = 0;
        while (cnt < 30 && rankIndex(str(cur->rankIdx)) < want) {
            cur = &nodes[D[cur->parentTaxId]];
            ++cnt;
        }
        if (cnt == 30) return t;
        return cur->taxId;
    }
};

} // namespace

/* ------------------------------------------------------------------------------------------
 * Database handle
 * ---------------------------------------------------------------------------------------- */
struct orc_db {
    std::string dir;
    std::vector<uint16_t> diffIdx;
    std::vector<uint32_t> info;
    std::vector<orc_split> splits;
    Taxonomy tax;
    std::unordered_map<int32_t, int32_t> taxId2speciesId;
    /* db.parameters */
    int p_reduced = 0, p_accession = -1, p_skipRedundancy = 0, p_syncmer = 0, p_kmerFormat = -1;
    bool hasParams = false;
};

namespace {

template <class Tp> bool slurp(const std::string &path, std::vector<Tp> &out) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { g_err = "cannot open " + path; return false; }
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize((size_t) sz / sizeof(Tp));
    size_t got = fread(out.data(), sizeof(Tp), out.size(), f);
    fclose(f);
    if (got != out.size()) { g_err = "short read " + path; return false; }
    return true;
}

/* KmerMatcher.cpp:56-120 loadTaxIdList (the contamList=="" branch) */
void load_taxid_list(orc_db *db) {
    std::ifstream in(db->dir + "/taxID_list");
    std::string line;
    const Taxonomy &tx = db->tax;
    while (std::getline(in, line)) {
        if (line.empty()) continue;
        int32_t taxId = (int32_t) std::stoul(line);
        int32_t sp = tx.atRank(taxId, "species");
        const TaxNode *n = tx.node(taxId);
        if (!n) continue; /* the reference would dereference NULL here */
        if (taxId != n->taxId) db->taxId2speciesId[taxId] = sp;
        int guard = 0;
        while (n->taxId != sp && guard++ < 1000) {
            db->taxId2speciesId[n->taxId] = sp;
            const TaxNode *pn = tx.node(n->parentTaxId);
            if (!pn || pn == n) break;
            n = pn;
        }
        db->taxId2speciesId[sp] = sp;
    }
}

} // namespace

extern "C" {

const char *orc_last_error(void) { return g_err.c_str(); }

void orc_default_params(orc_params *p) { /* src/workflow/classify.cpp:10-37 */
    p->kmer_format = 1; p->syncmer = 0; p->smer_len = 5; p->seq_mode = 2;
    p->min_cons_cnt = 4; p->min_cons_cnt_euk = 9; p->min_score = 0.f; p->min_sp_score = 0.f;
    p->tie_ratio = 0.95f; p->skip_redundancy = 0; p->accession_level = 0; p->threads = 1;
}

orc_db *orc_db_open(const char *dbdir) {
    orc_db *db = new orc_db();
    db->dir = dbdir;
    if (!slurp(db->dir + "/diffIdx", db->diffIdx) || !slurp(db->dir + "/info", db->info) ||
        !slurp(db->dir + "/split", db->splits) || !db->tax.load(db->dir + "/taxonomyDB")) {
        delete db;
        return nullptr;
    }
    load_taxid_list(db);
    std::ifstream pf(db->dir + "/db.parameters"); /* src/commons/common.cpp:88-133 */
    if (pf.is_open()) {
        db->hasParams = true;
        std::string line;
        while (std::getline(pf, line)) {
            size_t tab = line.find('\t');
            if (tab == std::string::npos) continue;
            std::string k = line.substr(0, tab), v = line.substr(tab + 1);
            if (k == "Reduced_alphabet") db->p_reduced = atoi(v.c_str());
            else if (k == "Accession_level") db->p_accession = atoi(v.c_str());
            else if (k == "Skip_redundancy") { if (v == "1") db->p_skipRedundancy = 1; }
            else if (k == "Syncmer") { if (v == "1") db->p_syncmer = 1; }
            else if (k == "Kmer_format") db->p_kmerFormat = atoi(v.c_str());
            /* "S-mer_len" is never written by the builder (it writes "Syncmer_len"), so
             * --smer-len stays at its command-line value */
            else if (k == "S-mer_len") { /* handled by caller if ever present */ }
        }
    }
    return db;
}

void orc_db_close(orc_db *db) { delete db; }

void orc_db_apply_parameters(const orc_db *db, orc_params *p) {
    if (!db->hasParams) return;
    if (db->p_skipRedundancy) p->skip_redundancy = 1;
    if (db->p_syncmer && !p->syncmer) p->syncmer = 1;
    if (db->p_kmerFormat > 0) p->kmer_format = db->p_kmerFormat;
    if (db->p_accession == 0 && p->accession_level == 1) p->accession_level = 0;
    if (db->p_accession == 1 && p->accession_level == 0) p->accession_level = 2;
}

size_t orc_db_num_kmers(const orc_db *db) { return db->info.size(); }
size_t orc_db_num_words(const orc_db *db) { return db->diffIdx.size(); }

/* KmerMatcher.h:282-297 getNextTargetKmer applied over the whole stream */
size_t orc_db_decode(const orc_db *db, uint64_t *values, uint32_t *infos, size_t cap) {
    uint64_t cur = 0, diff = 0;
    size_t n = 0;
    for (size_t w = 0; w < db->diffIdx.size(); ++w) {
        uint16_t frag = db->diffIdx[w];
        if (frag & 0x8000u) {
            diff |= (frag & 0x7FFFu);
            cur += diff;
            diff = 0;
            if (n < cap) {
                values[n] = cur;
                if (infos) infos[n] = n < db->info.size() ? db->info[n] : 0;
            }
            ++n;
        } else {
            diff |= frag;
            diff <<= 15;
        }
    }
    return n;
}

int32_t orc_tax_max_taxid(const orc_db *db) { return db->tax.maxTaxID; }
int32_t orc_tax_lca(const orc_db *db, int32_t a, int32_t b) { return db->tax.lca(a, b); }
int32_t orc_tax_parent(const orc_db *db, int32_t t) { const TaxNode *n = db->tax.node(t); return n ? n->parentTaxId : 0; }
int32_t orc_tax_species_of(const orc_db *db, int32_t t) {
    auto it = db->taxId2speciesId.find(t);
    return it == db->taxId2speciesId.end() ? 0 : it->second;
}
int32_t orc_tax_original(const orc_db *db, int32_t t) { return db->tax.original(t); }
const char *orc_tax_rank(const orc_db *db, int32_t t) { const TaxNode *n = db->tax.node(t); return n ? db->tax.str(n->rankIdx) : ""; }
const char *orc_tax_name(const orc_db *db, int32_t t) { const TaxNode *n = db->tax.node(t); return n ? db->tax.str(n->nameIdx) : ""; }
int32_t orc_tax_is_ancestor(const orc_db *db, int32_t anc, int32_t child) { return db->tax.isAncestor(anc, child) ? 1 : 0; }
int32_t orc_tax_eukaryota(const orc_db *db) { return db->tax.eukaryota; }
int32_t orc_tax_at_rank(const orc_db *db, int32_t t, const char *rank) { return db->tax.atRank(t, rank); }

} // extern "C"

/* The remaining stages live in separate files of the same translation unit. */
#include "orc_extract.inc"
#include "orc_match.inc"
#include "orc_score.inc"
#include "orc_io.inc"
#include "orc_mask.inc"
#include "orc_build.inc"
