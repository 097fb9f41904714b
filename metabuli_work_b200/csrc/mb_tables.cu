// mb_tables.cu — host-side construction of the translation / Hamming tables (uploaded once per context).
#include "mb_common.cuh"

#include <string.h>

namespace mb {

void build_tables(Tables &t) {
    // ASCII -> base code.  The reference first maps a character through the `atcg` string
    // (src/commons/common.cpp:13-23: IUPAC letters collapse onto one base, everything else is '.')
    // and then takes (c & 14) >> 1 (src/commons/GeneticCode.h:6): A0 C1 T2 G3, '.' and 'N' -> 7.
    memset(t.nuc, 4, sizeof(t.nuc));
    const char *from = "ABCDGHKMRSTUWY";
    const char *to   = "AGCGGTGCACTGAT";
    for (int i = 0; from[i]; ++i) {
        u8 code = (u8) (((unsigned) to[i] & 14u) >> 1);
        t.nuc[(unsigned char) from[i]] = code;
        t.nuc[(unsigned char) (from[i] | 0x20)] = code;
    }
    // Standard genetic code over the alphabet ARNDCQEGHILKMFPSTWYVX (X = stop), GeneticCode.h:34-193.
    // aaOf[b0][b1][b2] with our base codes; written as the usual TCAG-ordered table.
    const char *alphabet = "ARNDCQEGHILKMFPSTWYVX";
    const char *tcag = "FFLLSSSSYYXXCCXWLLLLPPPPHHQQRRRRIIIMTTTTNNKKSSRRVVVVAAAADDEEGGGG";
    const int tcag2code[4] = {2, 1, 0, 3};
    u8 aaOf[4][4][4], cidOf[4][4][4];
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b)
            for (int c = 0; c < 4; ++c) {
                const char letter = tcag[a * 16 + b * 4 + c];
                const int x = tcag2code[a], y = tcag2code[b], z = tcag2code[c];
                aaOf[x][y][z] = (u8) (strchr(alphabet, letter) - alphabet);
                cidOf[x][y][z] = (u8) z;         // codon id defaults to the third base
            }
    // Amino acids with six codons and the stops use ids 4..7 for the extra codons (GeneticCode.h:40-180)
    cidOf[0][3][3] = 4; cidOf[0][3][0] = 5;      // Arg AGG, AGA
    cidOf[2][2][3] = 4; cidOf[2][2][0] = 5;      // Leu TTG, TTA
    cidOf[0][3][2] = 6; cidOf[0][3][1] = 7;      // Ser AGT, AGC
    cidOf[2][3][0] = 5;                          // stop TGA
    for (int i = 0; i < 64; ++i) {
        const int b0 = i & 3, b1 = (i >> 2) & 3, b2 = (i >> 4) & 3;
        t.codonF[i] = (u8) (aaOf[b0][b1][b2] | (cidOf[b0][b1][b2] << 5));
        const int r0 = b2 ^ 2, r1 = b1 ^ 2, r2 = b0 ^ 2; // complement: A<->T, C<->G is code ^ 2
        t.codonR[i] = (u8) (aaOf[r0][r1][r2] | (cidOf[r0][r1][r2] << 5));
    }
    // hammingLookup (KmerMatcher.h:66-70)
    static const u8 H[8][8] = {
        {0, 1, 1, 1, 2, 1, 3, 3}, {1, 0, 1, 1, 2, 2, 3, 2}, {1, 1, 0, 1, 2, 2, 2, 3}, {1, 1, 1, 0, 1, 2, 3, 3},
        {2, 2, 2, 1, 0, 1, 4, 4}, {1, 2, 2, 2, 1, 0, 4, 4}, {3, 3, 2, 3, 4, 4, 0, 1}, {3, 2, 3, 3, 4, 4, 1, 0}};
    for (int w = 0; w < 4; ++w) t.ham4[w] = 0;
    for (int a = 0; a < 8; ++a)
        for (int b = 0; b < 8; ++b) {
            const int idx = a * 8 + b;
            t.ham4[idx >> 4] |= (u64) H[a][b] << ((idx & 15) * 4);
        }
    for (int qp = 0; qp < 64; ++qp)
        for (int tp = 0; tp < 64; ++tp) {
            const int c0 = H[qp & 7][tp & 7], c1 = H[qp >> 3][tp >> 3];
            const int v0 = c0 & 3, v1 = c1 & 3;
            t.lut2[qp << 6 | tp] = (u8) (c0 + c1);
            t.lutp[qp << 6 | tp] = (u8) ((v0 | v1 << 2) | ((v1 | v0 << 2) << 4));
        }
}

} // namespace mb
