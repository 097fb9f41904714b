#!/usr/bin/env python3
"""Seeded synthetic inputs for the golden fixtures (genomes, NCBI-style taxdump, reads).

Used only by tests/golden/make_golden.py / make_golden2.py, which build databases from them (oracle/synth.py) and feed
both to oracle/_ref/ref_harness (the reference's own classes).  Nothing here is product code.

Layout written under <outdir>:
  genomes/ACC_xxx.fna, fasta.list, acc2taxid.tsv, taxdump/{nodes,names,merged}.dmp,
  reads_1.fq, reads_2.fq (paired 2x150), long.fa (long reads)
"""
import argparse
import os

import numpy as np

BASES = np.frombuffer(b"ACGT", dtype=np.uint8)


def mutate(rng, seq, rate):
    """Substitute a fraction `rate` of positions with a different base."""
    out = seq.copy()
    n = len(seq)
    k = rng.binomial(n, rate)
    pos = rng.choice(n, size=k, replace=False)
    out[pos] = (out[pos] + rng.integers(1, 4, size=k)) % 4
    return out


def write_fasta(path, name, seq_idx):
    s = BASES[seq_idx].tobytes().decode()
    with open(path, "w") as f:
        f.write(f">{name}\n")
        for i in range(0, len(s), 80):
            f.write(s[i:i + 80] + "\n")


def revcomp(idx):
    # A0 C1 G2 T3 -> complement = 3 - x
    return (3 - idx)[::-1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("outdir")
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--families", type=int, default=2)
    ap.add_argument("--genera", type=int, default=2)
    ap.add_argument("--species", type=int, default=2)
    ap.add_argument("--strains", type=int, default=2, help="genomes per species (strain nodes under the species)")
    ap.add_argument("--genome-len", type=int, default=100_000)
    ap.add_argument("--pairs", type=int, default=2000)
    ap.add_argument("--long-reads", type=int, default=40)
    ap.add_argument("--euk-families", type=int, default=0,
                    help="the last N families hang under a second superkingdom named Eukaryota (Taxonomer.cpp:497-500: "
                         "minConsCntEuk applies to their species); does not change the random stream")
    args = ap.parse_args()
    rng = np.random.default_rng(args.seed)
    out = args.outdir
    os.makedirs(f"{out}/genomes", exist_ok=True)
    os.makedirs(f"{out}/taxdump", exist_ok=True)

    # ---- taxonomy: root(1) > Bacteria(2, superkingdom) > phylum(3) > class(4) > order(5) > families...
    nodes = [(1, 1, "no rank", "root"), (2, 1, "superkingdom", "Bacteria"), (3, 2, "phylum", "Synthphyla"),
             (4, 3, "class", "Synthclassia"), (5, 4, "order", "Synthorderales")]
    if args.euk_families:
        nodes += [(6, 1, "superkingdom", "Eukaryota"), (7, 6, "phylum", "Eukphyla"), (8, 7, "class", "Eukclassia"),
                  (9, 8, "order", "Eukorderales")]
    next_id = [10]

    def new_node(parent, rank, name):
        t = next_id[0]
        next_id[0] += 1
        nodes.append((t, parent, rank, name))
        return t

    genomes = []  # (accession, taxid, seq)
    root_seq = rng.integers(0, 4, size=args.genome_len, dtype=np.int64)
    for f in range(args.families):
        fam = new_node(9 if f >= args.families - args.euk_families else 5, "family", f"Famaceae{f}")
        fam_seq = mutate(rng, root_seq, 0.25)
        for g in range(args.genera):
            gen = new_node(fam, "genus", f"Genus{f}{g}")
            gen_seq = mutate(rng, fam_seq, 0.10)
            for s in range(args.species):
                sp = new_node(gen, "species", f"Genus{f}{g} species{s}")
                sp_seq = mutate(rng, gen_seq, 0.06)
                for k in range(args.strains):
                    st = new_node(sp, "no rank", f"Genus{f}{g} species{s} strain{k}")
                    st_seq = mutate(rng, sp_seq, 0.01)
                    acc = f"SYN_{len(genomes):06d}"
                    genomes.append((acc, st, st_seq))

    with open(f"{out}/taxdump/nodes.dmp", "w") as fn:
        for t, p, rank, _ in nodes:
            fn.write(f"{t}\t|\t{p}\t|\t{rank}\t|\t\t|\t0\t|\n")
    with open(f"{out}/taxdump/names.dmp", "w") as fn:
        for t, _, _, name in nodes:
            fn.write(f"{t}\t|\t{name}\t|\t\t|\tscientific name\t|\n")
    open(f"{out}/taxdump/merged.dmp", "w").close()

    with open(f"{out}/fasta.list", "w") as fl, open(f"{out}/acc2taxid.tsv", "w") as fa:
        fa.write("accession\taccession.version\ttaxid\tgi\n")
        for acc, t, seq in genomes:
            path = os.path.abspath(f"{out}/genomes/{acc}.fna")
            write_fasta(path, f"{acc}.1 synthetic genome", seq)
            fl.write(path + "\n")
            fa.write(f"{acc}\t{acc}.1\t{t}\t0\n")

    # ---- paired reads: 2x150 (some shorter / ragged), 0.5% substitutions, a few with N, lower case,
    # too-short mates and random (unclassifiable) reads
    def to_str(idx):
        return BASES[idx].tobytes().decode()

    with open(f"{out}/reads_1.fq", "w") as f1, open(f"{out}/reads_2.fq", "w") as f2:
        for i in range(args.pairs):
            kind = i % 50
            acc, t, seq = genomes[rng.integers(0, len(genomes))]
            insert = int(rng.integers(300, 500))
            start = int(rng.integers(0, len(seq) - insert))
            frag = mutate(rng, seq[start:start + insert], 0.005)
            if rng.random() < 0.5:
                frag = revcomp(frag)
            l1, l2 = 150, 150
            if kind == 7:
                l1, l2 = int(rng.integers(26, 150)), int(rng.integers(26, 150))   # ragged lengths, all L mod 3
            if kind == 11:
                l1 = int(rng.integers(1, 26))                                      # mate 1 too short
            if kind == 13:
                l2 = int(rng.integers(1, 26))                                      # mate 2 too short
            r1 = to_str(frag[:l1])
            r2 = to_str(revcomp(frag)[:l2])
            if kind == 17:                                                         # random read pair
                r1 = to_str(rng.integers(0, 4, size=150))
                r2 = to_str(rng.integers(0, 4, size=150))
            if kind == 19:                                                         # ambiguous bases
                r1 = list(r1)
                for p in rng.choice(len(r1), size=min(3, len(r1)), replace=False):
                    r1[p] = "N"
                r1 = "".join(r1)
                r2 = r2[:40] + "NNNN" + r2[44:]
            if kind == 23:                                                         # soft-masked lower case + IUPAC
                r1 = r1.lower()
                r2 = r2[:70] + "R" + r2[71:100] + "y" + r2[101:]
            name = f"read{i:06d}_{acc}"
            f1.write(f"@{name}/1\n{r1}\n+\n{'I' * len(r1)}\n")
            f2.write(f"@{name}/2\n{r2}\n+\n{'I' * len(r2)}\n")

    # ---- long reads: 2-8 kb, 5% errors (substitutions, insertions, deletions 2:1:2)
    with open(f"{out}/long.fa", "w") as fl:
        for i in range(args.long_reads):
            acc, t, seq = genomes[rng.integers(0, len(genomes))]
            L = int(rng.integers(2000, 8000))
            start = int(rng.integers(0, len(seq) - L))
            frag = seq[start:start + L]
            res = []
            for b in frag:
                r = rng.random()
                if r < 0.02:
                    res.append((b + rng.integers(1, 4)) % 4)
                elif r < 0.03:
                    res.append(b)
                    res.append(rng.integers(0, 4))
                elif r < 0.05:
                    continue
                else:
                    res.append(b)
            arr = np.array(res, dtype=np.int64)
            if rng.random() < 0.5:
                arr = revcomp(arr)
            fl.write(f">long{i:04d}_{acc}\n{to_str(arr)}\n")


if __name__ == "__main__":
    main()
