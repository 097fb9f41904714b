"""Synthetic database builder — TEST INFRASTRUCTURE ONLY (never imported by metabuli_work_b200/).

Builds a Metabuli database directory (diffIdx, info, split, taxonomyDB, taxID_list, db.parameters)
from the genomes + taxdump written by tests/golden/make_inputs.py, WITHOUT `metabuli build`
(its Prodigal gene finder needs a training blob that is missing from the reference tree and the
reference's cmake build system is off limits):

  * taxonomyDB is written by the reference's own TaxonomyWrapper (oracle/_ref/ref_mkdb);
  * target k-mers come from the oracle's scanner restatement applied to pseudo-genes (one strand
    and frame per ~1.5-kb block of each genome, which is the density a Prodigal-driven build
    gives: about one metamer per codon of the genome);
  * (value, species) de-duplication with LCA of the tax IDs follows IndexCreator.h:546-578 and
    the sort order Kmer.h:77-87;
  * the files are written by orc_write_db_files (IndexCreator.cpp:811-886) and db.parameters as
    in IndexCreator.cpp:1245-1266.

The reference reader (KmerMatcher inside oracle/_ref/ref_harness) then consumes these files, which
pins the format restatement.
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import orc  # noqa: E402

REF_MKDB = os.path.join(HERE, "_ref", "ref_mkdb")


def read_fasta_seq(path):
    with open(path, "rb") as f:
        lines = f.read().split(b"\n")
    return b"".join(l.strip() for l in lines if l and not l.startswith(b">"))


def load_nodes(taxdump):
    """nodes.dmp -> {taxid: (parent, rank)} in ORIGINAL ids."""
    nodes = {}
    with open(os.path.join(taxdump, "nodes.dmp")) as f:
        for line in f:
            c = line.split("\t|\t")
            nodes[int(c[0])] = (int(c[1]), c[2].strip())
    return nodes


def target_kmers(p, seq, block=1500, seed=0):
    """All six frames through the oracle scanner, then one (strand, frame) kept per block."""
    b = orc.Batch([seq])
    pp = orc.default_params(kmer_format=p.kmer_format, syncmer=p.syncmer, smer_len=p.smer_len, seq_mode=3)
    k = orc.extract(pp, b)
    pos = (k["qinfo"] & np.uint64(0xFFFFFFFF)).astype(np.int64)
    frame = (k["qinfo"] >> np.uint64(61)).astype(np.int64)
    blk = pos // block
    rng = np.random.default_rng(seed)
    chosen = rng.integers(0, 6, size=int(blk.max()) + 1 if len(blk) else 1)
    keep = chosen[blk] == frame
    return k["value"][keep]


def target_records(indir, dbdir, syncmer=0, kmer_format=2, smer_len=5, seed=7, max_genomes=None):
    """The raw target k-mers of the genomes under indir as IndexCreator's buffer holds them before the sort: value, external
    tax id, internal tax id, internal species id per k-mer (+ the genomes' tax ids, the external->internal map and an lca
    function over external ids).  dbdir receives taxonomyDB / tax_map.tsv (ref_mkdb)."""
    os.makedirs(dbdir, exist_ok=True)
    subprocess.check_call([REF_MKDB, os.path.join(indir, "taxdump"), dbdir], stdout=subprocess.DEVNULL)
    ext2int = {}
    with open(os.path.join(dbdir, "tax_map.tsv")) as f:
        for line in f:
            a, b = line.split()
            ext2int[int(a)] = int(b)
    nodes = load_nodes(os.path.join(indir, "taxdump"))

    def lineage(t):
        out = [t]
        while nodes[t][0] != t:
            t = nodes[t][0]
            out.append(t)
        return out

    def species_of(t):
        for x in lineage(t):
            if nodes[x][1] == "species":
                return x
        return t

    def lca(a, b):
        la = lineage(a)
        sb = set(lineage(b))
        for x in la:
            if x in sb:
                return x
        return 1

    p = orc.default_params(kmer_format=kmer_format, syncmer=syncmer, smer_len=smer_len)
    acc2tax = {}
    with open(os.path.join(indir, "acc2taxid.tsv")) as f:
        next(f)
        for line in f:
            c = line.split("\t")
            acc2tax[c[0]] = int(c[2])
    vals, taxs, sps = [], [], []
    genome_tax = []
    with open(os.path.join(indir, "fasta.list")) as f:
        paths = [l.strip() for l in f if l.strip()]
    if max_genomes:
        paths = paths[:max_genomes]
    for gi, path in enumerate(paths):
        acc = os.path.basename(path).split(".")[0]
        t = acc2tax[acc]
        genome_tax.append(t)
        v = target_kmers(p, read_fasta_seq(path), seed=seed * 1000003 + gi)
        vals.append(v)
        taxs.append(np.full(len(v), t, dtype=np.int64))
        sps.append(np.full(len(v), species_of(t), dtype=np.int64))
    v = np.concatenate(vals)
    t = np.concatenate(taxs)
    s = np.concatenate(sps)
    # internal ids decide the order (Kmer.h:77-87 compares internal species / tax ids)
    to_int = np.vectorize(lambda x: ext2int[x])
    uniq_t = np.unique(t)
    t_int = dict(zip(uniq_t.tolist(), to_int(uniq_t).tolist()))
    uniq_s = np.unique(s)
    s_int = dict(zip(uniq_s.tolist(), to_int(uniq_s).tolist()))
    ti = np.array([t_int[x] for x in uniq_t])[np.searchsorted(uniq_t, t)]
    si = np.array([s_int[x] for x in uniq_s])[np.searchsorted(uniq_s, s)]
    return v, t, ti, si, genome_tax, ext2int, lca


def build_db(indir, dbdir, syncmer=0, kmer_format=2, smer_len=5, split_num=4096, seed=7, skip_redundancy=1):
    """indir: output of make_inputs.py. Returns number of DB k-mers."""
    v, t, ti, si, genome_tax, ext2int, lca = target_records(indir, dbdir, syncmer, kmer_format, smer_len, seed)
    order = np.lexsort((ti, si, v))
    v, t, ti, si = v[order], t[order], ti[order], si[order]
    # group by (value, species): one entry each, tax id = LCA of the group's tax ids
    new_grp = np.ones(len(v), dtype=bool)
    new_grp[1:] = (v[1:] != v[:-1]) | (si[1:] != si[:-1])
    starts = np.flatnonzero(new_grp)
    ends = np.append(starts[1:], len(v))
    out_v = v[starts]
    out_t = ti[starts].copy()
    multi = np.flatnonzero((ends - starts) > 1)
    for g in multi:
        a = int(t[starts[g]])
        for j in range(starts[g] + 1, ends[g]):
            if int(t[j]) != a:
                a = lca(a, int(t[j]))
        out_t[g] = ext2int[a]
    # final order (value, species, taxid): groups are already unique per (value, species)
    out_v = np.ascontiguousarray(out_v, dtype=np.uint64)
    out_t = np.ascontiguousarray(out_t, dtype=np.uint32)
    if not skip_redundancy:
        # old databases (Skip_redundancy 0) flag redundant k-mers in bit 31 of `info`; the matcher masks it off
        # (KmerMatcher.cpp:204-205,381).  Every third entry is flagged here.
        out_t[::3] |= np.uint32(1 << 31)
    rc = orc.lib().orc_write_db_files(dbdir.encode(), orc._ptr(out_v), orc._ptr(out_t), len(out_v), split_num, kmer_format)
    assert rc == 0
    with open(os.path.join(dbdir, "taxID_list"), "w") as f:
        for x in sorted({ext2int[x] for x in genome_tax}):
            f.write(f"{x}\n")
    with open(os.path.join(dbdir, "db.parameters"), "w") as f:
        f.write("DB_name\tsynthetic\nCreation_date\t2026-01-01\n")
        f.write("Metabuli commit used to create the DB\tsynthetic\n")
        f.write("Reduced_alphabet\t0\nAccession_level\t0\nMask_mode\t0\nMask_prob\t0.900000\n")
        f.write(f"Skip_redundancy\t{1 if skip_redundancy else 0}\n")
        f.write(f"Syncmer\t{syncmer}\n")
        if syncmer:
            f.write(f"Syncmer_len\t{smer_len}\n")
        if kmer_format != 1:  # legacy DBs carry no Kmer_format line (classify.cpp:13, updateDB.cpp:16)
            f.write(f"Kmer_format\t{kmer_format}\n")
    return len(out_v)


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("indir")
    ap.add_argument("dbdir")
    ap.add_argument("--syncmer", type=int, default=0)
    ap.add_argument("--kmer-format", type=int, default=2)
    ap.add_argument("--split-num", type=int, default=4096)
    a = ap.parse_args()
    n = build_db(a.indir, a.dbdir, a.syncmer, a.kmer_format, split_num=a.split_num)
    print(f"{a.dbdir}: {n} k-mers")
