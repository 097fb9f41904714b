"""bench_workload.py writes the bench database straight in the on-disk format; its writer must agree byte for byte
with the oracle's restatement of IndexCreator::writeTargetFilesAndSplits / getDiffIdx (which the reference reader pins)."""
import os
import sys

import numpy as np
import pytest
import torch

import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.mark.parametrize("n,split", [(50000, 64), (5000, 4096), (300000, 4096)])
def test_db_writer_matches_oracle_writer(tmp_path, n, split):
    import bench_workload as W
    rng = np.random.default_rng(n)
    aa = np.sort(rng.integers(0, 1 << 40, size=n // 3, dtype=np.uint64))
    v = np.unique((rng.choice(aa, size=n) << np.uint64(24)) | rng.integers(0, 1 << 24, size=n, dtype=np.uint64))
    v = np.sort(np.concatenate([v, v[::7]]))                      # repeated values: one entry per (value, species)
    t = rng.integers(1, 100, size=len(v)).astype(np.uint32)
    d1, d2 = str(tmp_path / "a"), str(tmp_path / "b")
    os.makedirs(d1), os.makedirs(d2)
    assert orc.lib().orc_write_db_files(d1.encode(), orc._ptr(v), orc._ptr(t), len(v), split, 2) == 0

    class Tax:
        genome_tax = np.array([5, 6])
    db = {"values": torch.from_numpy(v.view(np.int64)), "info": torch.from_numpy(t.astype(np.int64))}
    W.write_db_dir(d2, db, Tax(), split_num=split)
    for f in ("diffIdx", "info", "split"):
        assert open(os.path.join(d1, f), "rb").read() == open(os.path.join(d2, f), "rb").read(), f


def test_taxonomy_tables_follow_reference_rules():
    import bench
    import bench_workload as W
    tax = W.Taxonomy(*bench.taxonomy_shape(100))
    assert len(tax.strain) == 100 and len(tax.species) == 100
    for st in tax.strain[:10]:
        sp = tax.parent[st]
        assert tax.rank[sp] == "species" and tax.species_rank_of[st] == sp and tax.species_of[st] == sp
    assert tax.species_rank_of[tax.genus[0]] == tax.genus[0]      # above species: the taxon itself (TaxonomyWrapper.cpp:479-498)


def test_fastq_text_of_the_bench_reads():
    """bench.py's input-pipeline leg feeds mb_stage_fastq with this text: four lines per record, fixed-width names,
    the mate's 150 bases, '+', 150 quality characters; parsed back it must give the packed reads."""
    import bench
    rng = np.random.default_rng(3)
    n = 37
    bases = rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=n * 300).astype(np.uint8)
    for mate in (0, 1):
        text = bytes(bench.fastq_text(bases, n, mate))
        lines = text.split(b"\n")
        assert lines[-1] == b"" and len(lines) == 4 * n + 1
        for i in range(n):
            h, s, p, q = lines[4 * i:4 * i + 4]
            assert h == b"@r%09d" % i and p == b"+" and q == b"I" * 150
            assert s == bytes(bases[(2 * i + mate) * 150:(2 * i + mate + 1) * 150])


@pytest.mark.parametrize("L", [3000, 3001, 3002])
def test_torch_db_kmers_equal_oracle_extraction(L):
    """bench_workload.syncmer_metamers (torch, builds the bench DB without the library under test) yields exactly the
    k-mer values the reference-pinned oracle extracts from the same genome read as one long read (syncmer, format 2)."""
    import bench_workload as W
    rng = np.random.default_rng(L)
    codes = rng.integers(0, 4, size=2 * L, dtype=np.uint8)
    starts = np.array([0, L, 2 * L], dtype=np.int64)
    value, pos, g, frame = W.syncmer_metamers(torch.from_numpy(codes), starts, torch.device("cpu"))
    got = np.sort(value.numpy().view(np.uint64))
    p = orc.default_params(kmer_format=2, syncmer=1, smer_len=5, seq_mode=3)
    seqs = [W.CODE2ASCII[codes[:L]].tobytes(), W.CODE2ASCII[codes[L:]].tobytes()]
    k = orc.extract(p, orc.Batch(seqs))
    want = np.sort(k["value"])
    assert len(got) == len(want) and np.array_equal(got, want)
    assert int(g.max()) == 1 and int(frame.max()) == 5
