"""Pins the CPU oracle (oracle/mb_oracle.cpp) to outputs of the REFERENCE itself.

tests/golden/data was produced by tests/golden/make_golden.py from oracle/_ref/ref_harness, i.e. the
reference's own KmerScanner / SyncmerScanner / OldMetamerScanner / KmerMatcher / Taxonomer classes
compiled from /root/reference.  Every stage of the oracle must reproduce it bit for bit: the sorted
query k-mer buffer, the sorted match buffer (sha256) and every per-read field incl. fp32 score bits.
"""
import numpy as np
import pytest

import helpers
import orc

CASES = ["plain_paired", "sync_paired", "legacy_paired", "plain_long", "sync_long", "plain_single",
         # round 2 (tests/golden/make_golden2.py): Eukaryota branch, redundancy-bit mask, s = 4 / 6, legacy x long reads
         "euk_paired", "euk_long", "redund_paired", "sync4_paired", "sync6_paired", "sync4_long", "legacy_long"]


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_reference(golden, case):
    meta = golden.cases[case]
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"], **({"smer_len": meta["smer_len"]} if "smer_len" in meta else {}))
    names, s1, s2 = golden.reads(case)
    assert len(names) == meta["n_reads"]
    b = orc.Batch(s1, s2)
    k, m, res, ids, cnt = orc.classify(db, p, b, threads=2)
    assert len(k) == meta["n_kmers"]
    assert helpers.digest(k) == meta["kmers_sha256"]
    assert len(m) == meta["n_matches"]
    assert helpers.digest(m) == meta["matches_sha256"]
    assert helpers.check_results_against_golden(golden.results(case), res, ids, cnt) == []
    db.close()


MASK_CASES = ["mask_paired", "mask_paired_p05", "mask_long"]


def _mates(b):
    return [bytes(b.seqs[int(o):int(o) + int(l)]).decode("latin-1") for o, l in zip(b.off, b.len)]


@pytest.mark.parametrize("case", MASK_CASES)
def test_oracle_mask_matches_reference(golden, case):
    """--mask 1 (tests/golden/make_golden3.py): the masked mates equal what the reference's
    SeqIterator::maskLowComplexityRegions wrote (SeqIterator.cpp:154-177), and the masked batch classifies like the
    reference's own KmerExtractor (with masking) + KmerMatcher + Taxonomer did."""
    meta = golden.cases[case]
    names, s1, s2 = golden.reads(case)
    b = orc.Batch(s1, s2)
    masked, probs = orc.mask(b, meta["mask_prob"], want_probs=True)
    want = open(helpers.DATA + "/" + case + ".masked.txt").read().split("\n")[:-1]
    got = [m for i, m in enumerate(_mates(masked)) if s2 is not None or i % 2 == 0]
    assert got == want
    assert sum(m.count("N") for m in got) > sum(m.count("N") for m in _mates(b))     # the case masks something
    assert float(probs.max()) <= 1.0 and float(probs.min()) >= 0.0
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"])
    k, m, res, ids, cnt = orc.classify(db, p, masked, threads=2)
    assert helpers.digest(k) == meta["kmers_sha256"] and helpers.digest(m) == meta["matches_sha256"]
    assert helpers.check_results_against_golden(golden.results(case), res, ids, cnt) == []
    db.close()


def test_mask_likelihood_ratios_are_the_reference_matrix():
    """mask_lr.bin = ref_harness --dump-lr: ProbabilityMatrix(NucleotideMatrix("nucleotide.out")) and aa2num as the
    reference computes them at run time; the oracle (and csrc/mb_mask.cuh) hold the two values and the letter map."""
    raw = open(helpers.DATA + "/mask_lr.bin", "rb").read()
    lr = np.frombuffer(raw[:200], dtype=np.float64).reshape(5, 5)
    codes = np.frombuffer(raw[200:456], dtype=np.uint8)
    assert raw[456] == 4
    match, mismatch = orc.lib().orc_mask_lr(1), orc.lib().orc_mask_lr(0)
    for i in range(5):
        for j in range(5):
            assert lr[i, j] == (match if i == j and i < 4 else mismatch)
    # letter map: mask a one-letter read with every byte value; 'N' comes back exactly for code 4
    for c in range(1, 256):
        if c in (10, 13):
            continue
        b = orc.Batch([bytes([c]) * 3])
        out = _mates(orc.mask(b, 1.0))[0]
        assert (out == "NNN") == (codes[c] == 4), c        # (byte 255 keeps code 0: the reference's mapping loop stops at 254)


def _sha(b):
    import hashlib
    return hashlib.sha256(bytes(b)).hexdigest()


@pytest.mark.parametrize("splits", [64, 4096])
def test_oracle_build_index_matches_reference(golden, tmp_path, splits):
    """DB-build inner loop (tests/golden/make_golden4.py): the files the oracle writes for the committed target records
    equal the files written by the REFERENCE's own sort + IndexCreator::filterKmers<DB_CREATION> +
    IndexCreator::writeTargetFilesAndSplits (oracle/_ref/ref_buildidx), by digest; the split table with 64 entries byte
    for byte."""
    import json
    want = json.load(open(helpers.DATA + "/build_digests.json"))
    rec = np.load(helpers.DATA + "/build_input.npz")
    assert len(rec["value"]) == want["records"]
    db = orc.Db(golden.dbdir("plain_paired"))
    n = orc.build_index(db, rec["value"], rec["taxid"], rec["species"], str(tmp_path), split_num=splits)
    db.close()
    w = want[f"splits{splits}"]
    assert n == w["db_kmers"]
    for f in ("diffIdx", "info", "split"):
        assert _sha(open(tmp_path / f, "rb").read()) == w[f + "_sha256"], f
    if splits == 64:
        assert open(tmp_path / "split", "rb").read() == open(helpers.DATA + "/build_split64.bin", "rb").read()
    if splits == 64:
        ev, et, es = helpers.build_edge_records(rec["taxid"], rec["species"])
        db = orc.Db(golden.dbdir("plain_paired"))
        d = str(tmp_path / "edge")
        assert orc.build_index(db, ev, et, es, d, split_num=64) == want["edge_splits64"]["db_kmers"]
        db.close()
        for f in ("diffIdx", "info", "split"):
            assert _sha(open(d + "/" + f, "rb").read()) == want["edge_splits64"][f + "_sha256"], ("edge", f)
    # the built files are a database: decoding gives a sorted stream of n k-mers
    import shutil
    for f in ("taxonomyDB", "taxID_list", "db.parameters"):
        shutil.copy(golden.dbdir("plain_paired") + "/" + f, tmp_path / f)
    built = orc.Db(str(tmp_path))
    v, t = built.decode()
    assert len(v) == n and np.all(v[1:] >= v[:-1])
    built.close()


def test_oracle_thread_invariance(golden):
    """KmerMatcher splits the query list per thread (KmerMatcher.cpp:166-194); the result must not depend on it."""
    case = "plain_paired"
    db = orc.Db(golden.dbdir(case))
    names, s1, s2 = golden.reads(case)
    b = orc.Batch(s1, s2)
    outs = []
    for t in (1, 3, 8):
        p = db.params(seq_mode=2, threads=t)
        k = orc.sort_kmers(orc.extract(p, b))
        outs.append(orc.sort_matches(orc.match(db, p, k)))
    assert all(np.array_equal(outs[0], o) for o in outs[1:])


def test_decode_is_sorted_and_counts(golden):
    """validatedb invariant (src/util/validateDatabase.cpp:98-130): #terminated deltas == #info entries."""
    for d in ("plain_paired", "sync_paired", "legacy_paired"):
        db = orc.Db(golden.dbdir(d))
        v, t = db.decode()
        assert len(v) == db.num_kmers()
        assert np.all(v[1:] >= v[:-1])
        db.close()
