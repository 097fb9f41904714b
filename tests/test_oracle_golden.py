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
