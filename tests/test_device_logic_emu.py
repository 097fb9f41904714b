"""CPU checks of the per-element DEVICE logic (MB_HD functions in metabuli_work_b200/csrc/*.cuh),
instantiated for the host by tests/emu/emu.cpp, against the oracle.  These run without a GPU; the
kernels themselves (tiling, scans, atomics) are covered by the -m gpu tests."""
import ctypes as C

import numpy as np
import pytest

import helpers
import orc

CASES = ["plain_paired", "sync_paired", "legacy_paired", "plain_long", "sync_long", "plain_single"]
RES_DT = np.dtype([("classification", "<i4"), ("score", "<f4"), ("is_classified", "<i4"), ("query_length", "<i4"),
                   ("hamming_dist", "<i4"), ("taxcnt_begin", "<i4"), ("taxcnt_end", "<i4"), ("n_matches", "<i4")])


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emu():
    return helpers.emu_lib()


@pytest.mark.parametrize("case", CASES)
def test_window_extraction_logic(golden, emu, case):
    meta = golden.cases[case]
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"])
    _, s1, s2 = golden.reads(case)
    b = orc.Batch(s1, s2)
    want = orc.sort_kmers(orc.extract(p, b))
    out = np.zeros(len(want) + 1024, dtype=orc.KMER_DT)
    n = emu.emu_extract(p.kmer_format, p.syncmer, p.smer_len, p.seq_mode, _ptr(b.seqs), _ptr(b.off), _ptr(b.len), b.n, _ptr(out), len(out))
    assert n == len(want)
    got = orc.sort_kmers(out[:n].copy())
    assert np.array_equal(got, want)
    db.close()


@pytest.mark.parametrize("case", ["plain_paired", "legacy_paired", "sync_long"])
def test_hamming_tables(golden, emu, case):
    """hamming / rightEndHamming of every oracle match recomputed by the device functions."""
    meta = golden.cases[case]
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"])
    _, s1, s2 = golden.reads(case)
    k = orc.sort_kmers(orc.extract(p, orc.Batch(s1, s2)))
    m = orc.match(db, p, k)
    order = np.argsort(k["qinfo"], kind="stable")
    idx = order[np.searchsorted(k["qinfo"][order], m["qinfo"])]
    q = np.ascontiguousarray(k["value"][idx])
    t = np.ascontiguousarray(m["dnaEncoding"].astype(np.uint64))
    frame = (m["qinfo"] >> np.uint64(61)).astype(np.int64)
    rev = np.ascontiguousarray(((frame < 3) != (p.kmer_format == 2)).astype(np.uint8))
    s = np.zeros(len(m), dtype=np.uint8)
    pk = np.zeros(len(m), dtype=np.uint16)
    emu.emu_hamming(_ptr(q), _ptr(t), _ptr(rev), len(m), _ptr(s), _ptr(pk))
    assert np.array_equal(s, m["hamming"])
    assert np.array_equal(pk, m["rightEndHamming"])
    s2, pk2 = np.zeros_like(s), np.zeros_like(pk)
    emu.emu_hamming_lut(_ptr(q), _ptr(t), _ptr(rev), len(m), _ptr(s2), _ptr(pk2))
    assert np.array_equal(s2, m["hamming"])
    assert np.array_equal(pk2, m["rightEndHamming"])
    db.close()


def test_hamming_two_codon_tables_exhaustive(emu):
    """The two-codon tables of the match kernel equal the cell-by-cell functions on random 24-bit pairs, both orientations."""
    rng = np.random.default_rng(3)
    n = 200000
    q = rng.integers(0, 1 << 24, size=n, dtype=np.uint64)
    t = rng.integers(0, 1 << 24, size=n, dtype=np.uint64)
    for r in (0, 1):
        rev = np.full(n, r, dtype=np.uint8)
        s1, p1 = np.zeros(n, dtype=np.uint8), np.zeros(n, dtype=np.uint16)
        s2, p2 = np.zeros(n, dtype=np.uint8), np.zeros(n, dtype=np.uint16)
        emu.emu_hamming(_ptr(q), _ptr(t), _ptr(rev), n, _ptr(s1), _ptr(p1))
        emu.emu_hamming_lut(_ptr(q), _ptr(t), _ptr(rev), n, _ptr(s2), _ptr(p2))
        assert np.array_equal(s1, s2) and np.array_equal(p1, p2)


def test_hamming_lut_quirk(emu):
    """HAMMING_LUT7 holds 1<<14 for query codon ids 4,5 against target ids 6,7 (KmerMatcher.h:148-157)."""
    q = np.array([4 << 21, 5 << 21, 6 << 21], dtype=np.uint64)
    t = np.array([6 << 21, 7 << 21, 4 << 21], dtype=np.uint64)
    rev = np.zeros(3, dtype=np.uint8)
    s = np.zeros(3, dtype=np.uint8)
    pk = np.zeros(3, dtype=np.uint16)
    emu.emu_hamming(_ptr(q), _ptr(t), _ptr(rev), 3, _ptr(s), _ptr(pk))
    assert list(s) == [4, 4, 4]
    assert list(pk) == [1 << 14, 1 << 14, 0]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("min_sp", [0.0, 0.9])
@pytest.mark.parametrize("fast", [0, 1])
def test_score_read_logic(golden, emu, case, min_sp, fast):
    meta = golden.cases[case]
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"], min_sp_score=min_sp)
    _, s1, s2 = golden.reads(case)
    b = orc.Batch(s1, s2)
    k, m, res, ids, cnt = orc.classify(db, p, b)
    parent, depth, sp_of, sp_rank, mx, node0, euk = helpers.dense_taxonomy(db)
    out = np.zeros(b.n, dtype=RES_DT)
    tid = np.zeros(len(m) + 16, dtype=np.int32)
    tcn = np.zeros(len(m) + 16, dtype=np.int32)
    emu.emu_set_fast(fast)
    used = emu.emu_score(_ptr(m), len(m), _ptr(b.len), b.n, p.seq_mode, p.kmer_format, p.syncmer, p.smer_len, p.min_cons_cnt,
                         p.min_cons_cnt_euk, p.min_score, p.min_sp_score, p.tie_ratio, _ptr(parent), _ptr(depth), _ptr(sp_of),
                         _ptr(sp_rank), mx, node0, euk, _ptr(out), _ptr(tid), _ptr(tcn), len(tid))
    assert used >= 0
    bad = []
    for i in range(b.n):
        a, o = res[i], out[i]
        tc_a = helpers.taxcnt_string(ids, cnt, int(a["taxcnt_begin"]), int(a["taxcnt_end"]))
        tc_o = helpers.taxcnt_string(tid, tcn, int(o["taxcnt_begin"]), int(o["taxcnt_end"]))
        if not (a["classification"] == o["classification"] and a["is_classified"] == o["is_classified"]
                and np.float32(a["score"]).view(np.uint32) == np.float32(o["score"]).view(np.uint32)
                and a["query_length"] == o["query_length"] and tc_a == tc_o):
            bad.append(i)
    assert bad == []
    if fast and meta["seq_mode"] != 3:
        assert emu.emu_fast_fallbacks() * 5 < b.n        # the streaming form holds the typical short read
    emu.emu_set_fast(0)
    db.close()


@pytest.mark.parametrize("case", ["plain_paired", "sync_paired", "sync_long"])
def test_sort_key_recoded(golden, emu, case):
    """The re-coded 24-bit sort key (format 2) is monotone over ascending metamers of queries AND of the database (the
    partition of k_match relies on it), fits 24 bits and equals the numpy restatement used by the GPU tests."""
    meta = golden.cases[case]
    db = orc.Db(golden.dbdir(case))
    p = db.params(seq_mode=meta["seq_mode"])
    assert p.kmer_format == 2
    _, s1, s2 = golden.reads(case)
    k = orc.sort_kmers(orc.extract(p, orc.Batch(s1, s2)))
    dbv, _ = db.decode()
    for values in (np.ascontiguousarray(k["value"]), np.ascontiguousarray(dbv)):
        out = np.zeros(len(values), dtype=np.uint64)
        emu.emu_sort_t(1, 0, _ptr(values), len(values), _ptr(out))
        assert np.all(out[1:] >= out[:-1])
        assert int(out.max()) < (1 << 24)
        assert np.array_equal(out, helpers.kmer_sort_key(values, 21, 2))
        # it must separate at least as well as the leading 25 raw bits... and never split equal amino-acid parts
        aa = values >> np.uint64(24)
        same = aa[1:] == aa[:-1]
        assert np.all(out[1:][same] == out[:-1][same])
    # plain leading-bits mode
    v = np.ascontiguousarray(k["value"])
    out = np.zeros(len(v), dtype=np.uint64)
    emu.emu_sort_t(0, 8, _ptr(v), len(v), _ptr(out))
    assert np.array_equal(out, v >> np.uint64(32))
    db.close()


@pytest.mark.parametrize("s", [3, 4, 5, 6, 7])
def test_syncmer_predicate(emu, s):
    """is_syncmer (compile-time forms for s = 4, 5, 6 and the loop) against the definition of SyncmerScanner.h:36-102:
    the LEFTMOST minimum of the 8-s+1 amino-acid s-mers sits at offset 0 or 8-s.  Low-entropy residues force ties."""
    rng = np.random.default_rng(s)
    n = 20000
    aa = np.zeros(n, dtype=np.uint64)
    for i in range(8):
        hi = 21 if i % 2 == 0 else 3
        aa = (aa << np.uint64(5)) | rng.integers(0, hi, n).astype(np.uint64)
    aa[: n // 4] = (aa[: n // 4] & np.uint64(0x1F)) * np.uint64(0x0842108421)      # one residue repeated eight times
    values = np.ascontiguousarray((aa << np.uint64(24)) | rng.integers(0, 1 << 24, n).astype(np.uint64))
    out = np.zeros(n, dtype=np.uint8)
    emu.emu_is_syncmer(_ptr(values), n, s, _ptr(out))
    last = 8 - s
    mask = np.uint64((1 << (5 * s)) - 1)
    smers = np.stack([(aa >> np.uint64(5 * (last - i))) & mask for i in range(last + 1)], axis=1)
    arg = np.argmin(smers, axis=1)                  # numpy returns the first minimum
    want = ((arg == 0) | (arg == last)).astype(np.uint8)
    assert np.array_equal(out, want)
