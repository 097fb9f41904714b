"""GPU parity tests (run on the B200 box with -m gpu): every stage of the CUDA path, called through the
C ABI, against the oracle on the same inputs, and against the reference-generated golden fixtures."""
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

import helpers
import orc

pytestmark = pytest.mark.gpu

CASES = ["plain_paired", "sync_paired", "legacy_paired", "plain_long", "sync_long", "plain_single",
         # round 2 (tests/golden/make_golden2.py): Eukaryota branch, redundancy-bit mask, s = 4 / 6, legacy x long reads
         "euk_paired", "euk_long", "redund_paired", "sync4_paired", "sync6_paired", "sync4_long", "legacy_long"]


@pytest.fixture(scope="module")
def mb():
    import metabuli_work_b200.binding as b
    return b


def canon_k(k):
    return orc.sort_kmers(np.ascontiguousarray(k).view(orc.KMER_DT).copy())


def as_orc_matches(m):
    out = np.zeros(len(m), dtype=orc.MATCH_DT)
    for f in ("qinfo", "targetId", "speciesId", "dnaEncoding", "rightEndHamming", "hamming"):
        out[f] = m[f]
    return out


def compare_results(res_gpu, ids_gpu, cnt_gpu, res_o, ids_o, cnt_o):
    bad = []
    for i in range(len(res_o)):
        a, o = res_o[i], res_gpu[i]
        tc_a = helpers.taxcnt_string(ids_o, cnt_o, int(a["taxcnt_begin"]), int(a["taxcnt_end"]))
        tc_o = helpers.taxcnt_string(ids_gpu, cnt_gpu, int(o["taxcnt_begin"]), int(o["taxcnt_end"]))
        if not (a["classification"] == o["classification"] and a["is_classified"] == o["is_classified"]
                and np.float32(a["score"]).view(np.uint32) == np.float32(o["score"]).view(np.uint32)
                and a["query_length"] == o["query_length"] and tc_a == tc_o):
            bad.append(i)
    return bad


def run_all_stages(mb, dbdir, seq_mode, s1, s2, sort_bits=0, smer_len=0, **kw):
    """Returns dict of GPU intermediates and the oracle's, after asserting stage-by-stage parity."""
    db = orc.Db(dbdir)
    if smer_len:
        kw["smer_len"] = smer_len
    p = db.params(seq_mode=seq_mode, **kw)
    batch = orc.Batch(s1, s2)
    ok, om, ores, oids, ocnt = orc.classify(db, p, batch, threads=4)
    clf = mb.Classifier(dbdir, seq_mode=seq_mode, min_score=kw.get("min_score", 0.0), min_sp_score=kw.get("min_sp_score", 0.0))
    try:
        if sort_bits:
            clf.set_params(kmer_sort_bits=sort_bits)
        if smer_len:
            clf.set_params(smer_len=smer_len)            # --smer-len (the DB's Syncmer_len line is never read: common.cpp:97-127)
        # K3 tap: decode of the resident stream
        v, _ = db.decode()
        assert np.array_equal(clf.decode_db(), v)
        bases, offsets, lens = mb.pack_reads(s1, s2)
        clf.stage(bases, offsets, lens)
        n = clf.extract()
        assert n == len(ok)
        assert np.array_equal(canon_k(clf.kmers()), ok)                      # K1
        clf.sort_kmers()
        ks = clf.kmers()
        t = helpers.kmer_sort_key(ks["value"], sort_bits, int(p.kmer_format))
        assert np.all(t[1:] >= t[:-1])                                        # K2: ordered by the sort's key (leading AA bits / re-coded)
        assert np.array_equal(canon_k(ks), ok)
        nm = clf.match()
        assert nm == len(om)
        gm = orc.sort_matches(as_orc_matches(clf.matches()))
        assert np.array_equal(gm, om)                                         # K3+K4
        clf.sort_matches()
        assert np.array_equal(as_orc_matches(clf.matches()), om)              # K5: exact compareMatches order
        used = clf.score()
        res, ids, cnt = clf.fetch(used)
        assert compare_results(res, ids, cnt, ores, oids, ocnt) == []         # K6
        # the one-call path with host buffers
        res2, ids2, cnt2 = clf.classify(bases, offsets, lens)
        assert compare_results(res2, ids2, cnt2, ores, oids, ocnt) == []
        t = clf.timings()
        assert t["kernel_launches"] > 0 and t["n_kmers"] == len(ok) and t["n_matches"] == len(om)
        return res, ids, cnt
    finally:
        clf.close()
        db.close()


@pytest.mark.parametrize("case", CASES)
def test_golden_case_all_stages(golden, mb, case):
    meta = golden.cases[case]
    _, s1, s2 = golden.reads(case)
    res, ids, cnt = run_all_stages(mb, golden.dbdir(case), meta["seq_mode"], s1, s2, smer_len=meta.get("smer_len", 0))
    # and directly against the reference-generated fixture
    assert helpers.check_results_against_golden(golden.results(case), res, ids, cnt) == []


@pytest.mark.parametrize("case", ["mask_paired", "mask_paired_p05", "mask_long"])
def test_mask_on_device(golden, mb, tmp_path, case):
    """--mask 1 (csrc/mb_mask.cuh): the masked batch is byte-identical to what the reference's
    SeqIterator::maskLowComplexityRegions produced (tests/golden/data/<case>.masked.txt, written by ref_harness), the
    stages that follow give the reference's results, and `metabuli classify --mask 1` writes the reference's files."""
    meta = golden.cases[case]
    _, s1, s2 = golden.reads(case)
    want = open(os.path.join(helpers.DATA, case + ".masked.txt"), "rb").read().split(b"\n")[:-1]
    bases, offsets, lens = mb.pack_reads(s1, s2)
    clf = mb.Classifier(golden.dbdir(case), seq_mode=meta["seq_mode"])
    try:
        clf.set_params(mask_mode=1, mask_prob=meta["mask_prob"])
        clf.stage(bases, offsets, lens)
        clf.mask_reads()
        clf.mask_reads()                                    # idempotent per staged batch
        gb, go, gl = clf.staged_reads(len(bases))
        got = [bytes(gb[int(o):int(o) + int(l)]) for o, l in zip(go, gl)]
        if s2 is None:
            got = got[0::2]
        assert got == want
        # the oracle's restatement agrees too (same bytes), then the rest of the path on the masked batch
        ob = orc.mask(orc.Batch(s1, s2), meta["mask_prob"])
        assert bytes(ob.seqs[:len(bases)]) == bytes(gb)
        n = clf.extract()
        assert n == meta["n_kmers"]
        assert helpers.digest(canon_k(clf.kmers())) == meta["kmers_sha256"]
        clf.sort_kmers()
        assert clf.match() == meta["n_matches"]
        clf.sort_matches()
        assert helpers.digest(as_orc_matches(clf.matches())) == meta["matches_sha256"]
        res, ids, cnt = clf.fetch(clf.score())
        assert helpers.check_results_against_golden(golden.results(case), res, ids, cnt) == []
        # one call from host buffers: masking happens inside
        res2, ids2, cnt2 = clf.classify(bases, offsets, lens)
        assert helpers.check_results_against_golden(golden.results(case), res2, ids2, cnt2) == []
    finally:
        clf.close()
    # the command line, device parser and host reader
    out = str(tmp_path)
    rep = os.path.join(helpers.DATA, "reports")
    files = [os.path.join(helpers.DATA, f) for f in meta["reads"]]
    flags = ["--seq-mode", str(meta["seq_mode"]), "--mask", "1", "--mask-prob", str(meta["mask_prob"])]
    for job, env in ((case, {}), (case + "_h", {"MB_HOST_PARSE": "1"})):
        subprocess.check_call([mb.CLI, "classify"] + flags + files + [golden.dbdir(case), out, job], env=dict(os.environ, **env))
        for f in ("classifications.tsv", "report.tsv"):
            assert open(os.path.join(out, job + "_" + f), "rb").read() == open(os.path.join(rep, case + "_" + f), "rb").read(), (job, f)


@pytest.mark.parametrize("splits", [64, 4096])
def test_build_index_on_device(golden, mb, tmp_path, splits):
    """DB-build inner loop (csrc/mb_build.cuh, mb_build_index): diffIdx / info / split for the committed target records are
    the files the REFERENCE's own sort + filterKmers<DB_CREATION> + writeTargetFilesAndSplits wrote (digests from
    oracle/_ref/ref_buildidx, tests/golden/make_golden4.py), and equal the oracle's byte for byte; edge cases: one record,
    all records equal, fewer k-mers than splits."""
    import hashlib
    import json
    want = json.load(open(os.path.join(helpers.DATA, "build_digests.json")))[f"splits{splits}"]
    rec = np.load(os.path.join(helpers.DATA, "build_input.npz"))
    clf = mb.Classifier(golden.dbdir("plain_paired"), seq_mode=2)
    db = orc.Db(golden.dbdir("plain_paired"))
    try:
        words, info, spl = clf.build_index(rec["value"], rec["taxid"], rec["species"], split_num=splits)
        assert len(info) == want["db_kmers"] and 2 * len(words) == want["diffIdx_bytes"]
        assert hashlib.sha256(words.tobytes()).hexdigest() == want["diffIdx_sha256"]
        assert hashlib.sha256(info.tobytes()).hexdigest() == want["info_sha256"]
        assert hashlib.sha256(spl.tobytes()).hexdigest() == want["split_sha256"]

        def same_as_oracle(v, t, s, k, tag):
            d = str(tmp_path / tag)
            n = orc.build_index(db, v, t, s, d, split_num=k)
            w, i, sp = clf.build_index(v, t, s, split_num=k)
            assert len(i) == n, tag
            assert w.tobytes() == open(d + "/diffIdx", "rb").read(), tag
            assert i.tobytes() == open(d + "/info", "rb").read(), tag
            assert sp.tobytes() == open(d + "/split", "rb").read(), tag
        v, t, s = rec["value"], rec["taxid"], rec["species"]
        same_as_oracle(v[:1], t[:1], s[:1], splits, "one")
        same_as_oracle(np.repeat(v[:1], 1000), t[:1000], np.repeat(s[:1], 1000), splits, "all_equal")
        same_as_oracle(v[:40], t[:40], s[:40], splits, "few")
        # values far apart (five-word deltas) and a giant amino-acid group swallowing several split thresholds: the digests
        # are the reference's (13 of the 63 split entries are used: thresholds inside the giant group share one)
        ev, et, es = helpers.build_edge_records(t, s)
        same_as_oracle(ev, et, es, 64, "wide_and_giant")
        w, i, sp = clf.build_index(ev, et, es, split_num=64)
        edge = json.load(open(os.path.join(helpers.DATA, "build_digests.json")))["edge_splits64"]
        assert (hashlib.sha256(w.tobytes()).hexdigest(), hashlib.sha256(i.tobytes()).hexdigest(), hashlib.sha256(sp.tobytes()).hexdigest()) == \
            (edge["diffIdx_sha256"], edge["info_sha256"], edge["split_sha256"])
        assert int((sp[:, 0] != 0).sum()) == edge["used_split_entries"]
    finally:
        clf.close()
        db.close()


@pytest.mark.parametrize("bits", [24, 31, 32, 40])
def test_kmer_sort_bits(golden, mb, bits):
    """The join result does not depend on how many leading amino-acid bits the query sort orders."""
    _, s1, s2 = golden.reads("sync_paired")
    run_all_stages(mb, golden.dbdir("sync_paired"), 2, s1, s2, sort_bits=bits)
    _, sl, _ = golden.reads("plain_long")
    run_all_stages(mb, golden.dbdir("plain_long"), 3, sl, None, sort_bits=bits)


def test_min_sp_score_branch(golden, mb):
    _, s1, s2 = golden.reads("plain_paired")
    run_all_stages(mb, golden.dbdir("plain_paired"), 2, s1, s2, min_sp_score=0.9)


def test_min_score_branch(golden, mb):
    _, s1, s2 = golden.reads("sync_paired")
    run_all_stages(mb, golden.dbdir("sync_paired"), 2, s1, s2, min_score=0.3)


def test_edge_reads(golden, mb):
    """Empty batch, reads shorter than one k-mer, all-N reads, N every few bases, lower case, ragged mates."""
    rng = np.random.default_rng(5)
    _, g1, g2 = golden.reads("plain_paired")
    s1 = [b"", b"ACGT", b"N" * 150, g1[0], g1[1].lower(), g1[2][:26], g1[3][:27], g1[4][:28], g1[5][:29], g1[6]]
    s2 = [b"", g2[1], g2[2], b"ACGTAC", g2[1], g2[2][:150], g2[3][:31], g2[4][:30], b"", bytes(g2[6][:75]) + b"NNN" + bytes(g2[6][78:])]
    x = bytearray(g1[7])
    for i in range(0, len(x), 29):
        x[i] = ord("N")
    s1.append(bytes(x))
    s2.append(g2[7])
    run_all_stages(mb, golden.dbdir("plain_paired"), 2, s1, s2)
    run_all_stages(mb, golden.dbdir("sync_paired"), 2, s1, s2)
    run_all_stages(mb, golden.dbdir("legacy_paired"), 1, s1, None)
    clf = mb.Classifier(golden.dbdir("plain_paired"), seq_mode=2)
    bases, offsets, lens = mb.pack_reads([], [])
    res, ids, cnt = clf.classify(bases, offsets, lens)
    assert len(res) == 0
    clf.close()


def _craft_db(golden, tmp, values, taxids):
    src = golden.dbdir("plain_paired")
    os.makedirs(tmp, exist_ok=True)
    for f in ("taxonomyDB", "taxID_list", "db.parameters"):
        shutil.copy(os.path.join(src, f), tmp)
    v = np.ascontiguousarray(values, dtype=np.uint64)
    t = np.ascontiguousarray(taxids, dtype=np.uint32)
    assert orc.lib().orc_write_db_files(tmp.encode(), orc._ptr(v), orc._ptr(t), len(v), 64, 2) == 0
    return tmp


def test_skewed_groups(golden, mb, tmp_path):
    """Amino-acid groups longer than the 512-word overhang and longer than a whole chunk (streaming path),
    groups straddling block boundaries, and the 'last k-mer of the DB is never a candidate' rule."""
    rng = np.random.default_rng(11)
    src_db = orc.Db(golden.dbdir("plain_paired"))
    base_v, base_t = src_db.decode()
    src_db.close()
    taxa = np.unique(base_t)
    _, s1, s2 = golden.reads("plain_paired")
    p = orc.default_params(kmer_format=2, seq_mode=2)
    qk = orc.extract(p, orc.Batch(s1[:200], s2[:200]))
    aa_pool = np.unique(qk["value"] >> np.uint64(24))
    vals = [base_v]
    for aa, size in zip(rng.choice(aa_pool, 6, replace=False), (600, 1500, 4000, 9000, 300, 2600)):
        dna = rng.choice(1 << 24, size=size, replace=False).astype(np.uint64)
        vals.append((np.uint64(aa) << np.uint64(24)) | dna)
    v = np.unique(np.concatenate(vals))
    t = rng.choice(taxa, size=len(v)).astype(np.uint32)
    # make the LAST DB k-mer share its AA with a query so that the exclusion rule matters
    last_aa = int(aa_pool[-1])
    tail = (np.uint64(last_aa) << np.uint64(24)) | np.arange(5, dtype=np.uint64)
    keep = (v >> np.uint64(24)) < np.uint64(last_aa)
    v = np.concatenate([v[keep], tail])
    t = np.concatenate([t[keep], rng.choice(taxa, size=5).astype(np.uint32)])
    dbdir = _craft_db(golden, str(tmp_path / "skew"), v, t)
    run_all_stages(mb, dbdir, 2, s1[:200], s2[:200])


def test_reads_with_thousands_of_matches(golden, mb, tmp_path):
    """Short reads whose match lists exceed every in-kernel segment buffer (> 1024 matches: k_sort_segments_huge, one CTA per
    such read) next to ordinary reads: the k-mers of three read pairs are in the DB under every species of the fixture taxonomy."""
    rng = np.random.default_rng(12)
    src_db = orc.Db(golden.dbdir("plain_paired"))
    base_v, base_t = src_db.decode()
    src_db.close()
    taxa = np.unique(base_t)
    _, s1, s2 = golden.reads("plain_paired")
    p = orc.default_params(kmer_format=2, seq_mode=2)
    hot = orc.extract(p, orc.Batch([s1[3], s1[40], s1[77]], [s2[3], s2[40], s2[77]]))["value"]
    hot = np.unique(hot)
    # a database holds a value once per SPECIES (IndexCreator::filterKmers): the added taxa belong to different species
    src_db = orc.Db(golden.dbdir("plain_paired"))
    by_species = {}
    for x in taxa.tolist():
        by_species.setdefault(int(orc.lib().orc_tax_species_of(src_db.h, int(x))), int(x))
    src_db.close()
    by_species.pop(0, None)
    chosen = np.array(sorted(by_species.values()), dtype=np.uint32)
    assert len(chosen) >= 8
    extra_v = np.repeat(hot, len(chosen))
    extra_t = np.tile(chosen, len(hot)).astype(np.uint32)
    # (the base DB's own entries for these values would repeat a species: drop them)
    base_keep = ~np.isin(base_v, hot)
    base_v, base_t = base_v[base_keep], base_t[base_keep]
    v = np.concatenate([base_v, extra_v])
    t = np.concatenate([base_t.astype(np.uint32), extra_t])
    order = np.lexsort((t, v))
    v, t = v[order], t[order]
    keep = np.ones(len(v), dtype=bool)
    keep[1:] = (v[1:] != v[:-1]) | (t[1:] != t[:-1])
    dbdir = _craft_db(golden, str(tmp_path / "hot"), v[keep], t[keep])
    db = orc.Db(dbdir)
    k, m, res, ids, cnt = orc.classify(db, db.params(seq_mode=2), orc.Batch(s1[:120], s2[:120]), threads=4)
    db.close()
    seq = (m["qinfo"] >> np.uint64(32)) & np.uint64((1 << 29) - 1)
    per_read = np.bincount(seq.astype(np.int64))
    assert per_read.max() > 2048 and (per_read > 1024).sum() >= 3          # the case really has huge segments
    run_all_stages(mb, dbdir, 2, s1[:120], s2[:120])


def test_long_read_sort_orders_dna_ties(golden, mb):
    """--seq-mode 3 sorts the full compareMatches key; the radix passes leave the 24 DNA bits out and k_fix_dna_ties orders
    the runs that agree above them.  Real matches, each repeated up to six times with other DNA bits (so that nearly every
    record sits in such a run), shuffled: the order must equal compareMatches' (KmerMatcher.cpp:1149-1166)."""
    _, sl, _ = golden.reads("plain_long")
    bases, offsets, lens = mb.pack_reads(sl, None)
    clf = mb.Classifier(golden.dbdir("plain_long"), seq_mode=3)
    try:
        clf.stage(bases, offsets, lens)
        clf.extract()
        clf.sort_kmers()
        clf.match()
        m = clf.matches()
        rng = np.random.default_rng(21)
        parts = [m]
        for k in range(5):
            c = m[rng.random(len(m)) < 0.6].copy()
            c["dnaEncoding"] = rng.integers(0, 1 << 24, len(c), dtype=np.uint32)
            parts.append(c)
        big = np.concatenate(parts)
        big = big[rng.permutation(len(big))]
        clf.set_matches(big)
        clf.sort_matches()
        got = as_orc_matches(clf.matches())
        want = orc.sort_matches(as_orc_matches(big))
        assert len(got) == len(want) > 3 * len(m)
        assert np.array_equal(got, want)
    finally:
        clf.close()


def test_async_result_fetch(golden, mb):
    """mb_fetch_results_async / mb_fetch_wait: the same arrays as mb_fetch_results, also when the next batch is started
    before the wait (its scoring must not overwrite results that are still travelling)."""
    import torch
    _, s1, s2 = golden.reads("sync_paired")
    half = len(s1) // 2
    a = mb.pack_reads(s1[:half], s2[:half])
    b = mb.pack_reads(s1[half:], s2[half:])
    clf = mb.Classifier(golden.dbdir("sync_paired"), seq_mode=2)
    try:
        def pinned(n, dt):
            t = torch.zeros(n * np.dtype(dt).itemsize, dtype=torch.uint8).pin_memory()
            return t, t.numpy().view(dt)
        want = []
        for bases, offsets, lens in (a, b):
            res, ids, cnt = clf.classify(bases, offsets, lens)
            want.append((res.copy(), ids.copy(), cnt.copy()))
        keep, got = [], []
        for bases, offsets, lens in (a, b):
            clf.stage(bases, offsets, lens)
            used = clf.classify_staged()
            tr, res = pinned(len(lens) // 2, mb.RESULT_DT)
            ti, ids = pinned(used + 1, np.int32)
            tc, cnt = pinned(used + 1, np.int32)
            keep.append((tr, ti, tc))
            clf.fetch_async(res, ids, cnt)              # no wait: the next batch starts while these travel
            got.append((res, ids, cnt, used))
        clf.fetch_wait()
        for (wr, wi, wc), (res, ids, cnt, used) in zip(want, got):
            assert compare_results(res, ids, cnt, wr, wi, wc) == []
    finally:
        clf.close()


def test_medium_random(mb, tmp_path):
    """A larger seeded case built on the box: 54 genomes x 60 kb, 20k pairs + long reads (multi-tile sort,
    hundreds of DB blocks), plain and syncmer."""
    import synth
    work = str(tmp_path / "med")
    subprocess.check_call([sys.executable, os.path.join(helpers.HERE, "golden", "make_inputs.py"), work, "--seed", "77", "--families", "3",
                           "--genera", "3", "--species", "3", "--strains", "2", "--genome-len", "60000", "--pairs", "20000",
                           "--long-reads", "30"])
    _, s1 = orc.read_fastx(os.path.join(work, "reads_1.fq"))
    _, s2 = orc.read_fastx(os.path.join(work, "reads_2.fq"))
    _, sl = orc.read_fastx(os.path.join(work, "long.fa"))
    for sync in (0, 1):
        dbdir = os.path.join(work, f"db{sync}")
        synth.build_db(work, dbdir, syncmer=sync, kmer_format=2)
        run_all_stages(mb, dbdir, 2, s1, s2)
        run_all_stages(mb, dbdir, 3, sl, None)


def test_cli_files_byte_identical(golden, mb, tmp_path):
    """`metabuli classify` outputs vs the files the REFERENCE's own Reporter wrote for the same inputs
    (tests/golden/data/reports, made by tests/golden/make_golden2.py from oracle/_ref/ref_harness --report;
    Reporter.cpp:38-83,146-253), byte for byte: classifications, report, Krona chart, --lineage 1."""
    case = "plain_paired"
    data = helpers.DATA
    rep = os.path.join(data, "reports")
    out = str(tmp_path)

    def rd(p):
        with open(p, "rb") as f:
            return f.read()
    prelude = os.path.join(out, "prelude.html")
    with open(prelude, "wb") as f:
        f.write(b"<html><!-- stand-in for krona_prelude.html -->\n")
    r1, r2 = os.path.join(data, "reads_1.fq"), os.path.join(data, "reads_2.fq")
    subprocess.check_call([mb.CLI, "classify", r1, r2, golden.dbdir(case), out, "job", "--batch-reads", "150", "--krona-prelude", prelude])
    # (plain FASTQ is parsed on the GPU by default; MB_HOST_PARSE=1 forces the host reader that FASTA / gzip input takes)
    subprocess.check_call([mb.CLI, "classify", r1, r2, golden.dbdir(case), out, "jobh", "--batch-reads", "150"], env=dict(os.environ, MB_HOST_PARSE="1"))
    for f in ("classifications.tsv", "report.tsv"):
        b = rd(os.path.join(rep, case + "_" + f))
        assert rd(os.path.join(out, "job_" + f)) == b, f
        assert rd(os.path.join(out, "jobh_" + f)) == b, f
    assert rd(os.path.join(out, "job_krona.html")) == rd(prelude) + rd(os.path.join(rep, case + "_krona_tail.html"))
    # multi-line FASTQ (sequence and quality wrapped at 70 columns, after a first record that looks plain) is not what the
    # device parser takes: the command notices on the first chunk and falls back to the host reader
    def wrap(path_in, path_out):
        with open(path_in, "rb") as f, open(path_out, "wb") as o:
            lines = f.read().split(b"\n")
            for i in range(0, len(lines) - 1, 4):
                h, sq, p, q = lines[i:i + 4]
                if i == 0:                          # the first record stays on four lines: the sniffing sees plain FASTQ
                    o.write(h + b"\n" + sq + b"\n" + p + b"\n" + q + b"\n")
                    continue
                o.write(h + b"\n" + b"\n".join(sq[k:k + 70] for k in range(0, len(sq), 70)) + b"\n" + p + b"\n" +
                        b"\n".join(q[k:k + 70] for k in range(0, len(q), 70)) + b"\n")
    wrap(r1, os.path.join(out, "w1.fq"))
    wrap(r2, os.path.join(out, "w2.fq"))
    log = subprocess.check_output([mb.CLI, "classify", os.path.join(out, "w1.fq"), os.path.join(out, "w2.fq"), golden.dbdir(case), out, "jobw"])
    assert b"using the host reader" in log
    assert rd(os.path.join(out, "jobw_classifications.tsv")) == rd(os.path.join(rep, case + "_classifications.tsv"))
    # --lineage 1: one more column (TaxonomyWrapper::taxLineage2)
    subprocess.check_call([mb.CLI, "classify", r1, r2, golden.dbdir(case), out, "jobl1", "--lineage", "1"])
    got = rd(os.path.join(out, "jobl1_classifications.tsv"))
    assert got == rd(os.path.join(rep, case + "_lineage_classifications.tsv"))
    assert b"\ts_" in got or b";s_" in got            # a species-level lineage entry is present
    # the other databases / modes through the command line: Eukaryota taxonomy, redundancy bits, --smer-len, legacy x long reads
    for c, extra in (("euk_paired", []), ("redund_paired", []), ("sync4_paired", ["--smer-len", "4"]), ("sync6_paired", ["--smer-len", "6"])):
        subprocess.check_call([mb.CLI, "classify", r1, r2, golden.dbdir(c), out, c] + extra)
        for f in ("classifications.tsv", "report.tsv"):
            assert rd(os.path.join(out, c + "_" + f)) == rd(os.path.join(rep, c + "_" + f)), (c, f)
    for c in ("legacy_long", "euk_long"):
        subprocess.check_call([mb.CLI, "classify", "--seq-mode", "3", os.path.join(data, "long.fa"), golden.dbdir(c), out, c])
        for f in ("classifications.tsv", "report.tsv"):
            assert rd(os.path.join(out, c + "_" + f)) == rd(os.path.join(rep, c + "_" + f)), (c, f)
    # gzip input: inflated by the reader threads, the text parsed on the device like a plain file (several chunks), and
    # through the host reader
    import gzip
    gz = []
    for src in (r1, r2):
        dst = os.path.join(out, os.path.basename(src) + ".gz")
        with open(src, "rb") as fi, gzip.open(dst, "wb") as fo:
            fo.write(fi.read())
        gz.append(dst)
    for job, env, extra in (("gz", {}, []), ("gz_chunks", {}, ["--batch-reads", "150"]), ("gz_host", {"MB_HOST_PARSE": "1"}, [])):
        log = subprocess.run([mb.CLI, "classify"] + gz + [golden.dbdir(case), out, job] + extra, env=dict(os.environ, **env), check=True,
                             capture_output=True, text=True).stdout
        assert ("parsed on the GPU" in log) == (job != "gz_host")
        for f in ("classifications.tsv", "report.tsv"):
            assert rd(os.path.join(out, job + "_" + f)) == rd(os.path.join(rep, case + "_" + f)), (job, f)
    # several GPUs in one process (host/mb_host_multi.cpp): DB shards per context, reads split over them, k-mers and matches
    # exchanged through peer stores.  Listing device 0 more than once runs the same code on one GPU.
    for c, mode, gl, extra in (("plain_paired", 2, "0,0", ["--batch-reads", "150"]), ("sync_paired", 2, "0,0,0", []),
                               ("legacy_long", 3, "0,0", []), ("euk_paired", 2, "0,0,0,0", ["--batch-reads", "97"])):
        inp = [r1, r2] if mode == 2 else ["--seq-mode", "3", os.path.join(data, "long.fa")]
        subprocess.check_call([mb.CLI, "classify"] + inp + [golden.dbdir(c), out, "mg_" + c, "--gpu-list", gl] + extra)
        for f in ("classifications.tsv", "report.tsv"):
            assert rd(os.path.join(out, "mg_" + c + "_" + f)) == rd(os.path.join(rep, c + "_" + f)), (c, gl, f)
    # shard re-balancing forced after every batch (thresholds lowered through the test knobs): the shards are reloaded with
    # moved boundaries in the middle of the run and the output must not change
    log = subprocess.run([mb.CLI, "classify", r1, r2, golden.dbdir("sync_paired"), out, "mg_rebal", "--gpu-list", "0,0,0", "--batch-reads", "97"],
                         env=dict(os.environ, MB_REBALANCE_MIN_READS="1", MB_REBALANCE_THRESHOLD="0"), check=True, capture_output=True, text=True)
    assert "re-balanced" in log.stdout
    for f in ("classifications.tsv", "report.tsv"):
        assert rd(os.path.join(out, "mg_rebal_" + f)) == rd(os.path.join(rep, "sync_paired_" + f)), f
    # receive buffers that start far too small: both exchanges overflow, everybody re-allocates and the step is redone
    log = subprocess.run([mb.CLI, "classify", r1, r2, golden.dbdir("sync_paired"), out, "mg_small", "--gpu-list", "0,0,0"],
                         env=dict(os.environ, MB_EXCHANGE_CAP0="64"), check=True, capture_output=True, text=True)
    assert "regrown" in log.stdout + log.stderr
    for f in ("classifications.tsv", "report.tsv"):
        assert rd(os.path.join(out, "mg_small_" + f)) == rd(os.path.join(rep, "sync_paired_" + f)), f
    # long-read mode through the library entry point
    mb.classify_files(os.path.join(data, "long.fa"), None, golden.dbdir("sync_long"), out, "jobl", seq_mode=3)
    assert rd(os.path.join(out, "jobl_classifications.tsv")) == rd(os.path.join(rep, "sync_long_classifications.tsv"))
    assert rd(os.path.join(out, "jobl_report.tsv")) == rd(os.path.join(rep, "sync_long_report.tsv"))


def test_sharded_db_matches_union(golden, mb):
    """DB cut at split boundaries into 3 self-contained shards: the union of the shards' matches equals
    the whole-DB match list, including the last-k-mer rule on the last shard only."""
    case = "plain_paired"
    _, s1, s2 = golden.reads(case)
    bases, offsets, lens = mb.pack_reads(s1, s2)
    whole = mb.Classifier(golden.dbdir(case), seq_mode=2)
    whole.stage(bases, offsets, lens)
    whole.extract()
    whole.sort_kmers()
    kmers = whole.kmers()
    # queries INSIDE the boundary amino-acid groups whose DNA part is below / above the group's first DB k-mer: they are
    # compared with the whole group (KmerMatcher.cpp:275-450), so they must be routed by their amino-acid part alone
    probe = mb.Classifier(golden.dbdir(case), seq_mode=2, shard=1, n_shards=3)
    b3 = probe.shard_bounds()
    probe.close()
    assert np.all(b3 & np.uint64(0xFFFFFF) == 0)
    extra = np.zeros(4 * len(b3), dtype=kmers.dtype)
    for i, bd in enumerate(b3):
        for j, dna in enumerate((0, 1, 0x924924, 0xFFFFFF)):
            extra[4 * i + j]["value"] = bd | np.uint64(dna)
            extra[4 * i + j]["qinfo"] = (np.uint64(7 + i) << np.uint64(32)) | np.uint64(3 * j) | (np.uint64(j % 6) << np.uint64(61))
    kmers = np.concatenate([kmers, extra])
    whole.set_kmers(kmers)
    whole.sort_kmers()
    kmers = whole.kmers()
    whole.match()
    want = orc.sort_matches(as_orc_matches(whole.matches()))
    db_o = orc.Db(golden.dbdir(case))
    assert np.array_equal(want, orc.sort_matches(orc.match(db_o, db_o.params(seq_mode=2), orc.sort_kmers(kmers.view(orc.KMER_DT).copy()))))
    db_o.close()
    hit = np.isin(want["qinfo"], extra["qinfo"])
    assert hit.any(), "the boundary-group probes must produce matches, or the case proves nothing"
    whole.close()
    parts = []
    total_k = 0
    for g in range(3):
        sh = mb.Classifier(golden.dbdir(case), seq_mode=2, shard=g, n_shards=3)
        total_k += sh.db_kmers()
        bounds = sh.shard_bounds()
        assert len(bounds) == 2
        owner = np.searchsorted(bounds, kmers["value"], side="right")
        sh.set_kmers(kmers[owner == g])
        sh.sort_kmers()
        sh.match()
        parts.append(as_orc_matches(sh.matches()))
        sh.close()
    got = orc.sort_matches(np.concatenate(parts))
    assert np.array_equal(got, want)
    db = orc.Db(golden.dbdir(case))
    assert total_k == db.num_kmers()
    db.close()


def test_partition_kernels(golden, mb):
    case = "plain_paired"
    _, s1, s2 = golden.reads(case)
    bases, offsets, lens = mb.pack_reads(s1, s2)
    clf = mb.Classifier(golden.dbdir(case), seq_mode=2)
    clf.stage(bases, offsets, lens)
    clf.extract()
    k0 = clf.kmers()
    qs = np.quantile(k0["value"].astype(np.float64), [0.25, 0.5, 0.75]).astype(np.uint64)
    counts = clf.partition_kmers(qs)
    k1 = clf.kmers()
    owner0 = np.searchsorted(qs, k0["value"], side="right")
    assert np.array_equal(counts, np.bincount(owner0, minlength=4).astype(np.uint64))
    # stable partition
    want = np.concatenate([k0[owner0 == g] for g in range(4)])
    assert np.array_equal(k1, want)
    clf.sort_kmers()
    clf.match()
    m0 = clf.matches()
    rpr = (len(s1) + 2) // 3
    counts = clf.partition_matches(rpr, 3)
    m1 = clf.matches()
    home = ((m0["qinfo"] >> np.uint64(32)) & np.uint64(0x1FFFFFFF)).astype(np.int64) - 1
    home //= rpr
    assert np.array_equal(counts, np.bincount(home, minlength=3).astype(np.uint64))
    assert np.array_equal(m1, np.concatenate([m0[home == g] for g in range(3)]))
    clf.close()


@pytest.mark.parametrize("case,world", [("plain_paired", 3), ("sync_paired", 2), ("legacy_paired", 2)])
def test_routed_producers_one_gpu(golden, mb, case, world):
    """The fused multi-GPU form on ONE GPU: `world` contexts (one DB shard each, a slice of the reads each) wired to each
    other's receive buffers by plain device pointers — the same kernels write the same "peer" stores, only the NVLink
    hop is missing.  k-mers must arrive at the shard that owns them, matches at their read's home context, and every
    context's results must equal the single-context oracle run of its reads."""
    meta = golden.cases[case]
    dbdir = golden.dbdir(case)
    _, s1, s2 = golden.reads(case)
    n = len(s1)
    per = (n + world - 1) // world
    db = orc.Db(dbdir)
    p = db.params(seq_mode=meta["seq_mode"])
    ok, om, ores, oids, ocnt = orc.classify(db, p, orc.Batch(s1, s2), threads=4)
    db.close()
    ctxs = [mb.Classifier(dbdir, seq_mode=meta["seq_mode"], shard=g, n_shards=world) for g in range(world)]
    try:
        bounds = ctxs[0].shard_bounds()
        assert len(bounds) == world - 1
        mem = [c.peer_alloc(world * (len(ok) + 64), world * (len(om) + 64)) for c in ctxs]      # a region per source rank
        for g, c in enumerate(ctxs):
            c.peer_attach(g, mem, bounds, per)
        emitted = 0
        for g, c in enumerate(ctxs):
            a, b = g * per, min(n, (g + 1) * per)
            bases, offsets, lens = mb.pack_reads(s1[a:b], None if s2 is None else s2[a:b])
            c.set_read_id_base(g * per)
            c.stage(bases, offsets, lens)
            emitted += c.extract_routed()
        assert emitted == len(ok)
        owner = np.searchsorted(bounds, ok["value"], side="right")
        m_emitted = 0
        for g, c in enumerate(ctxs):
            got, over = c.take_routed_kmers()
            assert not over and got == int((owner == g).sum())
            assert np.array_equal(canon_k(c.kmers()), ok[owner == g])         # the right k-mers reached shard g
            c.sort_kmers()
            m_emitted += c.match_routed()
        assert m_emitted == len(om)
        home = (((om["qinfo"] >> np.uint64(32)) & np.uint64(0x1FFFFFFF)).astype(np.int64) - 1) // per
        for g, c in enumerate(ctxs):
            got, over = c.take_routed_matches()
            assert not over and got == int((home == g).sum())
            assert np.array_equal(orc.sort_matches(as_orc_matches(c.matches())), om[home == g])
            c.sort_matches()
            used = c.score()
            res, ids, cnt = c.fetch(used)
            a, b = g * per, min(n, (g + 1) * per)
            sub = ores[a:b]
            assert compare_results(res, ids, cnt, sub, oids, ocnt) == []
        # a receive buffer that is too small is reported, not overrun
        small = [c.peer_alloc(world * (len(ok) + 64), 8) for c in ctxs]
        for g, c in enumerate(ctxs):
            c.peer_attach(g, small, bounds, per)
        for c in ctxs:
            c.extract_routed()
        for c in ctxs:
            c.take_routed_kmers()
            c.sort_kmers()
            c.match_routed()
        assert any(c.take_routed_matches()[1] for c in ctxs)
    finally:
        for c in ctxs:
            c.close()


def test_prefetch_overlap(golden, mb):
    """mb_prefetch_reads: the next batch is copied on the second stream while the current one is classified; the
    following stage() with the same arrays adopts it.  Results must equal the plain path, also when a prefetch is
    never claimed or is replaced."""
    case = "sync_paired"
    _, s1, s2 = golden.reads(case)
    half = len(s1) // 2
    A = mb.pack_reads(s1[:half], s2[:half])
    B = mb.pack_reads(s1[half:], s2[half:])
    clf = mb.Classifier(golden.dbdir(case), seq_mode=2)
    try:
        want = {}
        for name, (b, o, l) in (("A", A), ("B", B)):
            want[name] = clf.classify(b, o, l)
        clf.stage(*A)
        clf.prefetch(*B)                        # overlaps the work on A
        ra = clf.fetch(clf.classify_staged())
        clf.stage(*B)                           # adopted, no second copy
        clf.prefetch(*A)
        rb = clf.fetch(clf.classify_staged())
        clf.prefetch(*B)                        # replaces the unclaimed prefetch of A
        clf.stage(*A)                           # not the prefetched batch: ordinary copy
        ra2 = clf.fetch(clf.classify_staged())
        clf.stage(*B)                           # still pending, adopted now
        rb2 = clf.fetch(clf.classify_staged())
        for got, name in ((ra, "A"), (rb, "B"), (ra2, "A"), (rb2, "B")):
            assert compare_results(*got, *want[name]) == [], name           # (taxCnt ranges are handed out by atomics)
        # a malformed layout is caught by the device-side check
        b, o, l = A
        o2 = o.copy()
        o2[1], o2[2] = o[2], o[1]
        with pytest.raises(mb.MetabuliError):
            clf.stage(b, o2, l)
    finally:
        clf.close()


def _fastq_text(names, seqs, eol=b"\n", trailing=True, comment=True):
    recs = []
    for k, (n, s) in enumerate(zip(names, seqs)):
        hdr = b"@" + n + ((b" extra/%d comment" % k) if comment and k % 2 == 0 else b"")
        recs.append(hdr + eol + bytes(s) + eol + b"+" + (n if k % 3 == 0 else b"") + eol + b"I" * len(s))
    t = eol.join(recs)
    return t + (eol if trailing else b"")


@pytest.mark.parametrize("variant", ["plain", "crlf", "no_trailing_newline", "single_end", "chunked"])
def test_fastq_parsed_on_device(golden, mb, variant):
    """mb_stage_fastq (input pipeline on the device): raw FASTQ text -> the same staged batch the host packer
    produces (bases, offsets, lengths), the read names, whole-record chunk accounting, malformed-record errors, and the
    same classification results as the host-parsed path."""
    _, g1, g2 = golden.reads("plain_paired")
    s1 = [bytes(x) for x in g1[:300]] + [b"", b"ACGTN", b"acgtacgtacgtacgtacgtacgtacgtacgt"]
    s2 = [bytes(x) for x in g2[:300]] + [b"ACGT", b"", b"TTTTGGGGCCCCAAAATTTTGGGGCCCCAAAA"]
    names = [b"read_%d/x" % i for i in range(len(s1))]
    paired = variant != "single_end"
    eol = b"\r\n" if variant == "crlf" else b"\n"
    t1 = _fastq_text(names, s1, eol, trailing=variant != "no_trailing_newline")
    t2 = _fastq_text(names, s2, eol, trailing=variant != "no_trailing_newline", comment=False) if paired else None
    clf = mb.Classifier(golden.dbdir("plain_paired"), seq_mode=2 if paired else 1)
    try:
        want_b, want_o, want_l = mb.pack_reads(s1, s2 if paired else None)
        clf.stage(want_b, want_o, want_l)
        used = clf.classify_staged()
        want_res, want_ids, want_cnt = clf.fetch(used)
        if variant == "chunked":
            # the texts are cut in the middle of a record: only whole records are taken and the consumed bytes say where
            # the next chunk starts; a read limit cuts earlier
            cut1, cut2 = len(t1) * 2 // 3, len(t2) // 2
            n, c1, c2, nb = clf.stage_fastq(t1[:cut1], t2[:cut2], final=False)
            assert 0 < n < len(s1) and t1[c1:c1 + 1] == b"@" and t2[c2:c2 + 1] == b"@"
            assert t1[:c1].count(b"\n") == 4 * n and t2[:c2].count(b"\n") == 4 * n
            b0, o0, l0 = clf.staged_reads(nb)
            n2, d1, d2, nb2 = clf.stage_fastq(t1[c1:], t2[c2:], final=True)
            assert n + n2 == len(s1) and c1 + d1 == len(t1) and c2 + d2 == len(t2)
            b1, o1, l1 = clf.staged_reads(nb2)
            assert np.array_equal(np.concatenate([b0, b1]), want_b) and np.array_equal(np.concatenate([l0, l1]), want_l)
            assert np.array_equal(np.concatenate([o0, o1 + np.uint64(nb)]), want_o)
            n3, e1, e2, _ = clf.stage_fastq(t1, t2, max_reads=17)
            assert n3 == 17 and t1[:e1].count(b"\n") == 68 and t2[:e2].count(b"\n") == 68
            return
        n, c1, c2, nb = clf.stage_fastq(t1, t2)
        assert n == len(s1) and c1 == len(t1) and (not paired or c2 == len(t2)) and nb == len(want_b)
        got_b, got_o, got_l = clf.staged_reads(nb)
        assert np.array_equal(got_l, want_l) and np.array_equal(got_o, want_o) and np.array_equal(got_b, want_b)
        assert clf.fastq_names(t1) == names
        used = clf.classify_staged()
        res, ids, cnt = clf.fetch(used)
        assert compare_results(res, ids, cnt, want_res, want_ids, want_cnt) == []
        # the same text prefetched on the copy stream and adopted by the next stage call
        a1 = np.frombuffer(t1, dtype=np.uint8).copy()
        a2 = None if t2 is None else np.frombuffer(t2, dtype=np.uint8).copy()
        clf.prefetch_fastq(a1, a2)
        n, c1, c2, nb = clf.stage_fastq(a1, a2)
        assert n == len(s1) and nb == len(want_b)
        got_b, got_o, got_l = clf.staged_reads(nb)
        assert np.array_equal(got_l, want_l) and np.array_equal(got_o, want_o) and np.array_equal(got_b, want_b)
        used = clf.classify_staged()
        res, ids, cnt = clf.fetch(used)
        assert compare_results(res, ids, cnt, want_res, want_ids, want_cnt) == []
        # malformed records are reported, not classified
        bad = t1.replace(b"+", b"-", 1)
        with pytest.raises(mb.MetabuliError):
            clf.stage_fastq(bad, t2)
        lines = t1.split(eol)
        lines[3] = lines[3][:-1]                       # quality shorter than the sequence
        with pytest.raises(mb.MetabuliError):
            clf.stage_fastq(eol.join(lines), t2)
        n0, _, _, nb0 = clf.stage_fastq(b"", b"" if paired else None)
        assert n0 == 0 and nb0 == 0
    finally:
        clf.close()


@pytest.mark.parametrize("extra", [[], ["--mask"]])
def test_config1_sized_run_equals_reference_classes(tmp_path, extra):
    """BASELINE configs[0] at its stated size (100 genomes x 200 kb, 100 000 read pairs) through the whole product path
    (DB directory written in the on-disk format, C++ host, CUDA stages) and through the reference's own classes
    (oracle/_ref/ref_harness: KmerMatcher, Taxonomer, ips4o sort; with --mask also its KmerExtractor + tantan) on the
    same files: every read's is_classified, taxID, query length, fp32 score bits and taxID:count list must agree."""
    import json
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(os.path.join(root, "oracle", "_ref", "ref_harness")):
        pytest.skip("oracle/_ref/ref_harness was not built (it is compiled where /root/reference exists and travels with the repo)")
    cmd = [sys.executable, os.path.join(root, "bench.py"), "--genomes", "100", "--pairs", "100000", "--len-min", "200000", "--len-max", "200000",
           "--steps", "1", "--warmup", "3", "--no-fastq", "--workdir", str(tmp_path)] + extra
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=root)
    assert out.returncode == 0, out.stderr[-3000:]
    line = json.loads([ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["check"] == {"against": "ref_harness", "reads": 100000, "mismatches": 0}
    assert line["counts"]["classified"] > 90000 and line["gpu_launches"] > 0
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["same_config"] is True
