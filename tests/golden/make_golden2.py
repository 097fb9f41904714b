#!/usr/bin/env python3
"""Round-2 additions to the golden fixtures (run in the build container only; ADDS to tests/golden/data, the six round-1
cases stay byte-identical):

  * new databases + cases for branches no round-1 fixture reached:
      db_euk      second superkingdom "Eukaryota" over half of the families  -> minConsCntEuk (Taxonomer.cpp:497-500)
      db_redund   db.parameters Skip_redundancy 0, bit 31 set in `info`       -> redundancy mask (KmerMatcher.cpp:204-205,381)
      db_sync4/6  syncmer DBs with --smer-len 4 and 6                          -> SyncmerScanner with s != 5
      legacy_long legacy format-1 DB x --seq-mode 3                            -> OldMetamerScanner + long-read scoring
  * for EVERY case (old and new) the files the REFERENCE's own Reporter writes (Reporter.cpp:38-83,146-253, compiled into
    oracle/_ref/ref_harness): <case>_classifications.tsv, <case>_report.tsv, <case>_lineage_classifications.tsv
    (--lineage 1) and <case>_krona_tail.html = JOB_krona.html without the static prelude (a 224-KB third-party
    resource of the reference that is not copied into this repository).
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import orc  # noqa: E402
import synth  # noqa: E402

PRELUDE = "/root/reference/lib/mmseqs/data/resources/krona_prelude.html"

NEW_CASES = [  # name, db, seq_mode, reads, extra harness flags
    ("euk_paired", "db_euk", 2, ("reads_1.fq", "reads_2.fq"), []),
    ("euk_long", "db_euk", 3, ("long.fa",), []),
    ("redund_paired", "db_redund", 2, ("reads_1.fq", "reads_2.fq"), []),
    ("sync4_paired", "db_sync4", 2, ("reads_1.fq", "reads_2.fq"), ["--smer-len", "4"]),
    ("sync6_paired", "db_sync6", 2, ("reads_1.fq", "reads_2.fq"), ["--smer-len", "6"]),
    ("sync4_long", "db_sync4", 3, ("long.fa",), ["--smer-len", "4"]),
    ("legacy_long", "db_legacy", 3, ("long.fa",), []),
]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def harness(data, work, name, db, mode, reads, flags, extra):
    prefix = os.path.join(work, name)
    cmd = [orc.REF_HARNESS, os.path.join(data, db), prefix, str(mode), "2"] + [os.path.join(data, r) for r in reads] + flags + extra
    out = subprocess.check_output(cmd).decode().strip().splitlines()[-1]
    return prefix, json.loads(out)


def main():
    data = os.path.join(HERE, "data")
    work = "/tmp/mb_golden_work2"
    shutil.rmtree(work, ignore_errors=True)
    gen = [sys.executable, os.path.join(HERE, "make_inputs.py")]
    common = ["--seed", "1234", "--genome-len", "12000", "--pairs", "400", "--long-reads", "8"]
    subprocess.check_call(gen + [work + "/plain"] + common)
    subprocess.check_call(gen + [work + "/euk"] + common + ["--euk-families", "1"])
    for f in ("reads_1.fq", "reads_2.fq", "long.fa"):        # same random stream: the reads are the round-1 reads
        assert open(os.path.join(work, "plain", f), "rb").read() == open(os.path.join(data, f), "rb").read(), f
        assert open(os.path.join(work, "euk", f), "rb").read() == open(os.path.join(data, f), "rb").read(), f
    for db in ("db_euk", "db_redund", "db_sync4", "db_sync6"):
        shutil.rmtree(os.path.join(data, db), ignore_errors=True)
    synth.build_db(work + "/euk", os.path.join(data, "db_euk"), syncmer=0, kmer_format=2, split_num=64)
    synth.build_db(work + "/plain", os.path.join(data, "db_redund"), syncmer=0, kmer_format=2, split_num=64, skip_redundancy=0)
    synth.build_db(work + "/plain", os.path.join(data, "db_sync4"), syncmer=1, kmer_format=2, smer_len=4, split_num=64)
    synth.build_db(work + "/plain", os.path.join(data, "db_sync6"), syncmer=1, kmer_format=2, smer_len=6, split_num=64)

    with open(os.path.join(data, "digests.json")) as f:
        digests = json.load(f)
    for name, db, mode, reads, flags in NEW_CASES:
        prefix, info = harness(data, work, name, db, mode, reads, flags, ["--dump"])
        k = orc.sort_kmers(np.fromfile(prefix + ".kmers.bin", dtype=orc.KMER_DT))
        m = np.fromfile(prefix + ".matches.bin", dtype=orc.MATCH_DT)
        shutil.copy(prefix + ".results.tsv", os.path.join(data, name + ".results.tsv"))
        digests[name] = {"db": db, "seq_mode": mode, "reads": list(reads), "n_reads": info["reads"],
                         "n_kmers": int(len(k)), "n_matches": int(len(m)),
                         "kmers_sha256": digest(k), "matches_sha256": digest(m)}
        if flags:
            digests[name]["smer_len"] = int(flags[1])
        print(name, digests[name]["n_kmers"], digests[name]["n_matches"])

    # the reference Reporter's files for every case
    prelude_len = os.path.getsize(PRELUDE)
    rep = os.path.join(data, "reports")
    shutil.rmtree(rep, ignore_errors=True)
    os.makedirs(rep)
    for name in sorted(digests):
        d = digests[name]
        flags = ["--smer-len", str(d["smer_len"])] if "smer_len" in d else []
        out = os.path.join(work, "rep_" + name)
        os.makedirs(out, exist_ok=True)
        harness(data, work, name, d["db"], d["seq_mode"], d["reads"], flags, ["--report", out, "plain"])
        harness(data, work, name, d["db"], d["seq_mode"], d["reads"], flags, ["--report", out, "lineage", "--lineage"])
        shutil.copy(os.path.join(out, "plain_classifications.tsv"), os.path.join(rep, name + "_classifications.tsv"))
        shutil.copy(os.path.join(out, "plain_report.tsv"), os.path.join(rep, name + "_report.tsv"))
        shutil.copy(os.path.join(out, "lineage_classifications.tsv"), os.path.join(rep, name + "_lineage_classifications.tsv"))
        assert open(os.path.join(out, "plain_report.tsv"), "rb").read() == open(os.path.join(out, "lineage_report.tsv"), "rb").read()
        krona = open(os.path.join(out, "plain_krona.html"), "rb").read()
        assert krona[:prelude_len] == open(PRELUDE, "rb").read()
        with open(os.path.join(rep, name + "_krona_tail.html"), "wb") as f:
            f.write(krona[prelude_len:])
    with open(os.path.join(data, "digests.json"), "w") as f:
        json.dump(digests, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
