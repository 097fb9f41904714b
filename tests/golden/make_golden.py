#!/usr/bin/env python3
"""Regenerates the golden fixtures under tests/golden/data (run in the build container only).

  1. make_inputs.py  -> seeded genomes / taxdump / reads
  2. oracle/synth.py -> three databases (format 2 plain, format 2 syncmer, legacy format 1);
                        taxonomyDB comes from the reference's TaxonomyWrapper (oracle/_ref/ref_mkdb)
  3. oracle/_ref/ref_harness (the REFERENCE's KmerScanner/SyncmerScanner/KmerMatcher/Taxonomer
     classes compiled from /root/reference) -> per-read results + sha256 of the sorted k-mer and
     match buffers for each case

Only the reads, the DB directories, results.tsv and digests.json are kept; they are what
tests/test_oracle_golden.py pins the oracle with (and the GPU tests then compare with the oracle).
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

CASES = [  # name, db, seq_mode, reads
    ("plain_paired", "db_plain", 2, ("reads_1.fq", "reads_2.fq")),
    ("sync_paired", "db_sync", 2, ("reads_1.fq", "reads_2.fq")),
    ("legacy_paired", "db_legacy", 2, ("reads_1.fq", "reads_2.fq")),
    ("plain_long", "db_plain", 3, ("long.fa",)),
    ("sync_long", "db_sync", 3, ("long.fa",)),
    ("plain_single", "db_plain", 1, ("reads_1.fq",)),
]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    data = os.path.join(HERE, "data")
    work = "/tmp/mb_golden_work"
    shutil.rmtree(work, ignore_errors=True)
    shutil.rmtree(data, ignore_errors=True)
    os.makedirs(data)
    subprocess.check_call([sys.executable, os.path.join(HERE, "make_inputs.py"), work, "--seed", "1234",
                           "--genome-len", "12000", "--pairs", "400", "--long-reads", "8"])
    for f in ("reads_1.fq", "reads_2.fq", "long.fa"):
        shutil.copy(os.path.join(work, f), data)
    synth.build_db(work, os.path.join(data, "db_plain"), syncmer=0, kmer_format=2, split_num=64)
    synth.build_db(work, os.path.join(data, "db_sync"), syncmer=1, kmer_format=2, split_num=64)
    synth.build_db(work, os.path.join(data, "db_legacy"), syncmer=0, kmer_format=1, split_num=64)
    digests = {}
    for name, db, mode, reads in CASES:
        prefix = os.path.join(work, name)
        cmd = [orc.REF_HARNESS, os.path.join(data, db), prefix, str(mode), "2"] + [os.path.join(data, r) for r in reads] + ["--dump"]
        out = subprocess.check_output(cmd).decode().strip().splitlines()[-1]
        info = json.loads(out)
        k = orc.sort_kmers(np.fromfile(prefix + ".kmers.bin", dtype=orc.KMER_DT))
        m = np.fromfile(prefix + ".matches.bin", dtype=orc.MATCH_DT)
        shutil.copy(prefix + ".results.tsv", os.path.join(data, name + ".results.tsv"))
        digests[name] = {"db": db, "seq_mode": mode, "reads": list(reads), "n_reads": info["reads"],
                         "n_kmers": int(len(k)), "n_matches": int(len(m)),
                         "kmers_sha256": digest(k), "matches_sha256": digest(m)}
        print(name, digests[name]["n_kmers"], digests[name]["n_matches"])
    with open(os.path.join(data, "digests.json"), "w") as f:
        json.dump(digests, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
