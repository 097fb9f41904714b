#!/usr/bin/env python3
"""Round-2, second batch of golden fixtures (run in the build container only; ADDS to tests/golden/data):

  * for EVERY existing case the reference's own KmerExtractor (QueryIndexer::indexQueryFile +
    KmerExtractor::extractQueryKmers, compiled from src/commons/KmerExtractor.cpp: ref_harness --ref-extractor) must give
    the committed k-mer / match digests and per-read results — the read loop restated in the harness is no longer the only
    witness of the extraction;
  * `--mask 1` (tantan low-complexity masking, SeqIterator.cpp:154-177): reads with implanted tandem repeats / homopolymers
    (mask_1.fq, mask_2.fq, mask_long.fa), the masked sequences the reference's SeqIterator::maskLowComplexityRegions
    produces for them (mask_*.masked.txt, --dump-masked), the likelihood-ratio matrix and letter codes it works with
    (mask_lr.bin, --dump-lr), and the classification results / Reporter files of the cases
        mask_paired (db_sync, --mask 1), mask_paired_p05 (db_sync, --mask 1 --mask-prob 0.5), mask_long (db_plain, --mask 1)
    run through the reference extractor.
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

PRELUDE = "/root/reference/lib/mmseqs/data/resources/krona_prelude.html"
MASK_CASES = [  # name, db, seq_mode, reads, flags
    ("mask_paired", "db_sync", 2, ("mask_1.fq", "mask_2.fq"), ["--mask", "1"]),
    ("mask_paired_p05", "db_sync", 2, ("mask_1.fq", "mask_2.fq"), ["--mask", "1", "--mask-prob", "0.5"]),
    ("mask_long", "db_plain", 3, ("mask_long.fa",), ["--mask", "1"]),
]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run(cmd):
    return json.loads(subprocess.check_output(cmd).decode().strip().splitlines()[-1])


def implant(rng, seq):
    """A low-complexity stretch somewhere in the read: homopolymer, short tandem repeat, or a diverged repeat."""
    s = bytearray(seq)
    kind = rng.integers(0, 5)
    L = len(s)
    if L < 24:
        return bytes(s)
    n = int(rng.integers(12, min(90, L - 4)))
    at = int(rng.integers(0, L - n))
    if kind == 0:
        unit = bytes([b"ACGT"[rng.integers(0, 4)]])
    else:
        unit = bytes(b"ACGT"[rng.integers(0, 4)] for _ in range(int(rng.integers(2, 9 if kind < 4 else 30))))
    rep = (unit * (n // len(unit) + 1))[:n]
    rep = bytearray(rep)
    if kind >= 3:                       # a few substitutions inside the repeat
        for _ in range(max(1, n // 15)):
            rep[int(rng.integers(0, n))] = b"ACGT"[rng.integers(0, 4)]
    s[at:at + n] = rep
    if rng.random() < 0.15:             # ambiguity codes and lower case travel through the letter mapping
        p = int(rng.integers(0, L))
        s[p] = b"NRYKMSWnacgtX-"[rng.integers(0, 14)]
    return bytes(s)


def main():
    data = os.path.join(HERE, "data")
    work = "/tmp/mb_golden_work3"
    shutil.rmtree(work, ignore_errors=True)
    os.makedirs(work)
    with open(os.path.join(data, "digests.json")) as f:
        digests = json.load(f)

    # 1. the reference's own extractor reproduces every committed case
    for name in sorted(digests):
        d = digests[name]
        if name.startswith("mask_"):
            continue
        flags = ["--smer-len", str(d["smer_len"])] if "smer_len" in d else []
        prefix = os.path.join(work, name)
        run([orc.REF_HARNESS, os.path.join(data, d["db"]), prefix, str(d["seq_mode"]), "2"] + [os.path.join(data, r) for r in d["reads"]] +
            flags + ["--dump", "--ref-extractor"])
        k = orc.sort_kmers(np.fromfile(prefix + ".kmers.bin", dtype=orc.KMER_DT))
        m = np.fromfile(prefix + ".matches.bin", dtype=orc.MATCH_DT)
        assert digest(k) == d["kmers_sha256"] and digest(m) == d["matches_sha256"], name
        assert open(prefix + ".results.tsv", "rb").read() == open(os.path.join(data, name + ".results.tsv"), "rb").read(), name
        print("reference KmerExtractor ==", name)

    # 2. reads with low-complexity stretches
    rng = np.random.default_rng(4321)
    def fastq(path):
        rec = open(os.path.join(data, path)).read().split("\n")
        return [(rec[i], rec[i + 1]) for i in range(0, len(rec) - 3, 4)]
    for src, dst in (("reads_1.fq", "mask_1.fq"), ("reads_2.fq", "mask_2.fq")):
        with open(os.path.join(data, dst), "w") as f:
            for j, (h, s) in enumerate(fastq(src)[:200]):
                s2 = implant(rng, s.encode()).decode() if j % 4 != 3 else s
                f.write(f"{h}\n{s2}\n+\n{'I' * len(s2)}\n")
    fa = open(os.path.join(data, "long.fa")).read().split(">")[1:]
    with open(os.path.join(data, "mask_long.fa"), "w") as f:
        for rec in fa:
            h, s = rec.split("\n", 1)
            s = s.replace("\n", "").encode()
            for _ in range(6):
                at = int(rng.integers(0, len(s) - 200))
                s = s[:at] + implant(rng, s[at:at + 200]) + s[at + 200:]
            f.write(f">{h}\n{s.decode()}\n")

    # 3. what the reference does with them
    subprocess.check_call([orc.REF_HARNESS, "--dump-lr", os.path.join(data, "mask_lr.bin")])
    prelude_len = os.path.getsize(PRELUDE)
    rep = os.path.join(data, "reports")
    for name, db, mode, reads, flags in MASK_CASES:
        prefix = os.path.join(work, name)
        files = [os.path.join(data, r) for r in reads]
        base = [orc.REF_HARNESS, os.path.join(data, db), prefix, str(mode), "2"] + files + flags + ["--ref-extractor"]
        info = run(base + ["--dump", "--dump-masked", os.path.join(data, name + ".masked.txt")])
        k = orc.sort_kmers(np.fromfile(prefix + ".kmers.bin", dtype=orc.KMER_DT))
        m = np.fromfile(prefix + ".matches.bin", dtype=orc.MATCH_DT)
        shutil.copy(prefix + ".results.tsv", os.path.join(data, name + ".results.tsv"))
        digests[name] = {"db": db, "seq_mode": mode, "reads": list(reads), "n_reads": info["reads"], "n_kmers": int(len(k)),
                         "n_matches": int(len(m)), "kmers_sha256": digest(k), "matches_sha256": digest(m), "mask": 1,
                         "mask_prob": float(flags[3]) if len(flags) > 2 else 0.9}
        # the same reads without masking must differ (else the case tests nothing)
        info0 = run([orc.REF_HARNESS, os.path.join(data, db), prefix + "_nomask", str(mode), "2"] + files + ["--ref-extractor"])
        assert info0["kmers"] > info["kmers"], (name, info0["kmers"], info["kmers"])
        out = os.path.join(work, "rep_" + name)
        os.makedirs(out, exist_ok=True)
        run(base + ["--report", out, "plain"])
        run(base + ["--report", out, "lineage", "--lineage"])
        for f in ("classifications.tsv", "report.tsv"):
            shutil.copy(os.path.join(out, "plain_" + f), os.path.join(rep, name + "_" + f))
        shutil.copy(os.path.join(out, "lineage_classifications.tsv"), os.path.join(rep, name + "_lineage_classifications.tsv"))
        krona = open(os.path.join(out, "plain_krona.html"), "rb").read()
        with open(os.path.join(rep, name + "_krona_tail.html"), "wb") as f:
            f.write(krona[prelude_len:])
        print(name, "k-mers", info["kmers"], "of", info0["kmers"], "unmasked; matches", info["matches"])
    with open(os.path.join(data, "digests.json"), "w") as f:
        json.dump(digests, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
