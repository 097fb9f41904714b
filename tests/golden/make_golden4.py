#!/usr/bin/env python3
"""DB-build inner loop fixtures (run in the build container only; ADDS to tests/golden/data):

  build_input.npz   raw target k-mers (value, internal taxId, internal speciesId), unsorted, of the first genomes of the
                    fixture genome set (strains of one species share k-mers, so (value, species) duplicates with different
                    taxa are present, plus a block of exact duplicates and k-mers shared ACROSS species added below)
  build_digests.json  sha256 / sizes of the diffIdx, info and split files that the REFERENCE's own
                    SORT_PARALLEL(compareTargetKmer) -> IndexCreator::filterKmers<DB_CREATION> ->
                    IndexCreator::writeTargetFilesAndSplits wrote for those records (oracle/_ref/ref_buildidx, taxonomy of
                    db_plain), for split counts 64 and 4096 — identical for 1 and 4 threads
  build_split64.bin the split file itself (64 entries)
The oracle's orc_build_index must write the same three files (asserted here and in tests/test_oracle_golden.py)."""
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
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402

REF_BUILDIDX = os.path.join(ROOT, "oracle", "_ref", "ref_buildidx")


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def main():
    data = os.path.join(HERE, "data")
    work = "/tmp/mb_golden_work4"
    shutil.rmtree(work, ignore_errors=True)
    subprocess.check_call([sys.executable, os.path.join(HERE, "make_inputs.py"), work + "/plain", "--seed", "1234", "--genome-len", "12000",
                           "--pairs", "400", "--long-reads", "8"])
    v, t, ti, si, genome_tax, ext2int, lca = synth.target_records(work + "/plain", work + "/taxdb", syncmer=0, kmer_format=2, max_genomes=24)
    rng = np.random.default_rng(99)
    # exact duplicates (same value, species, taxon) and k-mers shared across species (same value, other species)
    dup = rng.choice(len(v), 3000, replace=False)
    other = rng.choice(len(v), 3000, replace=False)
    donor = rng.choice(len(v), 3000, replace=False)
    v = np.concatenate([v, v[dup], v[other]])
    ti = np.concatenate([ti, ti[dup], ti[donor]])
    si = np.concatenate([si, si[dup], si[donor]])
    perm = rng.permutation(len(v))
    v, ti, si = v[perm].astype(np.uint64), ti[perm].astype(np.int32), si[perm].astype(np.int32)
    np.savez_compressed(os.path.join(data, "build_input.npz"), value=v, taxid=ti, species=si)
    rec = np.zeros(len(v), dtype=[("v", "<u8"), ("t", "<i4"), ("s", "<i4")])
    rec["v"], rec["t"], rec["s"] = v, ti, si
    rec.tofile(os.path.join(work, "input.bin"))
    taxdir = os.path.join(data, "db_plain")
    # the taxonomy written for these genomes is the fixture taxonomy (same taxdump, same ref_mkdb)
    assert open(os.path.join(work, "taxdb", "taxonomyDB"), "rb").read() == open(os.path.join(taxdir, "taxonomyDB"), "rb").read()
    out = {"records": int(len(v))}
    db = orc.Db(taxdir)
    for splits in (64, 4096):
        dirs = []
        for threads in (1, 4):
            d = os.path.join(work, f"ref_s{splits}_t{threads}")
            os.makedirs(d)
            info = json.loads(subprocess.check_output([REF_BUILDIDX, taxdir, os.path.join(work, "input.bin"), d, str(splits), str(threads)])
                              .decode().strip().splitlines()[-1])
            dirs.append(d)
        for f in ("diffIdx", "info", "split"):
            assert open(os.path.join(dirs[0], f), "rb").read() == open(os.path.join(dirs[1], f), "rb").read(), (splits, f)
        od = os.path.join(work, f"orc_s{splits}")
        n = orc.build_index(db, v, ti, si, od, split_num=splits)
        assert n == info["selected"], (n, info)
        for f in ("diffIdx", "info", "split"):
            assert open(os.path.join(od, f), "rb").read() == open(os.path.join(dirs[0], f), "rb").read(), ("oracle != reference", splits, f)
        out[f"splits{splits}"] = {"db_kmers": int(info["selected"]), "diffIdx_sha256": sha(os.path.join(dirs[0], "diffIdx")),
                                  "info_sha256": sha(os.path.join(dirs[0], "info")), "split_sha256": sha(os.path.join(dirs[0], "split")),
                                  "diffIdx_bytes": os.path.getsize(os.path.join(dirs[0], "diffIdx"))}
        if splits == 64:
            shutil.copy(os.path.join(dirs[0], "split"), os.path.join(data, "build_split64.bin"))
        print("splits", splits, "records", len(v), "->", info["selected"], "DB k-mers; oracle == reference (1 and 4 threads)")
    # edge records (helpers.build_edge_records: five-word deltas, a giant amino-acid group over several split thresholds)
    ev, et, es = helpers.build_edge_records(ti, si)
    erec = np.zeros(len(ev), dtype=rec.dtype)
    erec["v"], erec["t"], erec["s"] = ev, et, es
    erec.tofile(os.path.join(work, "edge.bin"))
    d = os.path.join(work, "ref_edge")
    os.makedirs(d)
    info = json.loads(subprocess.check_output([REF_BUILDIDX, taxdir, os.path.join(work, "edge.bin"), d, "64", "2"]).decode().strip().splitlines()[-1])
    od = os.path.join(work, "orc_edge")
    assert orc.build_index(db, ev, et, es, od, split_num=64) == info["selected"]
    for f in ("diffIdx", "info", "split"):
        assert open(os.path.join(od, f), "rb").read() == open(os.path.join(d, f), "rb").read(), ("oracle != reference (edge)", f)
    nsplit = int((np.fromfile(os.path.join(d, "split"), dtype=np.uint64).reshape(-1, 3)[:, 0] != 0).sum())
    out["edge_splits64"] = {"db_kmers": int(info["selected"]), "diffIdx_sha256": sha(os.path.join(d, "diffIdx")), "info_sha256": sha(os.path.join(d, "info")),
                            "split_sha256": sha(os.path.join(d, "split")), "used_split_entries": nsplit}
    print("edge records", len(ev), "->", info["selected"], "DB k-mers,", nsplit, "split entries used of 63; oracle == reference")
    db.close()
    with open(os.path.join(data, "build_digests.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
