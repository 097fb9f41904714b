#!/usr/bin/env python3
"""DB-build inner loop at bench scale: the raw target k-mers of the configs[1] genomes (before the sort)
go through mb_build_index on the GPU; the result must equal, byte for byte, the diffIdx / info / split files of the DB
directory the bench wrote (built by bench_workload.py with torch from the same records) — and, for a bounded sample of
the records, the files written by the REFERENCE's own sort + filterKmers<DB_CREATION> + writeTargetFilesAndSplits
(oracle/_ref/ref_buildidx) on the host cores, which is also the timed CPU baseline.
usage: python scripts/gpu_build_bench.py [--genomes 10000] [--sample 50000000] > line.json"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genomes", type=int, default=10000)
    ap.add_argument("--sample", type=int, default=50_000_000)
    ap.add_argument("--workdir", default="/tmp/mb_bench")
    a = ap.parse_args()
    import torch
    import bench
    import bench_workload as W
    import metabuli_work_b200.binding as mb
    device = torch.device("cuda:0")
    sys.argv = ["bench.py", "--genomes", str(a.genomes), "--workdir", a.workdir]
    ba = bench.parse_args()
    dbdir, meta = bench.prepare_db(ba, device, 0, 1)
    # the raw records, exactly as bench_workload.build_db collects them before its sort
    base = os.path.join(ba.workdir, bench.workload_key(ba))
    tax = W.Taxonomy(*bench.taxonomy_shape(ba.genomes))
    ext2int = {}
    with open(os.path.join(dbdir, "tax_map.tsv")) as f:
        for line in f:
            e, i = line.split()
            ext2int[int(e)] = int(i)
    tax.use_internal_ids(ext2int)
    codes, starts = W.make_genomes(tax, ba.seed, ba.len_min, ba.len_max, device)
    gtax_of = torch.from_numpy(tax.genome_tax).to(device)
    gsp_of = torch.from_numpy(tax.genome_species).to(device)
    vs, ts, ss = [], [], []
    n = len(starts) - 1
    for g0 in range(0, n, 200):
        g1 = min(n, g0 + 200)
        x, y = int(starts[g0]), int(starts[g1])
        value, pos, g_rel, frame = W.syncmer_metamers(codes[x:y], starts[g0:g1 + 1] - x, device, syncmer=bool(ba.syncmer), kmer_format=ba.kmer_format)
        genome = g_rel + g0
        keep = frame == ((genome * 7919 + (pos // 1500) * 104729) >> 3) % 6
        vk, gk = value[keep], genome[keep]
        vs.append(vk.cpu().numpy().view(np.uint64))
        ts.append(gtax_of[gk].to(torch.int32).cpu().numpy())
        ss.append(gsp_of[gk].to(torch.int32).cpu().numpy())
    del codes
    torch.cuda.empty_cache()
    v, t, s = np.concatenate(vs), np.concatenate(ts), np.concatenate(ss)
    del vs, ts, ss
    print(f"[build] {len(v)} raw target k-mers of {n} genomes", file=sys.stderr)
    clf = mb.Classifier(dbdir, seq_mode=2)
    times = []
    for rep in range(3):
        t0 = time.perf_counter()
        words, info, splits = clf.build_index(v, t, s, split_num=4096)
        times.append(time.perf_counter() - t0)
    same = {f: bool(np.array_equal(np.fromfile(os.path.join(dbdir, f), dtype=dt), arr.reshape(-1)))
            for f, dt, arr in (("diffIdx", np.uint16, words), ("info", np.uint32, info), ("split", np.uint64, splits))}
    line = {"metric": "DB-build inner loop: target k-mers/s (sort + LCA de-duplication + delta encoding + split table)", "unit": "k-mers/s",
            "records": int(len(v)), "db_kmers": int(len(info)), "db_words": int(len(words)), "seconds_best": min(times), "seconds_all": times,
            "value": len(v) / min(times), "entry_point": "mb_build_index (host buffers in, host buffers out)",
            "equals_bench_db_files": same}
    # the reference's own code on a bounded sample: files must agree with the GPU's for the same sample, and its time is the baseline
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_buildidx")
    m = min(a.sample, len(v))
    if os.path.exists(harness) and m:
        rec = np.zeros(m, dtype=[("v", "<u8"), ("t", "<i4"), ("s", "<i4")])
        rec["v"], rec["t"], rec["s"] = v[:m], t[:m], s[:m]
        inp = os.path.join(base, "build_sample.bin")
        rec.tofile(inp)
        out = os.path.join(base, "build_ref")
        os.makedirs(out, exist_ok=True)
        threads = os.cpu_count() or 1
        t0 = time.perf_counter()
        subprocess.check_call([harness, dbdir, inp, out, "4096", str(threads)], stdout=subprocess.DEVNULL)
        ref_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        w2, i2, s2 = clf.build_index(v[:m], t[:m], s[:m], split_num=4096)
        gpu_s = time.perf_counter() - t0
        agree = {f: bool(np.array_equal(np.fromfile(os.path.join(out, f), dtype=dt), arr.reshape(-1)))
                 for f, dt, arr in (("diffIdx", np.uint16, w2), ("info", np.uint32, i2), ("split", np.uint64, s2))}
        line["reference_sample"] = {"records": int(m), "threads": threads, "reference_seconds": ref_s, "gpu_seconds": gpu_s,
                                    "reference_kmers_per_s": m / ref_s, "gpu_kmers_per_s": m / gpu_s, "files_equal": agree,
                                    "note": "reference time includes reading the 16-byte records from a file and writing the three files"}
        os.remove(inp)
    clf.close()
    print(json.dumps(line))


if __name__ == "__main__":
    main()
