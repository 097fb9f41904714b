#!/usr/bin/env python3
"""bench.py — reads/s of the classification hot path on BASELINE.json configs[1]:
1M synthetic 2x150 read pairs vs a 10k-genome synthetic SYNCMER database, on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W              (N>1: launched by torch.distributed.run)
  python bench.py --impl reference --gpus N --steps K --warmup W

One "step" = one pass of the hot path (extract -> sort -> match -> sort -> score) over one batch of
n_pairs read pairs per GPU.  `value` is measured with the reads already resident in HBM; `e2e` is the
same step through the public one-call entry point (mb_classify_batch) with pinned HOST buffers, so
the H2D copy of the reads and the D2H copy of the per-read results are inside the timed region.
The database is resident in HBM in its on-disk (compressed) form, as the product keeps it.

Multi-GPU (weak scaling): every rank is the home of its own n_pairs reads; the database is cut by
k-mer range into N shards (at `split` boundaries); query k-mers go to their owner GPU and matches
come back home with two all-to-alls (metabuli_work_b200/exchange.py).

The workload is generated on the box from a seed (bench_workload.py), written to disk as a real
Metabuli database directory and opened by the product's C++ host; it is cached under --workdir so
the reference arm and this arm, run back to back, use byte-identical inputs.

The oracle (oracle/) is used here only as the timed CPU baseline / reference arm (oracle/_ref/ref_harness) and as the checker:
by default the GPU results of the benched reads are compared with the per-read results ref_harness wrote ("check" in the line).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "classified reads/sec (read pairs, 2x150, syncmer DB)"
UNIT = "reads/s"


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5],
                    help="BASELINE.json configs (1-based): 2 = 1M 2x150 pairs vs 10k-genome syncmer DB (the driver's default); "
                         "3 = GTDB-scale taxonomy (130k species), DB sharded over the GPUs; 4 = ONT-like long reads, seq-mode 3; "
                         "5 = contigs vs non-syncmer DB (--legacy: legacy format-1 DB)")
    ap.add_argument("--legacy", action="store_true", help="config 5: legacy format-1 database (OldMetamerScanner)")
    ap.add_argument("--batch-reads", type=int, default=0, help="reads per library batch (0 = the config's default)")
    ap.add_argument("--genomes", type=int, default=0, help="genomes in the synthetic DB (0 = the config's: 10k, config 3: 130k)")
    ap.add_argument("--pairs", type=int, default=0, help="reads (pairs) per GPU per step (0 = the config's: 1M pairs / 1.25M pairs / 100k long reads / 500 contigs)")
    ap.add_argument("--len-min", type=int, default=0)
    ap.add_argument("--len-max", type=int, default=0)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--workdir", default=os.environ.get("MB_BENCH_WORKDIR", "/tmp/mb_bench"))
    ap.add_argument("--cpu-pairs", type=int, default=0, help="read pairs of the bounded CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--check", action="store_true", help="(kept for old command lines: the reference check is on by default)")
    ap.add_argument("--mask", action="store_true", help="--mask 1: tantan low-complexity masking of the reads before extraction (k_mask_mates)")
    ap.add_argument("--no-check", action="store_true", help="skip the comparison of the GPU results with ref_harness' results")
    ap.add_argument("--check-port", action="store_true", help="also compare 20000 reads with the oracle port (liborc)")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="multi-GPU: producers write into peer memory (fused) or bucket + NCCL all-to-all (nccl)")
    ap.add_argument("--no-rebalance", action="store_true", help="multi-GPU: keep the equal-count DB shards (default: after the warm-up the "
                    "shard boundaries are moved so that the owner-side time (k-mer sort + join) is equal, and the shards reloaded)")
    ap.add_argument("--no-fastq", action="store_true", help="skip the input-pipeline leg (FASTQ text parsed on the device)")
    ap.add_argument("--keep", action="store_true", help="keep the generated workload in --workdir")
    a = ap.parse_args()
    apply_config(a)
    return a


def apply_config(a):
    """Fills in the workload of the chosen BASELINE.json config (explicit flags win)."""
    c = {2: dict(label="configs[1]", genomes=10000, pairs=1_000_000, glen=(200_000, 500_000), kind="pairs", seq_mode=2, syncmer=1, fmt=2,
                 rlen=(150, 150), batch=0, cpu=1_000_000),
         3: dict(label="configs[2]", genomes=130_000, pairs=1_250_000, glen=(40_000, 80_000), kind="pairs", seq_mode=2, syncmer=1, fmt=2,
                 rlen=(150, 150), batch=0, cpu=100_000),
         4: dict(label="configs[3]", genomes=10000, pairs=100_000, glen=(200_000, 500_000), kind="long", seq_mode=3, syncmer=1, fmt=2,
                 rlen=(5_000, 20_000), batch=50_000, cpu=5_000),
         5: dict(label="configs[4]", genomes=10000, pairs=500, glen=(200_000, 500_000), kind="contigs", seq_mode=3, syncmer=0,
                 fmt=1 if a.legacy else 2, rlen=(50_000, 500_000), batch=0, cpu=50)}[a.config]
    a.label = c["label"]
    a.genomes = a.genomes or c["genomes"]
    a.pairs = a.pairs or c["pairs"]
    a.len_min = a.len_min or c["glen"][0]
    a.len_max = a.len_max or c["glen"][1]
    a.kind, a.seq_mode, a.syncmer, a.kmer_format, a.rlen = c["kind"], c["seq_mode"], c["syncmer"], c["fmt"], c["rlen"]
    a.batch_reads = a.batch_reads or c["batch"] or a.pairs
    if not a.cpu_pairs:
        a.cpu_pairs = min(a.pairs, c["cpu"])
    a.metric = {"pairs": METRIC, "long": "classified reads/sec (ONT-like long reads 5-20 kb, 5% errors, seq-mode 3, syncmer DB)",
                "contigs": "classified reads/sec (contigs 50-500 kb, seq-mode 3, non-syncmer DB" + (", legacy format 1)" if a.legacy else ")")}[a.kind]


# ---------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------
def taxonomy_shape(n_genomes):
    """families x genera x species x strains with F*G*S*T == n_genomes (10k -> 10 x 10 x 10 x 10)."""
    if n_genomes >= 100_000:                  # GTDB-scale: one genome per species (129,671 species in GTDB R226)
        return 50, 52, n_genomes // 2600, 1
    if n_genomes >= 10000:
        return 10, 10, 10, n_genomes // 1000
    if n_genomes >= 1000:
        return 5, 5, 4, n_genomes // 100
    if n_genomes >= 100:
        return 5, 5, 4, max(1, n_genomes // 100)
    return 2, 2, 2, max(1, n_genomes // 8)


def workload_key(a):
    return f"g{a.genomes}_l{a.len_min}-{a.len_max}_s{a.seed}" + ("" if (a.syncmer, a.kmer_format) == (1, 2) else f"_syn{a.syncmer}_f{a.kmer_format}")


def prepare_db(a, device, rank, world):
    """Rank 0 builds the database directory (if not cached); returns its path."""
    import torch
    import bench_workload as W
    base = os.path.join(a.workdir, workload_key(a))
    dbdir = os.path.join(base, "db")
    marker = os.path.join(base, "db.done")
    if rank == 0 and not os.path.exists(marker):
        os.makedirs(dbdir, exist_ok=True)
        t0 = time.time()
        tax = W.Taxonomy(*taxonomy_shape(a.genomes))
        W.write_taxdump(tax, os.path.join(base, "taxdump"))
        mkdb = os.path.join(ROOT, "oracle", "_ref", "ref_mkdb")
        if not os.path.exists(mkdb):
            raise SystemExit("oracle/_ref/ref_mkdb is missing: run __graft_entry__.build() where /root/reference exists "
                             "(taxonomyDB is written by the reference's own TaxonomyWrapper)")
        subprocess.check_call([mkdb, os.path.join(base, "taxdump"), dbdir], stdout=subprocess.DEVNULL)
        ext2int = {}
        with open(os.path.join(dbdir, "tax_map.tsv")) as f:
            for line in f:
                e, i = line.split()
                ext2int[int(e)] = int(i)
        tax.use_internal_ids(ext2int)
        codes, starts = W.make_genomes(tax, a.seed, a.len_min, a.len_max, device)
        log(f"genomes: {len(starts) - 1}, {int(starts[-1]) / 1e9:.2f} Gbp in {time.time() - t0:.1f}s")
        db = W.build_db(tax, codes, starts, device, log=log, syncmer=bool(a.syncmer), kmer_format=a.kmer_format)
        n_bases = int(starts[-1])
        del codes
        torch.cuda.empty_cache()
        n_words = W.write_db_dir(dbdir, db, tax, syncmer=a.syncmer, kmer_format=a.kmer_format)
        meta = {"genomes": len(starts) - 1, "bases": n_bases, "db_kmers": int(len(db["values"])), "db_words": int(n_words),
                "species": len(tax.species)}
        with open(os.path.join(base, "db.json"), "w") as f:
            json.dump(meta, f)
        del db
        torch.cuda.empty_cache()
        open(marker, "w").close()
        log(f"database written to {dbdir} in {time.time() - t0:.1f}s: {meta}")
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    with open(os.path.join(base, "db.json")) as f:
        meta = json.load(f)
    return dbdir, meta


def prepare_reads(a, device, rank):
    """Reads of home rank `rank` (cached as .npy): bases uint8 ASCII, offsets u64[2n], lens u32[2n]."""
    import bench_workload as W
    base = os.path.join(a.workdir, workload_key(a))
    tag = f"{a.kind}_p{a.pairs}_r{rank}"
    path = os.path.join(base, f"reads_{tag}.npy")
    n = a.pairs
    if os.path.exists(path) and (a.kind == "pairs" or os.path.exists(path[:-4] + "_lens.npy")):
        bases = np.load(path)
        rl = None if a.kind == "pairs" else np.load(path[:-4] + "_lens.npy")
    else:
        import torch
        tax = W.Taxonomy(*taxonomy_shape(a.genomes))
        codes, starts = W.make_genomes(tax, a.seed, a.len_min, a.len_max, device)
        if a.kind == "pairs":
            bases, _, _, _ = W.make_reads(codes, starts, n, a.seed * 1000 + 17 + rank, device)
            rl = None
        else:
            err = dict(sub=0.02, ins=0.01, dele=0.02) if a.kind == "long" else dict(sub=0.001, ins=0.0, dele=0.0)
            bases, _, lens = W.make_long_reads(codes, starts, n, a.seed * 1000 + 29 + rank, device, a.rlen[0], a.rlen[1],
                                               chunk=2000 if a.kind == "long" else 50, **err)
            rl = lens[0::2].copy()
        del codes
        torch.cuda.empty_cache()
        tmp = path + f".tmp{os.getpid()}.npy"
        np.save(tmp, bases)
        os.replace(tmp, path)
        if rl is not None:
            np.save(path[:-4] + "_lens.npy", rl)
    if a.kind == "pairs":
        offsets = np.arange(2 * n, dtype=np.uint64) * np.uint64(150)
        lens = np.full(2 * n, 150, dtype=np.uint32)
    else:
        lens = np.zeros(2 * n, dtype=np.uint32)
        lens[0::2] = rl
        ends = np.cumsum(rl.astype(np.uint64))
        offsets = np.zeros(2 * n, dtype=np.uint64)
        offsets[0::2][1:] = ends[:-1]
        offsets[1::2] = ends
    return bases, offsets, lens


# ---------------------------------------------------------------------------------------------------
# clocks (B200_PROFILING.md: sampled DURING the timed region)
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "samples": len(sm),
                "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's own classes (oracle/_ref/ref_harness), else the oracle port
# ---------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def write_sample_files(a, bases, offsets, lens, n_reads):
    """The first n_reads reads of rank 0 as files for ref_harness: two FASTQ files (pairs) or one FASTA file (long reads)."""
    import bench_workload as W
    base = os.path.join(a.workdir, workload_key(a))
    if a.kind == "pairs":
        p1, p2 = os.path.join(base, f"sample{n_reads}_1.fq"), os.path.join(base, f"sample{n_reads}_2.fq")
        if not (os.path.exists(p1) and os.path.exists(p2)):
            for mate, p in ((0, p1), (1, p2)):
                tmp = p + f".tmp{os.getpid()}"
                fastq_text(bases[:n_reads * 300], n_reads, mate).tofile(tmp)
                os.replace(tmp, p)
        return [p1, p2]
    p = os.path.join(base, f"sample_{a.kind}_p{a.pairs}_{n_reads}.fa")
    if not os.path.exists(p):
        tmp = p + f".tmp{os.getpid()}"
        W.write_fasta(tmp, bases, offsets, lens, n_reads)
        os.replace(tmp, p)
    return [p]


def ref_results_path(a, n_pairs):
    return os.path.join(a.workdir, workload_key(a), f"ref_{a.kind}_p{a.pairs}_n{n_pairs}{'_mask' if a.mask else ''}.results.tsv")


def run_cpu_reference(a, dbdir, reads, n_pairs, reps, skip):
    """-> dict(kind, cores, reads_per_s, s_per_rep, phases, sample, results_tsv).  The reference's own classes
    (oracle/_ref/ref_harness) on all host cores; times extract+sort+match+sort+score (no file parse); the per-read results of
    the first repetition are kept as results_tsv (what the default-on check compares the GPU results with)."""
    threads = host_threads()
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    base = os.path.join(a.workdir, workload_key(a))
    bases, offsets, lens = reads
    if os.path.exists(harness):
        files = write_sample_files(a, bases, offsets, lens, n_pairs)
        # warm the page cache for the DB files the matcher re-reads on every call
        for f in ("diffIdx", "info"):
            with open(os.path.join(dbdir, f), "rb") as fh:
                while fh.read(1 << 26):
                    pass
        prefix = ref_results_path(a, n_pairs)[:-len(".results.tsv")]
        cmd = [harness, dbdir, prefix, str(a.seq_mode), str(threads)] + files + ["--reps", str(reps + skip), "--skip-reps", str(skip)]
        if a.mask:      # masking lives in the reference's own KmerExtractor (which parses the files inside the timed extraction)
            cmd += ["--mask", "1", "--ref-extractor"]
        t0 = time.time()
        out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if out.returncode != 0:
            raise RuntimeError(f"ref_harness failed ({out.returncode}): {out.stderr[-2000:]}")
        js = None
        for ln in out.stdout.splitlines():
            if ln.startswith('{"harness"'):
                js = json.loads(ln)
        if js is None:
            raise RuntimeError("ref_harness printed no JSON line")
        ph = {k: js[k] for k in ("extract_s", "sort_kmers_s", "match_s", "sort_matches_s", "score_s")}
        s_rep = sum(ph.values())
        whole = js["reads"] == a.pairs
        return {"kind": "reference", "cores": threads, "reads_per_s": js["reads"] / s_rep, "s_per_rep": s_rep, "phases": ph,
                "kmers": js["kmers"], "matches": js["matches"], "wall_s": time.time() - t0, "reads": js["reads"],
                "results_tsv": prefix + ".results.tsv", "same_config": whole,
                "sample": (f"all {js['reads']} reads ({a.kind}) of rank 0" if whole else f"first {js['reads']} of the {a.pairs} reads ({a.kind}) of rank 0") +
                          f" vs the full DB; reference KmerMatcher/Taxonomer classes (oracle/_ref/ref_harness), {threads} OpenMP threads, "
                          f"mean of {reps} reps after {skip} warm-up, extract+sort+match+sort+score, file parsing excluded"}
    # oracle port
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    db = orc.Db(dbdir)
    p = db.params(seq_mode=a.seq_mode, threads=threads)
    m1 = [bases[int(offsets[2 * i]):int(offsets[2 * i]) + int(lens[2 * i])].tobytes() for i in range(n_pairs)]
    m2 = [bases[int(offsets[2 * i + 1]):int(offsets[2 * i + 1]) + int(lens[2 * i + 1])].tobytes() for i in range(n_pairs)] if a.kind == "pairs" else None
    batch = orc.Batch(m1, m2)
    times = []
    for r in range(reps + skip):
        t0 = time.time()
        orc.classify(db, p, batch, threads=threads)
        times.append(time.time() - t0)
    db.close()
    s_rep = float(np.mean(times[skip:]))
    return {"kind": "port", "cores": threads, "reads_per_s": n_pairs / s_rep, "s_per_rep": s_rep, "phases": {}, "wall_s": sum(times),
            "reads": n_pairs, "results_tsv": None, "same_config": n_pairs == a.pairs,
            "sample": f"first {n_pairs} of the {a.pairs} read pairs of rank 0 vs the full DB; oracle port (oracle/liborc.so), {threads} threads"}


def check_against_reference(tsv, h_res, h_ids, h_cnt, limit=None):
    """Default-on parity record of the benched config: the per-read results ref_harness wrote (the reference's own
    KmerMatcher / Taxonomer on the same DB directory and reads) vs the GPU results of the same reads — internal tax ID,
    is_classified, query length, fp32 score BITS and the taxID:count list."""
    n = bad = 0
    first_bad = []
    with open(tsv) as f:
        for i, line in enumerate(f):
            if limit is not None and i >= limit:
                break
            c = line.rstrip("\n").split("\t")
            g = h_res[i]
            tg = "".join(f"{t}:{k} " for t, k in zip(h_ids[g["taxcnt_begin"]:g["taxcnt_end"]].tolist(), h_cnt[g["taxcnt_begin"]:g["taxcnt_end"]].tolist()))
            ok = (int(c[0]) == int(g["is_classified"]) and int(c[2]) == int(g["classification"]) and int(c[4]) == int(g["query_length"])
                  and int(c[5], 16) == int(np.float32(g["score"]).view(np.uint32)) and c[6] == tg)
            n += 1
            if not ok:
                bad += 1
                if len(first_bad) < 5:
                    first_bad.append(i)
    out = {"against": "ref_harness", "reads": n, "mismatches": bad}
    if first_bad:
        out["first_mismatching_reads"] = first_bad
    return out


def workload_text(a, meta):
    what = {"pairs": f"{a.pairs} synthetic 2x150 read pairs", "long": f"{a.pairs} synthetic ONT-like long reads ({a.rlen[0]}-{a.rlen[1]} nt, 5% errors)",
            "contigs": f"{a.pairs} synthetic contigs ({a.rlen[0]}-{a.rlen[1]} nt, 0.1% substitutions)"}[a.kind]
    dbk = ("syncmer" if a.syncmer else "non-syncmer") + (" legacy format-1" if a.kmer_format == 1 else "")
    return (f"{a.label}: {what} per GPU vs {meta['genomes']}-genome synthetic {dbk} DB ({meta['db_kmers']} k-mers, {meta['species']} species), "
            f"seq-mode {a.seq_mode}" + (", s=5" if a.syncmer else "") + (", --mask 1" if a.mask else ""))


def cpu_sample_pairs(a):
    return min(a.cpu_pairs, a.pairs)


# ---------------------------------------------------------------------------------------------------
# roofline bookkeeping: ALGORITHMIC bytes per launch of every kernel of the step
# ---------------------------------------------------------------------------------------------------
def algorithmic_bytes(t, sizeof_match=24):
    """t: mb_phase_times of one step (counts of the batch this GPU processed) -> {kernel: bytes per launch}."""
    nb, nk, nm, nr = t["n_bases"], t["n_kmers"], t["n_matches"], t["n_reads"]
    dbw, dbk = t["db_words"], t["db_kmers"]
    return {
        "k_tile_first_mate": 8 * (nb // 2048 + 1),
        "k_extract<count>": nb,                                   # 1 B/base read
        "k_scan_counts": 12 * (nb // 2048 + 1),
        "k_extract": nb + 16 * nk,                                # bases read + 16 B/k-mer written (single pass)
        "k_mask_mates": 3 * nb + 8 * nb + nb // 2,               # bases twice read + once written, a float per base written + read, scales
        "k_sort_segments": 2 * sizeof_match * nm,                 # one read + one write of the grouped matches
        "k_score_fast": sizeof_match * nm + 16 * nr,
        "k_score_finish": sizeof_match * nm // 2 + 48 * nr,       # the best species' matches (about half) + results
        "k_radix_hist<kmer>": 16 * nk,
        "k_radix_scan_hist<kmer>": 5 * 256 * 16,
        "k_radix_pass<kmer>": 32 * nk,                            # one read + one write of the records
        "k_match_partition": 40 * (dbw // 448 + 1),          # a 32-byte unit read + an 8-byte slice start written per unit (<= one per 448 words)
        "k_match": 2 * dbw + 4 * dbk + 16 * nk + sizeof_match * nm,   # stream + info + queries + matches
        "k_radix_hist<match>": sizeof_match * nm,
        "k_radix_scan_hist<match>": 16 * 256 * 16,
        "k_radix_pass<match>": 2 * sizeof_match * nm,
        "k_segment_matches": sizeof_match * nm + 16 * nr,
        "k_segment_max": 16 * nr,
        "k_score": 0,                                             # leftover reads only (about 0.5 %): not rated
        "k_sort_segments_long": 0,
    }


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """dram bytes per launch from the committed ncu capture of this workload (profiles/traffic.json), else None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        with open(p) as f:
            ks = json.load(f).get("kernels", {})
        for name, v in ks.items():                   # ncu names carry template arguments: match on the base name
            if name == kernel or name.split("<")[0] == kernel.split("<")[0] and (("kmer" in name) == ("kmer" in kernel)):
                return int(v["dram_bytes_per_launch"])
    return None


def fastq_text(bases, n_pairs, mate):
    """FASTQ text of one mate file of the synthetic reads: '@r<9 digits>' / 150 bases / '+' / 150 x 'I' per record."""
    L = 150
    w = 12 + L + 1 + 2 + L + 1
    rec = np.empty((n_pairs, w), dtype=np.uint8)
    rec[:, 0] = ord("@"); rec[:, 1] = ord("r")
    idx = np.arange(n_pairs, dtype=np.int64)
    for d in range(9):
        rec[:, 10 - d] = (idx // 10 ** d % 10 + ord("0")).astype(np.uint8)
    rec[:, 11] = 10
    seq = bases.reshape(n_pairs, 2, L)[:, mate, :]
    rec[:, 12:12 + L] = seq
    rec[:, 12 + L] = 10
    rec[:, 13 + L] = ord("+")
    rec[:, 14 + L] = 10
    rec[:, 15 + L:15 + 2 * L] = ord("I")
    rec[:, 15 + 2 * L] = 10
    return rec.reshape(-1)


def fastq_leg(a, torch, clf, bases, n_pairs, h_res, h_ids, h_cnt, stream):
    """mb_stage_fastq on the bench reads written as two FASTQ texts in pinned host memory: parse-only time (text in,
    staged batch out) and the whole step from raw text to results on the host."""
    t1 = torch.from_numpy(fastq_text(bases, n_pairs, 0)).pin_memory()
    t2 = torch.from_numpy(fastq_text(bases, n_pairs, 1)).pin_memory()
    a1, a2 = t1.numpy(), t2.numpy()
    nbytes = a1.nbytes + a2.nbytes
    reps = max(2, a.steps)
    n, c1, c2, nb = clf.stage_fastq(a1, a2)                    # warm-up (allocations)
    if n != n_pairs or nb != len(bases):
        raise RuntimeError(f"device parser staged {n} reads / {nb} bases, expected {n_pairs} / {len(bases)}")
    got_b, _, _ = clf.staged_reads(nb)
    if not np.array_equal(got_b, bases):
        raise RuntimeError("device parser packed different bases than the generator")
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    for _ in range(reps):
        clf.stage_fastq(a1, a2)
    torch.cuda.synchronize()
    parse_ms = (time.perf_counter() - w0) * 1e3 / reps
    # whole step from raw text: the text of the next step is copied on the copy stream while this one is classified
    clf.prefetch_fastq(a1, a2)
    torch.cuda.synchronize()
    w0 = time.perf_counter()
    for _ in range(reps):
        clf.stage_fastq(a1, a2)
        clf.prefetch_fastq(a1, a2)
        clf.classify_staged()
        clf.fetch_into(h_res, h_ids, h_cnt)
    torch.cuda.synchronize()
    full_ms = (time.perf_counter() - w0) * 1e3 / reps
    return {"entry_point": "mb_stage_fastq + mb_prefetch_fastq(next) (+ mb_classify_staged + mb_fetch_results)", "text_bytes_per_step": int(nbytes),
            "parse_ms": round(parse_ms, 3), "parse_text_gbs": round(nbytes / parse_ms / 1e6, 1),
            "parse_note": "pinned host text -> staged batch on the device: H2D of the raw text, line index, record checks, packing; host wall clock",
            "from_text_ms_per_step": round(full_ms, 3), "from_text_reads_per_s": round(n_pairs / (full_ms / 1e3), 1)}


# ---------------------------------------------------------------------------------------------------
# the B200 arm
# ---------------------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run --nproc-per-node N")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    import metabuli_work_b200.binding as mb
    from metabuli_work_b200.exchange import DeviceEngine, ShardedClassifier

    dbdir, meta = prepare_db(a, device, rank, world)
    bases, offsets, lens = prepare_reads(a, device, rank)
    n_pairs = a.pairs
    t0 = time.time()
    clf = mb.Classifier(dbdir, device=local, seq_mode=a.seq_mode, shard=rank, n_shards=world)
    log(f"rank {rank}: shard {rank}/{world} resident ({clf.db_kmers()} k-mers, {clf.db_words()} words) in {time.time() - t0:.1f}s")
    clf.set_profile(True)
    if a.mask:
        clf.set_params(mask_mode=1, mask_prob=0.9)
    sharded = ShardedClassifier(DeviceEngine(clf, device), clf.shard_bounds(), n_pairs, rank, world, fused=a.exchange == "fused") if world > 1 else None

    # Library batches: the whole step in one batch (configs 2, 3, 5), or several (config 4: 100k long reads hold a billion
    # k-mers; the reference splits such inputs as well, QueryIndexer.cpp:62-67).  With several batches every batch is staged
    # from pinned host memory inside the timed region of BOTH legs.
    multi = a.batch_reads < n_pairs
    if multi and world > 1:
        raise SystemExit("configs with several batches per step run on one GPU")
    batches = []
    if multi:
        for s0 in range(0, n_pairs, a.batch_reads):
            s1 = min(n_pairs, s0 + a.batch_reads)
            b0 = int(offsets[2 * s0])
            b1 = int(offsets[2 * s1 - 1]) + int(lens[2 * s1 - 1])
            tb = torch.from_numpy(bases[b0:b1].copy()).pin_memory()
            to = torch.from_numpy((offsets[2 * s0:2 * s1] - np.uint64(b0)).view(np.int64).copy()).pin_memory()
            tl = torch.from_numpy(lens[2 * s0:2 * s1].view(np.int32).copy()).pin_memory()
            batches.append({"n": s1 - s0, "tb": tb, "b": tb.numpy(), "o": to.numpy().view(np.uint64), "l": tl.numpy().view(np.uint32),
                            "keep": (to, tl)})

    # pinned host buffers (inputs and results of the end-to-end leg)
    pin_bases = torch.from_numpy(bases).pin_memory()
    pin_off = torch.from_numpy(offsets.view(np.int64)).pin_memory()
    pin_len = torch.from_numpy(lens.view(np.int32)).pin_memory()
    h_bases = pin_bases.numpy()
    h_off = pin_off.numpy().view(np.uint64)
    h_len = pin_len.numpy().view(np.uint32)
    pin_res = torch.zeros(n_pairs * mb.RESULT_DT.itemsize, dtype=torch.uint8).pin_memory()
    tc_cap = 8 * n_pairs if a.kind == "pairs" else max(64 * n_pairs, 1 << 22)      # (internal taxID, count) entries of one step
    pin_ids = torch.zeros(tc_cap, dtype=torch.int32).pin_memory()
    pin_cnt = torch.zeros(tc_cap, dtype=torch.int32).pin_memory()
    h_res = pin_res.numpy().view(mb.RESULT_DT)
    h_ids, h_cnt = pin_ids.numpy(), pin_cnt.numpy()
    # a second set of page-locked result arrays: the results of step i travel (mb_fetch_results_async) while step i+1 runs
    pin_res2, pin_ids2, pin_cnt2 = torch.zeros_like(pin_res).pin_memory(), torch.zeros_like(pin_ids).pin_memory(), torch.zeros_like(pin_cnt).pin_memory()
    res_sets = [(h_res, h_ids, h_cnt), (pin_res2.numpy().view(mb.RESULT_DT), pin_ids2.numpy(), pin_cnt2.numpy())]

    stream = torch.cuda.ExternalStream(clf.stream(), device=device)

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize(device)

    def max_over_ranks(x):
        if world == 1:
            return x
        import torch.distributed as dist
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    step_counts = {}

    def staged_step():
        if not multi:
            return sharded.classify_staged() if sharded else clf.classify_staged()
        used = 0
        step_counts.clear()
        for bt in batches:
            clf.stage(bt["b"], bt["o"], bt["l"], bases_ptr=bt["tb"].data_ptr())
            used += clf.classify_staged()
            t = clf.timings()
            for k in ("n_bases", "n_kmers", "n_matches", "n_reads", "kernel_launches"):
                step_counts[k] = step_counts.get(k, 0) + t[k]
            for k, (ms, n) in clf.kernel_times().items():
                a_ms, a_n = step_counts.setdefault("kt", {}).get(k, (0.0, 0))
                step_counts["kt"][k] = (a_ms + ms, a_n + n)
        return used

    # end-to-end step: the reads of the step come from pinned host memory, the results go back to host memory.  Two
    # pinned input buffers alternate, and the copy of the NEXT step's reads (mb_prefetch_reads, second stream) overlaps
    # this step's device work; every step still copies its own inputs and reads its own results.
    pin_in = [(pin_bases, h_bases, h_off, h_len)]
    pb2, po2, pl2 = pin_bases.clone().pin_memory(), pin_off.clone().pin_memory(), pin_len.clone().pin_memory()
    pin_in.append((pb2, pb2.numpy(), po2.numpy().view(np.uint64), pl2.numpy().view(np.uint32)))
    e2e_i = [0]

    def e2e_step_multi():
        used = 0
        for i, bt in enumerate(batches):
            nx = batches[(i + 1) % len(batches)]
            clf.stage(bt["b"], bt["o"], bt["l"], bases_ptr=bt["tb"].data_ptr())
            clf.prefetch(nx["b"], nx["o"], nx["l"], bases_ptr=nx["tb"].data_ptr())
            u = clf.classify_staged()
            if "res" not in bt or len(bt["ids"]) < u:
                bt["res"] = np.zeros(bt["n"], dtype=mb.RESULT_DT)
                bt["ids"] = np.zeros(max(2 * u, 1024), dtype=np.int32)
                bt["cnt"] = np.zeros(max(2 * u, 1024), dtype=np.int32)
            clf.fetch_into(bt["res"], bt["ids"], bt["cnt"])
            used += u
        return used

    e2e_pending = [False]         # one GPU: results of the last finished step not yet sent to the host
    e2e_host = {"stage": 0.0, "prefetch": 0.0, "classify": 0.0, "fetch": 0.0, "n": 0}     # MB_BENCH_HOST_TIMES=1: host seconds per call

    def e2e_step():
        if multi:
            return e2e_step_multi()
        cur = pin_in[e2e_i[0] & 1]
        nxt = pin_in[(e2e_i[0] + 1) & 1]
        e2e_i[0] += 1
        if os.environ.get("MB_BENCH_HOST_TIMES") and not sharded:
            t0 = time.perf_counter()
            clf.stage(cur[1], cur[2], cur[3], bases_ptr=cur[0].data_ptr())
            t1 = time.perf_counter()
            clf.prefetch(nxt[1], nxt[2], nxt[3], bases_ptr=nxt[0].data_ptr())
            t2 = time.perf_counter()
            used = clf.classify_staged()
            t3 = time.perf_counter()
            clf.fetch_wait()
            clf.fetch_async(*res_sets[e2e_i[0] & 1])
            t4 = time.perf_counter()
            for k, v in (("stage", t1 - t0), ("prefetch", t2 - t1), ("classify", t3 - t2), ("fetch", t4 - t3)):
                e2e_host[k] += v
            e2e_host["n"] += 1
            return used
        if sharded:
            sharded.stage(cur[1], cur[2], cur[3], bases_ptr=cur[0].data_ptr())
            clf.prefetch(nxt[1], nxt[2], nxt[3], bases_ptr=nxt[0].data_ptr())
            used = sharded.classify_staged()
        else:
            # one GPU: the PREVIOUS step's results start their trip after this step's extraction — at the very start of a step
            # the small synchronous reads of staging and extraction would share the host link with them (0.7 ms measured)
            clf.stage(cur[1], cur[2], cur[3], bases_ptr=cur[0].data_ptr())
            clf.prefetch(nxt[1], nxt[2], nxt[3], bases_ptr=nxt[0].data_ptr())
            clf.extract()
            if e2e_pending[0]:
                clf.fetch_async(*res_sets[(e2e_i[0] + 1) & 1])
            clf.sort_kmers()
            clf.match()
            clf.sort_matches()
            used = clf.score()
            clf.fetch_wait()                               # (long done) — the previous step's results are on the host
            e2e_pending[0] = True                          # this step's results leave during the next step, or after the loop
            return used
        clf.fetch_wait()                                   # the previous step's results have arrived
        clf.fetch_async(*res_sets[e2e_i[0] & 1])           # this step's results travel under the next step's device work
        return used

    # ---- device-resident leg: reads staged once, K passes of the device stages ----
    if sharded:
        sharded.stage(h_bases, h_off, h_len, bases_ptr=pin_bases.data_ptr())
    elif not multi:
        clf.stage(h_bases, h_off, h_len, bases_ptr=pin_bases.data_ptr())
    for _ in range(a.warmup):
        staged_step()
    # Equal k-mer counts per shard are not equal work (shards that begin with six-codon amino acids have larger groups:
    # 21-40 ms of join per shard at configs[2]).  The owner-side wall time of the warm-up (k-mer sort + join) is gathered,
    # the boundaries are moved along the split table so that it comes out equal (piecewise-constant density), and every
    # rank reloads its shard; at most two rounds.
    cut_fractions, rebalance_log = None, []
    if sharded and not a.no_rebalance:
        import torch.distributed as dist
        for _round in range(2):
            t_own = [None] * world
            dist.all_gather_object(t_own, float(sharded.stats["phases_ms"]["sort_match"]))
            rebalance_log.append([round(x, 2) for x in t_own])
            if max(t_own) <= float(os.environ.get("MB_REBALANCE_THRESHOLD", "1.06")) * (sum(t_own) / world):     # (test knob: 0 forces a round)
                break
            moved = mb.rebalanced_cuts(t_own, cut_fractions)
            if not np.all(np.diff(np.concatenate([[0.0], moved, [1.0]])) > 1e-4):      # degenerate times: keep the shards as they are
                break
            cut_fractions = moved
            sharded.close()
            del sharded
            clf.close()
            barrier()
            clf = mb.Classifier(dbdir, device=local, seq_mode=a.seq_mode, shard=rank, n_shards=world, cut_fractions=cut_fractions)
            clf.set_profile(True)
            if a.mask:
                clf.set_params(mask_mode=1, mask_prob=0.9)
            stream = torch.cuda.ExternalStream(clf.stream(), device=device)      # the timing events go on the new context's stream
            sharded = ShardedClassifier(DeviceEngine(clf, device), clf.shard_bounds(), n_pairs, rank, world, fused=a.exchange == "fused")
            sharded.stage(h_bases, h_off, h_len, bases_ptr=pin_bases.data_ptr())
            for _ in range(a.warmup):
                staged_step()
        if cut_fractions is not None:
            log(f"rank {rank}: shards re-balanced, boundaries at {[round(float(x), 4) for x in cut_fractions]}; owner-side ms per round {rebalance_log}")
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(a.steps):
        used = staged_step()
    ev1.record(stream)
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    t_step = clf.timings()
    ktimes = clf.kernel_times()
    if multi:                                  # counts and kernel times of the whole step = sums over its batches
        t_step = dict(t_step, **{k: v for k, v in step_counts.items() if k != "kt"})
        ktimes = step_counts["kt"]
    launches_per_step = t_step["kernel_launches"]
    stats = dict(sharded.stats) if sharded else {}
    if sharded and world > 1:
        # every rank's wall-clock split and leftover-scorer load of the last timed step: where an imbalance comes from
        import torch.distributed as dist
        mine = dict(stats["phases_ms"], k_match=round(ktimes.get("k_match", (0, 0))[0], 2), k_score=round(ktimes.get("k_score", (0, 0))[0], 2),
                    k_sort_segments=round(ktimes.get("k_sort_segments", (0, 0))[0], 2), kmers_owned=stats["kmers_owned"], matches_owned=stats["matches_owned"])
        allr = [None] * world
        dist.all_gather_object(allr, mine)
        stats["by_rank"] = allr

    # ---- end-to-end leg: host buffers in, host results out, every step ----
    for _ in range(max(1, a.warmup // 2)):
        e2e_step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    w0 = time.perf_counter()
    for _ in range(a.steps):
        used = e2e_step()
    if e2e_pending[0]:                                     # one GPU: the last step's results leave now
        clf.fetch_async(*res_sets[e2e_i[0] & 1])
        e2e_pending[0] = False
    clf.fetch_wait()                                       # the last step's results are on the host: inside the timed region
    if e2e_host["n"]:
        log("e2e host ms per step: " + ", ".join(f"{k} {1e3 * v / e2e_host['n']:.3f}" for k, v in e2e_host.items() if k != "n"))
    e1.record(stream)
    barrier()
    wall_e2e = time.perf_counter() - w0
    if not multi:
        h_res, h_ids, h_cnt = res_sets[e2e_i[0] & 1]       # where the last step's results went
    ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), wall_e2e * 1e3))
    clocks = sampler.stop()
    if multi:
        classified = int(sum(int(bt["res"]["is_classified"].sum()) for bt in batches))
        h_res, h_ids, h_cnt = batches[0]["res"], batches[0]["ids"], batches[0]["cnt"]     # what the check compares: the first batch
    else:
        classified = int(h_res["is_classified"].sum())
    t_e2e = clf.timings()

    ms_step = ms_total / a.steps
    value = world * n_pairs / (ms_step / 1e3)
    e2e_value = world * n_pairs / (ms_e2e / a.steps / 1e3)
    h2d = len(h_bases) + h_off.nbytes + h_len.nbytes
    d2h = h_res.nbytes + 8 * int(used)

    # ---- roofline of the dominant kernel of the step (per-launch CUDA-event times of the last timed step) ----
    peak, peak_src = measured_peak()
    alg = algorithmic_bytes(t_step)
    if multi:                                  # the counts are sums over the step's batches: per launch = per batch
        alg = {k: v // len(batches) for k, v in alg.items()}
    table = {}
    for k, (ms, n) in ktimes.items():
        if n == 0:
            continue
        per = ms / n
        table[k] = {"ms": round(ms, 4), "launches": n, "alg_bytes_per_launch": int(alg.get(k, 0)),
                    "gbs": round(alg.get(k, 0) / (per * 1e-3) / 1e9, 1) if per > 0 else None}
    dom = max(table, key=lambda k: table[k]["ms"]) if table else None
    roofline = None
    if dom:
        ach = table[dom]["gbs"]
        roofline = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                    "traffic": ncu_traffic(dom), "peak_source": peak_src, "launch_ms": round(table[dom]["ms"] / table[dom]["launches"], 4),
                    "launches_per_step": table[dom]["launches"], "alg_bytes_per_launch": table[dom]["alg_bytes_per_launch"]}
    if "k_match" in table:
        m = table["k_match"]
        match_phase = {"kernel": "k_match", "achieved": m["gbs"], "peak": peak, "unit": "GB/s", "frac": round((m["gbs"] or 0) / peak, 4),
                       "launch_ms": round(m["ms"] / m["launches"], 4), "alg_bytes_per_launch": m["alg_bytes_per_launch"],
                       "traffic": ncu_traffic("k_match")}
    else:
        match_phase = None

    out = {
        "metric": a.metric, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {"workload": workload_text(a, meta),
                   "batches_per_step": max(1, len(batches)),
                   "pairs_per_gpu": n_pairs, "db_kmers": meta["db_kmers"], "db_words": meta["db_words"], "genomes": meta["genomes"],
                   "db_sharding": (f"{world} k-mer-range shard(s), one per GPU" + (f"; boundaries moved to equalise the owner-side time "
                                   f"(fractions of the split table {[round(float(x), 4) for x in cut_fractions]}, owner-side ms per "
                                   f"round {rebalance_log})" if cut_fractions is not None else "")) if world > 1 else "whole DB on one GPU",
                   "exchange": (a.exchange if world > 1 else None),
                   "l2": "inputs larger than L2 (k-mer buffer and DB stream are GBs vs 126 MB L2); no flush needed",
                   "seed": a.seed},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": round(ms_e2e / a.steps, 3), "entry_point": "mb_stage_reads + mb_prefetch_reads(next) + mb_classify_staged + mb_fetch_results_async / mb_fetch_wait (pinned host buffers; "
                               "every step copies its own reads in and its own results out, both overlapped with neighbouring steps)"},
        "gpu_launches": int(launches_per_step) * a.steps,
        "roofline": roofline,
        "match_phase": match_phase,
        "kernels": table,
        "phases_ms": {k: round(t_step[k], 3) for k in ("extract_ms", "sort_kmers_ms", "match_ms", "sort_matches_ms", "score_ms")},
        "e2e_phases_ms": {k: round(t_e2e[k], 3) for k in ("h2d_ms", "extract_ms", "sort_kmers_ms", "match_ms", "sort_matches_ms", "score_ms", "d2h_ms")},
        "counts": {"kmers": t_step["n_kmers"], "matches": t_step["n_matches"], "reads": t_step["n_reads"], "classified": classified,
                   "match_retries": t_step["match_retries"], **stats},
    }

    # ---- input pipeline: the same reads as raw FASTQ text, parsed and packed on the device ----
    if world == 1 and not a.no_fastq and a.kind == "pairs" and not multi:
        try:
            out["input_pipeline"] = fastq_leg(a, torch, clf, bases, n_pairs, h_res, h_ids, h_cnt, stream)
        except Exception as ex:  # reported, never required
            out["input_pipeline"] = {"error": str(ex)}

    if rank == 0:
        n_cpu = cpu_sample_pairs(a)
        tsv = ref_results_path(a, n_cpu)
        if not a.no_cpu_baseline and world == 1:
            try:
                cb = run_cpu_reference(a, dbdir, (bases, offsets, lens), n_cpu, reps=2, skip=1)
                out["cpu_baseline"] = {"value": round(cb["reads_per_s"], 1), "unit": UNIT, "cores": cb["cores"], "kind": cb["kind"],
                                       "sample": cb["sample"], "same_config": cb["same_config"], "phases_s": cb["phases"],
                                       "wall_s": round(cb["wall_s"], 1)}
            except Exception as ex:  # the baseline is reported, never required
                out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "reference", "sample": f"failed: {ex}"}
        if not a.no_check:
            try:
                if not os.path.exists(tsv):          # N > 1, or the baseline leg was skipped: one untimed pass of the harness
                    run_cpu_reference(a, dbdir, (bases, offsets, lens), n_cpu, reps=1, skip=0)
                out["check"] = check_against_reference(tsv, h_res, h_ids, h_cnt, limit=len(h_res))
            except Exception as ex:
                out["check"] = {"against": "ref_harness", "error": str(ex)}
        if a.check_port:
            out["check_port"] = check_against_oracle(a, dbdir, bases, h_res, h_ids, h_cnt, clf, mb, world)
    clf.close()
    barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def check_against_oracle(a, dbdir, bases, h_res, h_ids, h_cnt, clf, mb, world):
    """--check: the first pairs of rank 0 through the oracle (CPU restatement) vs the GPU results of the same reads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    n = min(a.pairs, 20000)
    db = orc.Db(dbdir)
    p = db.params(seq_mode=2)
    b = bases[:n * 300].reshape(n, 2, 150)
    batch = orc.Batch([x[0].tobytes() for x in b], [x[1].tobytes() for x in b])
    _, om, ores, oids, ocnt = orc.classify(db, p, batch, threads=host_threads())
    db.close()
    bad = 0
    for i in range(n):
        g, o = h_res[i], ores[i]
        tg = list(zip(h_ids[g["taxcnt_begin"]:g["taxcnt_end"]].tolist(), h_cnt[g["taxcnt_begin"]:g["taxcnt_end"]].tolist()))
        to = list(zip(oids[o["taxcnt_begin"]:o["taxcnt_end"]].tolist(), ocnt[o["taxcnt_begin"]:o["taxcnt_end"]].tolist()))
        if not (g["classification"] == o["classification"] and g["is_classified"] == o["is_classified"]
                and np.float32(g["score"]).view(np.uint32) == np.float32(o["score"]).view(np.uint32) and tg == to):
            bad += 1
    return {"reads_compared": n, "mismatches": bad, "oracle_matches": int(len(om)), "oracle_classified": int(ores["is_classified"].sum())}


# ---------------------------------------------------------------------------------------------------
# the reference arm: the reference's own CPU classes on the same workload, bounded sample per step
# ---------------------------------------------------------------------------------------------------
def run_reference(a):
    """The reference's own CPU implementation of the path (oracle/_ref/ref_harness: its KmerMatcher, Taxonomer, scanners and
    ips4o sort, all host cores) on this arm's config.  Nothing of metabuli_work_b200 is imported here: the workload
    generator is plain torch (bench_workload.py)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))) if torch.cuda.is_available() else None
    base = os.path.join(a.workdir, workload_key(a))
    if device is None and not os.path.exists(os.path.join(base, "db.done")):
        print(json.dumps({"impl": "reference", "unavailable": "the synthetic workload is generated on the GPU and no CUDA device is visible"}))
        return
    if device is not None:
        torch.cuda.set_device(device)
    dbdir, meta = prepare_db(a, device, 0, 1)
    reads = prepare_reads(a, device, 0)
    n = cpu_sample_pairs(a)
    cb = run_cpu_reference(a, dbdir, reads, n, reps=a.steps, skip=a.warmup)
    value = cb["reads_per_s"]
    out = {
        "impl": "reference", "metric": a.metric, "value": round(value, 1), "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": round(cb["s_per_rep"] * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": workload_text(a, meta),
                   "pairs_per_gpu": a.pairs, "db_kmers": meta["db_kmers"], "db_words": meta["db_words"], "genomes": meta["genomes"],
                   "step": f"one pass of the reference hot path over {cb['reads']} read pairs (whole DB scanned)",
                   "same_config": cb["same_config"], "seed": a.seed},
        "cpu_baseline": {"value": round(value, 1), "unit": UNIT, "cores": cb["cores"], "kind": cb["kind"], "sample": cb["sample"],
                         "same_config": cb["same_config"], "phases_s": cb["phases"]},
        "e2e": {"value": round(value, 1), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def main():
    a = parse_args()
    try:
        if a.impl == "reference":
            run_reference(a)
        else:
            run_b200(a)
    finally:
        if not a.keep and os.environ.get("MB_BENCH_KEEP", "1") == "0":
            import shutil
            shutil.rmtree(os.path.join(a.workdir, workload_key(a)), ignore_errors=True)


if __name__ == "__main__":
    main()
