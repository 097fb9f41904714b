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
    ap.add_argument("--genomes", type=int, default=10000, help="genomes in the synthetic DB (configs[1]: 10k)")
    ap.add_argument("--pairs", type=int, default=1_000_000, help="read pairs per GPU per step (configs[1]: 1M)")
    ap.add_argument("--len-min", type=int, default=200_000)
    ap.add_argument("--len-max", type=int, default=500_000)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--workdir", default=os.environ.get("MB_BENCH_WORKDIR", "/tmp/mb_bench"))
    ap.add_argument("--cpu-pairs", type=int, default=0, help="read pairs of the bounded CPU sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--check", action="store_true", help="(kept for old command lines: the reference check is on by default)")
    ap.add_argument("--no-check", action="store_true", help="skip the comparison of the GPU results with ref_harness' results")
    ap.add_argument("--check-port", action="store_true", help="also compare 20000 reads with the oracle port (liborc)")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="multi-GPU: producers write into peer memory (fused) or bucket + NCCL all-to-all (nccl)")
    ap.add_argument("--no-fastq", action="store_true", help="skip the input-pipeline leg (FASTQ text parsed on the device)")
    ap.add_argument("--keep", action="store_true", help="keep the generated workload in --workdir")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------
def taxonomy_shape(n_genomes):
    """families x genera x species x strains with F*G*S*T == n_genomes (10k -> 10 x 10 x 10 x 10)."""
    if n_genomes >= 10000:
        return 10, 10, 10, n_genomes // 1000
    if n_genomes >= 1000:
        return 5, 5, 4, n_genomes // 100
    if n_genomes >= 100:
        return 5, 5, 4, max(1, n_genomes // 100)
    return 2, 2, 2, max(1, n_genomes // 8)


def workload_key(a):
    return f"g{a.genomes}_l{a.len_min}-{a.len_max}_s{a.seed}"


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
        db = W.build_db(tax, codes, starts, device, log=log)
        n_words = W.write_db_dir(dbdir, db, tax)
        meta = {"genomes": len(starts) - 1, "bases": int(starts[-1]), "db_kmers": int(len(db["values"])), "db_words": int(n_words),
                "species": len(tax.species)}
        with open(os.path.join(base, "db.json"), "w") as f:
            json.dump(meta, f)
        del db, codes
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
    path = os.path.join(base, f"reads_p{a.pairs}_r{rank}.npy")
    if os.path.exists(path):
        bases = np.load(path)
    else:
        import torch
        tax = W.Taxonomy(*taxonomy_shape(a.genomes))
        codes, starts = W.make_genomes(tax, a.seed, a.len_min, a.len_max, device)
        bases, _, _, _ = W.make_reads(codes, starts, a.pairs, a.seed * 1000 + 17 + rank, device)
        del codes
        torch.cuda.empty_cache()
        tmp = path + f".tmp{os.getpid()}.npy"
        np.save(tmp, bases)
        os.replace(tmp, path)
    n = a.pairs
    offsets = np.arange(2 * n, dtype=np.uint64) * np.uint64(150)
    lens = np.full(2 * n, 150, dtype=np.uint32)
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


def write_sample_fastq(a, bases, n_pairs):
    base = os.path.join(a.workdir, workload_key(a))
    p1, p2 = os.path.join(base, f"sample{n_pairs}_1.fq"), os.path.join(base, f"sample{n_pairs}_2.fq")
    if not (os.path.exists(p1) and os.path.exists(p2)):
        for mate, p in ((0, p1), (1, p2)):
            tmp = p + f".tmp{os.getpid()}"
            fastq_text(bases[:n_pairs * 300], n_pairs, mate).tofile(tmp)
            os.replace(tmp, p)
    return p1, p2


def ref_results_path(a, n_pairs):
    return os.path.join(a.workdir, workload_key(a), f"ref_p{a.pairs}_n{n_pairs}.results.tsv")


def run_cpu_reference(a, dbdir, bases, n_pairs, reps, skip):
    """-> dict(kind, cores, reads_per_s, s_per_rep, phases, sample, results_tsv).  The reference's own classes
    (oracle/_ref/ref_harness) on all host cores; times extract+sort+match+sort+score (no file parse); the per-read results of
    the first repetition are kept as results_tsv (what the default-on check compares the GPU results with)."""
    threads = host_threads()
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    base = os.path.join(a.workdir, workload_key(a))
    if os.path.exists(harness):
        p1, p2 = write_sample_fastq(a, bases, n_pairs)
        # warm the page cache for the DB files the matcher re-reads on every call
        for f in ("diffIdx", "info"):
            with open(os.path.join(dbdir, f), "rb") as fh:
                while fh.read(1 << 26):
                    pass
        prefix = ref_results_path(a, n_pairs)[:-len(".results.tsv")]
        cmd = [harness, dbdir, prefix, "2", str(threads), p1, p2, "--reps", str(reps + skip), "--skip-reps", str(skip)]
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
                "sample": (f"all {js['reads']} read pairs of rank 0" if whole else f"first {js['reads']} of the {a.pairs} read pairs of rank 0") +
                          f" vs the full DB; reference KmerMatcher/Taxonomer classes (oracle/_ref/ref_harness), {threads} OpenMP threads, "
                          f"mean of {reps} reps after {skip} warm-up, extract+sort+match+sort+score, file parsing excluded"}
    # oracle port
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    db = orc.Db(dbdir)
    p = db.params(seq_mode=2, threads=threads)
    b = bases[:n_pairs * 300].reshape(n_pairs, 2, 150)
    batch = orc.Batch([x[0].tobytes() for x in b], [x[1].tobytes() for x in b])
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


def check_against_reference(tsv, h_res, h_ids, h_cnt):
    """Default-on parity record of the benched config: the per-read results ref_harness wrote (the reference's own
    KmerMatcher / Taxonomer on the same DB directory and reads) vs the GPU results of the same reads — internal tax ID,
    is_classified, query length, fp32 score BITS and the taxID:count list."""
    n = bad = 0
    first_bad = []
    with open(tsv) as f:
        for i, line in enumerate(f):
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


def cpu_sample_pairs(a):
    if a.cpu_pairs:
        return min(a.cpu_pairs, a.pairs)
    return min(a.pairs, 1_000_000)


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
        "k_sort_segments": 2 * sizeof_match * nm,                 # one read + one write of the grouped matches
        "k_score_fast": sizeof_match * nm + 16 * nr,
        "k_score_finish": sizeof_match * nm // 2 + 48 * nr,       # the best species' matches (about half) + results
        "k_radix_hist<kmer>": 16 * nk,
        "k_radix_scan_hist<kmer>": 5 * 256 * 16,
        "k_radix_pass<kmer>": 32 * nk,                            # one read + one write of the records
        "k_match_partition": 8 * (dbw // 2048 + 1),
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
    clf = mb.Classifier(dbdir, device=local, seq_mode=2, shard=rank, n_shards=world)
    log(f"rank {rank}: shard {rank}/{world} resident ({clf.db_kmers()} k-mers, {clf.db_words()} words) in {time.time() - t0:.1f}s")
    clf.set_profile(True)
    sharded = ShardedClassifier(DeviceEngine(clf, device), clf.shard_bounds(), n_pairs, rank, world, fused=a.exchange == "fused") if world > 1 else None

    # pinned host buffers (inputs and results of the end-to-end leg)
    pin_bases = torch.from_numpy(bases).pin_memory()
    pin_off = torch.from_numpy(offsets.view(np.int64)).pin_memory()
    pin_len = torch.from_numpy(lens.view(np.int32)).pin_memory()
    h_bases = pin_bases.numpy()
    h_off = pin_off.numpy().view(np.uint64)
    h_len = pin_len.numpy().view(np.uint32)
    pin_res = torch.zeros(n_pairs * mb.RESULT_DT.itemsize, dtype=torch.uint8).pin_memory()
    tc_cap = 8 * n_pairs
    pin_ids = torch.zeros(tc_cap, dtype=torch.int32).pin_memory()
    pin_cnt = torch.zeros(tc_cap, dtype=torch.int32).pin_memory()
    h_res = pin_res.numpy().view(mb.RESULT_DT)
    h_ids, h_cnt = pin_ids.numpy(), pin_cnt.numpy()

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

    def staged_step():
        return sharded.classify_staged() if sharded else clf.classify_staged()

    # end-to-end step: the reads of the step come from pinned host memory, the results go back to host memory.  Two
    # pinned input buffers alternate, and the copy of the NEXT step's reads (mb_prefetch_reads, second stream) overlaps
    # this step's device work; every step still copies its own inputs and reads its own results.
    pin_in = [(pin_bases, h_bases, h_off, h_len)]
    pb2, po2, pl2 = pin_bases.clone().pin_memory(), pin_off.clone().pin_memory(), pin_len.clone().pin_memory()
    pin_in.append((pb2, pb2.numpy(), po2.numpy().view(np.uint64), pl2.numpy().view(np.uint32)))
    e2e_i = [0]

    def e2e_step():
        cur = pin_in[e2e_i[0] & 1]
        nxt = pin_in[(e2e_i[0] + 1) & 1]
        e2e_i[0] += 1
        if sharded:
            sharded.stage(cur[1], cur[2], cur[3], bases_ptr=cur[0].data_ptr())
            clf.prefetch(nxt[1], nxt[2], nxt[3], bases_ptr=nxt[0].data_ptr())
            used = sharded.classify_staged()
        else:
            clf.stage(cur[1], cur[2], cur[3], bases_ptr=cur[0].data_ptr())
            clf.prefetch(nxt[1], nxt[2], nxt[3], bases_ptr=nxt[0].data_ptr())
            used = clf.classify_staged()
        clf.fetch_into(h_res, h_ids, h_cnt)
        return used

    # ---- device-resident leg: reads staged once, K passes of the device stages ----
    if sharded:
        sharded.stage(h_bases, h_off, h_len, bases_ptr=pin_bases.data_ptr())
    else:
        clf.stage(h_bases, h_off, h_len, bases_ptr=pin_bases.data_ptr())
    for _ in range(a.warmup):
        staged_step()
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
    launches_per_step = t_step["kernel_launches"]
    stats = dict(sharded.stats) if sharded else {}

    # ---- end-to-end leg: host buffers in, host results out, every step ----
    for _ in range(max(1, a.warmup // 2)):
        e2e_step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    w0 = time.perf_counter()
    for _ in range(a.steps):
        used = e2e_step()
    e1.record(stream)
    barrier()
    wall_e2e = time.perf_counter() - w0
    ms_e2e = max_over_ranks(max(e0.elapsed_time(e1), wall_e2e * 1e3))
    clocks = sampler.stop()
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
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic",
        "config": {"workload": f"configs[1]: {n_pairs} synthetic 2x150 read pairs per GPU vs {meta['genomes']}-genome synthetic syncmer DB "
                               f"({meta['db_kmers']} k-mers, {meta['species']} species), seq-mode 2, s=5",
                   "pairs_per_gpu": n_pairs, "db_kmers": meta["db_kmers"], "db_words": meta["db_words"], "genomes": meta["genomes"],
                   "db_sharding": f"{world} k-mer-range shard(s), one per GPU" if world > 1 else "whole DB on one GPU",
                   "exchange": (a.exchange if world > 1 else None),
                   "l2": "inputs larger than L2 (k-mer buffer and DB stream are GBs vs 126 MB L2); no flush needed",
                   "seed": a.seed},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": round(ms_e2e / a.steps, 3), "entry_point": "mb_stage_reads + mb_prefetch_reads(next) + mb_classify_staged + mb_fetch_results (pinned host buffers)"},
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
    if world == 1 and not a.no_fastq:
        try:
            out["input_pipeline"] = fastq_leg(a, torch, clf, bases, n_pairs, h_res, h_ids, h_cnt, stream)
        except Exception as ex:  # reported, never required
            out["input_pipeline"] = {"error": str(ex)}

    if rank == 0:
        n_cpu = cpu_sample_pairs(a)
        tsv = ref_results_path(a, n_cpu)
        if not a.no_cpu_baseline and world == 1:
            try:
                cb = run_cpu_reference(a, dbdir, bases, n_cpu, reps=2, skip=1)
                out["cpu_baseline"] = {"value": round(cb["reads_per_s"], 1), "unit": UNIT, "cores": cb["cores"], "kind": cb["kind"],
                                       "sample": cb["sample"], "same_config": cb["same_config"], "phases_s": cb["phases"],
                                       "wall_s": round(cb["wall_s"], 1)}
            except Exception as ex:  # the baseline is reported, never required
                out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "reference", "sample": f"failed: {ex}"}
        if not a.no_check:
            try:
                if not os.path.exists(tsv):          # N > 1, or the baseline leg was skipped: one untimed pass of the harness
                    run_cpu_reference(a, dbdir, bases, n_cpu, reps=1, skip=0)
                out["check"] = check_against_reference(tsv, h_res, h_ids, h_cnt)
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
    bases, _, _ = prepare_reads(a, device, 0)
    n = cpu_sample_pairs(a)
    cb = run_cpu_reference(a, dbdir, bases, n, reps=a.steps, skip=a.warmup)
    value = cb["reads_per_s"]
    out = {
        "impl": "reference", "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": round(cb["s_per_rep"] * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": f"configs[1]: {a.pairs} synthetic 2x150 read pairs per GPU vs {meta['genomes']}-genome synthetic syncmer DB "
                               f"({meta['db_kmers']} k-mers, {meta['species']} species), seq-mode 2, s=5",
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
