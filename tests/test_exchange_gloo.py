"""The multi-GPU exchange logic (metabuli_work_b200/exchange.py) on CPU: world_size-2 and -3 `gloo` groups, with a
checker engine built on the oracle standing in for the device.  What is under test is the host-side plumbing:
read-id bases, bucketing by owner shard, the two all-to-alls with ragged counts, and that scoring at the home rank
sees exactly the matches of its reads.  The union over ranks must equal the single-process oracle run."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEngine:
    """Same interface as exchange.DeviceEngine, computed by the oracle (checker side only)."""

    def __init__(self, dbdir, seq_mode):
        self.db = orc.Db(dbdir)
        self.p = self.db.params(seq_mode=seq_mode)
        self.base = 0

    def set_read_id_base(self, base):
        self.base = base

    def stage(self, bases, offsets, lens, bases_ptr=None):
        self.batch = bases                       # an orc.Batch
        self.n_reads = bases.n

    def extract(self):
        k = orc.extract(self.p, self.batch)
        k["qinfo"] += np.uint64(self.base) << np.uint64(32)
        self.k = k
        return len(k)

    def partition_kmers(self, bounds):
        owner = np.searchsorted(bounds, self.k["value"], side="right")
        order = np.argsort(owner, kind="stable")
        self.k = self.k[order]
        return np.bincount(owner, minlength=len(bounds) + 1).astype(np.uint64)

    def export_kmers(self):
        return torch.from_numpy(np.ascontiguousarray(self.k).view(np.int64).reshape(-1, 2).copy())

    def import_kmers(self, t):
        self.k = t.numpy().reshape(-1).view(orc.KMER_DT).copy()

    def sort_kmers(self):
        self.k = orc.sort_kmers(self.k)

    def match(self):
        # matches only occur inside an amino-acid group, and shards are cut at group boundaries: the whole DB
        # gives the shard's result for the k-mers this rank owns
        self.m = orc.match(self.db, self.p, self.k)
        return len(self.m)

    def partition_matches(self, reads_per_rank, world):
        seq = ((self.m["qinfo"] >> np.uint64(32)) & np.uint64(0x1FFFFFFF)).astype(np.int64)
        home = (seq - 1) // reads_per_rank
        order = np.argsort(home, kind="stable")
        self.m = self.m[order]
        return np.bincount(home, minlength=world).astype(np.uint64)

    def export_matches(self):
        return torch.from_numpy(np.ascontiguousarray(self.m).view(np.int64).reshape(-1, 3).copy())

    def import_matches(self, t):
        self.m = t.numpy().reshape(-1).view(orc.MATCH_DT).copy()

    def sort_matches(self):
        self.m = orc.sort_matches(self.m)

    def score(self):
        m = self.m.copy()
        m["qinfo"] -= np.uint64(self.base) << np.uint64(32)          # back to 1-based local read ids
        self.res, self.ids, self.cnt = orc.score(self.db, self.p, m, self.batch)
        return len(self.ids)

    def fetch(self, used):
        return self.res, self.ids, self.cnt

    # ---- fused form: on the device the producers write into peer memory; here an all-to-all stands in for the peer
    # stores, so that the control flow of ShardedClassifier (barriers, takes, overflow agreement, regrow) is what runs
    def peer_setup(self, group, rank, world, bounds, reads_per_rank, kmer_cap, match_cap):
        self.g = dict(group=group, rank=rank, world=world, bounds=np.asarray(bounds, dtype=np.uint64), rpr=reads_per_rank,
                      kmer_cap=kmer_cap, match_cap=match_cap)
        self.setups = getattr(self, "setups", 0) + 1

    def sync_tensor(self):
        return torch.zeros(1, dtype=torch.int64)

    def extract_routed(self):
        from metabuli_work_b200.exchange import _all_to_all_records
        n = self.extract()
        counts = self.partition_kmers(self.g["bounds"])
        self._arrived_k, _ = _all_to_all_records(self.export_kmers(), counts, 2, self.g["group"])
        return n

    def take_kmers(self):
        n = self._arrived_k.shape[0]
        if n > self.g["kmer_cap"]:
            return n, True
        self.import_kmers(self._arrived_k)
        return n, False

    def match_routed(self):
        from metabuli_work_b200.exchange import _all_to_all_records
        n = self.match()
        counts = self.partition_matches(self.g["rpr"], self.g["world"])
        self._arrived_m, _ = _all_to_all_records(self.export_matches(), counts, 3, self.g["group"])
        return n

    def take_matches(self):
        n = self._arrived_m.shape[0]
        if n > self.g["match_cap"]:
            return n, True
        self.import_matches(self._arrived_m)
        return n, False


def _worker(rank, world, port, dbdir, s1, s2, seq_mode, bounds, per_rank, outdir, fused):
    sys.path.insert(0, ROOT)
    from metabuli_work_b200.exchange import ShardedClassifier
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        a, b = rank * per_rank, min(len(s1), (rank + 1) * per_rank)
        batch = orc.Batch(s1[a:b], None if s2 is None else s2[a:b])
        eng = OracleEngine(dbdir, seq_mode)
        sc = ShardedClassifier(eng, bounds, per_rank, fused=fused, match_cap=16 if fused else None, kmer_cap=64 if fused else None)
        sc.stage(batch, None, batch.len)
        used = sc.classify_staged()
        res, ids, cnt = sc.fetch(used)
        tc = [helpers.taxcnt_string(ids, cnt, int(r["taxcnt_begin"]), int(r["taxcnt_end"])) for r in res]
        np.save(os.path.join(outdir, f"res{rank}.npy"), res)
        with open(os.path.join(outdir, f"tc{rank}.txt"), "w") as f:
            f.write("\n".join(tc))
        np.save(os.path.join(outdir, f"stats{rank}.npy"), np.array([sc.stats["kmers_extracted"], sc.stats["kmers_owned"],
                                                                  sc.stats["matches_owned"], sc.stats["matches_home"], sc.regrows]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,world,fused", [("plain_paired", 2, False), ("sync_paired", 3, False), ("sync_long", 2, False),
                                              ("plain_paired", 2, True), ("sync_paired", 3, True)])
def test_sharded_equals_single(golden, tmp_path, case, world, fused):
    meta = golden.cases[case]
    dbdir = golden.dbdir(case)
    _, s1, s2 = golden.reads(case)
    n = len(s1)
    per_rank = (n + world - 1) // world
    db = orc.Db(dbdir)
    p = db.params(seq_mode=meta["seq_mode"])
    v, _ = db.decode()
    # shard bounds: first value of an amino-acid group nearest to equal counts (what the split table encodes)
    bounds = []
    for g in range(1, world):
        i = g * len(v) // world
        while i < len(v) and (v[i] >> np.uint64(24)) == (v[i - 1] >> np.uint64(24)):
            i += 1
        bounds.append(v[min(i, len(v) - 1)] & np.uint64(0xFFFFFFFFFF000000))    # routed by the amino-acid part only
    bounds = np.array(bounds, dtype=np.uint64)
    k, m, res, ids, cnt = orc.classify(db, p, orc.Batch(s1, s2))
    db.close()
    port = 29600 + (os.getpid() % 300)
    mp.spawn(_worker, args=(world, port, dbdir, s1, s2, meta["seq_mode"], bounds, per_rank, str(tmp_path), fused), nprocs=world, join=True)
    got_res = np.concatenate([np.load(tmp_path / f"res{r}.npy") for r in range(world)])
    got_tc = []
    for r in range(world):
        txt = (tmp_path / f"tc{r}.txt").read_text()
        a, b = r * per_rank, min(n, (r + 1) * per_rank)
        got_tc += txt.split("\n") if b - a > 0 else []
    assert len(got_res) == n
    for f in ("classification", "is_classified", "query_length"):
        assert np.array_equal(got_res[f], res[f]), f
    assert np.array_equal(got_res["score"].view(np.uint32), res["score"].view(np.uint32))
    want_tc = [helpers.taxcnt_string(ids, cnt, int(r["taxcnt_begin"]), int(r["taxcnt_end"])) for r in res]
    assert got_tc == want_tc
    st = np.stack([np.load(tmp_path / f"stats{r}.npy") for r in range(world)])
    assert st[:, 0].sum() == st[:, 1].sum() == len(k)          # every k-mer reached exactly one owner
    assert st[:, 2].sum() == st[:, 3].sum() == len(m)          # every match came home
    assert (st[:, 1] > 0).all()                                # the shards really split the work
    if fused:
        assert (st[:, 4] >= 2).all()                           # the 64-record k-mer and the 16-record match buffers both overflowed and were regrown
