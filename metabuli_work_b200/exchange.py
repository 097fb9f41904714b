"""Multi-GPU classification: database sharded by k-mer range, two all-to-all exchanges per batch.

SURVEY.md §8(e).  The reference is single-node/OpenMP; its matcher already partitions the DB by the
`split` table into independent k-mer ranges (KmerMatcher.cpp:127-271), which is what shards here:

  home rank r   : stages its own reads, extracts k-mers with sequence IDs  r*reads_per_rank + 1 ...
                  and buckets them by owner shard (owner = number of shard bounds <= value)
  all-to-all #1 : 16-byte k-mer records go to the rank that holds their k-mer range
  owner rank    : sorts what it received, joins it against its resident shard -> matches
                  and buckets the matches by home rank ((sequenceID-1) // reads_per_rank)
  all-to-all #2 : 24-byte match records return to the read's home rank
  home rank     : sorts its matches (compareMatches order) and scores its reads

FUSED form (default on GPUs): the two bucketing passes and the two collectives disappear into the producing
kernels.  Every rank exposes fixed receive buffers through CUDA IPC; k_extract writes each k-mer straight into its
owner's buffer and k_match writes each match straight into its home rank's buffer (peer stores over NVLink, one remote
atomic reservation per run: mb_extract_kmers_routed / mb_match_kmers_routed).  What is left on the host is a barrier
between producer and consumer, carried by a one-element all-reduce that also agrees on buffer overflow.

This module is plumbing only: every compute step is a call into libmetabuli_b200.so through an
ENGINE object; the exchange is torch.distributed (NCCL over NVLink on the GPUs; gloo in the CPU
test-suite, where a checker engine built on the oracle stands in for the device).
"""
import numpy as np
import torch
import torch.distributed as dist


class DevicePtr:
    """Exposes a raw device pointer to torch through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr, n_words):
        self.__cuda_array_interface__ = {"shape": (n_words,), "typestr": "<i8", "data": (ptr, False), "version": 3, "strides": None}


class DeviceEngine:
    """Engine over an mb_ctx (binding.Context): records cross the exchange as int64 tensors on the GPU."""

    def __init__(self, ctx, device):
        self.ctx = ctx
        self.device = device
        self._keep = None

    def set_read_id_base(self, base):
        self.ctx.set_read_id_base(base)

    def stage(self, bases, offsets, lens, bases_ptr=None):
        self.ctx.stage(bases, offsets, lens, bases_ptr)

    def extract(self):
        return self.ctx.extract()

    def partition_kmers(self, bounds):
        return self.ctx.partition_kmers(bounds)

    def export_kmers(self):
        ptr, n = self.ctx.kmers_device()
        if n == 0:
            return torch.empty((0, 2), dtype=torch.int64, device=self.device)
        return torch.as_tensor(DevicePtr(ptr, n * 2), device=self.device).view(n, 2)

    def import_kmers(self, t):
        self._keep = t
        self.ctx.set_kmers_device(t.data_ptr(), t.shape[0])

    def sort_kmers(self):
        self.ctx.sort_kmers()

    def match(self):
        return self.ctx.match()

    def partition_matches(self, reads_per_rank, world):
        return self.ctx.partition_matches(reads_per_rank, world)

    def export_matches(self):
        ptr, n = self.ctx.matches_device()
        if n == 0:
            return torch.empty((0, 3), dtype=torch.int64, device=self.device)
        return torch.as_tensor(DevicePtr(ptr, n * 3), device=self.device).view(n, 3)

    def import_matches(self, t):
        self._keep = t
        self.ctx.set_matches_device(t.data_ptr(), t.shape[0])

    def sort_matches(self):
        self.ctx.sort_matches()

    def score(self):
        return self.ctx.score()

    def fetch(self, used):
        return self.ctx.fetch(used)

    # ---- fused form ----
    def peer_setup(self, group, rank, world, bounds, reads_per_rank, kmer_cap, match_cap):
        """Allocate this rank's receive side, trade CUDA IPC handles, map every peer's buffers, attach.  Collective."""
        from .binding import PeerMem
        self.peer_teardown(group)
        own = self.ctx.peer_alloc(kmer_cap, match_cap)
        mine = {"h": [self.ctx.ipc_export(p) for p in (own.kmers, own.matches, own.counters)], "kc": int(own.kmer_cap),
                "mc": int(own.match_cap)}
        everyone = [None] * world
        dist.all_gather_object(everyone, mine, group=group)
        peers, self._opened = [], []
        for r in range(world):
            if r == rank:
                peers.append(own)
                continue
            ptrs = [self.ctx.ipc_open(h) for h in everyone[r]["h"]]
            self._opened += ptrs
            peers.append(PeerMem(ptrs[0], ptrs[1], ptrs[2], everyone[r]["kc"], everyone[r]["mc"]))
        self.ctx.peer_attach(rank, peers, bounds, reads_per_rank)
        dist.barrier(group=group)

    def peer_teardown(self, group):
        if getattr(self, "_opened", None):
            for p in self._opened:
                self.ctx.ipc_close(p)
            self._opened = []
            dist.barrier(group=group)              # nobody frees a buffer a peer still maps

    def sync_tensor(self):
        return torch.zeros(1, dtype=torch.int64, device=self.device)

    def extract_routed(self):
        return self.ctx.extract_routed()

    def take_kmers(self):
        return self.ctx.take_routed_kmers()

    def match_routed(self):
        return self.ctx.match_routed()

    def take_matches(self):
        return self.ctx.take_routed_matches()


def _all_to_all_records(send, send_counts, width, group=None):
    """send: [n, width] int64 rows grouped by destination rank (send_counts rows each) -> rows received, recv_counts."""
    world = dist.get_world_size(group)
    sc = torch.as_tensor(np.asarray(send_counts, dtype=np.int64), device=send.device)
    rc = torch.empty_like(sc)
    dist.all_to_all_single(rc, sc, group=group)
    recv_counts = rc.cpu().tolist()
    recv = torch.empty((int(sum(recv_counts)), width), dtype=torch.int64, device=send.device)
    dist.all_to_all_single(recv, send.contiguous(), output_split_sizes=recv_counts,
                           input_split_sizes=[int(x) for x in send_counts], group=group)
    assert len(recv_counts) == world
    if recv.is_cuda:
        # the library consumes `recv` on its own stream: the collective must have landed
        torch.cuda.current_stream(recv.device).synchronize()
    return recv, recv_counts


class ShardedClassifier:
    """One per rank.  `engine` holds shard `rank` of the database; `bounds` are the first k-mer values of shards
    1..world-1 (binding.Classifier.shard_bounds())."""

    def __init__(self, engine, bounds, reads_per_rank, rank=None, world=None, group=None, fused=False, match_cap=None, kmer_cap=None):
        self.e = engine
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.bounds = np.ascontiguousarray(bounds, dtype=np.uint64)
        assert len(self.bounds) == self.world - 1
        self.reads_per_rank = int(reads_per_rank)
        self.stats = {}
        self.fused = bool(fused)
        self.kmer_cap = self.match_cap = 0
        self._match_cap0 = match_cap          # initial match receive capacity (records); default: half the k-mer capacity
        self._kmer_cap0 = kmer_cap            # initial k-mer receive capacity (records); default: 3 x the bases of the batch
        self.regrows = 0

    def _agree(self, value=0):
        """Barrier + agreement in one: max of `value` over the ranks."""
        t = self.e.sync_tensor()
        t.fill_(int(value))
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return int(t.item())

    def close(self):
        """Unmap the peers' buffers (collective): call on every rank before the engines' contexts are destroyed."""
        if self.fused and hasattr(self.e, "peer_teardown"):
            self.e.peer_teardown(self.group)

    def _setup(self, kmer_cap, match_cap):
        self.kmer_cap, self.match_cap = int(kmer_cap), int(match_cap)
        self.e.peer_setup(self.group, self.rank, self.world, self.bounds, self.reads_per_rank, self.kmer_cap, self.match_cap)

    def stage(self, bases, offsets, lens, bases_ptr=None):
        assert len(lens) // 2 <= self.reads_per_rank
        self.e.set_read_id_base(self.rank * self.reads_per_rank)
        self.e.stage(bases, offsets, lens, bases_ptr)
        if self.fused:
            # every window start emits at most two k-mers (one per strand): 2 x bases bounds what one rank produces; an
            # owner receives about that much too unless the query k-mers pile up in one shard (slack 1.5)
            # mates are packed back to back (mb_stage_reads checks it), so the base count is the buffer length: no pass over
            # lens (an oracle-backed test engine hands over its own batch object instead of a byte buffer)
            n_bases = len(bases) if hasattr(bases, "__len__") else int(np.asarray(lens, dtype=np.int64).sum())
            need = self._agree(self._kmer_cap0 if self._kmer_cap0 and not self.kmer_cap else max(3 * n_bases, 1 << 16))
            if need > self.kmer_cap:
                self._setup(need, max(self.match_cap, self._match_cap0 or need // 2))

    def classify_staged(self):
        """All device stages + both exchanges on the staged reads; returns the taxCnt entries used.
        Every engine call below returns after its device work has finished (counts come back to the host), so the
        wall-clock split in stats["phases_ms"] is a device-time split."""
        import time
        if self.fused:
            return self._classify_staged_fused()
        e = self.e
        t = [time.perf_counter()]
        lap = lambda: t.append(time.perf_counter())
        n_local = e.extract(); lap()
        counts = e.partition_kmers(self.bounds); lap()
        recv, _ = _all_to_all_records(e.export_kmers(), counts, 2, self.group); lap()
        e.import_kmers(recv)
        e.sort_kmers()
        n_matches_owner = e.match(); lap()
        mcounts = e.partition_matches(self.reads_per_rank, self.world); lap()
        mrecv, _ = _all_to_all_records(e.export_matches(), mcounts, 3, self.group); lap()
        e.import_matches(mrecv)
        e.sort_matches()
        used = e.score(); lap()
        names = ("extract", "partition_kmers", "exchange_kmers", "sort_match", "partition_matches", "exchange_matches", "sort_score")
        self.stats = {"kmers_extracted": int(n_local), "kmers_owned": int(recv.shape[0]), "matches_owned": int(n_matches_owner),
                      "matches_home": int(mrecv.shape[0]), "kmer_bytes_sent": int(n_local) * 16,
                      "match_bytes_sent": int(n_matches_owner) * 24,
                      "phases_ms": {n: round((t[i + 1] - t[i]) * 1e3, 3) for i, n in enumerate(names)}}
        return used

    def _classify_staged_fused(self):
        import time
        e = self.e
        t = [time.perf_counter()]
        lap = lambda: t.append(time.perf_counter())
        while True:
            n_local = e.extract_routed(); lap()             # k-mers land in their owners' receive buffers
            self._agree(); lap()                            # every producer is done
            n_owned, over = e.take_kmers()
            need = self._agree(n_owned if over else 0)      # (on overflow the count is the capacity that would have sufficed)
            if not need:
                break
            # query k-mers piled up in one shard beyond the slack (low-complexity or contaminated samples against an
            # equal-count split): everyone re-allocates with the worst demand and the extraction is redone, as on the match side
            del t[-2:]
            self.regrows += 1
            self._setup(need + need // 4, self.match_cap)
        e.sort_kmers()
        while True:
            n_matches_owner = e.match_routed(); lap()       # matches land in their home ranks' receive buffers
            self._agree(); lap()
            n_home, over = e.take_matches()
            need = self._agree(n_home if over else 0)
            if not need:
                break
            # some rank received more matches than its buffer holds: everyone re-allocates and the join is redone (the
            # sorted k-mers are still resident); the k-mer side is re-attached with the same capacity
            del t[-2:]
            self.regrows += 1
            self._setup(self.kmer_cap, need + need // 4)
        e.sort_matches()
        used = e.score(); lap()
        names = ("extract", "barrier_kmers", "sort_match", "barrier_matches", "sort_score")
        ph = {}
        for i, n in enumerate(names):
            ph[n] = round((t[i + 1] - t[i]) * 1e3, 3)
        self.stats = {"kmers_extracted": int(n_local), "kmers_owned": int(n_owned), "matches_owned": int(n_matches_owner),
                      "matches_home": int(n_home), "kmer_bytes_sent": int(n_local) * 16,
                      "match_bytes_sent": int(n_matches_owner) * 24, "phases_ms": ph}
        return used

    def fetch(self, used):
        return self.e.fetch(used)
