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

    def __init__(self, engine, bounds, reads_per_rank, rank=None, world=None, group=None):
        self.e = engine
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        self.bounds = np.ascontiguousarray(bounds, dtype=np.uint64)
        assert len(self.bounds) == self.world - 1
        self.reads_per_rank = int(reads_per_rank)
        self.stats = {}

    def stage(self, bases, offsets, lens, bases_ptr=None):
        assert len(lens) // 2 <= self.reads_per_rank
        self.e.set_read_id_base(self.rank * self.reads_per_rank)
        self.e.stage(bases, offsets, lens, bases_ptr)

    def classify_staged(self):
        """All device stages + both exchanges on the staged reads; returns the taxCnt entries used."""
        e = self.e
        n_local = e.extract()
        counts = e.partition_kmers(self.bounds)
        recv, _ = _all_to_all_records(e.export_kmers(), counts, 2, self.group)
        e.import_kmers(recv)
        e.sort_kmers()
        n_matches_owner = e.match()
        mcounts = e.partition_matches(self.reads_per_rank, self.world)
        mrecv, _ = _all_to_all_records(e.export_matches(), mcounts, 3, self.group)
        e.import_matches(mrecv)
        e.sort_matches()
        used = e.score()
        self.stats = {"kmers_extracted": int(n_local), "kmers_owned": int(recv.shape[0]), "matches_owned": int(n_matches_owner),
                      "matches_home": int(mrecv.shape[0]), "kmer_bytes_sent": int(n_local) * 16,
                      "match_bytes_sent": int(n_matches_owner) * 24}
        return used

    def fetch(self, used):
        return self.e.fetch(used)
