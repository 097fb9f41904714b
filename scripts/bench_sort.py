"""Microbenchmark of the query k-mer radix sort alone (tuning aid, not product): random 16-byte records."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import metabuli_work_b200.binding as mb
n = int(os.environ.get("N", 256_000_000))
ctx = mb.RawContext(device=0, kmer_format=2, syncmer=1, seq_mode=2)
ctx.set_profile(True)
g = torch.Generator(device="cuda"); g.manual_seed(1)
buf = torch.randint(-(1 << 62), 1 << 62, (n, 2), generator=g, device="cuda", dtype=torch.int64)
for it in range(3):
    ctx.set_kmers_device(buf.data_ptr(), n)
    ctx.n_reads = 0
    mb._cuda.mb_set_profile(ctx.ctx, 1)
    ctx.sort_kmers()
    kt = ctx.kernel_times()
print({k: (round(v[0] / v[1], 3), v[1]) for k, v in kt.items()}, "GB/s per pass:", round(32 * n / (kt["k_radix_pass<kmer>"][0] / kt["k_radix_pass<kmer>"][1] * 1e-3) / 1e9, 1))
