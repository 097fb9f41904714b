"""Routed producers with every "peer" on ONE GPU (two contexts, shards 0/2 and 1/2 of the bench DB): the compute cost of
routing without the NVLink hop, from the per-kernel event times.  usage: python scripts/gpu_routed_local.py [world]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench

world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
sys.argv = sys.argv[:1]
a = bench.parse_args()
device = torch.device("cuda:0")
import metabuli_work_b200.binding as mb
dbdir, meta = bench.prepare_db(a, device, 0, 1)
bases, offsets, lens = bench.prepare_reads(a, device, 0)
n = a.pairs
ctxs = [mb.Classifier(dbdir, device=0, seq_mode=2, shard=g, n_shards=world) for g in range(world)]
bounds = ctxs[0].shard_bounds()
kc = 3 * len(bases)
mem = [c.peer_alloc(kc, kc // 2) for c in ctxs]
for g, c in enumerate(ctxs):
    c.peer_attach(g, mem, bounds, n)
    c.set_profile(True)
    c.set_read_id_base(g * n)
    c.stage(bases, offsets, lens)
out = {}
for it in range(3):
    for c in ctxs:
        c.extract_routed()
    out["extract_routed"] = {k: round(v[0] / max(v[1], 1), 3) for k, v in ctxs[0].kernel_times().items() if v[1]}
    for c in ctxs:
        got, over = c.take_routed_kmers()
        assert not over
        c.sort_kmers()
    torch.cuda.synchronize()
    ctxs[0].match()
    torch.cuda.synchronize()
    out["match_plain_half_db"] = ctxs[0].kernel_times()["k_match"][0]
    for c in ctxs:
        c.match_routed()
        torch.cuda.synchronize()
    out["match_routed_half_db"] = ctxs[0].kernel_times()["k_match"][0]
    for c in ctxs:
        got, over = c.take_routed_matches()
        assert not over, got
        c.sort_matches()
        c.score()
    out["all"] = {k: round(v[0] / max(v[1], 1), 3) for k, v in ctxs[0].kernel_times().items() if v[1]}
    # plain path on context 0 for comparison
ctxs[0].extract()
out["extract_plain"] = {k: round(v[0] / max(v[1], 1), 3) for k, v in ctxs[0].kernel_times().items() if v[1]}
print(json.dumps(out, indent=1))
