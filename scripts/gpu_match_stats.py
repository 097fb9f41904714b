"""One-off workload statistics for the join design (not product): AA-group sizes seen by the queries,
exact-hit fraction, candidate pairs."""
import sys, os, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import metabuli_work_b200.binding as mb
a = bench.parse_args()
dev = torch.device("cuda", 0)
dbdir, meta = bench.prepare_db(a, dev, 0, 1)
bases, offsets, lens = bench.prepare_reads(a, dev, 0)
clf = mb.Classifier(dbdir, device=0, seq_mode=2)
dbv = torch.from_numpy(clf.decode_db().view(np.int64)).to(dev)
clf.stage(bases, offsets, lens)
nk = clf.extract()
buf = torch.empty((nk, 2), dtype=torch.int64, device=dev)
clf.export_kmers_device(buf.data_ptr(), nk)
q = buf[:, 0].contiguous()
# values are < 2^63 here? AA part uses 40 bits << 24 = 64 bits: top bit may be set -> compare as unsigned via xor
FLIP = -(1 << 63)
dbs = dbv ^ FLIP          # signed order == unsigned order
qs = q ^ FLIP
assert bool((dbs[1:] >= dbs[:-1]).all())
db_aa = (dbv >> 24) & ((1 << 40) - 1)
q_aa = (q >> 24) & ((1 << 40) - 1)
lo = torch.searchsorted(db_aa, q_aa, right=False)
hi = torch.searchsorted(db_aa, q_aa, right=True)
g = hi - lo
p = torch.searchsorted(dbs, qs, right=False)
exact = (p < len(dbs)) & (dbs[p.clamp(max=len(dbs) - 1)] == qs)
pe = torch.searchsorted(dbs, qs, right=True)
out = {"queries": nk, "with_group": int((g > 0).sum()), "exact": int(exact.sum()), "pairs_all": int(g.sum()),
       "pairs_nonexact": int(g[~exact].sum()), "exact_run_sum": int((pe - p)[exact].sum()),
       "nonexact_with_group": int(((g > 0) & ~exact).sum())}
edges = [0, 1, 2, 3, 5, 9, 17, 33, 65, 129, 257, 513, 1025, 1 << 30]
gg = g[(g > 0) & ~exact]
out["nonexact_group_hist"] = {f"<{edges[i+1]}": int(((gg >= edges[i]) & (gg < edges[i + 1])).sum()) for i in range(1, len(edges) - 1)}
ge = g[exact]
out["exact_group_hist"] = {f"<{edges[i+1]}": int(((ge >= edges[i]) & (ge < edges[i + 1])).sum()) for i in range(1, len(edges) - 1)}
r = (pe - p)[exact]
out["exact_run_hist"] = {f"<{edges[i+1]}": int(((r >= edges[i]) & (r < edges[i + 1])).sum()) for i in range(1, len(edges) - 1)}
# DB-side group sizes
ch = torch.ones(len(db_aa), dtype=torch.bool, device=dev); ch[1:] = db_aa[1:] != db_aa[:-1]
out["db_groups"] = int(ch.sum()); out["db_kmers"] = len(db_aa)
print(json.dumps(out))
