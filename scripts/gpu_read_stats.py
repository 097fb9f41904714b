"""One-off: distribution of matches per read in the bench workload (not product)."""
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
res, ids, cnt = clf.classify(bases, offsets, lens)
n = res["n_matches"]
print("reads", len(n), "mean", n.mean(), "max", n.max(), "p50/p90/p99/p99.9", np.percentile(n, [50, 90, 99, 99.9]))
top = np.argsort(n)[-10:]
print("top", n[top])
for t in top[-3:]:
    b = bases[int(offsets[2*t]):int(offsets[2*t])+150].tobytes()
    print(t, n[t], b[:80])
