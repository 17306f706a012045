"""Development probe: distribution of candidate rows per query on the bench workload."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import bench, orfanage_b200 as ob
d = bench.generate(dict(bench.C2), "c2")
w = ob.Workload(d + "/query.gtf", [d + "/template.gtf"], d + "/genome.fa")
b = w.pack(0, 1, layout="islands4")
with ob.Context(ob.Params.make(mode=bench.MODE, want_rows=False), device=0) as cx:
    r = cx.run(b).best
n = r["n_rows"]
print("queries", len(n), "rows", int(n.sum()), "max", int(n.max()), "mean %.2f" % n.mean())
for lo in (1, 2, 4, 8, 16, 32, 64, 128, 256):
    print("n_rows >= %d: %d queries, %d rows" % (lo, int((n >= lo).sum()), int(n[n >= lo].sum())))
