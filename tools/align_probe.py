"""Development probe: one orfx_align_batch call on a C5-law sample (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench
import orfanage_b200 as ob

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
aa, off, pq, pt = bench.c5_pairs(n, 20261104)
if len(sys.argv) > 2:   # keep pairs shorter than this (steady-state captures without the titin tail)
    import numpy as np
    keep = np.nonzero((off[pq + 1] - off[pq]) < int(sys.argv[2]))[0]
    pq, pt = pq[keep], pt[keep]
    n = len(pq)
lib = ob.load_device_library(os.environ["ORFX_LIB"]) if os.environ.get("ORFX_LIB") else None   # a variant build of liborfx.so
with ob.Context(ob.Params.make(), device=0, lib=lib) as cx:
    cx.align(aa, off, pq[:64], pt[:64])
    best = None
    for _ in range(3):
        out, cells, ms = cx.align(aa, off, pq, pt)
        best = ms if best is None else min(best, ms)
    ms = best
    print(os.environ.get("ORFX_LIB", "liborfx.so"), end=" ")
    print("pairs %d cells %.4g ms %.3f GCUPS %.1f" % (n, cells, ms, cells / ms / 1e6))
