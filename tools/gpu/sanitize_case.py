#!/usr/bin/env python
"""A compact workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck): every kernel of the library once
or twice, on inputs small enough for a 10-100x slowdown, each result still checked against the CPU oracle.

  compute-sanitizer --tool racecheck python tools/gpu/sanitize_case.py

Covers: the whole path on a batch with titin-class (heavy) bundles, --overhang flanks, strand '.' loci and the alignment
stage enabled; the heavy-bundle kernels' global-memory branches (orfx_debug_set); template cleaning; pair scoring; and the
alignment kernel on short, multi-stripe and block-cooperative (phase A: warps handing stripe edges to each other through
shared-memory progress counters) pairs."""
import ctypes
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import orfanage_b200 as ob   # noqa: E402


def same(got, want, what):
    assert got.n_queries == want.n_queries and got.n_rows == want.n_rows, what
    assert got.best.tobytes() == want.best.tobytes(), what
    for k in want.rows:
        assert got.rows[k].tobytes() == want.rows[k].tobytes(), (what, k)


def device_context(params, oracle):
    """The CUDA library; ORFX_SANITIZE_DRY=1 (no GPU at hand) runs the script's own logic against the oracle twice."""
    if os.environ.get("ORFX_SANITIZE_DRY") == "1":
        return ob.Context(params, lib=oracle, prefix="orfx_oracle_")
    return ob.Context(params, device=0)


def main():
    dry = os.environ.get("ORFX_SANITIZE_DRY") == "1"
    d = "/tmp/orfx_sanitize"
    os.makedirs(d, exist_ok=True)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", "77", "--loci", "24", "--chroms", "2", "--chrom-len", "1500000",
                    "--queries", "6", "--templates", "3", "--lower-frac", "0.3", "--iupac-rate", "0.001", "--titin-frac", "0.15", "--dot-frac", "0.2",
                    "--edge-loci"], check=True, stderr=subprocess.DEVNULL)
    oracle = ctypes.CDLL(os.path.join(ROOT, "oracle", "liborfx_oracle.so"))
    w = ob.Workload(d + "/query.gtf", [d + "/template.gtf"], d + "/genome.fa", overhang=60)
    batch = w.pack(layout="islands_t")
    for pk in (dict(mode="LONGEST_MATCH,BEST", align=True), dict(mode="ALL", non_aug=True)):
        params = ob.Params.make(overhang=60, **pk)
        with ob.Context(params, lib=oracle, prefix="orfx_oracle_") as ox:
            want = ox.run(batch)
        for segs, pairs in ((9216, 8192), (0, 0)):
            with device_context(params, oracle) as cx:
                if not dry:
                    cx.debug_set("hb_max_segs", segs)
                    cx.debug_set("hb_max_pairs", pairs)
                same(cx.run(batch), want, "path %s hb %d/%d" % (pk, segs, pairs))
    with device_context(ob.Params.make(mode="BEST", overhang=60), oracle) as cx, ob.Context(ob.Params.make(mode="BEST", overhang=60), lib=oracle, prefix="orfx_oracle_") as ox:
        b4 = w.pack(layout="islands4")
        g, o = cx.score_pairs(b4), ox.score_pairs(b4)
        assert all(g[k].tobytes() == o[k].tobytes() for k in o)
        # cleaning batch = the queries with made-up CDS bounds (first exon .. last exon, cut to whole codons on the device side)
        a = {k: v.copy() for k, v in b4.arrays.items()}
        a["tx_kind"] = np.where(a["tx_kind"] == 0, 0, 2).astype(np.uint8)
        q = np.nonzero(a["tx_kind"] == 0)[0]
        a["tx_cds_start"] = np.zeros_like(a["tx_cds_start"]); a["tx_cds_end"] = np.zeros_like(a["tx_cds_end"])
        a["tx_cds_start"][q] = a["segs"][a["tx_seg_off"][q], 0] + 3
        a["tx_cds_end"][q] = a["segs"][a["tx_seg_off"][q + 1] - 1, 1] - 2
        a["seq_format"], a["seq_bases"] = np.array([1], np.int32), np.array([b4.struct.seq_bytes], np.int64)
        cb = ob.Batch(**a)
        for rescue in (False, True):
            gc, oc = cx.clean(cb, rescue), ox.clean(cb, rescue)
            assert all((x == y).all() for x, y in zip(gc, oc)), rescue
        # alignment: short pairs, multi-stripe pairs, and long pairs that take whole CTAs (more of them than one round)
        rng = np.random.default_rng(3)
        seqs, pq, pt = [], [], []
        for L, T in [(30, 40), (64, 64), (65, 200), (300, 280), (700, 650), (1500, 400)] + [(1100 + 37 * i, 4000 + 11 * i) for i in range(6)]:
            qa = rng.integers(0, 20, size=L).astype(np.uint8)
            ta = np.resize(qa, T).copy()
            m = rng.random(T) < 0.2
            ta[m] = rng.integers(0, 24, size=int(m.sum()))
            pq.append(len(seqs)); seqs.append(qa)
            pt.append(len(seqs)); seqs.append(ta)
        aa = np.concatenate(seqs)
        off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
        if dry:
            print("sanitize_case: dry run ok")
            return
        out, cells, ms = cx.align(aa, off, pq, pt)
        mat = np.ascontiguousarray(np.load(os.path.join(ROOT, "tests", "golden", "mat50.npy")), dtype=np.int8)
        n = len(pq)
        res = [np.zeros(n, dtype=np.int32) for _ in range(4)]
        pq_, pt_ = np.array(pq, np.int32), np.array(pt, np.int32)
        oracle.orfx_oracle_align_batch.restype = ctypes.c_int
        oracle.orfx_oracle_align_batch(ctypes.c_void_p(aa.ctypes.data), ctypes.c_void_p(off.ctypes.data), ctypes.c_void_p(pq_.ctypes.data),
                                       ctypes.c_void_p(pt_.ctypes.data), ctypes.c_int64(n), ctypes.c_void_p(mat.ctypes.data), 24, 12, 2, 0,
                                       *[ctypes.c_void_p(r.ctypes.data) for r in res])
        for i, k in enumerate(("max", "score", "mqe", "mte")):
            assert (out[k] == res[i]).all(), k
    w.close()
    print("sanitize_case: all results identical to the oracle")


if __name__ == "__main__":
    main()
