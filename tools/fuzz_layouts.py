#!/usr/bin/env python
"""Randomised check of the packer's upload formats, CPU only: for random synthetic workloads and parameters, every layout
orfxh_pack offers (whole windows, islands, 4-bit codes, 2-bit codes + exception list, segment dictionary, dictionary without
the per-segment arrays) must give the oracle the same rows and the same chosen ORFs, on the whole workload and on locus
shards.  The CUDA library reads the very same batches (tests/test_gpu_parity.py compares it with the oracle per layout).

  python tools/fuzz_layouts.py [--seeds 0:100]"""
import argparse
import ctypes
import os
import random
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import orfanage_b200 as ob   # noqa: E402

SYNTH = os.path.join(ROOT, "tools", "orfx_synth")
MODES = ["START_MATCH", "LONGEST", "LONGEST_MATCH", "BEST", "FIRST"]


def same(a, b):
    if a.n_queries != b.n_queries or a.n_rows != b.n_rows or a.best.tobytes() != b.best.tobytes():
        return False
    return all(a.rows[k].tobytes() == b.rows[k].tobytes() for k in a.rows)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", default="0:50")
    args = ap.parse_args()
    lo, _, hi = args.seeds.partition(":")
    oracle = ctypes.CDLL(os.path.join(ROOT, "oracle", "liborfx_oracle.so"))
    fails = cases = 0
    for seed in range(int(lo), int(hi)):
        rng = random.Random(seed)
        d = tempfile.mkdtemp(prefix="orfx_fuzzlay_%d_" % seed)
        n_tf = rng.choice([1, 2])
        synth = [SYNTH, "--out", d, "--seed", str(seed), "--loci", str(rng.choice([15, 40])), "--chroms", str(rng.choice([1, 3])),
                 "--chrom-len", str(rng.choice([400000, 1500000])), "--queries", str(rng.choice([3, 9])), "--templates", str(rng.choice([2, 5])),
                 "--template-files", str(n_tf), "--lower-frac", str(rng.choice([0, 0.4])), "--iupac-rate", str(rng.choice([0, 0.002, 0.02])),
                 "--titin-frac", str(rng.choice([0, 0.1]))]
        dot = rng.random() < 0.4
        if dot:
            synth += ["--dot-frac", "0.4"]
        if rng.random() < 0.4:
            synth += ["--edge-loci"]
        subprocess.run(synth, check=True, stderr=subprocess.DEVNULL)
        tpl = [os.path.join(d, "template.gtf")] if n_tf == 1 else [os.path.join(d, "template%d.gtf" % (i + 1)) for i in range(n_tf)]
        overhang = rng.choice([0, 0, 7, 90, 2000])
        spliced = overhang > 0 and rng.random() < 0.5
        w = ob.Workload(os.path.join(d, "query.gtf"), tpl, os.path.join(d, "genome.fa"), overhang=overhang, spliced_overhang=spliced)
        mode = "ALL" if rng.random() < 0.25 else ",".join(rng.sample(MODES, rng.choice([1, 2, 3])))
        params = ob.Params.make(mode=mode, overhang=overhang, spliced_overhang=spliced, non_aug=rng.random() < 0.3)
        n_shards = rng.choice([1, 1, 2, 3])
        layouts = ["islands", "islands4", "islands2", "islands2d"] + ([] if (dot and overhang > 0) else ["windows"])   # strand '.' flanks need islands
        with ob.Context(params, lib=oracle, prefix="orfx_oracle_") as ox:
            for shard in range(n_shards):
                want = None
                for layout in layouts:
                    b = w.pack(shard, n_shards, layout=layout)
                    got = ox.run(b)
                    cases += 1
                    if want is None:
                        want = got
                    elif not same(got, want):
                        fails += 1
                        print("FAIL seed %d shard %d/%d layout %s differs from %s | synth %s | overhang %d spliced %s mode %s"
                              % (seed, shard, n_shards, layout, layouts[0], " ".join(synth[3:]), overhang, spliced, mode))
                    if layout == "islands2d":
                        bare = b.without_segment_arrays()
                        cases += 1
                        if not same(ox.run(bare), want):
                            fails += 1
                            print("FAIL seed %d shard %d/%d dictionary-only batch differs | synth %s" % (seed, shard, n_shards, " ".join(synth[3:])))
        w.close()
        shutil.rmtree(d, ignore_errors=True)
    print("fuzz_layouts: %d batches over %d seeds, %d failures" % (cases, int(hi) - int(lo), fails))
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
