#!/usr/bin/env python
"""Generates the committed fixtures under tests/golden/ from the REFERENCE itself.

Run in the build container (where /root/reference exists and `make oracle` has produced
oracle/_ref/{orfanage_ref,orfcompare_ref} from the reference's own sources):

    python tests/golden/make_golden.py

Nothing here runs at test time on the GPU box: the tests only read the files this script wrote.

  mat50.npy, aa_table.npy   parsed from the literals in src/common.hpp:239-277
  codon_map.txt             the 180 codon_map facts of src/common.hpp:279, "NNNa" per line
  orfcompare_kat.tsv        oracle/_ref/orfcompare_ref on example/patl.chess3.gtf x example/patl.gencode.gtf
                            (the docs' worked example, docs/content/csvs/orfcompare_stats.csv)
  c1_batch.npz, c1_names.json, c1_ref.gtf, c1_ref.tsv
                            config 1: example/query.gtf x example/template.gtf on a random chr19 (36 Mb, seed
                            20261104), --mode BEST --cleanq; the packed batch + the reference's outputs
  synth_s<seed>_<tag>.{gtf,tsv}
                            reference outputs on seeded synthetic inputs (tools/orfx_synth regenerates the
                            inputs bit for bit from the seed)
"""
import json
import os
import re
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "orfanage_ref")
CMP_BIN = os.path.join(ROOT, "oracle", "_ref", "orfcompare_ref")
SYNTH = os.path.join(ROOT, "tools", "orfx_synth")

# (tag, CLI args) of the synthetic goldens; shared with tests/test_golden.py
SYNTH_SEEDS = [101, 102]
SYNTH_ARGS = ["--loci", "10", "--chroms", "1", "--chrom-len", "600000", "--queries", "6", "--templates", "3",
              "--template-files", "2", "--lower-frac", "0.3", "--iupac-rate", "0.0005", "--titin-frac", "0.1"]
SYNTH_RUNS = [
    ("default", ["--reference", "genome.fa", "--query", "query.gtf"]),
    ("best_cleanq", ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "BEST", "--cleanq"]),
    ("all", ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "ALL"]),
    ("lm_best_thr", ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "LONGEST_MATCH,BEST", "--lpi", "50",
                     "--ilpi", "60", "--mlpi", "70", "--minlen", "300"]),
    ("nonaug_longest", ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "LONGEST", "--non_aug"]),
    ("noref", ["--query", "query.gtf", "--mode", "LONGEST_MATCH,BEST"]),
    ("cleant", ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--mode", "LONGEST_MATCH,BEST"]),
    ("cleant_rescue", ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--rescue"]),
    # --overhang / --spliced_overhang (TX::extend_seq, Transcriptome.cpp:342-399)
    ("overhang300", ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "300"]),
    ("overhang60_spliced_all", ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "60", "--spliced_overhang", "--mode", "ALL"]),
]
CMP_RUNS = [   # orfcompare (reference: oracle/_ref/orfcompare_ref) on the same synthetic inputs
    ("t2_vs_t1_ref", ["--query", "template2.gtf", "--template", "template1.gtf", "--reference", "genome.fa"]),
    ("t1_vs_t2", ["--query", "template1.gtf", "--template", "template2.gtf"]),
    ("t1_vs_t1_id", ["--query", "template1.gtf", "--template", "template1.gtf", "--use_id"]),
]
C1_SEED = 20261104


def parse_constants():
    src = open(os.path.join(REF, "src", "common.hpp")).read()
    m = re.search(r"mat50\[\]\s*=\s*\{(.*?)\};", src, re.S)
    body = re.sub(r"//[^\n]*", "", m.group(1))
    vals = [int(x) for x in re.findall(r"-?\d+", body)]
    assert len(vals) == 24 * 24, len(vals)
    np.save(os.path.join(HERE, "mat50.npy"), np.array(vals, dtype=np.int8).reshape(24, 24))
    m = re.search(r"aa_table\[128\]\s*=\s*\{(.*?)\};", src, re.S)
    vals = [int(x) for x in re.findall(r"-?\d+", re.sub(r"//[^\n]*", "", m.group(1)))]
    assert len(vals) == 128, len(vals)
    np.save(os.path.join(HERE, "aa_table.npy"), np.array(vals, dtype=np.uint8))
    m = re.search(r"codon_map\s*=\s*\{(.*?)\};", src, re.S)
    pairs = re.findall(r"\{\s*\"([A-Z]{3})\"\s*,\s*'(.)'\s*\}", m.group(1))
    assert len(pairs) == 180, len(pairs)
    with open(os.path.join(HERE, "codon_map.txt"), "w") as f:
        for c, a in sorted(pairs):
            f.write(c + a + "\n")


def orfcompare_kat():
    out = os.path.join(HERE, "orfcompare_kat.tsv")
    subprocess.run([CMP_BIN, "--query", os.path.join(REF, "example", "patl.chess3.gtf"), "--template",
                    os.path.join(REF, "example", "patl.gencode.gtf"), "--output", out, "--threads", "1"], check=True,
                   capture_output=True)
    rows = [l.rstrip("\n").split("\t") for l in open(out)]
    # the docs' worked rows (docs/content/csvs/orfcompare_stats.csv, Examples 1-2)
    assert rows[1][:12] == ["CHS.8812.2", "ENST00000300146.10", "1566", "2313", "1566", "1566", "0", "0", "747", "92.5532",
                            "92.5532", "92.5532"], rows[1]
    assert rows[2][:12] == ["CHS.8812.5", "ENST00000300146.10", "2310", "2313", "2310", "2310", "0", "0", "3", "99.8703",
                            "99.8703", "99.8703"], rows[2]


def kat_compare_batch():
    """The docs' orfcompare example as a packed compare batch (orfx_score_pairs): CDS chains of example/patl.chess3.gtf
    (queries) and example/patl.gencode.gtf (template), one bundle, so that the known-answer rows can be checked on the
    GPU box, where /root/reference does not exist."""
    import re
    import numpy as np

    def cds_chains(path):
        tx = {}
        for l in open(path):
            f = l.rstrip("\n").split("\t")
            if len(f) < 9 or f[2] != "CDS":
                continue
            tid = re.search(r'transcript_id "([^"]+)"', f[8]).group(1)
            tx.setdefault(tid, dict(strand=f[6], segs=[]))["segs"].append((int(f[3]), int(f[4])))
        return tx
    q = cds_chains(os.path.join(REF, "example", "patl.chess3.gtf"))
    t = cds_chains(os.path.join(REF, "example", "patl.gencode.gtf"))
    rows = [l.rstrip("\n").split("\t") for l in open(os.path.join(HERE, "orfcompare_kat.tsv"))][1:]
    names, kind, strand, seg_off, segs, cs, ce = [], [], [], [0], [], [], []
    for tid, rec, k in [(r[1], t[r[1]], 1) for r in rows[:1]] + [(r[0], q[r[0]], 0) for r in rows]:
        names.append(tid); kind.append(k); strand.append(ord(rec["strand"]))
        ch = sorted(rec["segs"])
        segs += ch; seg_off.append(len(segs)); cs.append(ch[0][0]); ce.append(ch[-1][1])
    n = len(names)
    np.savez_compressed(os.path.join(HERE, "kat_compare_batch.npz"), seq=np.zeros(0, np.uint8), seq_off=np.zeros(2, np.int64),
                        bundle_start=np.array([min(cs)], np.int32), bundle_tx_off=np.array([0, n], np.int32),
                        tx_kind=np.array(kind, np.uint8), tx_strand=np.array(strand, np.uint8), tx_seg_off=np.array(seg_off, np.int32),
                        tx_cds_start=np.array(cs, np.int32), tx_cds_end=np.array(ce, np.int32), tx_tid_class=np.arange(n, dtype=np.int32),
                        segs=np.array(segs, np.int32), seg_phase=np.zeros(len(segs), np.int8), seg_seq_off=np.zeros(0, np.int32))
    json.dump({"tx_id": names}, open(os.path.join(HERE, "kat_compare_names.json"), "w"))


def config1():
    import orfanage_b200 as ob
    d = "/tmp/orfx_c1"
    os.makedirs(d, exist_ok=True)
    # genome only: one random chromosome named chr19 (chrom-offset 18), 36 Mb
    subprocess.run([SYNTH, "--out", d, "--seed", str(C1_SEED), "--loci", "0", "--chroms", "1", "--chrom-offset", "18",
                    "--chrom-len", "36000000"], check=True, capture_output=True)
    q = os.path.join(REF, "example", "query.gtf")
    t = os.path.join(REF, "example", "template.gtf")
    subprocess.run([REF_BIN, "--reference", "genome.fa", "--mode", "BEST", "--cleanq", "--threads", "1", "--query", q,
                    "--output", os.path.join(HERE, "c1_ref.gtf"), "--stats", os.path.join(HERE, "c1_ref.tsv"), t], cwd=d,
                   check=True, capture_output=True)
    w = ob.Workload(q, [t], os.path.join(d, "genome.fa"))
    b = w.pack()
    b.save(os.path.join(HERE, "c1_batch.npz"))
    names = [w.tx_id(i) for i in range(len(b.arrays["tx_kind"]))]
    json.dump({"tx_id": names, "mode": "BEST", "seed": C1_SEED}, open(os.path.join(HERE, "c1_names.json"), "w"), indent=1)
    w.close()


def synth_goldens():
    for seed in SYNTH_SEEDS:
        d = "/tmp/orfx_golden_s%d" % seed
        os.makedirs(d, exist_ok=True)
        subprocess.run([SYNTH, "--out", d, "--seed", str(seed)] + SYNTH_ARGS, check=True, capture_output=True)
        tpl = ["template1.gtf", "template2.gtf"]
        for tag, args in SYNTH_RUNS:
            base = os.path.join(HERE, "synth_s%d_%s" % (seed, tag))
            subprocess.run([REF_BIN] + args + ["--threads", "1", "--output", base + ".gtf", "--stats", base + ".tsv"] + tpl,
                           cwd=d, check=True, capture_output=True)
        for tag, args in CMP_RUNS:
            subprocess.run([os.path.join(os.path.dirname(REF_BIN), "orfcompare_ref")] + args + ["--threads", "1", "--output",
                           os.path.join(HERE, "cmp_s%d_%s.tsv" % (seed, tag))], cwd=d, check=True, capture_output=True)


if __name__ == "__main__":
    parse_constants()
    orfcompare_kat()
    kat_compare_batch()
    config1()
    synth_goldens()
    print("goldens written to", HERE)
