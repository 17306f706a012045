#!/usr/bin/env python
"""Randomised differential campaign, CPU only: the oracle-backed CLI (oracle/orfanage_oracle_cli: this repo's host code over
the CPU oracle) against the reference binary (oracle/_ref/orfanage_ref, built from the reference's own sources) on seeded
synthetic workloads with random generator settings and random option combinations.  GTF and stats TSV must be byte-identical
and the exit codes equal.  The CUDA library is compared with the same oracle in tests/ (-m gpu), so a divergence found here
is a divergence of the product.

  python tools/fuzz_parity.py [--seeds 200:240] [--runs-per-seed 6] [--keep]

Prints one line per failure with the command to reproduce it; exit code 1 if there was any."""
import argparse
import os
import random
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SYNTH = os.path.join(ROOT, "tools", "orfx_synth")
ORACLE_CLI = os.path.join(ROOT, "oracle", "orfanage_oracle_cli")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "orfanage_ref")
MODES = ["START_MATCH", "LONGEST", "LONGEST_MATCH", "BEST", "FIRST"]


def random_workload(rng, d, seed):
    n_tf = rng.choice([1, 1, 2, 3])
    args = [SYNTH, "--out", d, "--seed", str(seed), "--loci", str(rng.choice([20, 40, 80])), "--chroms", str(rng.choice([1, 2, 5])),
            "--chrom-len", str(rng.choice([400000, 1500000, 3000000])), "--queries", str(rng.choice([2, 6, 12])),
            "--templates", str(rng.choice([1, 3, 6])), "--template-files", str(n_tf), "--lower-frac", str(rng.choice([0, 0.3, 0.9])),
            "--iupac-rate", str(rng.choice([0, 0.0005, 0.01])), "--titin-frac", str(rng.choice([0, 0.02, 0.15]))]
    if rng.random() < 0.4:
        args += ["--dot-frac", str(rng.choice([0.1, 0.5]))]
    if rng.random() < 0.4:
        args += ["--edge-loci"]
    subprocess.run(args, check=True, stderr=subprocess.DEVNULL)
    tpl = ["template.gtf"] if n_tf == 1 else ["template%d.gtf" % (i + 1) for i in range(n_tf)]
    return args, tpl


def random_options(rng, tpl, overhang=True):
    a = []
    query = "query.gtf"
    templates = list(tpl)
    if len(tpl) > 1 and rng.random() < 0.2:       # a template file as the query: queries that carry a CDS
        query = templates.pop()
        a += [rng.choice(["--keep_all_cds", "--keep_cds_if_not_found", "--cleanq"])] if rng.random() < 0.7 else []
    has_ref = rng.random() < 0.85
    if has_ref:
        a += ["--reference", "genome.fa"]
    a += ["--query", query]
    if rng.random() < 0.8:
        if rng.random() < 0.2:
            a += ["--mode", "ALL"]
        else:
            k = rng.choice([1, 1, 2, 2, 3, 5])
            a += ["--mode", ",".join(rng.sample(MODES, k))]
    for flag, p in (("--non_aug", 0.25), ("--use_id", 0.15), ("--cleanq", 0.15)):
        if rng.random() < p and flag not in a:
            a += [flag]
    if not has_ref:                      # the reference asserts on --cleanq without --reference (orfanage.cpp:1050)
        a = [x for x in a if x != "--cleanq"]
    for flag, vals, p in (("--lpi", [0, 30, 80, 100], 0.25), ("--ilpi", [0, 50, 90], 0.25), ("--mlpi", [0, 60, 95], 0.25),
                          ("--minlen", [0, 3, 150, 600], 0.3)):
        if rng.random() < p:
            a += [flag, str(rng.choice(vals))]
    if has_ref:
        if rng.random() < 0.35:
            a += ["--cleant"] + (["--rescue"] if rng.random() < 0.6 else [])
        if overhang and rng.random() < 0.45:
            a += ["--overhang", str(rng.choice([1, 2, 3, 7, 30, 89, 300, 2000, 60000]))]
            if rng.random() < 0.5:
                a += ["--spliced_overhang"]
    return a, templates


def run(binary, d, args, tag, templates, extra=()):
    cmd = [binary] + list(args) + list(extra) + ["--output", tag + ".gtf", "--stats", tag + ".tsv"] + templates
    for f in (tag + ".gtf", tag + ".tsv"):
        if os.path.exists(os.path.join(d, f)):
            os.remove(os.path.join(d, f))
    try:
        p = subprocess.run(cmd, cwd=d, capture_output=True, timeout=600)
        rc, err = p.returncode, p.stderr.decode(errors="replace")
    except subprocess.TimeoutExpired:
        rc, err = -999, "timeout"

    def rd(name):
        path = os.path.join(d, name)
        return open(path, "rb").read() if os.path.exists(path) else b""
    return rc, rd(tag + ".gtf"), rd(tag + ".tsv"), err


def fuzz_orfcompare(seeds, runs_per_seed):
    """orfcompare: two annotation files of one workload against each other."""
    cmp_ref = os.path.join(ROOT, "oracle", "_ref", "orfcompare_ref")
    fails = runs = 0
    for seed in seeds:
        rng = random.Random(seed)
        d = tempfile.mkdtemp(prefix="orfx_fuzzcmp_%d_" % seed)
        n_tf = rng.choice([2, 3])
        synth = [SYNTH, "--out", d, "--seed", str(seed), "--loci", str(rng.choice([20, 60, 150])), "--chroms", str(rng.choice([1, 3])),
                 "--chrom-len", str(rng.choice([400000, 3000000])), "--queries", str(rng.choice([2, 8])), "--templates", str(rng.choice([2, 5, 9])),
                 "--template-files", str(n_tf), "--titin-frac", str(rng.choice([0, 0.05])), "--lower-frac", "0.3"]
        if rng.random() < 0.4:
            synth += ["--dot-frac", "0.3"]
        if rng.random() < 0.3:
            synth += ["--edge-loci"]
        subprocess.run(synth, check=True, stderr=subprocess.DEVNULL)
        files = ["template%d.gtf" % (i + 1) for i in range(n_tf)] + ["query.gtf"]
        for k in range(runs_per_seed):
            q, t = rng.sample(files, 2) if rng.random() < 0.9 else (files[0], files[0])
            opts = ["--query", q, "--template", t] + (["--use_id"] if rng.random() < 0.3 else []) + (["--reference", "genome.fa"] if rng.random() < 0.5 else [])
            pr = subprocess.run([cmp_ref] + opts + ["--threads", "1", "--output", "cmp_ref.tsv"], cwd=d, capture_output=True)
            po = subprocess.run([ORACLE_CLI, "orfcompare"] + opts + ["--output", "cmp_or.tsv"], cwd=d, capture_output=True)
            runs += 1
            if pr.returncode < 0:
                print("note: reference orfcompare aborted (rc %d) seed %d: %s" % (pr.returncode, seed, " ".join(opts)))
                continue

            def rd(n):
                path = os.path.join(d, n)
                return open(path, "rb").read() if os.path.exists(path) else b""
            if po.returncode != pr.returncode or (pr.returncode == 0 and rd("cmp_ref.tsv") != rd("cmp_or.tsv")):
                fails += 1
                print("FAIL orfcompare seed %d run %d (rc %d vs reference %d)\n  synth: %s\n  opts: %s\n  stderr(ours): %s\n  stderr(ref): %s"
                      % (seed, k, po.returncode, pr.returncode, " ".join(synth[1:]), " ".join(opts), po.stderr.decode(errors="replace")[-300:],
                         pr.stderr.decode(errors="replace")[-300:]))
        shutil.rmtree(d, ignore_errors=True)
    print("fuzz_parity (orfcompare): %d runs over %d seeds, %d failures" % (runs, len(seeds), fails))
    return fails


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--orfcompare", action="store_true", help="fuzz the orfcompare front-end instead of orfanage")
    ap.add_argument("--seeds", default="200:220")
    ap.add_argument("--runs-per-seed", type=int, default=6)
    ap.add_argument("--no-overhang", action="store_true", help="never pass --overhang (every other option still varies)")
    ap.add_argument("--keep", action="store_true", help="keep the directories of failing cases")
    args = ap.parse_args()
    lo, _, hi = args.seeds.partition(":")
    seeds = range(int(lo), int(hi or int(lo) + 1))
    if not (os.path.exists(ORACLE_CLI) and os.path.exists(REF_BIN)):
        sys.exit("build first: make oracle (needs /root/reference for oracle/_ref)")
    if args.orfcompare:
        sys.exit(1 if fuzz_orfcompare(seeds, args.runs_per_seed) else 0)
    fails = runs = aborted = bytes_cmp = 0
    for seed in seeds:
        rng = random.Random(seed)
        d = tempfile.mkdtemp(prefix="orfx_fuzz_%d_" % seed)
        synth_args, tpl = random_workload(rng, d, seed)
        bad_here = False
        for k in range(args.runs_per_seed):
            opts, templates = random_options(rng, tpl, not args.no_overhang)
            rc_r, gtf_r, tsv_r, err_r = run(REF_BIN, d, opts, "ref", templates, ["--threads", "1"])
            rc_o, gtf_o, tsv_o, err_o = run(ORACLE_CLI, d, opts, "cand", templates)
            runs += 1
            if rc_r < 0 and rc_r != -999:      # the reference itself died on a signal (its own assertion): nothing to match
                aborted += 1
                print("note: reference aborted (rc %d) seed %d: %s | %s" % (rc_r, seed, " ".join(opts), err_r.strip().splitlines()[-1:] or ""))
                continue
            what = None
            if rc_o != rc_r:
                what = "exit code %d vs reference %d" % (rc_o, rc_r)
            elif rc_r == 0 and gtf_o != gtf_r:
                what = "GTF differs"
            elif rc_r == 0 and tsv_o != tsv_r:
                what = "stats differ"
            bytes_cmp += len(gtf_r) + len(tsv_r)
            if what:
                fails += 1
                bad_here = True
                print("FAIL seed %d run %d: %s\n  synth: %s\n  opts: %s %s\n  stderr(ours): %s\n  stderr(ref): %s"
                      % (seed, k, what, " ".join(synth_args[1:]), " ".join(opts), " ".join(templates), err_o.strip()[-300:], err_r.strip()[-300:]))
        if bad_here and args.keep:
            print("  kept:", d)
        else:
            shutil.rmtree(d, ignore_errors=True)
    print("fuzz_parity: %d runs over %d seeds, %d failures, %d reference aborts, %.1f MB compared" % (runs, len(seeds), fails, aborted, bytes_cmp / 1e6))
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
