"""Full-size byte-parity pins: runs the reference's own binary (oracle/_ref/orfanage_ref --threads 1) on the seeded
full-size workloads below and records the SHA-256 of its GTF and stats TSV in tests/golden/fullsize_hashes.json.
The GPU suite (tests/test_gpu_fullsize.py) regenerates the same workloads on the box (tools/orfx_synth is a pure
function of its flags), runs the CUDA CLI with the same options and compares the digests; with ORFX_PARITY_DIRECT=1
it also runs the reference there and cmp's the files.  Run here (the container that has /root/reference):

    python tests/golden/make_fullsize_hashes.py [case ...]

About 5 minutes of single-thread CPU per case (the reference at --threads 1: its output order is nondeterministic above 1)."""
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
OUT = os.path.join(ROOT, "tests", "golden", "fullsize_hashes.json")

# name -> (generator flags, template files, CLI options)
CASES = {
    # BASELINE.json configs[1] exactly as bench.py generates it
    "c2": (dict(seed=20261104, loci=20000, chroms=24, chrom_len=125000000, queries=13, templates=8, template_files=1, titin=0.01),
           ["--mode", "LONGEST_MATCH,BEST"]),
    # configs[2]-shaped: three template files (MANE / GENCODE / CHESS-like, heavy dedup across files) + the three thresholds
    "c3": (dict(seed=20261105, loci=5000, chroms=6, chrom_len=125000000, queries=13, templates=4, template_files=3, titin=0.01),
           ["--mode", "LONGEST_MATCH,BEST", "--lpi", "50", "--ilpi", "60", "--mlpi", "70"]),
    # configs[3]-shaped: ~100 query isoforms per locus; titin-class loci of 100-360 exons then hold 10-36 k segments, above
    # the heavy-bundle kernels' shared-memory stage (9 216 segments / 8 192 pairs)
    "c4": (dict(seed=20261106, loci=1500, chroms=2, chrom_len=112500000, queries=100, templates=6, template_files=1, titin=0.02),
           ["--mode", "LONGEST_MATCH,BEST"]),
}


def synth_args(cfg, out):
    return ["--out", out, "--seed", str(cfg["seed"]), "--loci", str(cfg["loci"]), "--chroms", str(cfg["chroms"]), "--chrom-len", str(cfg["chrom_len"]),
            "--queries", str(cfg["queries"]), "--templates", str(cfg["templates"]), "--template-files", str(cfg["template_files"]),
            "--titin-frac", str(cfg["titin"])]


def template_names(cfg):
    return ["template.gtf"] if cfg["template_files"] == 1 else ["template%d.gtf" % (k + 1) for k in range(cfg["template_files"])]


def generate(name, base):
    cfg, _ = CASES[name]
    d = os.path.join(base, "fullsize_" + name)
    if not os.path.exists(os.path.join(d, ".done")):
        os.makedirs(d, exist_ok=True)
        subprocess.run([os.path.join(ROOT, "tools", "orfx_synth")] + synth_args(cfg, d), check=True, stderr=subprocess.DEVNULL)
        open(os.path.join(d, ".done"), "w").write("ok\n")
    return d


def cli_args(name, tag, extra=()):
    cfg, opts = CASES[name]
    return ["--reference", "genome.fa", "--query", "query.gtf", "--output", tag + ".gtf", "--stats", tag + ".tsv"] + opts + list(extra) + template_names(cfg)


def digest(path):
    h = hashlib.sha256()
    n = 0
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 22), b""):
            h.update(chunk)
            n += chunk.count(b"\n")
    return h.hexdigest(), n


def main():
    names = sys.argv[1:] or list(CASES)
    base = os.environ.get("ORFX_BENCH_DIR", "/tmp/orfx_bench")
    res = json.load(open(OUT)) if os.path.exists(OUT) else {}
    ref = os.path.join(ROOT, "oracle", "_ref", "orfanage_ref")
    for name in names:
        d = generate(name, base)
        t = time.time()
        subprocess.run([ref] + cli_args(name, "ref", ["--threads", "1"]), cwd=d, check=True, capture_output=True)
        g, gl = digest(os.path.join(d, "ref.gtf"))
        s, sl = digest(os.path.join(d, "ref.tsv"))
        res[name] = {"generator": CASES[name][0], "options": CASES[name][1], "gtf_sha256": g, "gtf_lines": gl, "tsv_sha256": s, "tsv_lines": sl,
                     "reference_seconds_threads1": round(time.time() - t, 1)}
        print(name, res[name], flush=True)
        json.dump(res, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
