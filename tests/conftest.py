"""Shared fixtures.  `-m "not gpu"` covers the oracle against the reference binary and the goldens, the
host logic and the C-ABI exports; `-m gpu` holds the parity tests proper (CUDA path vs oracle)."""
import ctypes
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "orfanage_ref")
ORACLE_CLI = os.path.join(ROOT, "oracle", "orfanage_oracle_cli")
ORACLE_LIB = os.path.join(ROOT, "oracle", "liborfx_oracle.so")
SYNTH = os.path.join(ROOT, "tools", "orfx_synth")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def native_build():
    """CPU-side targets are cheap: make sure they exist.  The CUDA library is built by
    __graft_entry__.build() / `make cuda` and only checked for presence here."""
    need = [ORACLE_CLI, ORACLE_LIB, SYNTH, os.path.join(ROOT, "orfanage_b200", "liborfx_host.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.run(["make", "-C", ROOT, "tools", "host", "oracle"], check=True, stdout=subprocess.DEVNULL)
    return True


_SYNTH_CACHE = {}


def synth_dir(tmp_root, seed, loci=60, chroms=2, chrom_len=3000000, queries=6, templates=3, template_files=1, lower=0.3,
              iupac=0.0005, titin=0.02, dot=0.0, edge=False):
    key = (seed, loci, chroms, chrom_len, queries, templates, template_files, lower, iupac, titin, dot, edge)
    if key in _SYNTH_CACHE:
        return _SYNTH_CACHE[key]
    d = os.path.join(str(tmp_root), "synth_%d_%d_%d_%d" % (seed, loci, template_files, len(_SYNTH_CACHE)))
    os.makedirs(d, exist_ok=True)
    subprocess.run([SYNTH, "--out", d, "--seed", str(seed), "--loci", str(loci), "--chroms", str(chroms), "--chrom-len",
                    str(chrom_len), "--queries", str(queries), "--templates", str(templates), "--template-files",
                    str(template_files), "--lower-frac", str(lower), "--iupac-rate", str(iupac), "--titin-frac", str(titin)]
                   + (["--dot-frac", str(dot)] if dot else []) + (["--edge-loci"] if edge else []),
                   check=True, stderr=subprocess.DEVNULL)
    _SYNTH_CACHE[key] = d
    return d


@pytest.fixture(scope="session")
def synth(tmp_path_factory):
    root = tmp_path_factory.mktemp("orfx")

    def make(seed, **kw):
        return synth_dir(root, seed, **kw)
    return make


def template_files(d):
    single = os.path.join(d, "template.gtf")
    if os.path.exists(single):
        return [single]
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.startswith("template") and f.endswith(".gtf"))


@pytest.fixture(scope="session")
def oracle_lib():
    return ctypes.CDLL(ORACLE_LIB)


def run_cli(binary, d, args, tag, templates=None):
    """Runs an orfanage-compatible CLI in directory d; returns (rc, gtf bytes, stats bytes, stderr)."""
    tpl = templates if templates is not None else [os.path.basename(t) for t in template_files(d)]
    cmd = [binary] + list(args) + ["--output", tag + ".gtf", "--stats", tag + ".tsv"] + tpl
    p = subprocess.run(cmd, cwd=d, capture_output=True)
    def rd(name):
        path = os.path.join(d, name)
        return open(path, "rb").read() if os.path.exists(path) else b""
    return p.returncode, rd(tag + ".gtf"), rd(tag + ".tsv"), p.stderr.decode(errors="replace")


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def write_overhang_kat(d, minus=False):
    """Two hand-derivable --overhang cases on an all-C genome (no codon of interest outside what is planted):
    Q1, exon 131-160 = CC ATG (GCT)x8 G with CT TAA right behind it: the stop search runs 5 bases into the 3' flank, the
       exon and the CDS grow to 165 (Transcriptome.cpp:418-429);
    Q2, exon 231-260 = (GCT)x9 TAA, no ATG, with ATG at 225-227 and AG at 215-216: the unguided start search finds the ATG
       6 bases into the 5' flank -> exon and CDS start at 225; with --spliced_overhang the flank is cut after the AG (14
       bases left), the ATG sits at position 8 of the flanked sequence and the exon grows by overhang - 8 = 12 -> exon
       starts at 219 while the CDS starts at 225 (cpp:461-474, as written).
    minus=True mirrors everything onto the reverse strand (reverse-complemented genome, coordinates 401 - x).
    Returns the expected (exon start, exon end, CDS start, CDS end) per transcript for plain and spliced runs."""
    g = ["C"] * 400
    L = 400

    def put(pos, s):
        for i, ch in enumerate(s):
            g[pos - 1 + i] = ch
    put(131, "CC" + "ATG" + "GCT" * 8 + "G")
    put(161, "CT" + "TAA")
    put(231, "GCT" * 9 + "TAA")
    put(225, "ATG")
    put(215, "AG")
    if minus:
        g = [{"A": "T", "C": "G", "G": "C", "T": "A"}[c] for c in reversed(g)]
    open(os.path.join(d, "genome.fa"), "w").write(">chr1\n" + "".join(g) + "\n")
    strand = "-" if minus else "+"

    def span(s, e):
        return (L + 1 - e, L + 1 - s) if minus else (s, e)

    def tx(tid, gid, s, e, cds=None):
        attrs = 'transcript_id "%s"; gene_id "%s";' % (tid, gid)
        s, e = span(s, e)
        rows = ["chr1\tKAT\ttranscript\t%d\t%d\t.\t%s\t.\t%s" % (s, e, strand, attrs), "chr1\tKAT\texon\t%d\t%d\t.\t%s\t.\t%s" % (s, e, strand, attrs)]
        if cds:
            rows.append("chr1\tKAT\tCDS\t%d\t%d\t.\t%s\t0\t%s" % (span(*cds) + (strand, attrs)))
        return "\n".join(rows) + "\n"
    q = [tx("Q1", "G1", 131, 160), tx("Q2", "G2", 231, 260)]
    t = [tx("T1", "G1", 133, 160, (133, 159)), tx("T2", "G2", 231, 260, (231, 260))]
    open(os.path.join(d, "query.gtf"), "w").write("".join(reversed(q) if minus else q))
    open(os.path.join(d, "template.gtf"), "w").write("".join(reversed(t) if minus else t))
    want = {False: {"Q1": (131, 165, 133, 165), "Q2": (225, 260, 225, 260)},
            True: {"Q1": (131, 165, 133, 165), "Q2": (219, 260, 225, 260)}}
    return {sp: {tid: span(v[0], v[1]) + span(v[2], v[3]) for tid, v in w.items()} for sp, w in want.items()}


def check_overhang_kat(binary, d, expected):
    for spliced, want in expected.items():
        args = ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "20"] + (["--spliced_overhang"] if spliced else [])
        rc, gtf, tsv, err = run_cli(binary, d, args, "kat%d" % spliced, ["template.gtf"])
        assert rc == 0, err
        got = {}
        for line in gtf.decode().splitlines():
            f = line.split("\t")
            if len(f) < 9 or f[2] not in ("exon", "CDS"):
                continue
            tid = f[8].split('"')[1]
            got.setdefault(tid, {})[f[2]] = (int(f[3]), int(f[4]))
        for tid, (es, ee, cs, ce) in want.items():
            assert got[tid]["exon"] == (es, ee) and got[tid]["CDS"] == (cs, ce), (spliced, tid, got[tid])


def write_edge_kat(d):
    """A transcript that ends on the last base of its chromosome (TX::extend_seq, Transcriptome.cpp:348-364: the clamp
    re-reads that base as a one-base 3' flank).  Exon 201-300 of a 300-base chromosome; ORF A, ATG at 203, runs to the end
    and stops short at "TA": the repeated last base completes TAA, the exon and the CDS grow to 301, the re-loaded sequence
    ends one base early (Bundle::get_nts, cpp:776-782) and the row ends as "missing_stop" keyed 203-301.  ORF B (213-224,
    another frame) is complete and is what gets reported.  Returns the expected stats keys in print order."""
    L = 300
    g = ["C"] * L

    def put(pos, s):
        for i, ch in enumerate(s):
            g[pos - 1 + i] = ch
    put(203, "ATG" + "GCC" * 31 + "TA")
    put(213, "ATGGCCGCCTAA")
    open(os.path.join(d, "genome.fa"), "w").write(">chr1\n" + "".join(g) + "\n")

    def tx(tid, s, e, cds=None):
        attrs = 'transcript_id "%s"; gene_id "G1";' % tid
        rows = ["chr1\tKAT\ttranscript\t%d\t%d\t.\t+\t.\t%s" % (s, e, attrs), "chr1\tKAT\texon\t%d\t%d\t.\t+\t.\t%s" % (s, e, attrs)]
        if cds:
            rows.append("chr1\tKAT\tCDS\t%d\t%d\t.\t+\t0\t%s" % (cds + (attrs,)))
        return "\n".join(rows) + "\n"
    open(os.path.join(d, "query.gtf"), "w").write(tx("Q1", 201, 300))
    open(os.path.join(d, "template.gtf"), "w").write(tx("T1", 201, 300, (203, 298)) + tx("T2", 201, 300, (213, 224)))
    return [("Q1.203-301", "T1", "missing_stop"), ("Q1.213-224", "T2", "best_gtf")]


def check_edge_kat(binary, d, expected):
    rc, gtf, tsv, err = run_cli(binary, d, ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "20", "--mode", "ALL"] +
                                (["--threads", "1"] if binary == REF_BIN else []), "edge", ["template.gtf"])
    assert rc == 0, err
    rows = [tuple(l.split("\t")[i] for i in (0, 1, 3)) for l in tsv.decode().splitlines()[1:]]
    assert rows == expected, rows
    return gtf, tsv
