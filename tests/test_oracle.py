"""CPU suite, part 1: the oracle (CPU restatement) is pinned against the reference.

  * byte parity of oracle/orfanage_oracle_cli with oracle/_ref/orfanage_ref (the reference's own sources
    compiled against the gclib stand-in) on seeded synthetic inputs over an option grid;
  * byte parity with the committed goldens (tests/golden/, written by make_golden.py from the reference);
  * config 1 at the C-ABI level: committed packed batch -> oracle library -> the reference's stats rows;
  * constants: codon table, mat50, aa_table against the literals parsed from src/common.hpp;
  * the docs' worked orfcompare rows reproduce from the reference build (pins the gclib stand-in).
"""
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import orfanage_b200 as ob
from conftest import GOLDEN, ORACLE_CLI, REF_BIN, ROOT, run_cli, template_files
from fmt import stats_lines
from golden.make_golden import SYNTH_ARGS, SYNTH_RUNS, SYNTH_SEEDS

OPTION_GRID = [
    ["--reference", "genome.fa", "--query", "query.gtf"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "BEST", "--cleanq"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "ALL"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "START_MATCH,FIRST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "LONGEST", "--non_aug"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "LONGEST_MATCH,BEST", "--lpi", "50", "--ilpi", "60",
     "--mlpi", "70", "--minlen", "300"],
    ["--query", "query.gtf", "--mode", "LONGEST_MATCH,BEST"],
    ["--query", "query.gtf", "--mode", "ALL", "--lpi", "80"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--use_id"],
    # template cleaning (Transcriptome::clean_cds / clean_short_orfs)
    ["--reference", "genome.fa", "--query", "query.gtf", "--cleant"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--rescue", "--mode", "LONGEST_MATCH,BEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--rescue", "--non_aug", "--mode", "BEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--minlen", "450", "--use_id"],
    # --overhang (TX::extend_seq and the exon growth of extend_to_stop / extend_to_start)
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "30"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "300", "--mode", "ALL"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "90", "--spliced_overhang", "--mode", "BEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "2000", "--spliced_overhang", "--non_aug", "--mode", "LONGEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "150", "--cleant", "--rescue"],
]
SELF_GRID = [   # query file = a template file (queries that carry a CDS)
    (["--reference", "genome.fa", "--query", "template2.gtf", "--keep_all_cds"], ["template1.gtf"]),
    (["--reference", "genome.fa", "--query", "template3.gtf", "--keep_cds_if_not_found", "--mode", "FIRST"], ["template1.gtf"]),
    (["--reference", "genome.fa", "--query", "template3.gtf", "--mode", "BEST"], ["template1.gtf"]),
]

needs_ref_bin = pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/orfanage_ref not built")


def compare_clis(cand, d, args, templates=None):
    rc_r, gtf_r, tsv_r, err_r = run_cli(REF_BIN, d, args + ["--threads", "1"], "ref", templates)
    rc_o, gtf_o, tsv_o, err_o = run_cli(cand, d, args, "cand", templates)
    assert rc_o == rc_r, (args, err_o[-400:], err_r[-400:])
    assert gtf_o == gtf_r, "GTF differs for %s" % args
    assert tsv_o == tsv_r, "stats TSV differs for %s" % args
    return len(gtf_r)


@needs_ref_bin
@pytest.mark.parametrize("seed", [1, 2])
def test_oracle_cli_matches_reference_binary(synth, seed):
    d = synth(seed, loci=60, template_files=3)
    total = 0
    for args in OPTION_GRID:
        total += compare_clis(ORACLE_CLI, d, args)
    for args, tpl in SELF_GRID:
        total += compare_clis(ORACLE_CLI, d, args, tpl)
    assert total > 100000   # the runs produced real output


@needs_ref_bin
def test_fuzz_parity_sample():
    """A slice of the randomised differential campaign (tools/fuzz_parity.py: random generator settings x random option
    combinations, oracle CLI vs reference binary, bytes and exit codes).  The full campaigns and what they found are in
    DESIGN.md §4c / profiles/r02_fuzz_parity.txt."""
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "--seeds", "200:212", "--runs-per-seed", "5"],
                       capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-2000:]
    assert "60 runs over 12 seeds, 0 failures" in p.stdout


@pytest.mark.parametrize("seed", SYNTH_SEEDS)
def test_oracle_cli_matches_goldens(tmp_path, seed):
    d = str(tmp_path)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", str(seed)] + SYNTH_ARGS, check=True,
                   capture_output=True)
    for tag, args in SYNTH_RUNS:
        rc, gtf, tsv, err = run_cli(ORACLE_CLI, d, args, "o_" + tag, ["template1.gtf", "template2.gtf"])
        assert rc == 0, err
        base = os.path.join(GOLDEN, "synth_s%d_%s" % (seed, tag))
        assert gtf == open(base + ".gtf", "rb").read(), tag
        assert tsv == open(base + ".tsv", "rb").read(), tag


def test_overhang_worked_examples(tmp_path):
    """--overhang by hand: stop in the 3' flank, ATG in the 5' flank, and the reference's
    `extend_len - starts[0]` growth after a splice cut — the oracle-backed CLI and, where built, the reference binary."""
    from conftest import check_overhang_kat, write_overhang_kat
    for minus in (False, True):   # and mirrored onto the reverse strand
        d = str(tmp_path / ("minus" if minus else "plus"))
        os.makedirs(d)
        expected = write_overhang_kat(d, minus)
        check_overhang_kat(ORACLE_CLI, d, expected)
        if os.path.exists(REF_BIN):
            check_overhang_kat(REF_BIN, d, expected)


DOT_EDGE_GRID = [
    ["--reference", "genome.fa", "--query", "query.gtf"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "ALL", "--non_aug"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--cleant", "--rescue", "--mode", "LONGEST_MATCH,BEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "30"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "300", "--mode", "ALL"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "90", "--spliced_overhang", "--mode", "BEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "2000", "--spliced_overhang", "--non_aug", "--mode", "LONGEST"],
    ["--reference", "genome.fa", "--query", "query.gtf", "--overhang", "40", "--cleant", "--rescue", "--spliced_overhang", "--mode", "ALL"],
]


def dot_edge_dir(synth, seed):
    """Loci written with strand '.' (read as reverse-strand sequence, with the reference's `strand == '-'` / `strand == '+'`
    asymmetries: both --overhang flanks from the low side, 5' CDS extensions landing on the 3' end) and loci whose main
    isoform touches the first / last base of its chromosome (one-base flank that re-reads the edge base)."""
    return synth(seed, loci=120, chroms=6, chrom_len=800000, queries=8, templates=3, template_files=2, dot=0.3, edge=True)


@needs_ref_bin
@pytest.mark.parametrize("seed", [71, 72])
def test_oracle_cli_matches_reference_on_strand_dot_and_chromosome_edges(synth, seed):
    d = dot_edge_dir(synth, seed)
    strands = {l.split("\t")[6] for l in open(os.path.join(d, "query.gtf")) if "\ttranscript\t" in l}
    assert strands == {"+", "-", "."}
    for args in DOT_EDGE_GRID:
        compare_clis(ORACLE_CLI, d, args)


def test_chromosome_edge_worked_example(tmp_path):
    """The one-base flank of a transcript that ends on its chromosome's last base, by hand (conftest.write_edge_kat)."""
    from conftest import check_edge_kat, write_edge_kat
    d = str(tmp_path)
    expected = write_edge_kat(d)
    got = check_edge_kat(ORACLE_CLI, d, expected)
    if os.path.exists(REF_BIN):
        assert check_edge_kat(REF_BIN, d, expected) == got


def test_config1_oracle_library_matches_reference_rows(oracle_lib):
    batch = ob.Batch.load(os.path.join(GOLDEN, "c1_batch.npz"))
    names = json.load(open(os.path.join(GOLDEN, "c1_names.json")))["tx_id"]
    with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as cx:
        res = cx.run(batch)
    want = open(os.path.join(GOLDEN, "c1_ref.tsv")).read().splitlines()[1:]
    assert stats_lines(res, names) == want
    # the chosen ORFs are the CDS rows of the reference's GTF
    gtf = open(os.path.join(GOLDEN, "c1_ref.gtf")).read()
    for b in res.best:
        tid = names[b["q_tx"]]
        cds = [(int(m.group(1)), int(m.group(2))) for m in
               re.finditer(r"\tCDS\t(\d+)\t(\d+)\t[^\n]*transcript_id \"%s\";" % re.escape(tid), gtf)]
        if b["t_tx"] < 0:
            assert not cds
        else:
            assert (min(c[0] for c in cds), max(c[1] for c in cds)) == (b["cds_start"], b["cds_end"])


def test_codon_table_matches_reference_literal():
    want = open(os.path.join(GOLDEN, "codon_map.txt")).read().split()
    hdr = open(os.path.join(ROOT, "include", "orfx_codon_table.h")).read()
    got = re.findall(r"\"([A-Z]{3}[A-Z.])\"", hdr)
    assert sorted(got) == sorted(want) and len(got) == 180
    if os.path.exists("/root/reference/src/common.hpp"):   # the golden itself is current
        src = open("/root/reference/src/common.hpp").read()
        pairs = re.findall(r"\{\s*\"([A-Z]{3})\"\s*,\s*'(.)'\s*\}", src)
        assert sorted(c + a for c, a in pairs) == sorted(want)


def test_mat50_and_aa_table_match_reference_literals():
    mat = np.load(os.path.join(GOLDEN, "mat50.npy"))
    hdr = open(os.path.join(ROOT, "include", "orfx_mat50.h")).read()   # the literal the device library and the oracle share
    body = hdr[hdr.index("#define ORFX_MAT50_LITERAL"):]
    body = body[body.index("{") + 1:body.index("}")]
    got = np.array([int(x) for x in re.findall(r"-?\d+", body)], dtype=np.int8).reshape(24, 24)
    assert (got == mat).all()
    assert (mat == mat.T).all()
    table = np.load(os.path.join(GOLDEN, "aa_table.npy"))
    all_bytes = bytes(range(128))
    assert (ob.encode_protein(all_bytes) == table).all()


def test_orfcompare_known_answers():
    rows = [l.rstrip("\n").split("\t") for l in open(os.path.join(GOLDEN, "orfcompare_kat.tsv"))]
    # docs/content/csvs/orfcompare_stats.csv, Examples 1 and 2
    assert rows[1][:12] == ["CHS.8812.2", "ENST00000300146.10", "1566", "2313", "1566", "1566", "0", "0", "747", "92.5532",
                            "92.5532", "92.5532"]
    assert rows[2][:12] == ["CHS.8812.5", "ENST00000300146.10", "2310", "2313", "2310", "2310", "0", "0", "3", "99.8703",
                            "99.8703", "99.8703"]
    cmp_bin = os.path.join(ROOT, "oracle", "_ref", "orfcompare_ref")
    if os.path.exists(cmp_bin) and os.path.exists("/root/reference/example/patl.chess3.gtf"):
        out = "/tmp/orfx_kat.tsv"
        subprocess.run([cmp_bin, "--query", "/root/reference/example/patl.chess3.gtf", "--template",
                        "/root/reference/example/patl.gencode.gtf", "--output", out, "--threads", "1"], check=True,
                       capture_output=True)
        assert open(out).read() == open(os.path.join(GOLDEN, "orfcompare_kat.tsv")).read()


# ---- orfcompare -------------------------------------------------------------------
CMP_REF = os.path.join(ROOT, "oracle", "_ref", "orfcompare_ref")


def kat_rows():
    return [l.rstrip("\n").split("\t")[:12] for l in open(os.path.join(GOLDEN, "orfcompare_kat.tsv"))][1:]


def test_orfcompare_oracle_library_reproduces_the_docs_rows(oracle_lib):
    from fmt import compare_lines
    batch = ob.Batch.load(os.path.join(GOLDEN, "kat_compare_batch.npz"))
    names = json.load(open(os.path.join(GOLDEN, "kat_compare_names.json")))["tx_id"]
    with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as ox:
        assert compare_lines(ox.score_pairs(batch), names) == kat_rows()


@pytest.mark.parametrize("seed", SYNTH_SEEDS)
def test_orfcompare_oracle_cli_matches_goldens(tmp_path, seed):
    from golden.make_golden import CMP_RUNS
    d = str(tmp_path)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", str(seed)] + SYNTH_ARGS, check=True,
                   capture_output=True)
    for tag, args in CMP_RUNS:
        p = subprocess.run([ORACLE_CLI, "orfcompare"] + args + ["--output", "o_" + tag + ".tsv"], cwd=d, capture_output=True)
        assert p.returncode == 0, p.stderr.decode()[-400:]
        assert open(os.path.join(d, "o_" + tag + ".tsv"), "rb").read() == open(os.path.join(GOLDEN, "cmp_s%d_%s.tsv" % (seed, tag)), "rb").read(), tag


@pytest.mark.skipif(not os.path.exists(CMP_REF), reason="oracle/_ref/orfcompare_ref not built")
def test_orfcompare_oracle_cli_matches_reference_binary(synth):
    d = synth(3, loci=150, template_files=3, queries=8)
    total = 0
    for args in (["--query", "template2.gtf", "--template", "template1.gtf", "--reference", "genome.fa"],
                 ["--query", "template3.gtf", "--template", "template2.gtf"],
                 ["--query", "template1.gtf", "--template", "template3.gtf", "--use_id", "--reference", "genome.fa"]):
        subprocess.run([CMP_REF] + args + ["--threads", "1", "--output", "cmp_ref.tsv"], cwd=d, check=True, capture_output=True)
        subprocess.run([ORACLE_CLI, "orfcompare"] + args + ["--output", "cmp_or.tsv"], cwd=d, check=True, capture_output=True)
        a, b = open(os.path.join(d, "cmp_ref.tsv"), "rb").read(), open(os.path.join(d, "cmp_or.tsv"), "rb").read()
        assert a == b, args
        total += a.count(b"\n")
    assert total > 500
