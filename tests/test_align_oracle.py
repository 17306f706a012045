"""CPU suite, part 3: the alignment oracle.

ksw2 is absent from /root/reference (un-vendored submodule, no call site): the scalar restatement in oracle/orfx_oracle.cpp is
"parity unpinned" against ksw2 itself.  What can be pinned is pinned here: (1) the restatement against an independent
three-matrix Gotoh extension DP written straight from the published recurrence; (2) the in-path alignment of the oracle
(which protein, which template, which encoding) against a Python restatement of SEQ::translate + aa_table + mat50 on the
reference's own literals (tests/golden/codon_map.txt, aa_table.npy, mat50.npy)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import orfanage_b200 as ob
from conftest import GOLDEN, template_files

MAT = np.load(os.path.join(GOLDEN, "mat50.npy")).astype(np.int64)
AA_TABLE = np.load(os.path.join(GOLDEN, "aa_table.npy"))
NEG = -0x40000000


def gotoh_extz(q, t, gapo=12, gape=2):
    """Three explicit matrices, H / E / F, global extension from (0,0) (ksw2's extz: SURVEY.md §8a-A1):
         H(i,j)   = max{H(i-1,j-1) + s(t_i,q_j), E(i,j), F(i,j)}
         E(i+1,j) = max{H(i,j) - gapo, E(i,j)} - gape,   F(i,j+1) = max{H(i,j) - gapo, F(i,j)} - gape
       borders H(-1,-1) = 0, H(-1,j) = H(j,-1) = -(gapo + (j+1) gape); E / F of a border cell open from that border.
       Returns (max(0, max H), H(tl-1,ql-1), max_i H(i,ql-1), max_j H(tl-1,j))."""
    ql, tl = len(q), len(t)
    if ql == 0 or tl == 0:
        return 0, NEG, NEG, NEG
    H = np.full((tl + 1, ql + 1), NEG, dtype=np.int64)   # index +1: row / column -1 are the borders
    E = np.full((tl + 1, ql + 1), NEG, dtype=np.int64)
    F = np.full((tl + 1, ql + 1), NEG, dtype=np.int64)
    H[0, 0] = 0
    for j in range(1, ql + 1):
        H[0, j] = -(gapo + j * gape)
    for i in range(1, tl + 1):
        H[i, 0] = -(gapo + i * gape)
    for i in range(1, tl + 1):
        for j in range(1, ql + 1):
            E[i, j] = max(H[i - 1, j] - gapo, E[i - 1, j]) - gape     # gap along the target: from the cell above
            F[i, j] = max(H[i, j - 1] - gapo, F[i, j - 1]) - gape     # gap along the query: from the cell on the left
            H[i, j] = max(H[i - 1, j - 1] + MAT[t[i - 1], q[j - 1]], E[i, j], F[i, j])
    body = H[1:, 1:]
    return max(0, int(body.max())), int(body[-1, -1]), int(body[:, -1].max()), int(body[-1, :].max())


def oracle_align(lib, q, t, gapo=12, gape=2):
    lib.orfx_oracle_align.restype = C.c_int
    lib.orfx_oracle_align.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    o = [C.c_int32() for _ in range(4)]
    q = np.ascontiguousarray(q, dtype=np.uint8)
    t = np.ascontiguousarray(t, dtype=np.uint8)
    mat = np.ascontiguousarray(MAT, dtype=np.int8)
    lib.orfx_oracle_align(q.ctypes.data, len(q), t.ctypes.data, len(t), mat.ctypes.data, 24, gapo, gape, *[C.byref(x) for x in o])
    return tuple(x.value for x in o)


def test_scalar_restatement_against_an_independent_gotoh(oracle_lib):
    rng = np.random.default_rng(11)
    n = 0
    for gapo, gape in ((12, 2), (10, 1), (0, 1), (5, 3), (60, 4)):
        for _ in range(24):
            ql, tl = int(rng.integers(1, 70)), int(rng.integers(1, 70))
            q = rng.integers(0, 24, size=ql).astype(np.uint8)
            t = q[:tl].copy() if rng.random() < 0.5 and tl <= ql else rng.integers(0, 24, size=tl).astype(np.uint8)
            if len(t) < tl:
                t = np.concatenate([t, rng.integers(0, 24, size=tl - len(t)).astype(np.uint8)])
            assert oracle_align(oracle_lib, q, t, gapo, gape) == gotoh_extz(q, t, gapo, gape), (gapo, gape, ql, tl)
            n += 1
    assert n == 120
    s = ob.encode_protein("MKTAYIAKQRQISFVKSHFSRQLEERLGLIEVQAPILSRVG")
    assert oracle_align(oracle_lib, s, s)[1] == int(MAT[s, s].sum())           # identical strings: the diagonal
    assert oracle_align(oracle_lib, s, np.concatenate([s[:20], s[25:]]))[1] == int(MAT[s, s].sum() - MAT[s[20:25], s[20:25]].sum() - (12 + 5 * 2))


def read_fasta(path):
    seqs, name = {}, None
    for line in open(path):
        if line.startswith(">"):
            name = line[1:].split()[0]
            seqs[name] = []
        else:
            seqs[name].append(line.strip())
    return {k: "".join(v) for k, v in seqs.items()}


COMP = str.maketrans("AaCcTtGgUuMmRrWwSsYyKkVvHhDdBbNnXx", "TtGgAaCcAaKkYyWwSsRrMmBbDdHhVvNnXx")


def spliced(genome, segs, strand):
    s = "".join(genome[a - 1:b] for a, b in segs)
    return (s if strand == "+" else s.translate(COMP)[::-1]).upper()


def test_in_path_alignment_of_the_oracle_against_a_python_restatement(synth, oracle_lib):
    """Every scored row: protein of the row's ORF (SEQ::translate of the genomic span cds_start..cds_end on the spliced
    query, stop codon included) against the protein of its template (its spliced CDS chain from the first base), both
    through aa_table, aligned with mat50, gap open 12 / extend 2 -- recomputed here from the GTFs and the FASTA."""
    d = synth(81, loci=40, lower=0.4, iupac=0.002, titin=0.0)
    codon = {}
    for line in open(os.path.join(GOLDEN, "codon_map.txt")):
        if len(line.strip()) == 4:
            codon[line[:3]] = line[3]
    genome = read_fasta(os.path.join(d, "genome.fa"))
    chrom = {}
    for f in [os.path.join(d, "query.gtf")] + template_files(d):
        for line in open(f):
            c = line.split("\t")
            if len(c) > 8 and c[2] == "transcript":
                chrom[re.search(r'transcript_id "([^"]+)"', c[8]).group(1)] = c[0]
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    batch = w.pack(layout="islands_t")
    a = batch.arrays
    with ob.Context(ob.Params.make(mode="LONGEST_MATCH,BEST", align=True), lib=oracle_lib, prefix="orfx_oracle_") as ox:
        res = ox.run(batch)
    rows = res.rows

    def chain(tx):
        return [(int(s), int(e)) for s, e in a["segs"][a["tx_seg_off"][tx]:a["tx_seg_off"][tx + 1]]]

    def protein(nt):
        aa = [codon.get(nt[i:i + 3], "\0") for i in range(0, len(nt) - 2, 3)]
        return np.array([AA_TABLE[ord(c)] if ord(c) < 128 else 23 for c in aa], dtype=np.uint8)
    scored = np.nonzero(rows["qlen"] > 0)[0]
    assert len(scored) > 150
    assert not rows["aln_max"][rows["qlen"] == 0].any() and not rows["aln_score"][rows["qlen"] == 0].any()
    rng = np.random.default_rng(3)
    strands = set()
    for r in rng.choice(scored, size=40, replace=False):
        qx, tx = int(rows["q_tx"][r]), int(rows["t_tx"][r])
        g = genome[chrom[w.tx_id(qx)]]
        strand = chr(a["tx_strand"][qx])
        strands.add(strand)
        cs, ce = int(rows["cds_start"][r]), int(rows["cds_end"][r])
        qcds = [(max(s, cs), min(e, ce)) for s, e in chain(qx) if e >= cs and s <= ce]
        qp = protein(spliced(g, qcds, strand))
        tp = protein(spliced(g, chain(tx), chr(a["tx_strand"][tx])))
        assert len(qp) * 3 == rows["qlen"][r] and qp[-1] == 23          # the ORF ends with its stop codon ('.' -> residue 23)
        want = gotoh_extz(qp, tp) if len(qp) * len(tp) < 40000 else oracle_align(oracle_lib, qp, tp)
        got = tuple(int(rows[k][r]) for k in ("aln_max", "aln_score", "aln_mqe", "aln_mte"))
        assert got == want, (w.tx_id(qx), w.tx_id(tx), got, want)
    assert strands == {"+", "-"}
    w.close()
