"""GPU suite: protein alignment kernel (orfx_align_batch) against the scalar CPU restatement of the
ksw2 extension recurrence (oracle/orfx_oracle.cpp: orfx_oracle_align — parity UNPINNED against ksw2
itself) and against hand-computable cases."""
import ctypes as C
import os

import numpy as np
import pytest

import orfanage_b200 as ob
from conftest import GOLDEN, ORACLE_CLI, run_cli, template_files

pytestmark = pytest.mark.gpu

MAT = np.load(os.path.join(GOLDEN, "mat50.npy"))
# background amino-acid frequencies (order ARNDCQEGHILKMFPSTWYV), proteome-like
FREQ = np.array([8.25, 5.53, 4.06, 5.45, 1.37, 3.93, 6.75, 7.07, 2.27, 5.96, 9.66, 5.84, 2.42, 3.86, 4.70, 6.56, 5.34, 1.08,
                 2.92, 6.87])
FREQ = FREQ / FREQ.sum()


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


def mutate(rng, s, sub, n_indel):
    s = s.copy()
    m = rng.random(len(s)) < sub
    s[m] = rng.choice(20, size=int(m.sum()), p=FREQ)
    for _ in range(n_indel):
        p = int(rng.integers(0, max(len(s), 1)))
        k = int(rng.integers(1, 12))
        if rng.random() < 0.5:
            s = np.concatenate([s[:p], s[p + k:]])
        else:
            s = np.concatenate([s[:p], rng.choice(20, size=k, p=FREQ).astype(np.uint8), s[p:]])
    return s if len(s) else np.array([0], dtype=np.uint8)


def make_pairs(rng, lengths):
    seqs, pq, pt = [], [], []
    for L in lengths:
        a = rng.choice(20, size=int(L), p=FREQ).astype(np.uint8)
        b = mutate(rng, a, rng.uniform(0.05, 0.4), int(rng.integers(0, 4))).astype(np.uint8)
        pq.append(len(seqs)); seqs.append(a)
        pt.append(len(seqs)); seqs.append(b)
    off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    return seqs, np.concatenate(seqs), off, np.array(pq, dtype=np.int32), np.array(pt, dtype=np.int32)


@pytest.fixture(scope="module")
def gpu():
    cx = ob.Context(ob.Params.make(), device=0)
    yield cx
    cx.close()


def test_alignment_matches_scalar_restatement(gpu, oracle_lib):
    rng = np.random.default_rng(2026)
    lengths = list(rng.integers(1, 40, size=40)) + list(np.exp(rng.normal(np.log(400), 0.9, size=150)).clip(50, 4000).astype(int)) \
        + [255, 256, 257, 511, 512, 513, 1024, 2048, 3000]
    seqs, aa, off, pq, pt = make_pairs(rng, lengths)
    out, cells, ms = gpu.align(aa, off, pq, pt)
    assert cells == sum(len(seqs[q]) * len(seqs[t]) for q, t in zip(pq, pt))
    for i, (q, t) in enumerate(zip(pq, pt)):
        want = oracle_align(oracle_lib, seqs[q], seqs[t])
        got = (int(out["max"][i]), int(out["score"][i]), int(out["mqe"][i]), int(out["mte"][i]))
        assert got == want, "pair %d (qlen %d, tlen %d): %s != %s" % (i, len(seqs[q]), len(seqs[t]), got, want)


def test_alignment_other_gap_costs_and_wildcards(gpu, oracle_lib):
    rng = np.random.default_rng(7)
    seqs, aa, off, pq, pt = make_pairs(rng, [30, 333, 700, 1500])
    aa = aa.copy()
    aa[rng.integers(0, len(aa), size=40)] = rng.integers(20, 24, size=40)   # B Z X *
    seqs = [aa[off[i]:off[i + 1]] for i in range(len(off) - 1)]
    for gapo, gape in [(12, 2), (10, 1), (5, 3), (0, 1)]:
        out, _, _ = gpu.align(aa, off, pq, pt, gapo, gape)
        for i, (q, t) in enumerate(zip(pq, pt)):
            want = oracle_align(oracle_lib, seqs[q], seqs[t], gapo, gape)
            assert (int(out["max"][i]), int(out["score"][i]), int(out["mqe"][i]), int(out["mte"][i])) == want


def test_alignment_hand_computable_cases(gpu):
    # identical strings: score = sum of the diagonal entries; a single k-residue gap costs gapo + k*gape
    s = ob.encode_protein("MKTAYIAKQRQISFVKSHFSRQLEERLGLIEVQAPILSRVGDGTQDNLSGAEKAVQVKVKALPDAQFEVVHSLAKWKR")
    diag = int(MAT[s, s].sum())
    gapped = np.concatenate([s[:30], s[35:]])   # 5 residues deleted from the target
    aa = np.concatenate([s, s, gapped])
    off = np.array([0, len(s), 2 * len(s), 2 * len(s) + len(gapped)], dtype=np.int64)
    out, _, _ = gpu.align(aa, off, [0, 0], [1, 2])
    assert out["score"][0] == diag and out["max"][0] == diag and out["mqe"][0] == diag and out["mte"][0] == diag
    assert out["score"][1] == diag - int(MAT[s[30:35], s[30:35]].sum()) - (12 + 5 * 2)


def test_alignment_titin_class_pair(gpu, oracle_lib):
    """One long pair (block-striped path, int32 scores beyond the int16 range)."""
    rng = np.random.default_rng(99)
    seqs, aa, off, pq, pt = make_pairs(rng, [9000])
    out, cells, ms = gpu.align(aa, off, pq, pt)
    want = oracle_align(oracle_lib, seqs[0], seqs[1])
    assert (int(out["max"][0]), int(out["score"][0]), int(out["mqe"][0]), int(out["mte"][0])) == want
    assert want[0] > 32767


def test_alignment_empty_inputs(gpu):
    out, cells, _ = gpu.align(np.zeros(0, dtype=np.uint8), np.array([0], dtype=np.int64), [], [])
    assert cells == 0 and len(out["max"]) == 0
    aa = np.array([3, 4, 5], dtype=np.uint8)
    out, _, _ = gpu.align(aa, np.array([0, 3, 3], dtype=np.int64), [0], [1])   # empty target
    assert out["max"][0] == 0


def test_alignment_many_long_pairs_share_a_cta(gpu, oracle_lib):
    """More long pairs than CTAs, of very different shapes (1 to 40 stripes, some finishing before their predecessor),
    plus short ones: the long pairs of a CTA are pipelined into each other without a CTA barrier, so result slots,
    progress counters and edge buffers are reused while neighbouring pairs are still in flight."""
    rng = np.random.default_rng(4242)
    ql = np.concatenate([rng.integers(1024, 2600, size=380), rng.integers(2600, 10000, size=90), [1024, 1025, 4096, 4097, 256 * 16, 256 * 16 + 1]])
    rng.shuffle(ql)
    seqs, pq, pt = [], [], []
    for L in ql:
        a = rng.choice(20, size=int(L), p=FREQ).astype(np.uint8)
        tl = int(max(4194304 // int(L) + 1, rng.integers(300, 5000)))       # >= 4 Mi cells: taken by the block-cooperative phase
        b = mutate(rng, np.resize(a, tl), rng.uniform(0.05, 0.4), int(rng.integers(0, 4))).astype(np.uint8)
        pq.append(len(seqs)); seqs.append(a)
        pt.append(len(seqs)); seqs.append(b)
    for L in rng.integers(20, 900, size=300):                                # phase B work for the same launch
        a = rng.choice(20, size=int(L), p=FREQ).astype(np.uint8)
        pq.append(len(seqs)); seqs.append(a)
        pt.append(len(seqs)); seqs.append(mutate(rng, a, 0.2, 2).astype(np.uint8))
    aa = np.concatenate(seqs)
    off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    pq, pt = np.array(pq, dtype=np.int32), np.array(pt, dtype=np.int32)
    out, cells, ms = gpu.align(aa, off, pq, pt)
    n = len(pq)
    res = [np.zeros(n, dtype=np.int32) for _ in range(4)]
    mat = np.ascontiguousarray(MAT, dtype=np.int8)
    oracle_lib.orfx_oracle_align_batch.restype = C.c_int
    oracle_lib.orfx_oracle_align_batch(C.c_void_p(aa.ctypes.data), C.c_void_p(off.ctypes.data), C.c_void_p(pq.ctypes.data),
                                       C.c_void_p(pt.ctypes.data), C.c_int64(n), C.c_void_p(mat.ctypes.data), 24, 12, 2, 0,
                                       *[C.c_void_p(r.ctypes.data) for r in res])
    for i, name in enumerate(("max", "score", "mqe", "mte")):
        bad = np.nonzero(out[name] != res[i])[0]
        assert len(bad) == 0, "%s differs for %d pairs, first: pair %d (qlen %d, tlen %d)" % (
            name, len(bad), bad[0], len(seqs[pq[bad[0]]]), len(seqs[pt[bad[0]]]))
    out2, _, _ = gpu.align(aa, off, pq, pt)                                  # and again: no state left behind
    assert all((out[k] == out2[k]).all() for k in out)


def test_alignment_padded_columns_never_reach_the_results(gpu, oracle_lib):
    """Query lengths around every stripe boundary (64 / 128 / 256 columns and their multiples) at the largest gap costs the
    kernel accepts (gapo + gape = 64): the virtual columns right of the query's end score -128 and must stay below
    everything that max / mte / mqe / score look at."""
    rng = np.random.default_rng(64)
    lengths = [b + d for b in (64, 128, 192, 256, 320, 384, 512, 768, 1024) for d in (-2, -1, 0, 1, 2)] + [1, 2, 3, 31, 32, 33]
    seqs, pq, pt = [], [], []
    for L in lengths:
        a = rng.choice(24, size=int(L)).astype(np.uint8)
        for tl in (1, 7, 33, 200):
            pq.append(len(seqs)); seqs.append(a)
            pt.append(len(seqs)); seqs.append(np.resize(mutate(rng, a, 0.3, 2), tl).astype(np.uint8))
    aa = np.concatenate(seqs)
    off = np.concatenate([[0], np.cumsum([len(x) for x in seqs])]).astype(np.int64)
    for gapo, gape in ((60, 4), (32, 32), (12, 2), (0, 64)):
        out, _, _ = gpu.align(aa, off, pq, pt, gapo, gape)
        for i, (q, t) in enumerate(zip(pq, pt)):
            want = oracle_align(oracle_lib, seqs[q], seqs[t], gapo, gape)
            assert (int(out["max"][i]), int(out["score"][i]), int(out["mqe"][i]), int(out["mte"][i])) == want, (gapo, gape, len(seqs[q]), len(seqs[t]))
    with pytest.raises(ob.OrfxError):
        gpu.align(aa, off, pq[:2], pt[:2], 60, 5)


def test_alignment_true_titin_class_pair(gpu, oracle_lib):
    """BASELINE.json configs[4]'s upper end: one 33 k x 33 k pair (1.1 G cells, ~130 stripes of the block-cooperative path)
    and one 35 k x 30 k pair against the scalar restatement."""
    rng = np.random.default_rng(35000)
    a = rng.choice(20, size=33000, p=FREQ).astype(np.uint8)
    b = mutate(rng, a, 0.25, 3).astype(np.uint8)
    c = rng.choice(20, size=35000, p=FREQ).astype(np.uint8)
    d = mutate(rng, c[:30000], 0.1, 3).astype(np.uint8)
    seqs = [a, b, c, d]
    aa = np.concatenate(seqs)
    off = np.concatenate([[0], np.cumsum([len(x) for x in seqs])]).astype(np.int64)
    out, cells, ms = gpu.align(aa, off, [0, 2], [1, 3])
    for i, (q, t) in enumerate(((0, 1), (2, 3))):
        want = oracle_align(oracle_lib, seqs[q], seqs[t])
        assert (int(out["max"][i]), int(out["score"][i]), int(out["mqe"][i]), int(out["mte"][i])) == want
        assert want[0] > 100000


@pytest.mark.parametrize("seed,kw,overhang", [(91, dict(loci=60, lower=0.4, iupac=0.002), 0), (92, dict(loci=50, titin=0.1), 0),
                                              (93, dict(loci=60, queries=8, template_files=2, dot=0.3, edge=True, chroms=6, chrom_len=800000), 90)])
def test_in_path_alignment_matches_oracle(synth, oracle_lib, seed, kw, overhang):
    """params.align_enabled: every scored row's ORF protein against its template's protein (Finder::align, src/aln.hpp:27-41,
    build definition) -- translation on the device (translate.cu), pair list from the rows, the DP kernel on the library's
    stream -- against the oracle, which tests/test_align_oracle.py pins to an independent restatement.  Both strands, IUPAC
    and soft-masked bases, titin-class loci, --overhang flanks, strand '.'."""
    from test_gpu_parity import assert_same_results
    d = synth(seed, **kw)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=overhang)
    layouts = ["islands_t"] + (["windows"] if not kw.get("dot") else [])
    for layout in layouts:
        batch = w.pack(layout=layout)
        for pk in (dict(mode="LONGEST_MATCH,BEST"), dict(mode="ALL", non_aug=True, gapo=10, gape=1), dict(mode="BEST", lpi=50, ilpi=60, mlpi=70, minlen=300)):
            params = ob.Params.make(align=True, overhang=overhang, **pk)
            with ob.Context(params, device=0) as cx:
                got = cx.run(batch)
                again = cx.run(batch)
            with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
                want = ox.run(batch)
            assert "aln_max" in want.rows and (want.rows["aln_max"] > 0).sum() > 100
            assert_same_results(got, want, "in-path alignment, seed %d %s %s" % (seed, layout, pk))
            assert_same_results(again, want, "second batch on the same context")
            scored = want.rows["qlen"] > 0
            assert not want.rows["aln_score"][~scored].any()
    # the 4-bit base codes cannot feed the translation: refused, not approximated
    with ob.Context(ob.Params.make(align=True, overhang=overhang), device=0) as cx:
        with pytest.raises(ob.OrfxError):
            cx.run(w.pack(layout="islands4"))
    w.close()


def test_cli_pi_flag_adds_alignment_columns_and_leaves_the_gtf_alone(synth):
    d = synth(94, loci=80, template_files=2)
    base = ["--reference", "genome.fa", "--query", "query.gtf", "--mode", "LONGEST_MATCH,BEST"]
    rc0, gtf0, tsv0, err0 = run_cli(ob.cli_path(), d, base, "nopi")
    rc1, gtf1, tsv1, err1 = run_cli(ob.cli_path(), d, base + ["--pi", "1"], "pi")
    rc2, gtf2, tsv2, err2 = run_cli(ORACLE_CLI, d, base + ["--pi", "1"], "pi_oracle")
    assert rc0 == 0 and rc1 == 0 and rc2 == 0, (err0, err1, err2)
    assert gtf1 == gtf0 and gtf2 == gtf0
    assert tsv1 == tsv2
    l0, l1 = tsv0.decode().splitlines(), tsv1.decode().splitlines()
    assert l1[0] == l0[0] + "\taln_max\taln_score\taln_mqe\taln_mte" and len(l0) == len(l1)
    n = 0
    for x, y in zip(l0[1:], l1[1:]):
        if len(y.split("\t")) == 25:
            assert y.startswith(x + "\t")
            n += 1
        else:
            assert x == y     # the 20-column "-" rows of queries without a candidate
    assert n > 300
