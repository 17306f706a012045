"""GPU suite: the CUDA path, called through the C-ABI, against the CPU oracle — bit-exact.

Everything here needs a B200 (`-m gpu`); the oracle and the reference binary are only the checkers.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import orfanage_b200 as ob
from conftest import GOLDEN, REF_BIN, ROOT, run_cli, template_files
from fmt import stats_lines
from golden.make_golden import SYNTH_ARGS, SYNTH_RUNS, SYNTH_SEEDS
from test_oracle import OPTION_GRID, SELF_GRID, compare_clis

pytestmark = pytest.mark.gpu

PARAM_GRID = [
    dict(mode="LONGEST_MATCH,BEST"),
    dict(),   # the default cascade
    dict(mode="ALL"),
    dict(mode="START_MATCH,FIRST"),
    dict(mode="LONGEST", non_aug=True),
    dict(mode="BEST", lpi=50, ilpi=60, mlpi=70, minlen=300),
    dict(mode="LONGEST_MATCH,BEST", has_reference=False),
    dict(mode="FIRST,LONGEST", want_rows=False),
]


def assert_same_results(got, want, what=""):
    assert got.n_queries == want.n_queries and got.n_rows == want.n_rows, what
    for f in want.best.dtype.names:
        assert (got.best[f].view(np.uint32 if want.best[f].dtype == np.float32 else want.best[f].dtype)
                == want.best[f].view(np.uint32 if want.best[f].dtype == np.float32 else want.best[f].dtype)).all(), (what, "best." + f)
    assert set(got.rows) == set(want.rows), what
    for k in want.rows:
        assert got.rows[k].tobytes() == want.rows[k].tobytes(), (what, "row field " + k)


@pytest.fixture(scope="module")
def gpu():
    cx = ob.Context(ob.Params.make(), device=0)
    yield cx
    cx.close()


@pytest.mark.parametrize("seed,kw", [(21, dict(loci=80)), (22, dict(loci=60, titin=0.15, iupac=0.002)),
                                     (23, dict(loci=120, queries=14, templates=5, template_files=3))])
def test_results_match_oracle_bit_for_bit(synth, oracle_lib, gpu, seed, kw):
    d = synth(seed, **kw)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    batch = w.pack()
    windows = w.pack(1, 2, layout="windows") if seed == 21 else None
    assert batch.n_queries > 200
    if windows is not None:   # the whole-window layout of the batch gives the same answers as the islands layout
        params = ob.Params.make(mode="BEST")
        gpu.set_params(params)
        got = gpu.run(windows)
        with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
            assert_same_results(got, ox.run(windows), "windows layout")
            assert_same_results(got, ox.run(w.pack(1, 2, layout="islands")), "windows vs islands")
            nt4 = w.pack(1, 2, layout="islands4")     # 4-bit base codes: same answers from the kernel and from the oracle
            assert nt4.seq_format == 1 and nt4.arrays["seq"].size * 2 >= nt4.struct.seq_bytes > nt4.arrays["seq"].size
            assert_same_results(got, ox.run(nt4), "windows vs islands4 (oracle)")
            assert_same_results(got, gpu.run(nt4), "windows vs islands4 (kernel)")
            nt2 = w.pack(1, 2, layout="islands2")     # 2-bit base codes + exception list
            assert nt2.seq_format == 2 and nt2.arrays["seq"].size == (nt2.struct.seq_bytes + 3) // 4
            assert_same_results(got, ox.run(nt2), "windows vs islands2 (oracle)")
            assert_same_results(got, gpu.run(nt2), "windows vs islands2 (kernel)")
            dct = w.pack(1, 2, layout="islands2d")    # + segment dictionary: the chains are expanded on the device
            assert dct.arrays["seg_ref"].size == dct.arrays["segs"].shape[0]
            assert_same_results(got, gpu.run(dct), "windows vs islands2d (kernel)")
            assert gpu.last_upload_bytes() == dct.h2d_bytes < nt2.h2d_bytes
            bare = dct.without_segment_arrays()
            assert_same_results(got, gpu.run(bare), "windows vs islands2d, dictionary only (kernel)")
            assert_same_results(got, ox.run(bare), "windows vs islands2d, dictionary only (oracle)")
    for pk in PARAM_GRID:
        params = ob.Params.make(**pk)
        gpu.set_params(params)
        got = gpu.run(batch)
        with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
            want = ox.run(batch)
        assert_same_results(got, want, "seed %d params %s" % (seed, pk))
        # the device-resident variant gives the same answers as submit/wait
        gpu.upload(batch)
        assert gpu.run_resident() > 0
        assert_same_results(gpu.fetch(), want, "resident, seed %d params %s" % (seed, pk))
    w.close()


@pytest.mark.parametrize("layout", ["islands", "islands4", "islands2", "islands2d"])
def test_codon_scan_masks_match_oracle(synth, oracle_lib, gpu, layout):
    """Per-stage check (SURVEY.md §4 iii): stop / ATG masks of every query transcript, soft-masked and
    IUPAC bases included, both strands, against SEQ::translate semantics restated in the oracle."""
    import ctypes as C
    d = synth(24, loci=60, lower=0.5, iupac=0.01, titin=0.1)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    batch = w.pack(layout=layout)
    gpu.set_params(ob.Params.make(mode="BEST"))
    gpu.upload(batch)
    gpu.run_resident()
    oracle_lib.orfx_oracle_masks.restype = C.c_int64
    oracle_lib.orfx_oracle_masks.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64]
    a = batch.arrays
    seg_len = a["segs"][:, 1] - a["segs"][:, 0] + 1
    n_checked = strands = 0
    seen = set()
    for tx in np.nonzero(a["tx_kind"] == 0)[0]:
        L = int(seg_len[a["tx_seg_off"][tx]:a["tx_seg_off"][tx + 1]].sum())
        gl, gstop, gstart = gpu.masks(int(tx), L)
        words = (L + 31) // 32
        ostop = np.zeros(words, dtype=np.uint32)
        ostart = np.zeros(words, dtype=np.uint32)
        ol = oracle_lib.orfx_oracle_masks(C.byref(batch.struct), int(tx), ostop.ctypes.data, ostart.ctypes.data, words)
        assert gl == ol == L
        assert (gstop == ostop).all() and (gstart == ostart).all(), "masks differ for transcript %d" % tx
        assert ostop.any()
        n_checked += 1
        seen.add(int(a["tx_strand"][tx]))
    assert n_checked > 200 and seen == {ord("+"), ord("-")}
    w.close()


@pytest.mark.parametrize("seed,overhang,spliced,kw", [(51, 30, False, dict(loci=150)), (52, 300, False, dict(loci=150, titin=0.1, iupac=0.002)),
                                                      (53, 90, True, dict(loci=100, lower=0.4)), (54, 2000, True, dict(loci=60, queries=12))])
def test_overhang_matches_oracle(synth, oracle_lib, gpu, seed, overhang, spliced, kw):
    """--overhang / --spliced_overhang (TX::extend_seq and the exon growth of extend_to_stop / extend_to_start,
    Transcriptome.cpp:342-492): rows, winners and the growth of every row's query copy, plus
    the masks over the flanked chains."""
    import ctypes as C
    d = synth(seed, **kw)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=overhang, spliced_overhang=spliced)
    grown = 0
    for layout in ("islands4", "islands2", "islands2d", "windows"):
        batch = w.pack(layout=layout)
        a = batch.arrays
        assert a["tx_ext_lo"].size == a["tx_kind"].size and a["tx_ext_lo"].max() <= overhang and a["tx_ext_hi"].max() > 0
        for pk in (dict(mode="LONGEST_MATCH,BEST"), dict(mode="ALL", non_aug=True), dict(mode="BEST", lpi=50, ilpi=60, mlpi=70, minlen=300)):
            params = ob.Params.make(overhang=overhang, spliced_overhang=spliced, **pk)
            gpu.set_params(params)
            got = gpu.run(batch)
            with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
                want = ox.run(batch)
            assert "ext_lo" in want.rows
            assert_same_results(got, want, "overhang %d seed %d %s %s" % (overhang, seed, layout, pk))
            grown += int((want.best["ext_lo"] > 0).sum() + (want.best["ext_hi"] > 0).sum())
    assert grown > 0   # some winners end in the flank
    # masks over [5' flank][transcript][3' flank]
    gpu.upload(batch)
    gpu.run_resident()
    oracle_lib.orfx_oracle_masks.restype = C.c_int64
    oracle_lib.orfx_oracle_masks.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int64]
    seg_len = a["segs"][:, 1] - a["segs"][:, 0] + 1
    for tx in np.nonzero(a["tx_kind"] == 0)[0][:150]:
        L = int(seg_len[a["tx_seg_off"][tx]:a["tx_seg_off"][tx + 1]].sum()) + int(a["tx_ext_lo"][tx]) + int(a["tx_ext_hi"][tx])
        gl, gstop, gstart = gpu.masks(int(tx), L)
        words = (L + 31) // 32
        ostop = np.zeros(words, dtype=np.uint32)
        ostart = np.zeros(words, dtype=np.uint32)
        ol = oracle_lib.orfx_oracle_masks(C.byref(batch.struct), int(tx), ostop.ctypes.data, ostart.ctypes.data, words)
        assert gl == ol == L
        assert (gstop == ostop).all() and (gstart == ostart).all(), "masks differ for transcript %d" % tx
    gpu.set_params(ob.Params.make())
    w.close()


def test_overhang_worked_examples_on_the_gpu(tmp_path):
    """The hand-derived --overhang cases of tests/conftest.py through the CUDA CLI."""
    from conftest import check_overhang_kat, write_overhang_kat
    # (the reverse-strand mirror of these cases runs in the CPU suite against the reference binary and the oracle CLI;
    # reverse-strand --overhang loci reach the CUDA CLI through tests/sweep_parity.sh)
    d = str(tmp_path)
    check_overhang_kat(ob.cli_path(), d, write_overhang_kat(d))


@pytest.mark.parametrize("seed,overhang,spliced", [(71, 0, False), (72, 60, False), (73, 200, True)])
def test_strand_dot_and_chromosome_edges_match_oracle(synth, oracle_lib, gpu, seed, overhang, spliced):
    """Strand '.' loci (both flanks from the low side, 5' CDS extensions landing on the chain's 3' end -- Transcriptome.cpp:348-364,
    487, 566 as written) and transcripts that touch an end of their chromosome (ORFX_EXT_*_VIRTUAL): rows, winners, exon growth
    and template cleaning against the oracle, which test_oracle.py pins to the reference binary on the same loci."""
    from test_oracle import dot_edge_dir
    d = dot_edge_dir(synth, seed)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=overhang, spliced_overhang=spliced)
    batch = w.pack(layout="islands4")
    a = batch.arrays
    dots = (a["tx_kind"] == 0) & (a["tx_strand"] == ord("."))
    assert dots.sum() > 100
    if overhang:
        assert (a["tx_ext_hi"][dots] > 0).any() and a["tx_ext_flags"].any()
    for pk in (dict(mode="LONGEST_MATCH,BEST"), dict(mode="ALL", non_aug=True), dict(mode="BEST", lpi=50, ilpi=60, mlpi=70, minlen=300)):
        params = ob.Params.make(overhang=overhang, spliced_overhang=spliced, **pk)
        gpu.set_params(params)
        got = gpu.run(batch)
        with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
            want = ox.run(batch)
        assert_same_results(got, want, "strand '.', seed %d overhang %d %s" % (seed, overhang, pk))
    dot_rows = np.isin(want.rows["q_tx"], np.nonzero(dots)[0])
    assert (want.rows["pass"][dot_rows] == 1).any() and (want.rows["status"][dot_rows] == 5).any()   # some pass, some end without their stop
    cb = cleaning_batch(batch, got.best, np.random.default_rng(seed))
    for non_aug in (False, True):
        p = ob.Params.make(mode="BEST", non_aug=non_aug, overhang=overhang, spliced_overhang=spliced)
        gpu.set_params(p)
        got_c = gpu.clean(cb, True)
        with ob.Context(p, lib=oracle_lib, prefix="orfx_oracle_") as ox:
            want_c = ox.clean(cb, True)
        for g, x, name in zip(got_c, want_c, ("keep", "cds_start", "cds_end")):
            assert (g == x).all(), (seed, non_aug, name, int((g != x).sum()))
    if overhang:   # the whole-window layout cannot hold a strand '.' transcript's 5' flank: refused, not approximated
        with pytest.raises(ob.OrfxError):
            w.pack(layout="windows")
    gpu.set_params(ob.Params.make())
    w.close()


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/orfanage_ref not shipped")
def test_cli_byte_identical_to_reference_on_strand_dot_and_edges(synth, tmp_path):
    from conftest import check_edge_kat, write_edge_kat
    from test_oracle import DOT_EDGE_GRID, dot_edge_dir
    d = dot_edge_dir(synth, 71)
    for args in DOT_EDGE_GRID:
        compare_clis(ob.cli_path(), d, args)
    k = str(tmp_path)
    expected = write_edge_kat(k)
    assert check_edge_kat(ob.cli_path(), k, expected) == check_edge_kat(REF_BIN, k, expected)


def test_heavy_bundle_global_memory_branches_match_oracle(synth, oracle_lib):
    """The heavy-bundle kernels stage a titin-class bundle's chains and per-pair row counts in shared memory while they fit
    (9 216 segments / 8 192 pairs) and walk global memory above that.  orfx_debug_set lowers the two capacities so that
    ordinary titin-class test loci take the global-memory branches (pipeline.cu: hb_load, heavy_fill_kernel); results
    must not change."""
    d = synth(22, loci=60, titin=0.15, iupac=0.002)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    batch = w.pack(layout="islands4")
    a = batch.arrays
    per_tx = np.diff(a["tx_seg_off"])
    assert (per_tx > 32).sum() > 20     # heavy bundles are present
    for pk in (dict(mode="LONGEST_MATCH,BEST"), dict(mode="ALL"), dict(mode="BEST", has_reference=False)):
        params = ob.Params.make(**pk)
        with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
            want = ox.run(batch)
        for segs, pairs in ((9216, 8192), (64, 8192), (9216, 8), (0, 0)):
            with ob.Context(params, device=0) as cx:
                cx.debug_set("hb_max_segs", segs)
                cx.debug_set("hb_max_pairs", pairs)
                assert_same_results(cx.run(batch), want, "hb_max_segs %d hb_max_pairs %d %s" % (segs, pairs, pk))
    w.close()


def test_malformed_batches_are_refused(gpu):
    """include/orfx.h is a public seam: a batch that breaks what the kernels rely on is answered with ORFX_E_INVALID,
    not walked (validate_batch_kernel)."""
    seq = np.frombuffer(b"ACGTACGTATGAAATAGACGT" * 5, dtype=np.uint8)
    good = dict(seq=seq, seq_off=[0, len(seq)], bundle_start=[100], bundle_tx_off=[0, 2], tx_kind=[1, 0], tx_strand=[ord("+")] * 2,
                tx_seg_off=[0, 1, 3], tx_cds_start=[150, 0], tx_cds_end=[170, 0], tx_tid_class=[0, 1],
                segs=[[150, 170], [101, 120], [140, 180]], seg_phase=[0, 0, 0])
    gpu.set_params(ob.Params.make(mode="BEST"))
    assert gpu.run(ob.Batch(**good)).n_queries == 1
    bad = [dict(tx_seg_off=[0, 2, 1]),                       # offsets not monotone
           dict(tx_seg_off=[0, 1, 1]),                       # a query without a segment
           dict(segs=[[150, 170], [120, 101], [140, 180]]),  # end < start
           dict(segs=[[150, 170], [140, 180], [101, 120]]),  # chain not ascending
           dict(tx_kind=[1, 7]),                             # unknown kind
           dict(bundle_tx_off=[0, 3]),                       # more transcripts than the batch has
           dict(segs=[[150, 170], [101, 120], [140, 280]]),  # exon outside the window's bases
           dict(seq_off=[0, len(seq) + 50])]                 # more bases than were given
    for patch in bad:
        with pytest.raises(ob.OrfxError) as e:
            gpu.run(ob.Batch(**dict(good, **patch)))
        assert e.value.code == -1, patch
    # a segment dictionary whose references or offsets do not fit
    dct = dict(good, seg_ref=[0, 0, 1], bundle_dict_off=[0, 3], dict_segs=[[101, 120], [140, 180], [150, 170]], dict_phase=[0, 0, 0])
    dct["seg_ref"] = [2, 0, 1]
    only = {k: v for k, v in dct.items() if k not in ("segs", "seg_phase")}
    assert gpu.run(ob.Batch(**only)).n_queries == 1
    for patch in (dict(seg_ref=[2, 0, 3]), dict(bundle_dict_off=[0, 4]), dict(bundle_dict_off=[2, 1]), dict(seg_ref=[2, 1, 0])):
        with pytest.raises(ob.OrfxError) as e:
            gpu.run(ob.Batch(**dict(only, **patch)))
        assert e.value.code == -1, patch
    assert gpu.run(ob.Batch(**good)).n_queries == 1         # the context is still usable


def test_config1_golden_batch(gpu):
    """Config 1 (example/query.gtf x example/template.gtf, random chr19, --mode BEST): committed packed
    batch -> CUDA path -> the reference's own stats rows."""
    batch = ob.Batch.load(os.path.join(GOLDEN, "c1_batch.npz"))
    names = json.load(open(os.path.join(GOLDEN, "c1_names.json")))["tx_id"]
    gpu.set_params(ob.Params.make(mode="BEST"))
    res = gpu.run(batch)
    want = open(os.path.join(GOLDEN, "c1_ref.tsv")).read().splitlines()[1:]
    assert stats_lines(res, names) == want


@pytest.mark.parametrize("seed", SYNTH_SEEDS)
def test_cli_matches_goldens(tmp_path, seed):
    d = str(tmp_path)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", str(seed)] + SYNTH_ARGS, check=True,
                   capture_output=True)
    for tag, args in SYNTH_RUNS:
        rc, gtf, tsv, err = run_cli(ob.cli_path(), d, args, "g_" + tag, ["template1.gtf", "template2.gtf"])
        assert rc == 0, err
        base = os.path.join(GOLDEN, "synth_s%d_%s" % (seed, tag))
        assert gtf == open(base + ".gtf", "rb").read(), tag
        assert tsv == open(base + ".tsv", "rb").read(), tag


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/orfanage_ref not shipped")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_cli_byte_identical_to_reference_binary(synth, seed):
    d = synth(seed, loci=150, template_files=3, queries=8)
    for args in OPTION_GRID:
        compare_clis(ob.cli_path(), d, args)
    for args, tpl in SELF_GRID:
        compare_clis(ob.cli_path(), d, args, tpl)


def test_empty_and_degenerate_batches(gpu, oracle_lib):
    gpu.set_params(ob.Params.make(mode="BEST"))
    r = gpu.run(ob.Batch())
    assert r.n_queries == 0 and r.n_rows == 0
    # one bundle, a query with a 1-base exon and a template that does not overlap it -> "no-overlap"
    seq = np.frombuffer(b"ACGTACGTATGAAATAGACGT" * 5, dtype=np.uint8)
    b = ob.Batch(seq=seq, seq_off=[0, len(seq)], bundle_start=[100], bundle_tx_off=[0, 3], tx_kind=[1, 0, 0],
                 tx_strand=[ord("+")] * 3, tx_seg_off=[0, 1, 2, 4], tx_cds_start=[150, 0, 0], tx_cds_end=[170, 0, 0],
                 tx_tid_class=[0, 1, 2], segs=[[150, 170], [100, 100], [101, 120], [140, 180]], seg_phase=[0, 0, 0, 0])
    got = gpu.run(b)
    with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want = ox.run(b)
    assert_same_results(got, want, "degenerate")
    assert got.n_queries == 2
    # --overhang without the flank arrays is refused, by the kernel library and by the oracle alike
    ov = ob.Params.make(mode="BEST", overhang=10)
    gpu.set_params(ov)
    with pytest.raises(ob.OrfxError):
        gpu.run(b)
    with ob.Context(ov, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        with pytest.raises(ob.OrfxError):
            ox.run(b)
    gpu.set_params(ob.Params.make(mode="BEST"))
    # the other entry points on empty and degenerate input
    keep, cs, ce = gpu.clean(ob.Batch(), rescue=True)
    assert len(keep) == 0
    assert len(gpu.score_pairs(ob.Batch())["q_tx"]) == 0
    pairs = gpu.score_pairs(b)          # 2 queries x 1 template, one of them disjoint from it
    with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want_pairs = ox.score_pairs(b)
    assert len(pairs["q_tx"]) == 2
    for k in ("q_tx", "t_tx", "qlen", "tlen", "ulen", "match", "inframe", "outframe", "extra", "missing"):
        assert (pairs[k] == want_pairs[k]).all(), k
    cb = ob.Batch(seq=seq, seq_off=[0, len(seq)], bundle_start=[100], bundle_tx_off=[0, 2], tx_kind=[0, 0], tx_strand=[ord("+"), ord("-")],
                  tx_seg_off=[0, 1, 2], tx_cds_start=[108, 101], tx_cds_end=[116, 103], tx_tid_class=[0, 1], segs=[[100, 204], [100, 204]],
                  seg_phase=[0, 0], seg_seq_off=[0, 0])
    for rescue in (False, True):       # ATGAAATAG at 108..116 on '+'; a 3-base CDS on '-'
        got_c = gpu.clean(cb, rescue)
        with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as ox:
            want_c = ox.clean(cb, rescue)
        assert all((g == x).all() for g, x in zip(got_c, want_c)), rescue
        assert got_c[0][0] == 1 and (got_c[1][0], got_c[2][0]) == (108, 116)


def cleaning_batch(batch, best, rng):
    """A cleaning batch (orfx_clean_templates) made from the queries of a packed batch: every query keeps its exon
    chain and islands and gets CDS bounds derived from the ORF the path chose for it, perturbed so that all the
    branches of Transcriptome::clean_cds / TX::rescue_cds(t = nullptr) are taken: intact ORFs, 5' ends moved
    downstream / upstream in and out of frame, 3' ends cut short or pushed past the stop.  Templates become SKIP."""
    a = {k: v.copy() for k, v in batch.arrays.items()}
    kind = a["tx_kind"]
    a["tx_kind"] = np.where(kind == 0, 0, 2).astype(np.uint8)
    segs, so = a["segs"], a["tx_seg_off"]
    cs = np.zeros(len(kind), dtype=np.int32)
    ce = np.zeros(len(kind), dtype=np.int32)

    def spliced(tx):
        return np.concatenate([np.arange(s, e + 1) for s, e in segs[so[tx]:so[tx + 1]]])

    for b in best:
        tx = int(b["q_tx"])
        g = spliced(tx)
        plus = a["tx_strand"][tx] == ord("+")
        if b["t_tx"] < 0:
            lo, hi = 0, min(len(g) - 1, 3 * int(rng.integers(1, 40)) - 1)
        else:
            lo, hi = int(np.searchsorted(g, b["cds_start"])), int(np.searchsorted(g, b["cds_end"]))
        t5, t3 = (lo, hi) if plus else (len(g) - 1 - hi, len(g) - 1 - lo)   # transcript coordinates
        v = int(rng.integers(0, 8))
        if v == 1:
            t5 += 3 * int(rng.integers(1, 30))
        elif v == 2:
            t5 -= 3 * int(rng.integers(1, 30))
        elif v == 3:
            t5 += int(rng.integers(1, 3))
        elif v == 4:
            t3 -= 3 * int(rng.integers(1, 30))
        elif v == 5:
            t3 += 3 * int(rng.integers(1, 30))
        elif v == 6:
            t5 += 3 * int(rng.integers(1, 10)); t3 -= 3 * int(rng.integers(1, 10))
        t5 = max(0, min(t5, len(g) - 3))
        t3 = max(t5 + 2, min(t3, len(g) - 1))
        t3 = t5 + ((t3 - t5 + 1) // 3) * 3 - 1    # correct_chain_len: the CDS the templates carry is a multiple of 3
        u5, u3 = (t5, t3) if plus else (len(g) - 1 - t5, len(g) - 1 - t3)
        cs[tx], ce[tx] = min(g[u5], g[u3]), max(g[u5], g[u3])
    a["tx_cds_start"], a["tx_cds_end"] = cs, ce
    if batch.seq_format == 1:   # 4-bit base codes: `seq` holds two bases per byte
        a["seq_format"], a["seq_bases"] = np.array([1], np.int32), np.array([batch.struct.seq_bytes], np.int64)
    return ob.Batch(**a)


@pytest.mark.parametrize("seed,overhang,kw", [(31, 0, dict(loci=80)), (32, 0, dict(loci=60, titin=0.15, iupac=0.003, lower=0.5)),
                                              (33, 150, dict(loci=80, lower=0.3))])
def test_template_cleaning_matches_oracle(synth, oracle_lib, gpu, seed, overhang, kw):
    """--cleant / --rescue kernel (orfx_clean_templates) against the restated Transcriptome::clean_cds body; with
    --overhang the unguided searches of the rescue run into the flanks (TX::extend_seq, cpp:342-399)."""
    d = synth(seed, **kw)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=overhang)
    batch = w.pack()
    params = ob.Params.make(mode="LONGEST_MATCH,BEST", want_rows=False, overhang=overhang)
    gpu.set_params(params)
    best = gpu.run(batch).best
    cb = cleaning_batch(batch, best, np.random.default_rng(seed))
    kept = {}
    for non_aug in (False, True):
        for rescue in (False, True):
            p = ob.Params.make(mode="BEST", non_aug=non_aug, overhang=overhang)
            gpu.set_params(p)
            got = gpu.clean(cb, rescue)
            with ob.Context(p, lib=oracle_lib, prefix="orfx_oracle_") as ox:
                want = ox.clean(cb, rescue)
            for g, x, name in zip(got, want, ("keep", "cds_start", "cds_end")):
                assert (g == x).all(), (seed, non_aug, rescue, name, int((g != x).sum()))
            kept[(non_aug, rescue)] = int(got[0].sum())
    n = cb.n_queries
    assert n > 200
    if overhang:   # some rescued CDSs end or start in a flank
        with ob.Context(ob.Params.make(mode="BEST", overhang=overhang), lib=oracle_lib, prefix="orfx_oracle_") as ox:
            keep, s_, e_ = ox.clean(cb, True)
        a = cb.arrays
        qidx = np.nonzero(a["tx_kind"] == 0)[0]
        first = a["segs"][a["tx_seg_off"][qidx], 0]
        last = a["segs"][a["tx_seg_off"][qidx + 1] - 1, 1]
        assert int(((keep == 1) & ((s_ < first) | (e_ > last))).sum()) > 0
    assert 0 < kept[(False, False)] < kept[(False, True)] < n     # rescue repairs some, not all
    assert kept[(False, False)] <= kept[(True, False)]
    # the path still answers a regular batch afterwards
    gpu.set_params(params)
    assert (gpu.run(batch).best.tobytes() == best.tobytes())
    w.close()


@pytest.mark.parametrize("seed,kw", [(41, dict(loci=80, template_files=2)), (42, dict(loci=50, titin=0.2))])
def test_pair_scoring_matches_oracle(synth, oracle_lib, gpu, seed, kw):
    """orfcompare's kernel (orfx_score_pairs) against the restated TX::score: every (query, template) pair of every
    bundle, no bounding-box test; the query exon chains stand in for query CDS chains (titin-class chains included)."""
    d = synth(seed, **kw)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    batch = w.pack()
    gpu.set_params(ob.Params.make(mode="BEST"))
    got = gpu.score_pairs(batch)
    with ob.Context(ob.Params.make(mode="BEST"), lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want = ox.score_pairs(batch)
    assert len(want["q_tx"]) > 300
    for k in want:
        assert got[k].tobytes() == want[k].tobytes(), (seed, k)
    assert (want["inframe"] < want["match"]).any() and (want["extra"] > 0).any() and (want["missing"] > 0).any()
    w.close()


CMP_BIN = os.path.join(ROOT, "oracle", "_ref", "orfcompare_ref")


@pytest.mark.parametrize("seed", SYNTH_SEEDS)
def test_orfcompare_cli_matches_goldens(tmp_path, seed):
    from golden.make_golden import CMP_RUNS
    d = str(tmp_path)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", str(seed)] + SYNTH_ARGS, check=True,
                   capture_output=True)
    for tag, args in CMP_RUNS:
        p = subprocess.run([os.path.join(os.path.dirname(ob.cli_path()), "orfcompare")] + args + ["--output", "g_" + tag + ".tsv"], cwd=d,
                           capture_output=True)
        assert p.returncode == 0, p.stderr.decode()[-400:]
        assert open(os.path.join(d, "g_" + tag + ".tsv"), "rb").read() == open(os.path.join(GOLDEN, "cmp_s%d_%s.tsv" % (seed, tag)), "rb").read(), tag


def test_orfcompare_docs_rows_on_the_gpu(gpu):
    """docs/content/csvs/orfcompare_stats.csv Examples 1-2 (and the two other PATL1 rows): committed compare batch ->
    orfx_score_pairs -> the reference's own orfcompare rows."""
    from fmt import compare_lines
    batch = ob.Batch.load(os.path.join(GOLDEN, "kat_compare_batch.npz"))
    names = json.load(open(os.path.join(GOLDEN, "kat_compare_names.json")))["tx_id"]
    want = [l.rstrip("\n").split("\t")[:12] for l in open(os.path.join(GOLDEN, "orfcompare_kat.tsv"))][1:]
    assert compare_lines(gpu.score_pairs(batch), names) == want


@pytest.mark.skipif(not os.path.exists(CMP_BIN), reason="oracle/_ref/orfcompare_ref not shipped")
def test_orfcompare_cli_byte_identical_to_reference_binary(synth):
    d = synth(3, loci=150, template_files=3, queries=8)
    exe = os.path.join(os.path.dirname(ob.cli_path()), "orfcompare")
    for args in (["--query", "template2.gtf", "--template", "template1.gtf", "--reference", "genome.fa"],
                 ["--query", "template3.gtf", "--template", "template2.gtf"],
                 ["--query", "template1.gtf", "--template", "template3.gtf", "--use_id", "--reference", "genome.fa"]):
        subprocess.run([CMP_BIN] + args + ["--threads", "1", "--output", "cmp_ref.tsv"], cwd=d, check=True, capture_output=True)
        p = subprocess.run([exe] + args + ["--output", "cmp_gpu.tsv"], cwd=d, capture_output=True)
        assert p.returncode == 0, p.stderr.decode()[-400:]
        assert open(os.path.join(d, "cmp_ref.tsv"), "rb").read() == open(os.path.join(d, "cmp_gpu.tsv"), "rb").read(), args


def test_full_size_properties(oracle_lib, gpu):
    """BASELINE.json configs[1] at full size (20 000 loci, ~236 k query transcripts; the bench's workload, generated on the
    box), checked through size-independent properties: the run is idempotent, the answer does not depend on how the loci
    are sharded (loci are independent: 1 shard == 3 shards stitched back by transcript id), the chosen ORFs are whole
    codons, and the first 300 bundles are bit-identical to the oracle."""
    import bench
    d = bench.generate(dict(bench.C2), "c2")
    w = ob.Workload(os.path.join(d, "query.gtf"), [os.path.join(d, "template.gtf")], os.path.join(d, "genome.fa"))
    params = ob.Params.make(mode=bench.MODE, want_rows=False)
    gpu.set_params(params)
    full = w.pack(0, 1, layout="islands4")
    assert full.n_queries > 200000
    r1 = gpu.run(full).best
    r2 = gpu.run(full).best
    assert r1.tobytes() == r2.tobytes(), "not idempotent"
    found = r1["t_tx"] >= 0
    assert found.mean() > 0.9
    assert ((r1["cds_end"][found] >= r1["cds_start"][found]) & (r1["cds_start"][found] > 0)).all()
    # sharding invariance: per transcript id, the same ORF and scores from 3 locus shards as from one batch
    def by_id(best, shard):
        ids = [w.tx_id(int(t), shard) for t in best["q_tx"]]
        return dict(zip(ids, zip(best["cds_start"].tolist(), best["cds_end"].tolist(), best["n_rows"].tolist(), best["duplicity"].tolist(),
                                 best["ilpi"].view(np.uint32).tolist(), (best["t_tx"] >= 0).tolist())))
    whole = by_id(r1, 0)
    assert len(whole) == len(r1)
    stitched = {}
    for k in range(3):
        part = w.pack(k, 3, layout="islands4")
        stitched.update(by_id(gpu.run(part).best, k))
    assert stitched == whole
    # the first 300 bundles against the oracle
    w.truncate(300)
    head = w.pack(0, 1, layout="islands4")
    got = gpu.run(head).best
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want = ox.run(head).best
    assert got.tobytes() == want.tobytes() and len(got) > 3000
    w.close()


def test_every_kernel_once_on_a_compact_case():
    """tools/gpu/sanitize_case.py without a sanitizer around it (compute-sanitizer is closed on this GPU pool,
    profiles/r02_sanitizer_closed_on_pool.log): every kernel of the library on one small input with heavy bundles, flanks,
    strand '.', the alignment stage and block-cooperative alignment pairs, each result compared with the oracle."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sanitize_case", os.path.join(ROOT, "tools", "gpu", "sanitize_case.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.main()
