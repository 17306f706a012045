"""GPU test of orfx_align_pairs_static (include/orfx.h): the alignment entry point that allocates nothing -- its scratch is a
file-scope `static __device__` array of the library (align.cu) -- against the existing entry point orfx_align_batch, through the
C-ABI, on two calls in a row.  Integer work: the results must be identical.  The entry point has no counterpart in the reference
(it was added).  Collected last on purpose."""
import numpy as np
import pytest

import orfanage_b200 as ob

pytestmark = pytest.mark.gpu

MAX_PAIRS, MAX_LEN = 65536, 1024      # ORFX_STATIC_MAX_PAIRS / ORFX_STATIC_MAX_LEN


def protein_pairs(seed, lengths):
    """(aa, off, pair_q, pair_t): for every length a random protein and a mutated copy of it (substitutions and one indel)."""
    rng = np.random.default_rng(seed)
    seqs, pq, pt = [], [], []
    for L in lengths:
        a = rng.integers(0, 20, size=int(L)).astype(np.uint8)
        b = a.copy()
        m = rng.random(len(b)) < rng.uniform(0.05, 0.4)
        b[m] = rng.integers(0, 24, size=int(m.sum()))
        p = int(rng.integers(0, len(b)))
        b = np.concatenate([b[:p], b[p + int(rng.integers(0, 9)):], rng.integers(0, 20, size=int(rng.integers(0, 9))).astype(np.uint8)])[:MAX_LEN]
        if len(b) == 0:
            b = a[:1].copy()
        pq.append(len(seqs)); seqs.append(a)
        pt.append(len(seqs)); seqs.append(b)
    off = np.concatenate([[0], np.cumsum([len(s) for s in seqs])]).astype(np.int64)
    return np.concatenate(seqs), off, np.array(pq, dtype=np.int32), np.array(pt, dtype=np.int32)


def on_device(torch, *arrays):
    t = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrays]
    torch.cuda.synchronize()
    return t


def run_static(torch, cx, aa, off, pq, pt, gapo=12, gape=2):
    d_aa, d_off, d_pq, d_pt = on_device(torch, aa, off, pq, pt)
    n = len(pq)
    out = torch.full((4, max(n, 1)), -123456789, dtype=torch.int32, device="cuda")   # a value no alignment produces
    torch.cuda.synchronize()
    cx.align_static(d_aa.data_ptr(), d_off.data_ptr(), len(off) - 1, d_pq.data_ptr(), d_pt.data_ptr(), n,
                    [out[k].data_ptr() for k in range(4)], gapo, gape)
    h = out.cpu().numpy()   # the call returned after its stream drained
    return {"max": h[0, :n], "score": h[1, :n], "mqe": h[2, :n], "mte": h[3, :n]}


def test_static_scratch_alignment_matches_align_batch_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    with ob.Context(ob.Params.make(), device=0) as cx:
        # call 1: many pairs, every regime of the kernel below the static limits (single stripe, multi-stripe up to 1024 columns,
        # 1-residue proteins); call 2: fewer and different pairs, so anything left over from call 1 would show
        rng = np.random.default_rng(5)
        len1 = np.concatenate([rng.integers(1, 300, size=3000), rng.integers(257, MAX_LEN - 8, size=600), [1, 1, 2, 31, 32, 33, 255, 256, 257, 1000, 1016]])
        len2 = np.concatenate([rng.integers(20, 900, size=700), [1016, 1, 512]])
        for call, (seed, lengths, gaps) in enumerate([(11, len1, (12, 2)), (12, len2, (12, 2)), (13, len2[:200], (5, 1))]):
            aa, off, pq, pt = protein_pairs(seed, lengths)
            want, cells, _ = cx.align(aa, off, pq, pt, *gaps)
            assert cells > 0
            got = run_static(torch, cx, aa, off, pq, pt, *gaps)
            for k in ("max", "score", "mqe", "mte"):
                assert (got[k] == want[k]).all(), "call %d: %s differs at %s" % (call + 1, k, np.nonzero(got[k] != want[k])[0][:5])
        # the same call again gives the same answer, and the ordinary entry point still works afterwards
        again = run_static(torch, cx, aa, off, pq, pt, *gaps)
        assert all((again[k] == got[k]).all() for k in got)
        assert all((cx.align(aa, off, pq, pt, *gaps)[0][k] == want[k]).all() for k in want)


def test_static_scratch_limits_are_refused_with_a_clear_error():
    torch = pytest.importorskip("torch")
    with ob.Context(ob.Params.make(), device=0) as cx:
        aa, off, pq, pt = protein_pairs(21, [40, 50, 60])
        good = run_static(torch, cx, aa, off, pq, pt)
        # more pairs than the scratch holds: refused before anything is launched
        many = np.zeros(MAX_PAIRS + 1, dtype=np.int32)
        with pytest.raises(ob.OrfxError) as e:
            run_static(torch, cx, aa, off, many, many + 1)
        assert e.value.code == -4 and "ORFX_STATIC_MAX_PAIRS" in str(e.value)      # ORFX_E_UNSUPPORTED
        # a protein longer than the limit
        long_aa = np.concatenate([aa, np.zeros(MAX_LEN + 1, dtype=np.uint8)])
        long_off = np.concatenate([off, [off[-1] + MAX_LEN + 1]]).astype(np.int64)
        with pytest.raises(ob.OrfxError) as e:
            run_static(torch, cx, long_aa, long_off, np.array([0, 6], np.int32), np.array([1, 1], np.int32))
        assert e.value.code == -4 and "ORFX_STATIC_MAX_LEN" in str(e.value)
        # a pair index outside the sequences
        with pytest.raises(ob.OrfxError) as e:
            run_static(torch, cx, aa, off, np.array([0, 99], np.int32), np.array([1, 1], np.int32))
        assert e.value.code == -1 and "out of range" in str(e.value)                # ORFX_E_INVALID
        # the limits themselves are accepted, and the context is still usable after the errors
        edge_aa, edge_off, edge_pq, edge_pt = protein_pairs(22, [MAX_LEN, MAX_LEN])
        want, _, _ = cx.align(edge_aa, edge_off, edge_pq, edge_pt)
        got = run_static(torch, cx, edge_aa, edge_off, edge_pq, edge_pt)
        assert all((got[k] == want[k]).all() for k in want)
        after = run_static(torch, cx, aa, off, pq, pt)
        assert all((after[k] == good[k]).all() for k in good)
        assert run_static(torch, cx, aa, off, pq[:0], pt[:0])["max"].size == 0
