"""CPU suite, part 2: host logic and the C-ABI surface (no compute calls on the CUDA library)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import orfanage_b200 as ob
from conftest import ORACLE_LIB, ROOT, template_files


def declared_functions(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orfxh?_[a-z_0-9]+)\s*\(", src)))


def test_device_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(ob.device_library_path())
    names = declared_functions("orfx.h")
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "liborfx.so does not export %s" % n
    ob.load_device_library()
    assert lib.orfx_abi_version() == 7


def test_host_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(ob.host_library_path())
    names = [n for n in declared_functions("orfx_pack.h") if n.startswith("orfxh_")]
    assert len(names) >= 7
    for n in names:
        assert hasattr(lib, n), "liborfx_host.so does not export %s" % n


def test_no_cpu_fallback_without_a_gpu(synth):
    """The product must fail loudly, not compute on the CPU, when no CUDA device is usable."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(ob.OrfxError) as e:
        ob.Context(ob.Params.make())
    assert e.value.code == -3 and "no CPU path" in str(e.value)
    d = synth(5, loci=40)
    p = subprocess.run([ob.cli_path(), "--reference", "genome.fa", "--query", "query.gtf", "--output", "nogpu.gtf", "template.gtf"],
                       cwd=d, capture_output=True, text=True)
    assert p.returncode == 1 and "no CPU path" in p.stderr


def test_product_does_not_link_or_import_the_oracle():
    for lib in (ob.device_library_path(), ob.host_library_path(), ob.cli_path()):
        out = subprocess.run(["ldd", lib], capture_output=True, text=True).stdout
        assert "oracle" not in out
        syms = subprocess.run(["nm", "-D", lib], capture_output=True, text=True).stdout
        assert "orfx_oracle" not in syms
    for root, _, files in os.walk(os.path.join(ROOT, "orfanage_b200")):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f), errors="replace").read()
                assert "orfx_oracle" not in text and "liborfx_oracle" not in text, f


def test_workload_counts_and_batch_layout(synth):
    d = synth(5, loci=40)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    b = w.pack(layout="windows")
    a = b.arrays
    assert a["seg_seq_off"].size == 0
    assert w.count("active_bundles") == len(a["bundle_start"]) <= w.count("bundles")
    assert w.count("queries") == b.n_queries
    assert w.count("query_nt") == b.query_nt()
    # CSR invariants
    assert a["bundle_tx_off"][0] == 0 and a["bundle_tx_off"][-1] == len(a["tx_kind"])
    assert a["tx_seg_off"][0] == 0 and a["tx_seg_off"][-1] == a["segs"].shape[0]
    assert (np.diff(a["seq_off"]) > 0).all()
    segs = a["segs"]
    assert (segs[:, 0] <= segs[:, 1]).all()
    for x in range(len(a["tx_kind"])):
        s = segs[a["tx_seg_off"][x]:a["tx_seg_off"][x + 1]]
        assert (s[1:, 0] > s[:-1, 1]).all()   # ascending, disjoint
    # every transcript lies inside its bundle window
    for j in range(len(a["bundle_start"])):
        lo = a["bundle_start"][j]
        hi = lo + (a["seq_off"][j + 1] - a["seq_off"][j]) - 1
        for x in range(a["bundle_tx_off"][j], a["bundle_tx_off"][j + 1]):
            s = segs[a["tx_seg_off"][x]:a["tx_seg_off"][x + 1]]
            assert s[0, 0] >= lo and s[-1, 1] <= hi
    # window bytes are the FASTA bytes, case preserved
    fa = "".join(l.strip() for l in open(os.path.join(d, "genome.fa")) if not l.startswith(">"))
    j = 0
    lo = int(a["bundle_start"][j])
    n = int(a["seq_off"][1])
    win = bytes(a["seq"][:n]).decode()
    assert win in fa and any(c.islower() for c in fa)
    # islands layout: same chains, only the bases under query exons, every query segment finds its bytes
    chroms = {}
    name = None
    for l in open(os.path.join(d, "genome.fa")):
        if l.startswith(">"):
            name = l[1:].split()[0]
            chroms[name] = []
        else:
            chroms[name].append(l.strip())
    chroms = {k: "".join(v) for k, v in chroms.items()}
    bi = w.pack(layout="islands")
    ai = bi.arrays
    for k in ("segs", "tx_kind", "tx_seg_off", "bundle_tx_off", "bundle_start", "tx_cds_start", "seg_phase"):
        assert (ai[k] == np.array(a[k])).all(), k
    assert ai["seq"].size < a["seq"].size / 2 and ai["seg_seq_off"].size == ai["segs"].shape[0]
    n_checked = 0
    for j in range(len(ai["bundle_start"])):
        for x in range(ai["bundle_tx_off"][j], ai["bundle_tx_off"][j + 1]):
            if ai["tx_kind"][x] != 0:
                continue
            for k in range(ai["tx_seg_off"][x], ai["tx_seg_off"][x + 1]):
                s0, e0 = ai["segs"][k]
                o = int(ai["seq_off"][j]) + int(ai["seg_seq_off"][k])
                assert o + (e0 - s0 + 1) <= ai["seq_off"][j + 1]
                got = bytes(ai["seq"][o:o + (e0 - s0 + 1)]).decode()
                assert any(c[s0 - 1:e0] == got for c in chroms.values())
                n_checked += 1
    assert n_checked > 1000
    w.close()


def test_nt2_islands_decode_to_the_raw_islands(synth, oracle_lib):
    """ORFX_SEQ_NT2 (include/orfx.h): four bases per byte plus an exception list.  Decoded with numpy, the batch gives back the
    raw islands (case folded, non-IUPAC-pair letters as the 4-bit codes name them), bundles start on whole bytes, exception
    positions ascend, and the oracle answers the same on both batches."""
    d = synth(7, loci=50, lower=0.4, iupac=0.004, titin=0.1)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=45)
    raw = w.pack(1, 2, layout="islands")
    ar = {k: v.copy() for k, v in raw.arrays.items()}
    want_raw = None
    params = ob.Params.make(mode="ALL", overhang=45)
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want_raw = ox.run(raw)
    b2 = w.pack(1, 2, layout="islands2")
    a2 = b2.arrays
    assert b2.seq_format == 2 and a2["seq"].size == (b2.struct.seq_bytes + 3) // 4
    assert (a2["seq_off"] % 4 == 0).all() and a2["seq_exc_pos"].size == a2["seq_exc_code"].size > 0
    assert (np.diff(a2["seq_exc_pos"]) > 0).all()
    for k in ("segs", "tx_kind", "tx_seg_off", "bundle_tx_off", "seg_seq_off", "tx_ext_lo", "tx_ext_hi"):
        assert (a2[k] == ar[k]).all(), k
    codes = (a2["seq"][:, None] >> (np.arange(4, dtype=np.uint8) * 2)[None, :]) & 3
    letters = np.frombuffer(b"ACGT", dtype=np.uint8)[codes.reshape(-1)].copy()
    nt4_letter = np.frombuffer(b"NACUGRNNTNYNNNNN", dtype=np.uint8)
    letters[a2["seq_exc_pos"]] = nt4_letter[a2["seq_exc_code"]]
    known = set(b"ACGTURYN")
    n_checked = n_exc = 0
    for j in range(len(ar["bundle_start"])):
        n = int(ar["seq_off"][j + 1] - ar["seq_off"][j])
        r = bytes(ar["seq"][ar["seq_off"][j]:ar["seq_off"][j + 1]]).upper()
        got = bytes(letters[a2["seq_off"][j]:a2["seq_off"][j] + n])
        assert a2["seq_off"][j + 1] - a2["seq_off"][j] == (n + 3) // 4 * 4
        r = bytes(c if c in known else ord("N") for c in r)
        assert got == r, j
        n_checked += n
        n_exc += sum(c not in b"ACGT" for c in r)
    assert n_checked > 50000 and n_exc == a2["seq_exc_pos"].size
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        got = ox.run(b2)
    assert got.best.tobytes() == want_raw.best.tobytes()
    for k in want_raw.rows:
        assert got.rows[k].tobytes() == want_raw.rows[k].tobytes(), k
    # save / load keeps the format
    path = os.path.join(d, "b2.npz")
    b2.save(path)
    again = ob.Batch.load(path)
    assert again.seq_format == 2 and again.struct.seq_bytes == b2.struct.seq_bytes and again.struct.n_seq_exc == b2.struct.n_seq_exc
    w.close()


def test_segment_dictionary_expands_to_the_segment_arrays(synth, oracle_lib):
    """orfx_batch::seg_ref (include/orfx.h): per bundle the distinct (segment, phase, sequence offset) records and a 16-bit
    reference per segment.  Expanded with numpy it gives back segs / seg_phase / seg_seq_off; the oracle, which takes the
    chains from the dictionary alone, answers the same with and without the per-segment arrays; strand '.' loci with their
    private island copies (--overhang) and chromosome-edge loci included."""
    d = synth(8, loci=60, lower=0.3, iupac=0.002, titin=0.1, dot=0.2, edge=True)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=45)
    plain = w.pack(1, 2, layout="islands2")
    ap = {k: v.copy() for k, v in plain.arrays.items()}
    params = ob.Params.make(mode="ALL", overhang=45)
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        want = ox.run(plain)
    b = w.pack(1, 2, layout="islands2d")
    a = b.arrays
    assert a["seg_ref"].size == a["segs"].shape[0] > 1000 and a["dict_segs"].shape[0] == a["dict_phase"].size == a["dict_seq_off"].size
    assert a["dict_segs"].shape[0] < a["segs"].shape[0] / 2, "loci share exons: the dictionary is much smaller than the chains"
    for k in ("segs", "seg_phase", "seg_seq_off", "tx_seg_off", "bundle_tx_off", "seq", "tx_ext_lo", "tx_ext_hi"):
        assert (a[k] == ap[k]).all(), k
    seg_tx = np.repeat(np.arange(len(a["tx_kind"])), np.diff(a["tx_seg_off"]))
    seg_bundle = np.repeat(np.arange(len(a["bundle_tx_off"]) - 1), np.diff(a["bundle_tx_off"]))[seg_tx]
    n_rec = np.diff(a["bundle_dict_off"])
    assert (n_rec > 0).all() and n_rec.max() <= 65536 and a["bundle_dict_off"][-1] == a["dict_segs"].shape[0]
    assert (a["seg_ref"] < n_rec[seg_bundle]).all()
    at = a["bundle_dict_off"][seg_bundle] + a["seg_ref"].astype(np.int64)
    assert (a["dict_segs"][at] == a["segs"]).all() and (a["dict_phase"][at] == a["seg_phase"]).all() and (a["dict_seq_off"][at] == a["seg_seq_off"]).all()
    # records of a bundle are distinct
    rec = np.stack([np.repeat(np.arange(len(n_rec)), n_rec), a["dict_segs"][:, 0], a["dict_segs"][:, 1], a["dict_phase"], a["dict_seq_off"]], 1)
    assert len(np.unique(rec, axis=0)) == len(rec)
    assert b.h2d_bytes < plain.h2d_bytes * 0.9
    bare = b.without_segment_arrays()
    assert bare.arrays["segs"].size == 0 and bare.struct.n_segs == b.struct.n_segs and not bare.struct.segs
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as ox:
        for what, batch in (("with the arrays", b), ("dictionary only", bare)):
            got = ox.run(batch)
            assert got.best.tobytes() == want.best.tobytes(), what
            for k in want.rows:
                assert got.rows[k].tobytes() == want.rows[k].tobytes(), (what, k)
        # a reference outside its bundle's records is refused
        broken = {k: v.copy() for k, v in bare.arrays.items()}
        broken["seg_ref"][5] = 65535
        with pytest.raises(ob.OrfxError):
            ox.run(ob.Batch(**broken, seq_format=np.array([2], np.int32), seq_bases=np.array([b.struct.seq_bytes], np.int64)))
    w.close()


def test_layout_fuzz_sample():
    """A slice of tools/fuzz_layouts.py: random workloads, parameters and shards; every upload format of the packer gives the
    oracle the same rows and winners."""
    import subprocess
    import sys
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_layouts.py"), "--seeds", "0:12"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "0 failures" in p.stdout


def test_sharding_is_a_partition(synth, oracle_lib):
    """Shards are disjoint, cover every active bundle, and evaluating them separately gives the same
    per-query answers as one batch (loci are independent: SURVEY.md §8e)."""
    d = synth(6, loci=50)
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    params = ob.Params.make(mode="LONGEST_MATCH,BEST")
    with ob.Context(params, lib=oracle_lib, prefix="orfx_oracle_") as cx:
        whole = w.pack(0, 1)
        ids_whole = [w.tx_id(int(q)) for q in np.nonzero(whole.arrays["tx_kind"] == 0)[0]]
        res = cx.run(whole)
        want = {ids_whole[i]: (int(b["cds_start"]), int(b["cds_end"]), int(b["t_tx"] >= 0)) for i, b in enumerate(res.best)}
        got = {}
        starts = []
        for k in range(3):
            part = w.pack(k, 3)
            starts += list(part.arrays["bundle_start"])
            ids = {int(q): w.tx_id(int(q), k) for q in np.nonzero(part.arrays["tx_kind"] == 0)[0]}
            r = cx.run(part)
            for b in r.best:
                key = ids[int(b["q_tx"])]
                assert key not in got
                got[key] = (int(b["cds_start"]), int(b["cds_end"]), int(b["t_tx"] >= 0))
    assert len(starts) == len(whole.arrays["bundle_start"])
    assert got == want and len(got) > 100
    w.close()


WORKER = r'''
import os, sys, ctypes, json
sys.path.insert(0, {root!r})
import numpy as np, torch, torch.distributed as dist
import orfanage_b200 as ob
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
d = {d!r}
w = ob.Workload(d + "/query.gtf", [d + "/template.gtf"], d + "/genome.fa")
part = w.pack(rank, world)
lib = ctypes.CDLL({oracle!r})
with ob.Context(ob.Params.make(mode="LONGEST_MATCH,BEST"), lib=lib, prefix="orfx_oracle_") as cx:
    r = cx.run(part)
found = torch.tensor([int((r.best["t_tx"] >= 0).sum()), r.n_queries, int(r.best["cds_start"].astype(np.int64).sum())])
dist.all_reduce(found)          # results are gathered on the host; no collective on the data path
if rank == 0:
    json.dump(found.tolist(), open({out!r}, "w"))
dist.destroy_process_group()
'''


def test_two_rank_sharding_over_gloo(synth, tmp_path, oracle_lib):
    """The N>1 launch shape of bench.py (one process per shard, torch.distributed rendezvous, results
    reduced on the host) on world_size 2 / gloo."""
    d = synth(6, loci=50)
    out = str(tmp_path / "found.json")
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, d=d, oracle=ORACLE_LIB, out=out))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29531")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29531", str(script)], env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    import json
    found, nq, cs = json.load(open(out))
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"))
    with ob.Context(ob.Params.make(mode="LONGEST_MATCH,BEST"), lib=oracle_lib, prefix="orfx_oracle_") as cx:
        r = cx.run(w.pack())
    assert (found, nq, cs) == (int((r.best["t_tx"] >= 0).sum()), r.n_queries, int(r.best["cds_start"].astype(np.int64).sum()))
    w.close()


def _read_fasta(path):
    seqs, name = {}, None
    for line in open(path):
        if line.startswith(">"):
            name = line[1:].split()[0]
            seqs[name] = []
        else:
            seqs[name].append(line.strip())
    return {k: "".join(v) for k, v in seqs.items()}


@pytest.mark.parametrize("spliced", [False, True])
def test_overhang_flanks_of_the_packer(synth, spliced):
    """orfx_batch::tx_ext_lo / tx_ext_hi against a direct restatement of TX::extend_seq (Transcriptome.cpp:342-399): at most
    --overhang bases per side, clamped to the padded bundle window, cut before the first GT (3' side) / after the last AG (5'
    side) of the strand-corrected, upper-cased flank with --spliced_overhang; and the islands hold exactly those bases."""
    N = 75
    d = synth(61, loci=40, lower=0.4)
    genome = _read_fasta(os.path.join(d, "genome.fa"))
    w = ob.Workload(os.path.join(d, "query.gtf"), template_files(d), os.path.join(d, "genome.fa"), overhang=N, spliced_overhang=spliced)
    b = w.pack(layout="islands")
    a = b.arrays
    comp = str.maketrans("ACGTUacgtu", "TGCAAtgcaa")
    chrom_of = {}
    for line in open(os.path.join(d, "query.gtf")):
        f = line.split("\t")
        if len(f) > 8 and f[2] == "transcript":
            chrom_of[re.search(r'transcript_id "([^"]+)"', f[8]).group(1)] = f[0]
    checked = cut = 0
    for bi in range(len(a["bundle_start"])):
        x0, x1 = a["bundle_tx_off"][bi], a["bundle_tx_off"][bi + 1]
        bstart = int(a["bundle_start"][bi])
        for x in range(x0, x1):
            if a["tx_kind"][x] != 0:
                assert a["tx_ext_lo"][x] == 0 and a["tx_ext_hi"][x] == 0
                continue
            g = genome[chrom_of[w.tx_id(int(x))]]
            s0, s1 = a["tx_seg_off"][x], a["tx_seg_off"][x + 1]
            first, last = int(a["segs"][s0, 0]), int(a["segs"][s1 - 1, 1])
            plus = a["tx_strand"][x] == ord("+")
            for high in (False, True):
                lo, hi = (last + 1, min(last + N, len(g))) if high else (max(first - N, max(bstart, 1)), first - 1)
                flank = g[lo - 1:hi] if hi >= lo else ""
                want = len(flank)
                if spliced and flank:
                    f5 = (flank if plus else flank.translate(comp)[::-1]).upper()
                    if plus == high:                      # towards the stop
                        at = f5.find("GT")
                        want = len(flank) if at < 0 else at
                    else:
                        at = f5.rfind("AG")
                        want = len(flank) if at < 0 else len(flank) - (at + 2)
                    cut += want < len(flank)
                got = int(a["tx_ext_hi"][x] if high else a["tx_ext_lo"][x])
                assert got == want, (w.tx_id(int(x)), high, got, want)
                # the island bytes next to the terminal exon are the flank bases
                if got:
                    off = int(a["seq_off"][bi])
                    if high:
                        p0 = off + int(a["seg_seq_off"][s1 - 1]) + (last - int(a["segs"][s1 - 1, 0]) + 1)
                        assert bytes(a["seq"][p0:p0 + got]).decode() == g[last:last + got]
                    else:
                        p0 = off + int(a["seg_seq_off"][s0]) - got
                        assert bytes(a["seq"][p0:p0 + got]).decode() == g[first - 1 - got:first - 1]
                checked += 1
    assert checked > 400 and (cut > 100 if spliced else cut == 0)
    w.close()


def test_pair_sharding_partitions_and_balances():
    """C5 pairs over N GPUs: every pair lands in exactly one shard and the DP cells per shard are balanced."""
    rng = np.random.default_rng(5)
    lens = np.clip(np.exp(rng.normal(np.log(400.0), 0.9, 4000)), 50, 35000).astype(np.int64)
    lens[:6] = 33000                                      # a few titin-class sequences
    off = np.concatenate([[0], np.cumsum(lens)])
    pq, pt = rng.integers(0, 4000, 3000), rng.integers(0, 4000, 3000)
    pq[:3], pt[:3] = [0, 1, 2], [3, 4, 5]                 # ... paired with each other: the heaviest pairs
    cells = (off[pq + 1] - off[pq]) * (off[pt + 1] - off[pt])
    for n in (1, 2, 8):
        shards = ob.shard_pairs(off, pq, pt, n)
        assert len(shards) == n
        assert sorted(np.concatenate(shards).tolist()) == list(range(3000))
        load = np.array([cells[s].sum() for s in shards], dtype=np.float64)
        assert load.max() <= max(load.mean() * 1.05, load.mean() + cells.max())
    assert ob.shard_pairs(off, pq[:0], pt[:0], 4)[0].size == 0


def test_sliced_gtf_reader_gives_the_single_pass_result(tmp_path):
    """The GTF reader parses big files in concurrent slices; forced on a small file (cuts fall inside transcripts)
    it must reproduce the goldens written by the reference."""
    from conftest import GOLDEN, ORACLE_CLI, run_cli
    from golden.make_golden import SYNTH_ARGS, SYNTH_RUNS
    d = str(tmp_path)
    subprocess.run([os.path.join(ROOT, "tools", "orfx_synth"), "--out", d, "--seed", "101"] + SYNTH_ARGS, check=True, capture_output=True)
    old = os.environ.get("ORFX_GTF_SLICES")
    try:
        for n in ("3", "13"):
            os.environ["ORFX_GTF_SLICES"] = n
            for tag, args in SYNTH_RUNS[:2]:
                rc, gtf, tsv, err = run_cli(ORACLE_CLI, d, args, "s_" + tag, ["template1.gtf", "template2.gtf"])
                assert rc == 0, err
                assert gtf == open(os.path.join(GOLDEN, "synth_s101_%s.gtf" % tag), "rb").read(), (n, tag)
                assert tsv == open(os.path.join(GOLDEN, "synth_s101_%s.tsv" % tag), "rb").read(), (n, tag)
    finally:
        if old is None:
            os.environ.pop("ORFX_GTF_SLICES", None)
        else:
            os.environ["ORFX_GTF_SLICES"] = old
