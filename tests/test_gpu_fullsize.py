"""GPU suite: byte parity with the reference at full size.

The CUDA front-end (orfanage_b200/bin/orfanage) on the seeded full-size workloads of tests/golden/make_fullsize_hashes.py --
BASELINE.json configs[1] itself (20 000 loci, ~236 k evaluated queries), a configs[2]-shaped run (three template files +
--lpi / --ilpi / --mlpi) and a configs[3]-shaped run (~100 query isoforms per locus: titin-class bundles of 10-36 k segments,
above the heavy-bundle kernels' shared-memory stage) -- must write exactly the GTF and stats TSV the reference's own binary
writes (oracle/_ref/orfanage_ref --threads 1).  The reference's output is pinned by SHA-256 digests computed in the build
container (tests/golden/fullsize_hashes.json: the reference needs 1-5 minutes of one core per case); with
ORFX_PARITY_DIRECT=1 the reference also runs on the box and the files are cmp'ed directly."""
import hashlib
import json
import os
import subprocess
import sys
import time

import pytest

import orfanage_b200 as ob
from conftest import GOLDEN, REF_BIN, ROOT

sys.path.insert(0, GOLDEN)
import make_fullsize_hashes as fs   # noqa: E402

pytestmark = pytest.mark.gpu
HASHES = json.load(open(os.path.join(GOLDEN, "fullsize_hashes.json")))


def sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 22), b""):
            h.update(chunk)
    return h.hexdigest()


@pytest.mark.parametrize("case", ["c2", "c3", "c4"])
def test_full_size_output_is_byte_identical_to_the_reference(case):
    want = HASHES[case]
    assert want["generator"] == fs.CASES[case][0] and want["options"] == fs.CASES[case][1], "regenerate tests/golden/fullsize_hashes.json"
    base = os.environ.get("ORFX_BENCH_DIR", "/tmp/orfx_bench")
    d = fs.generate(case, base)
    t = time.time()
    p = subprocess.run([ob.cli_path()] + fs.cli_args(case, "gpu", ["--timing"]), cwd=d, capture_output=True, text=True)
    wall = time.time() - t
    assert p.returncode == 0, p.stderr[-800:]
    got = {"gtf_sha256": sha(os.path.join(d, "gpu.gtf")), "tsv_sha256": sha(os.path.join(d, "gpu.tsv"))}
    record = {"case": case, "cli_wall_s": round(wall, 2), "gtf_identical": got["gtf_sha256"] == want["gtf_sha256"],
              "tsv_identical": got["tsv_sha256"] == want["tsv_sha256"], "gtf_lines": want["gtf_lines"], "tsv_lines": want["tsv_lines"],
              "reference": "sha256 of oracle/_ref/orfanage_ref --threads 1 output (tests/golden/fullsize_hashes.json)",
              "timing": [l for l in p.stderr.splitlines() if l.startswith("[timing]")]}
    if os.environ.get("ORFX_PARITY_DIRECT") == "1" and os.path.exists(REF_BIN):
        t = time.time()
        subprocess.run([REF_BIN] + fs.cli_args(case, "ref", ["--threads", "1"]), cwd=d, check=True, capture_output=True)
        record["reference_wall_s"] = round(time.time() - t, 1)
        record["cmp_gtf"] = subprocess.run(["cmp", "-s", "ref.gtf", "gpu.gtf"], cwd=d).returncode == 0
        record["cmp_tsv"] = subprocess.run(["cmp", "-s", "ref.tsv", "gpu.tsv"], cwd=d).returncode == 0
        record["reference"] = "oracle/_ref/orfanage_ref --threads 1 run on this box, files compared with cmp"
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(record, open(os.path.join(out, "fullsize_parity_%s.json" % case), "w"), indent=1)
    assert record["gtf_identical"] and record["tsv_identical"], record
    assert record.get("cmp_gtf", True) and record.get("cmp_tsv", True), record


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs more than one GPU on the box")
@pytest.mark.parametrize("case", ["c4", "c2"])
def test_multi_gpu_cli_is_byte_identical_to_the_reference(case):
    """The product's own multi-GPU path -- shard_bundles (LPT) over --gpus 0,1,...,N-1, one host thread and one context per
    device, results merged back into bundle order (host/runner.cpp) -- writes the same bytes as the reference."""
    n = _n_gpus()
    want = HASHES[case]
    d = fs.generate(case, os.environ.get("ORFX_BENCH_DIR", "/tmp/orfx_bench"))
    gpus = ",".join(str(k) for k in range(n))
    t = time.time()
    p = subprocess.run([ob.cli_path()] + fs.cli_args(case, "mgpu", ["--timing", "--gpus", gpus]), cwd=d, capture_output=True, text=True)
    wall = time.time() - t
    assert p.returncode == 0, p.stderr[-800:]
    record = {"case": case, "gpus": gpus, "cli_wall_s": round(wall, 2), "gtf_identical": sha(os.path.join(d, "mgpu.gtf")) == want["gtf_sha256"],
              "tsv_identical": sha(os.path.join(d, "mgpu.tsv")) == want["tsv_sha256"], "timing": [l for l in p.stderr.splitlines() if l.startswith("[timing]")]}
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(record, open(os.path.join(out, "fullsize_parity_%s_%dgpu.json" % (case, n)), "w"), indent=1)
    assert record["gtf_identical"] and record["tsv_identical"], record
