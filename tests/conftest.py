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
              iupac=0.0005, titin=0.02):
    key = (seed, loci, chroms, chrom_len, queries, templates, template_files, lower, iupac, titin)
    if key in _SYNTH_CACHE:
        return _SYNTH_CACHE[key]
    d = os.path.join(str(tmp_root), "synth_%d_%d_%d_%d" % (seed, loci, template_files, len(_SYNTH_CACHE)))
    os.makedirs(d, exist_ok=True)
    subprocess.run([SYNTH, "--out", d, "--seed", str(seed), "--loci", str(loci), "--chroms", str(chroms), "--chrom-len",
                    str(chrom_len), "--queries", str(queries), "--templates", str(templates), "--template-files",
                    str(template_files), "--lower-frac", str(lower), "--iupac-rate", str(iupac), "--titin-frac", str(titin)],
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
