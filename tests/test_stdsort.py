"""stdsort.cuh replays libstdc++'s std::sort move for move (the reference's tie behaviour)."""
import os
import subprocess

from conftest import ROOT


def test_stdsort_replays_libstdcxx(tmp_path):
    exe = str(tmp_path / "stdsort_harness")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "stdsort_harness.cpp")], check=True)
    p = subprocess.run([exe], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.startswith("ok "), p.stdout
    assert int(p.stdout.split()[1]) > 5000
