"""The table the fast passes read instead of iterating libsvm's two-class coupling per row (csrc/hfe_solver.cuh) against the
exact procedure (svm.cpp:2035-2107, 2932-2941, restated in the same header and pinned to scikit-learn through the oracle
tests): built as a host program with nvcc, no GPU needed."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_coupling_table_matches_the_exact_solver(tmp_path):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not (os.path.exists(nvcc) or shutil.which(nvcc)):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "coupling_table_check")
    subprocess.run([nvcc, "-std=c++17", "-O2", "-I", os.path.join(ROOT, "machine-learning-replications_b200", "csrc"),
                    "-o", exe, os.path.join(ROOT, "tests", "csrc", "coupling_table_check.cu")], check=True,
                   capture_output=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    assert re.search(r"edges 6:", out), out                      # six sweep-count edges for eps = 0.005 / 2
    assert "ok 1" in out, out
    fit = float(re.search(r"max fit error ([0-9.e+-]+)", out).group(1))
    rnd = float(re.search(r"random check worst ([0-9.e+-]+)", out).group(1))
    assert fit <= 1e-9 and rnd <= 1e-9, out                      # ~3e-11 measured; the parity bar is 1e-6
