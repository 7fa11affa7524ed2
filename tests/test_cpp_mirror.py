"""The C++ host mirror (include/textblaster_b200.hpp): builds tests/cpp/test_mirror.cpp against the C-ABI
library and runs it.  Without a GPU the program has to stop with the library's "no CUDA device" error
(there is no CPU path); on a GPU box every reference unit-test case in it has to pass."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "textblaster_b200")


def _build(tmp_path):
    exe = str(tmp_path / "test_mirror")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "cpp", "test_mirror.cpp"), "-o", exe,
           "-L", LIBDIR, "-ltextblaster_b200", "-Wl,-rpath," + LIBDIR]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def _have_lib():
    return os.path.exists(os.path.join(LIBDIR, "libtextblaster_b200.so"))


@pytest.mark.skipif(not _have_lib(), reason="library not built (python -c 'import __graft_entry__ as g; g.build()')")
def test_mirror_compiles_and_fails_loudly_without_gpu(tmp_path):
    import torch

    exe = _build(tmp_path)
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 3
    assert "no CUDA device available" in r.stderr and "all checks passed" not in r.stdout


@pytest.mark.gpu
def test_mirror_reference_cases_on_gpu(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all checks passed" in r.stdout
