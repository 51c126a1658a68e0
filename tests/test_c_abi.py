"""The drop-in boundary used from plain C: tests/c_abi/ctrfe_c_smoke.c is compiled with gcc
against include/ctrfe.h and libctrfe.so (no torch, no Python in the process) and run."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c_abi", "ctrfe_c_smoke.c")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _compile(out):
    import __graft_entry__ as g
    g.build()
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None or not os.path.exists(os.path.join(CUDA, "include", "cuda_runtime_api.h")):
        pytest.skip("no C compiler / CUDA headers")
    libdir = os.path.join(ROOT, "paddle_infer_b200")
    cmd = [cc, "-std=c11", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(CUDA, "include"),
           SRC, "-o", out, "-L", libdir, "-l:libctrfe.so", "-L", os.path.join(CUDA, "lib64"), "-lcudart",
           f"-Wl,-rpath,{libdir}", f"-Wl,-rpath,{os.path.join(CUDA, 'lib64')}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_header_is_plain_c_and_the_caller_links(tmp_path):
    """include/ctrfe.h compiles as C11 with -Wall -Werror and the program links against the
    library (CPU check: nothing is run)."""
    _compile(str(tmp_path / "ctrfe_c_smoke"))


@pytest.mark.gpu
def test_plain_c_caller_runs(tmp_path):
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    exe = _compile(str(tmp_path / "ctrfe_c_smoke"))
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ctrfe C smoke: ok" in r.stdout
