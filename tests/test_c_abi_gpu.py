"""The drop-in boundary is a C ABI: a plain C program (tests/c_abi/abi_smoke.c; no Python, no torch)
is compiled against include/b200x.h, linked with libb200x.so and run."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ray-project-ray_b200")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _build(tmp_path):
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no C compiler")
    exe = str(tmp_path / "abi_smoke")
    cmd = [gcc, "-std=c11", "-Wall", "-Werror", "-o", exe, os.path.join(ROOT, "tests", "c_abi", "abi_smoke.c"),
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(CUDA, "include"), "-L" + PKG, "-lb200x",
           "-L" + os.path.join(CUDA, "lib64"), "-lcudart", "-Wl,-rpath," + PKG, "-Wl,-rpath," + os.path.join(CUDA, "lib64")]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return exe


def test_header_compiles_as_c_and_links(tmp_path):
    """include/b200x.h is valid C11 and every symbol the C caller uses resolves against libb200x.so (CPU-only check;
    without a GPU the program reports that the library has no CPU path and exits 4)."""
    exe = _build(tmp_path)
    import torch

    if not torch.cuda.is_available():
        r = subprocess.run([exe, "1000"], capture_output=True, text=True)
        assert r.returncode == 4 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_c_caller_sorts_and_gathers_through_the_abi(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe, "5000000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "abi_smoke ok" in r.stdout
