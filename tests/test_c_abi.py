"""The C ABI from C: tests/c_abi/abi_check.c includes include/zest_gpu.h as C99 (-Wall -Wextra -Werror -pedantic), links
libzest_gpu.so and calls it the way zest's Zig `extern` block would (no Zig toolchain exists here, so a
C caller is the closest stand-in for a non-Python binding).  Without a GPU it checks the host-only entry points and that
zg_init refuses with ZG_ERR_NO_DEVICE; with one it verifies and extracts a small xorb through a context and a group."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    exe = str(tmp_path / "abi_check")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c_abi", "abi_check.c"), "-o", exe, "-L" + os.path.join(ROOT, "zest_b200"),
                           "-lzest_gpu", "-Wl,-rpath," + os.path.join(ROOT, "zest_b200")])
    return exe


def test_c_caller_links_and_runs(tmp_path):
    out = subprocess.run([_build(tmp_path)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip() in ("no-device", "gpu-ok")


@pytest.mark.gpu
def test_c_caller_on_the_gpu(tmp_path):
    out = subprocess.run([_build(tmp_path)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip() == "gpu-ok"
