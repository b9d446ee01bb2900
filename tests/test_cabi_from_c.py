"""The boundary is a C ABI: include/parsec_b200.h must compile as plain C11 (and C++) and a C program linked against
libparsec_b200.so must be able to drive the host-only entry points.  No GPU involved."""
import os
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
INC = os.path.join(ROOT, "include")
LIBDIR = os.path.join(ROOT, "parsec-1_b200")
GCC = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else shutil.which("gcc")


@pytest.mark.skipif(GCC is None, reason="no C compiler")
def test_header_is_plain_c_and_cxx():
    hdr = os.path.join(INC, "parsec_b200.h")
    for args in (["-std=c11", "-x", "c"], ["-std=c++17", "-x", "c++"]):
        r = subprocess.run([GCC, "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only"] + args + [hdr],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr


@pytest.mark.skipif(GCC is None, reason="no C compiler")
def test_c_program_links_and_runs(tmp_path):
    exe = str(tmp_path / "cabi_host")
    r = subprocess.run([GCC, "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", INC, os.path.join(HERE, "c", "cabi_host.c"),
                        "-o", exe, "-L", LIBDIR, "-lparsec_b200", "-Wl,-rpath," + LIBDIR], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "C ABI ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.skipif(GCC is None, reason="no C compiler")
def test_c_program_runs_a_tiled_dgemm_on_the_gpu(tmp_path):
    exe = str(tmp_path / "cabi_gpu")
    r = subprocess.run([GCC, "-std=gnu11", "-O2", "-Wall", "-Wextra", "-Werror", "-I", INC, os.path.join(HERE, "c", "cabi_gpu.c"),
                        "-o", exe, "-L", LIBDIR, "-lparsec_b200", "-lm", "-Wl,-rpath," + LIBDIR], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    for args in (["256", "64"], ["384", "128"]):
        r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "C ABI gpu ok" in r.stdout, r.stdout + r.stderr
