"""The oracle is test infrastructure: nothing the product is made of may import, link, load or mention it, and the
product must fail loudly without a CUDA device instead of computing anything on the CPU."""
import os
import re
import subprocess

from conftest import load_hostapi

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "parsec-1_b200")


def test_product_sources_never_reference_the_oracle():
    offenders = []
    for base, _, files in os.walk(PKG):
        if os.path.basename(base) in ("build", "__pycache__"):
            continue
        for f in files:
            if not f.endswith((".cu", ".cuh", ".h", ".py", "Makefile")):
                continue
            text = open(os.path.join(base, f), errors="replace").read()
            if re.search(r"liboracle|pb2_oracle|\borc_[a-z]|oracle_lib|import oracle|oracle/", text):
                offenders.append(os.path.relpath(os.path.join(base, f), ROOT))
    # tools/ are measurement programs, not the library; they must not use the oracle either
    assert offenders == []
    hdr = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "parsec_b200.h")).read(), flags=re.S)   # code only
    assert not re.search(r"liboracle|pb2_oracle|orc_[a-z]", hdr)


def test_library_links_no_oracle_and_no_blas():
    lib = os.path.join(PKG, "libparsec_b200.so")
    out = subprocess.run(["ldd", lib], capture_output=True, text=True).stdout
    assert "oracle" not in out and "cublas" not in out.lower() and "blas" not in out.lower()
    dyn = subprocess.run(["nm", "-D", "--undefined-only", lib], capture_output=True, text=True).stdout
    assert "orc_" not in dyn and "cublas" not in dyn.lower()


def test_no_cpu_fallback_without_a_device():
    import ctypes as C
    H = load_hostapi()
    try:
        import torch
        if torch.cuda.is_available():
            return                      # on a GPU box the device exists; the CPU suite covers the other side
    except Exception:
        pass
    dev = C.c_void_p()
    assert H.lib().pb2_device_create(0, None, C.byref(dev)) != 0 and not dev.value
    try:
        H.Device(0)
        raise AssertionError("a device module came up without a CUDA device")
    except RuntimeError:
        pass
