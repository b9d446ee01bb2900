"""The built library really contains the Blackwell paths it claims: FP64 tensor-core DMMA, TMA loads and stores,
tcgen05.mma with and without CTA pairs, TMEM loads, mbarriers (SASS mnemonics per B200_PROFILING.md).  CPU only."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUOBJDUMP = shutil.which("cuobjdump") or ("/usr/local/cuda/bin/cuobjdump" if os.path.exists("/usr/local/cuda/bin/cuobjdump") else None)


@pytest.mark.skipif(CUOBJDUMP is None, reason="cuobjdump not found")
def test_library_sass_has_the_blackwell_instructions():
    lib = os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so")
    sass = subprocess.run([CUOBJDUMP, "-sass", lib], capture_output=True, text=True, timeout=300).stdout
    assert "sm_100a" in sass or "SM100" in sass.upper() or "EF_CUDA_SM100" in sass
    for mnemonic in ("DMMA.8x8x4",            # mma.sync.m8n8k4.f64: the FP64 tile kernels
                     "UTMALDG.2D",            # cp.async.bulk.tensor loads (TMA)
                     "UTMALDG.2D.2CTA",       # ... issued for a CTA pair
                     "UTMASTG.2D",            # TMA stores of C from the epilogue
                     "UTCHMMA",               # tcgen05.mma (bf16 / tf32)
                     "UTCHMMA.2CTA",          # tcgen05.mma.cta_group::2
                     "UTCBAR.2CTA.MULTICAST",  # tcgen05.commit multicast to both CTAs of a pair
                     "LDTM",                  # tcgen05.ld (TMEM -> registers)
                     "SYNCS.ARRIVE.TRANS64",  # mbarrier arrive.expect_tx
                     "USETMAXREG"):           # setmaxnreg in the warp-specialised DGEMM
        assert mnemonic in sass, mnemonic
