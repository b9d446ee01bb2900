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


@pytest.mark.skipif(CUOBJDUMP is None, reason="cuobjdump not found")
def test_every_fp64_tile_kernel_is_the_tma_dmma_code():
    """The transposed GEMM variants and the SYRK kernels are instantiations of the same tile code as dgemm_nn_tma: each of
    them must contain the TMA loads, the mbarrier waits and the DMMA instructions, and none of them a local-memory
    spill loop (a handful of spill instructions at most)."""
    import re
    lib = os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so")
    sass = subprocess.run([CUOBJDUMP, "-sass", lib], capture_output=True, text=True, timeout=300).stdout
    functions = {}
    for chunk in sass.split("Function : ")[1:]:
        name, _, body = chunk.partition("\n")
        functions[name.strip()] = body
    wanted = {"dgemm_nn_tma": 1, "dgemm_tma_trans": 3, "dsyrk_tma": 4, "dgemm_nn_tma_grouped": 2}
    for stem, count in wanted.items():
        # nvcc mangles the names as <length><identifier>: anchor on that so that dgemm_nn_tma does not also count the grouped kernel
        names = [n for n in functions if re.search(r"\d+%s(E|I)" % stem, n)]
        assert len(names) == count, (stem, names)
        for n in names:
            body = functions[n]
            assert body.count("DMMA.8x8x4") >= 128, n             # 32 per k-step x 4 k-steps per stage, unrolled
            assert "UTMALDG.2D" in body and "SYNCS.PHASECHK" in body and "USETMAXREG" in body, n
            assert body.count("STL") + body.count("LDL") <= 12, (n, body.count("STL"), body.count("LDL"))
    for stem in ("trtri_diag_blocks", "potf2_trtri"):
        assert any(stem in n for n in functions), stem
