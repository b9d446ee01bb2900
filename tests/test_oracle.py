"""The oracle against the reference's golden vectors and pinned behaviour (CPU only)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle_lib as O

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


# ---------------------------------------------------------------- zone allocator vs the reference
def _replay_zone(case, malloc, free, in_use):
    for op in case["ops"]:
        if op[0] == "malloc":
            _, size, want_off, want_use = op
            assert malloc(size) == want_off
        else:
            _, off, want_use = op
            free(off)
        assert in_use() == want_use


def test_zone_oracle_matches_reference_golden():
    cases = json.load(open(os.path.join(HERE, "golden", "zone_trace.json")))
    L = O.lib()
    for case in cases:
        z = L.orc_zone_init(case["nseg"], case["unit"])
        _replay_zone(case, lambda s: L.orc_zone_malloc(z, s), lambda o: L.orc_zone_free(z, o),
                     lambda: L.orc_zone_in_use(z))
        L.orc_zone_fini(z)


def test_zone_oracle_matches_reference_live():
    """When oracle/_ref (the reference's zone_malloc.c compiled in place) is present, fuzz both."""
    path = os.path.join(ROOT, "oracle", "_ref", "libparsec_refbits.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    ref = C.CDLL(path)
    ref.zone_malloc_init.restype = C.c_void_p
    ref.zone_malloc_init.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
    ref.zone_malloc.restype = C.c_void_p
    ref.zone_malloc.argtypes = [C.c_void_p, C.c_size_t]
    ref.zone_free.argtypes = [C.c_void_p, C.c_void_p]
    L = O.lib()
    base = 1 << 40
    rng = np.random.default_rng(7)
    for trial in range(20):
        nseg, unit = int(rng.integers(4, 300)), int(rng.choice([256, 4096, 524288]))
        zr, zo = ref.zone_malloc_init(base, nseg, unit), L.orc_zone_init(nseg, unit)
        live = []
        for _ in range(600):
            if live and rng.random() < 0.5:
                off = live.pop(int(rng.integers(len(live))))
                ref.zone_free(zr, base + off)
                assert L.orc_zone_free(zo, off) == 0
            else:
                size = int(rng.integers(1, 9 * unit))
                p = ref.zone_malloc(zr, size)
                want = -1 if not p else p - base
                assert L.orc_zone_malloc(zo, size) == want
                if want >= 0:
                    live.append(want)
        L.orc_zone_fini(zo)


def test_zone_free_of_unallocated_is_reported():
    L = O.lib()
    z = L.orc_zone_init(8, 4096)
    assert L.orc_zone_free(z, 4096) == -1
    off = L.orc_zone_malloc(z, 4096)
    assert off == 0 and L.orc_zone_free(z, off) == 0 and L.orc_zone_free(z, off) == -1
    assert L.orc_zone_malloc(z, 9 * 4096) == -1  # larger than the zone
    L.orc_zone_fini(z)


# ---------------------------------------------------------------- block-cyclic mapping
def test_grid_ranks_match_reference_golden():
    for g in json.load(open(os.path.join(HERE, "golden", "grid.json"))):
        d = O.bc(g["rank"], 4, 4, 64, 64, 0, 0, 64, 64, g["P"], g["Q"], g["kp"], g["kq"], g["ip"], g["jq"])
        assert (d.rrank, d.crank) == (g["rrank"], g["crank"])


@pytest.mark.parametrize("P,Q", [(1, 1), (1, 2), (2, 2), (2, 4)])
def test_block_cyclic_mapping(P, Q):
    N, NB = 2048, 256
    owners = {}
    for rank in range(P * Q):
        d = O.square(N, NB, P, Q, rank)
        assert (d.mt, d.nt, d.lmt, d.bsiz) == (8, 8, 8, NB * NB)
        seen = set()
        for n in range(d.nt):
            for m in range(d.mt):
                r = O.lib().orc_bc_rank_of(C.byref(d), m, n)
                assert r == (m % P) * Q + (n % Q)  # owner of C(m, n)
                assert O.lib().orc_bc_key(C.byref(d), m, n) == n * d.lmt + m
                if r == rank:
                    pos = O.lib().orc_bc_position(C.byref(d), m, n)
                    assert 0 <= pos < d.nb_local_tiles and pos not in seen
                    seen.add(pos)
                    owners[(m, n)] = rank
        assert len(seen) == d.nb_local_tiles == (8 // P) * (8 // Q)
    assert len(owners) == 64


def test_ragged_matrix_descriptor():
    d = O.bc(0, 40, 40, 250, 130, 0, 0, 250, 130)
    assert (d.mt, d.nt, d.lm, d.ln) == (7, 4, 280, 160)  # tile storage pads to whole tiles
    assert d.nb_local_tiles == 28


# ---------------------------------------------------------------- generator
def _py_lcg_tile(m, n, mb, nb, M, seed):
    """Independent numpy restatement of initialize_tile (dtd_test_simple_gemm.c:68-91)."""
    a, c, mask = 6364136223846793005, 1, (1 << 64) - 1
    out = np.empty((mb, nb), order="F")
    for col in range(nb):
        jump = m + (n + col) * M
        x, mul, add, k = seed, a, c, jump
        while k:
            if k & 1:
                x = (mul * x + add) & mask
            add = (add * (mul + 1)) & mask
            mul = (mul * mul) & mask
            k >>= 1
        for row in range(mb):
            out[row, col] = np.float32(0.5) - np.float32(x) * np.float32(5.4210108624275222e-20)
            x = (a * x + c) & mask
    return out


def test_generator_jump_is_n_steps():
    L = O.lib()
    x = 1789
    for n in range(70):
        assert L.orc_rnd64_jump(n, 1789) == x
        x = (6364136223846793005 * x + 1) & ((1 << 64) - 1)


@pytest.mark.parametrize("m,n,seed", [(0, 0, 1789), (3, 5, 1805), (7, 7, 1901)])
def test_generator_tile_bit_exact(m, n, seed):
    mb = nb = 16
    M = 128
    tile = np.empty((mb, nb), order="F")
    O.lib().orc_init_tile(tile.ctypes.data, m, n, mb, nb, M, mb, seed)
    want = _py_lcg_tile(m, n, mb, nb, M, seed)
    assert np.array_equal(tile, want)
    assert np.all(tile > -0.5 - 1e-7) and np.all(tile <= 0.5)
    assert np.array_equal(tile, tile.astype(np.float32).astype(np.float64))  # single-precision values


# ---------------------------------------------------------------- tile / tiled GEMM
def test_dgemm_tile_vs_numpy():
    rng = np.random.default_rng(0)
    for (mb, nb, kb) in [(1, 1, 1), (7, 5, 3), (64, 64, 64), (40, 40, 40), (33, 65, 17)]:
        A = np.asfortranarray(rng.standard_normal((mb, kb)))
        B = np.asfortranarray(rng.standard_normal((kb, nb)))
        Cm = np.asfortranarray(rng.standard_normal((mb, nb)))
        want = 1.5 * A @ B - 0.5 * Cm
        O.dgemm_tile(1.5, A, B, -0.5, Cm)
        assert np.allclose(Cm, want, rtol=0, atol=1e-12 * kb)


@pytest.mark.parametrize("m,n,k,alpha,beta", [(256, 256, 256, 1.0, 1.0), (100, 37, 53, -1.5, 0.5), (8, 4, 1, 1.0, 0.0),
                                              (7, 3, 9, 2.0, 1.0), (129, 130, 131, 1.0, 0.0), (1, 1, 1, 1.0, 1.0),
                                              (16, 8, 0, 1.0, 0.5), (64, 1, 64, 1.0, 1.0), (5, 64, 64, 0.0, 2.0)])
def test_register_blocked_tile_gemm_is_the_definition_bit_for_bit(m, n, k, alpha, beta):
    """orc_dgemm_tile packs 8 rows of A and keeps 8 x 4 sums in registers; every element is still the ascending-k sum with
    a separate multiply and add, so it must reproduce the one-column-at-a-time statement of the definition exactly."""
    rng = np.random.default_rng(m * 131 + n * 17 + k)
    kk = max(k, 1)
    a = np.asfortranarray(rng.uniform(-0.5, 0.5, (m, kk)))
    b = np.asfortranarray(rng.uniform(-0.5, 0.5, (kk, n)))
    c1 = np.asfortranarray(rng.uniform(-0.5, 0.5, (m, n)))
    c2 = c1.copy(order="F")
    L = O.lib()
    L.orc_dgemm_tile(m, n, k, alpha, a.ctypes.data, m, b.ctypes.data, kk, beta, c1.ctypes.data, m)
    L.orc_dgemm_tile_simple(m, n, k, alpha, a.ctypes.data, m, b.ctypes.data, kk, beta, c2.ctypes.data, m)
    assert np.array_equal(c1, c2)


def test_tiled_dgemm_config1_counts_versions_and_values():
    """BASELINE config 1: N=2048, NB=256 -> 512 GEMM + 64 READ_A + 64 READ_B tasks, C versions == KT == 8
    (the values the reference runtime produced in the survey probe)."""
    N, NB = 2048, 256
    d = O.square(N, NB)
    A, B, Cm = O.fill(d, 1789), O.fill(d, 1805), O.fill(d, 1901)
    C0 = O.tiled_to_dense(d, Cm)
    cnt, ver = O.tiled_dgemm(N, N, N, NB, NB, NB, A, B, Cm)
    assert (cnt.gemm, cnt.read_a, cnt.read_b) == (512, 64, 64)
    assert np.all(ver == 8)
    Ad, Bd = O.tiled_to_dense(d, A), O.tiled_to_dense(d, B)
    want = C0 + Ad @ Bd
    assert O.parity_norm(O.tiled_to_dense(d, Cm), want, Ad, Bd, N) < 10


def test_tiled_dgemm_ragged():
    M, N, K, mb = 250, 130, 90, 40
    dA, dB, dC = O.bc(0, mb, mb, M, K, 0, 0, M, K), O.bc(0, mb, mb, K, N, 0, 0, K, N), O.bc(0, mb, mb, M, N, 0, 0, M, N)
    A, B, Cm = O.fill(dA, 1), O.fill(dB, 2), O.fill(dC, 3)
    C0 = O.tiled_to_dense(dC, Cm)
    cnt, ver = O.tiled_dgemm(M, N, K, mb, mb, mb, A, B, Cm)
    assert cnt.gemm == 7 * 4 * 3 and np.all(ver == 3)
    want = C0 + O.tiled_to_dense(dA, A) @ O.tiled_to_dense(dB, B)
    assert np.allclose(O.tiled_to_dense(dC, Cm), want, atol=1e-12)


def test_tiled_dgemm_rejects_empty():
    z = np.zeros(1)
    assert O.lib().orc_tiled_dgemm(0, 4, 4, 2, 2, 2, z.ctypes.data, z.ctypes.data, z.ctypes.data, 1, None, None) == -4


def test_bf16_rounding_and_tile():
    L = O.lib()
    assert L.orc_f32_to_bf16_rn(1.0) == 0x3F80
    assert L.orc_f32_to_bf16_rn(np.float32(1.0 + 2 ** -8)) == 0x3F80      # tie -> even
    assert L.orc_f32_to_bf16_rn(np.float32(1.0 + 3 * 2 ** -8)) == 0x3F82  # tie -> even (up)
    rng = np.random.default_rng(1)
    mb = nb = kb = 32
    Af = rng.uniform(-0.5, 0.5, (mb, kb)).astype(np.float32)
    Bf = rng.uniform(-0.5, 0.5, (kb, nb)).astype(np.float32)
    to_bf16 = np.vectorize(lambda v: L.orc_f32_to_bf16_rn(float(v)), otypes=[np.uint16])
    Ab, Bb = np.asfortranarray(to_bf16(Af)), np.asfortranarray(to_bf16(Bf))
    back = lambda X: (X.astype(np.uint32) << 16).view(np.float32)
    Cm = np.zeros((mb, nb), dtype=np.float32, order="F")
    L.orc_gemm_tile_bf16(mb, nb, kb, 1.0, Ab.ctypes.data, mb, Bb.ctypes.data, kb, 0.0, Cm.ctypes.data, mb)
    want = back(Ab).astype(np.float64) @ back(Bb).astype(np.float64)
    assert np.allclose(Cm, want, rtol=1e-6, atol=1e-6)


# ---------------------------------------------------------------- coherency / version sequences
def test_c_tile_version_sequence_matches_appendix_c():
    """gpu.version goes 0,0,1,..,KT-1 and cpu.version 0,1,..,KT for one C tile."""
    KT = 4
    ev = O.trace(1, 1, KT)
    c = ev[ev["dc"] == 2]
    kinds = list(c["kind"])
    per_task = [O.EV_STAGE_IN, O.EV_POP, O.EV_EPILOG, O.EV_COMPLETE]
    assert kinds == [O.EV_RESERVE, O.EV_STAGE_IN, O.EV_PUSH_DONE, O.EV_POP, O.EV_EPILOG, O.EV_COMPLETE] + per_task * (KT - 1)
    # k = 0: reserve (INVALID, v0) -> stage-in with transfer, gpu.version overwritten by the host's 0
    r, s, p = c[0], c[1], c[2]
    assert (r["gpu_coh"], r["gpu_version"], r["cpu_coh"], r["owner_device"]) == (O.COH_INVALID, 0, O.COH_OWNED, 0)
    assert (s["transfer"], s["gpu_version"], s["cpu_coh"], s["owner_device"]) == (1, 0, O.COH_SHARED, 2)
    assert p["gpu_coh"] == O.COH_OWNED
    done = c[c["kind"] == O.EV_COMPLETE]
    assert list(done["gpu_version"]) == [0, 1, 2, 3]
    assert list(done["cpu_version"]) == [1, 2, 3, 4]
    stage = c[c["kind"] == O.EV_STAGE_IN]
    assert list(stage["transfer"]) == [1, 0, 0, 0]          # one H2D for the whole chain
    assert list(stage["gpu_version"]) == [0, 1, 2, 3]
    pops = c[c["kind"] == O.EV_POP]
    assert list(pops["transfer"]) == [0, 0, 0, 1]           # one D2H, at k == KT-1 (pushout)
    assert np.all(c[c["kind"] == O.EV_EPILOG]["gpu_coh"] == O.COH_SHARED)


def test_read_only_tiles_stay_version_zero_and_move_once():
    ev = O.trace(3, 2, 4)
    for dc, ntiles, readers in ((0, 3 * 4, 2), (1, 4 * 2, 3)):
        e = ev[ev["dc"] == dc]
        assert np.all(e["gpu_version"] == 0) and np.all(e["cpu_version"] == 0)
        assert (e["kind"] == O.EV_RESERVE).sum() == ntiles
        assert e[e["kind"] == O.EV_STAGE_IN]["transfer"].sum() == ntiles      # each tile crosses PCIe once
        assert (e["kind"] == O.EV_STAGE_IN).sum() == ntiles * readers
        assert (e["kind"] == O.EV_POP).sum() == ntiles * readers
        assert e[e["kind"] == O.EV_POP]["transfer"].sum() == 0
    assert (ev["kind"] == O.EV_COMPLETE).sum() == 3 * 2 * 4


def test_generator_matches_reference_initialize_tile_golden():
    """tests/golden/lcg_tiles.json was written by the reference's own initialize_tile() compiled in place."""
    for c in json.load(open(os.path.join(HERE, "golden", "lcg_tiles.json"))):
        tile = np.empty(c["mb"] * c["nb"])
        O.lib().orc_init_tile(tile.ctypes.data, c["m"], c["n"], c["mb"], c["nb"], c["M"], c["ld"], c["seed"])
        want = np.array([float.fromhex(v) for v in c["values"]])
        assert np.array_equal(tile, want), c


def test_oracle_matches_the_reference_runtime_golden():
    """tests/golden/dag_reference_runtime.json was recorded from the reference RUNTIME itself (unmodified libparsec.so,
    the dgemm_NN JDF through parsec-ptgpp, CPU bodies; tests/golden/make_dag_golden.py).  The oracle reproduces its
    per-class task counts, the final version of every tile's host copy, and C bit for bit (SHA-256)."""
    import hashlib
    golden = json.load(open(os.path.join(HERE, "golden", "dag_reference_runtime.json")))
    assert len(golden) >= 5
    for g in golden:
        M, N, K, NB, runs = g["M"], g["N"], g["K"], g["NB"], g["runs"]
        dA = O.bc(0, NB, NB, M, K, 0, 0, M, K)
        dB = O.bc(0, NB, NB, K, N, 0, 0, K, N)
        dC = O.bc(0, NB, NB, M, N, 0, 0, M, N)
        A, B, Cm = O.fill(dA, 1789), O.fill(dB, 1805), O.fill(dC, 1901)
        gemm = read_a = read_b = 0
        versions = np.zeros(dC.mt * dC.nt, dtype=np.int64)
        for _ in range(runs):
            cnt, ver = O.tiled_dgemm(M, N, K, NB, NB, NB, A, B, Cm)
            gemm, read_a, read_b = gemm + cnt.gemm, read_a + cnt.read_a, read_b + cnt.read_b
            versions += ver                                   # a second taskpool starts from the versions of the first
        assert {"read_a": read_a, "read_b": read_b, "gemm": gemm} == g["counts"]
        assert versions.tolist() == g["versions"]["C"]
        assert not any(g["versions"]["A"]) and not any(g["versions"]["B"])
        assert len(g["versions"]["A"]) == dA.mt * dA.nt and len(g["versions"]["B"]) == dB.mt * dB.nt
        assert hashlib.sha256(Cm.tobytes()).hexdigest() == g["c_sha256"]
