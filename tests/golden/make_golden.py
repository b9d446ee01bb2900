"""Regenerates the golden vectors from the REFERENCE's own code.

Run in the build container only (needs /root/reference): `make -C oracle ref` compiles the reference's
parsec/utils/zone_malloc.c and parsec/data_dist/matrix/grid_2Dcyclic.c from where they lie into
oracle/_ref/libparsec_refbits.so; this script drives that library and records its outputs:

  zone_trace.json  seeded malloc/free sequences -> offsets returned by the reference allocator
  grid.json        (rank, P, Q, kp, kq, ip, jq) -> (rrank, crank) from parsec_grid_2Dcyclic_init

The oracle restatement (oracle/pb2_oracle.c) and the product's HBM heap are then checked against these
files on any machine.
"""
import ctypes as C
import json
import os
import random

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libparsec_refbits.so"))
ref.zone_malloc_init.restype = C.c_void_p
ref.zone_malloc_init.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
ref.zone_malloc.restype = C.c_void_p
ref.zone_malloc.argtypes = [C.c_void_p, C.c_size_t]
ref.zone_free.argtypes = [C.c_void_p, C.c_void_p]
ref.zone_in_use.restype = C.c_size_t
ref.zone_in_use.argtypes = [C.c_void_p]
BASE = 1 << 40  # fake device base address; the allocator only does arithmetic on it


def zone_case(seed, nseg, unit, nops, sizes):
    rng = random.Random(seed)
    z = ref.zone_malloc_init(BASE, nseg, unit)
    live, ops = [], []
    for _ in range(nops):
        if live and (rng.random() < 0.45 or len(live) > nseg):
            i = rng.randrange(len(live))
            off = live.pop(i)
            ref.zone_free(z, BASE + off)
            ops.append(["free", off, int(ref.zone_in_use(z))])
        else:
            sz = rng.choice(sizes)
            p = ref.zone_malloc(z, sz)
            off = -1 if not p else p - BASE
            if off >= 0:
                live.append(off)
            ops.append(["malloc", sz, off, int(ref.zone_in_use(z))])
    return {"seed": seed, "nseg": nseg, "unit": unit, "ops": ops}


def main():
    KiB = 1024
    cases = [
        zone_case(1, 64, 512 * KiB, 400, [512 * KiB, 8 * 1024 * KiB, 2 * 1024 * KiB]),  # 512K/2M/8M tiles
        zone_case(2, 200, 512 * KiB, 1500, [8 * 1024 * KiB]),                             # NB=1024 fp64 only
        zone_case(3, 97, 4096, 2000, [1, 4096, 4097, 10000, 40000, 65536]),              # ragged sizes
        zone_case(4, 16, 512 * KiB, 300, [32 * 1024 * KiB, 8 * 1024 * KiB]),              # mostly failing
    ]
    json.dump(cases, open(os.path.join(HERE, "zone_trace.json"), "w"))

    class Grid(C.Structure):
        _fields_ = [(n, C.c_int) for n in "rank rows cols ip jq krows kcols crank rrank rloc cloc vp_p vp_q".split()]

    grids = []
    for (P, Q) in [(1, 1), (1, 2), (2, 2), (2, 4), (4, 2), (3, 5)]:
        for (kp, kq, ip, jq) in [(1, 1, 0, 0), (2, 3, 0, 0), (1, 1, 1, 1 % Q)]:
            for rank in range(P * Q):
                g = Grid()
                ref.parsec_grid_2Dcyclic_init(C.byref(g), rank, P, Q, kp, kq, ip % P, jq % Q)
                grids.append({"rank": rank, "P": P, "Q": Q, "kp": kp, "kq": kq, "ip": ip % P, "jq": jq % Q,
                              "rrank": g.rrank, "crank": g.crank})
    json.dump(grids, open(os.path.join(HERE, "grid.json"), "w"))
    print("wrote zone_trace.json (%d cases), grid.json (%d grids)" % (len(cases), len(grids)))


if __name__ == "__main__":
    main()


# ---------------------------------------------------------------------------------------------------------------
# ownership_trace.json: random operation sequences driven through the REFERENCE's parsec/data.c (attach / detach a
# device copy, start / end of an ownership transfer, and the field updates the device module makes around them:
# version bump, version overwrite, epilog, complete hook, reader release); the full data state after every step.
def ownership_cases():
    ref.ref_data_new.restype = C.c_void_p
    ref.ref_data_new.argtypes = [C.c_uint64]
    for name in ("ref_copy_attach_new", "ref_copy_detach", "ref_start_transfer", "ref_data_owner"):
        getattr(ref, name).argtypes = [C.c_void_p] + [C.c_int] * {"ref_copy_attach_new": 1, "ref_copy_detach": 1,
                                                                   "ref_start_transfer": 2, "ref_data_owner": 0}[name]
    ref.ref_end_transfer.argtypes = [C.c_void_p, C.c_int, C.c_int]
    ref.ref_end_transfer.restype = None
    ref.ref_copy_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    ref.ref_copy_state.restype = None
    ref.ref_copy_set.argtypes = [C.c_void_p] + [C.c_int] * 4
    ref.ref_copy_set.restype = None
    READ, WRITE = 4, 8
    DEVS = (0, 2, 3)

    def snapshot(d):
        out = [int(C.c_int8(ref.ref_data_owner(d)).value)]
        buf = (C.c_int * 4)()
        for dev in DEVS:
            ref.ref_copy_state(d, dev, buf)
            out.append(list(buf))
        return out

    cases = []
    for seed in range(24):
        rng = random.Random(1000 + seed)
        d = ref.ref_data_new(seed)
        steps, pending = [], []       # pending: (dev, access) transfers started and not yet ended
        for _ in range(150):
            st = snapshot(d)
            present = [dev for dev, s in zip(DEVS, st[1:]) if s[0]]
            gpus_absent = [dev for dev in (2, 3) if dev not in present]
            roll = rng.random()
            if gpus_absent and roll < 0.15:
                dev = rng.choice(gpus_absent)
                ref.ref_copy_attach_new(d, dev)
                op = ["attach", dev]
            elif pending and roll < 0.45:
                dev, access = pending.pop(rng.randrange(len(pending)))
                ref.ref_end_transfer(d, dev, access)
                op = ["end", dev, access]
            elif roll < 0.75:
                # a writer must own or be about to own: only start a WRITE where no other copy is OWNED by someone else
                dev = rng.choice(present)
                access = rng.choice([READ, READ, READ | WRITE, WRITE])
                owned_elsewhere = any(s[0] and s[1] == 1 and dv != dev for dv, s in zip(DEVS, st[1:]))
                if (access & WRITE) and owned_elsewhere:
                    access = READ
                r = ref.ref_start_transfer(d, dev, access)
                pending.append((dev, access))
                op = ["start", dev, access, r]
            elif roll < 0.9:
                dev = rng.choice(present)
                cur = st[1 + DEVS.index(dev)]
                kind = rng.choice(["version++", "version=host", "release_reader", "shared"])
                if kind == "version++":
                    ref.ref_copy_set(d, dev, -1, cur[3] + 1, -1)
                elif kind == "version=host":
                    ref.ref_copy_set(d, dev, -1, st[1][3], -1)
                elif kind == "release_reader":
                    ref.ref_copy_set(d, dev, -1, -1, max(0, cur[2] - 1))
                else:
                    ref.ref_copy_set(d, dev, 4, -1, -1)
                op = ["set", dev, kind]
            else:
                idle = [dev for dev in (2, 3) if dev in present and st[1 + DEVS.index(dev)][2] == 0
                        and not any(p[0] == dev for p in pending) and st[0] != dev]
                if not idle:
                    continue
                dev = rng.choice(idle)
                ref.ref_copy_detach(d, dev)
                op = ["detach", dev]
            steps.append({"op": op, "state": snapshot(d)})
        cases.append({"seed": seed, "steps": steps})
    json.dump(cases, open(os.path.join(HERE, "ownership_trace.json"), "w"))
    print("wrote ownership_trace.json (%d cases, %d steps)" % (len(cases), sum(len(c["steps"]) for c in cases)))


if __name__ == "__main__":
    ownership_cases()


# ---------------------------------------------------------------------------------------------------------------
# block_cyclic.json: parsec_matrix_block_cyclic_init + rank_of / data_key / data_of of the REFERENCE's
# two_dim_rectangle_cyclic.c, for every tile of a few (matrix, tiling, grid, rank) combinations.
def block_cyclic_cases():
    ref.ref_bc_new.restype = C.c_void_p
    ref.ref_bc_new.argtypes = [C.c_int] * 15
    ref.ref_bc_geometry.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    ref.ref_bc_rank_of.argtypes = [C.c_void_p, C.c_int, C.c_int]
    ref.ref_bc_data_key.restype = C.c_uint64
    ref.ref_bc_data_key.argtypes = [C.c_void_p, C.c_int, C.c_int]
    ref.ref_bc_tile_offset.restype = C.c_int64
    ref.ref_bc_tile_offset.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    cases = []
    plain = [(M, N, mb, nb, P, Q, 1, 1, 0, 0)
             for (M, N, mb, nb) in [(2048, 2048, 256, 256), (1000, 1000, 96, 96), (250, 130, 40, 40), (640, 384, 64, 128)]
             for (P, Q) in [(1, 1), (1, 2), (2, 2), (2, 4), (3, 2)]]
    # k-cyclic super-tiles (twoDBC_kcyclic_*) and a shifted process-grid origin (ip, jq)
    kcyc = [(M, N, mb, nb, P, Q, kp, kq, ip, jq)
            for (M, N, mb, nb) in [(1000, 1000, 96, 96), (640, 384, 64, 128)]
            for (P, Q, kp, kq, ip, jq) in [(1, 1, 2, 3, 0, 0), (2, 2, 2, 1, 0, 0), (2, 3, 2, 3, 0, 0), (3, 2, 3, 2, 1, 1),
                                           (2, 2, 1, 1, 1, 0)]]
    # PARSEC_MATRIX_LAPACK storage (parsec_matrix_block_cyclic_lapack_init): column-major local matrix, ld = mloc
    ref.ref_bc_lapack_new.restype = C.c_void_p
    ref.ref_bc_lapack_new.argtypes = [C.c_int] * 17
    lapack = [(M, N, mb, nb, P, Q, kp, kq, ip, jq, True)
              for (M, N, mb, nb) in [(1000, 1000, 96, 96), (250, 130, 40, 40)]
              for (P, Q, kp, kq, ip, jq) in [(1, 1, 1, 1, 0, 0), (2, 2, 1, 1, 0, 0), (2, 3, 2, 1, 0, 0)]]
    for spec in [c + (False,) for c in plain + kcyc] + lapack:
            (M, N, mb, nb, P, Q, kp, kq, ip, jq, is_lapack) = spec
            for rank in range(P * Q):
                dc = ref.ref_bc_new(rank, mb, nb, M, N, 0, 0, M, N, P, Q, kp, kq, ip, jq)
                mloc = nloc = 0
                if is_lapack:   # local matrix = the locally owned tile rows / columns, whole tiles (pad >= numroc)
                    g0 = (C.c_int * 14)()
                    ref.ref_bc_geometry(dc, g0)
                    mloc, nloc = max(g0[9] * mb, 1), max(g0[10] * nb, 1)
                    dc = ref.ref_bc_lapack_new(rank, mb, nb, M, N, 0, 0, M, N, P, Q, kp, kq, ip, jq, mloc, nloc)
                g = (C.c_int * 14)()
                ref.ref_bc_geometry(dc, g)
                geo = dict(zip("mt nt lmt lnt lm ln llm lln bsiz nb_elem_r nb_elem_c nb_local_tiles rrank crank".split(), list(g)))
                tiles = []
                for n in range(geo["nt"]):
                    for m in range(geo["mt"]):
                        owner = ref.ref_bc_rank_of(dc, m, n)
                        entry = [m, n, owner, int(ref.ref_bc_data_key(dc, m, n))]
                        if owner == rank:
                            key, nbe = C.c_uint64(), C.c_uint32()
                            off = ref.ref_bc_tile_offset(dc, m, n, C.byref(key), C.byref(nbe))
                            entry += [int(off), int(key.value), int(nbe.value)]
                        tiles.append(entry)
                cases.append({"M": M, "N": N, "mb": mb, "nb": nb, "P": P, "Q": Q, "kp": kp, "kq": kq, "ip": ip, "jq": jq,
                              "lapack": [mloc, nloc] if is_lapack else None, "rank": rank, "geo": geo, "tiles": tiles})
    json.dump(cases, open(os.path.join(HERE, "block_cyclic.json"), "w"))
    print("wrote block_cyclic.json (%d descriptors, %d tiles)" % (len(cases), sum(len(c["tiles"]) for c in cases)))


if __name__ == "__main__":
    block_cyclic_cases()


# ---------------------------------------------------------------------------------------------------------------
# lcg_tiles.json: tiles produced by the REFERENCE's own initialize_tile() (tests/dsl/dtd/dtd_test_simple_gemm.c:68-91,
# compiled in place through oracle/refshim/gen_wrapper.c), the synthetic-input generator of the path.
def lcg_cases():
    ref.ref_init_tile.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_uint]
    cases = []
    for (m, n, mb, nb, M, seed) in [(0, 0, 8, 8, 64, 1789), (3, 5, 16, 4, 1024, 1805), (31, 31, 8, 8, 32768, 1901),
                                    (7, 2, 5, 9, 250, 1789), (63, 63, 4, 4, 65536, 1805), (1, 0, 32, 2, 2048, 1901)]:
        buf = (C.c_double * (mb * nb))()
        assert ref.ref_init_tile(buf, m, n, mb, nb, M, mb, seed) == 0
        cases.append({"m": m, "n": n, "mb": mb, "nb": nb, "M": M, "ld": mb, "seed": seed,
                      "values": [float(v).hex() for v in buf]})
    json.dump(cases, open(os.path.join(HERE, "lcg_tiles.json"), "w"))
    print("wrote lcg_tiles.json (%d tiles)" % len(cases))


if __name__ == "__main__":
    lcg_cases()
