"""libpb2mpi.so (overlay/pb2mpi): the MPI datatype subset the reference's datatype layer is built on
(parsec/datatype/datatype_mpi.c), checked against an independent numpy model of the MPI-3.1 typemap rules:
size, lower bound, extent, and the byte stream MPI_Pack produces, for the constructors the reference uses for its tiles
(parsec/data_dist/matrix/matrixtypes.c: contiguous, vector, resized; upper / lower triangles through indexed) and for the
activation message (struct).  No GPU, one process."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "baseline", "_ref", "dist", "libpb2mpi.so")

MPI_BYTE, MPI_INT, MPI_DOUBLE, MPI_LB, MPI_UB = 1, 8, 15, 28, 29
SIZES = {MPI_BYTE: 1, MPI_INT: 4, MPI_DOUBLE: 8}


@pytest.fixture(scope="module")
def M():
    if not os.path.exists(LIB):
        pytest.skip("baseline/_ref/dist/libpb2mpi.so is not built (overlay/build_overlay.sh)")
    C.CDLL(os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so"), mode=C.RTLD_GLOBAL)
    m = C.CDLL(LIB)
    m.MPI_Type_create_hvector.argtypes = [C.c_int, C.c_int, C.c_long, C.c_int, C.POINTER(C.c_int)]
    m.MPI_Type_create_resized.argtypes = [C.c_int, C.c_long, C.c_long, C.POINTER(C.c_int)]
    return m


# ---- the model: a type is (list of (disp, len) in packing order, lb, ub)
def prim(t):
    return ([(0, SIZES[t])], 0, SIZES[t])


def place(out, old, disp, count):
    segs, lb, ub = old
    ext = ub - lb
    for c in range(count):
        out[0].extend((disp + c * ext + d, n) for d, n in segs)
    if count > 0:
        out[1].append(disp + lb + min(0, (count - 1) * ext))
        out[2].append(disp + ub + max(0, (count - 1) * ext))


def finish(out):
    return (out[0], min(out[1]) if out[1] else 0, max(out[2]) if out[2] else 0)


def m_contig(n, old):
    out = ([], [], [])
    place(out, old, 0, n)
    return finish(out)


def m_hvector(count, bl, stride, old):
    out = ([], [], [])
    for i in range(count):
        place(out, old, i * stride, bl)
    return finish(out)


def m_vector(count, bl, stride, old):
    return m_hvector(count, bl, stride * (old[2] - old[1]), old)


def m_indexed(bls, disps, old):
    out = ([], [], [])
    for bl, d in zip(bls, disps):
        place(out, old, d * (old[2] - old[1]), bl)
    return finish(out)


def m_resized(old, lb, ext):
    return (list(old[0]), lb, lb + ext)


def m_pack(buf, count, t):
    segs, lb, ub = t
    return b"".join(bytes(buf[i * (ub - lb) + d: i * (ub - lb) + d + n]) for i in range(count) for d, n in segs)


def check(M, handle, model, count=3):
    size, lb, ext = C.c_int(), C.c_long(), C.c_long()
    assert M.MPI_Type_commit(C.byref(C.c_int(handle))) == 0
    assert M.MPI_Type_size(handle, C.byref(size)) == 0 and size.value == sum(n for _, n in model[0])
    assert M.MPI_Type_get_extent(handle, C.byref(lb), C.byref(ext)) == 0
    assert (lb.value, ext.value) == (model[1], model[2] - model[1])
    lo = min([d for d, _ in model[0]] + [0])
    assert lo >= 0                                           # the cases below keep their pieces at or above the buffer
    span = (count - 1) * max(ext.value, 0) + max([d + n for d, n in model[0]] + [0]) + 8
    rng = np.random.default_rng(handle)
    buf = rng.integers(0, 256, span, dtype=np.uint8)
    out = np.zeros(size.value * count + 16, dtype=np.uint8)
    pos, psize = C.c_int(8), C.c_int()
    assert M.MPI_Pack_size(count, handle, 1, C.byref(psize)) == 0 and psize.value == size.value * count
    assert M.MPI_Pack(buf.ctypes.data, count, handle, out.ctypes.data, out.size, C.byref(pos), 1) == 0
    assert pos.value == 8 + size.value * count
    assert bytes(out[8:pos.value]) == m_pack(buf, count, model) and not out[:8].any()
    back = np.zeros_like(buf)                                 # unpack is the inverse on the bytes the type names
    pos = C.c_int(8)
    assert M.MPI_Unpack(out.ctypes.data, out.size, C.byref(pos), back.ctypes.data, count, handle, 1) == 0
    named = np.zeros(span, dtype=bool)
    for i in range(count):
        for d, n in model[0]:
            named[i * (model[2] - model[1]) + d: i * (model[2] - model[1]) + d + n] = True
    assert np.array_equal(back[named], buf[named]) and not back[~named].any()
    # the layout handed to the comm engine: same bytes, adjacent pieces merged
    M.PB2MPI_Type_layout.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    segs, nseg, e, s = C.c_void_p(), C.c_int(), C.c_int64(), C.c_int64()
    assert M.PB2MPI_Type_layout(handle, C.byref(segs), C.byref(nseg), C.byref(e), C.byref(s)) == 0
    arr = np.ctypeslib.as_array(C.cast(segs, C.POINTER(C.c_int64)), shape=(nseg.value, 2)) if nseg.value else np.zeros((0, 2))
    merged = []
    for d, n in model[0]:
        if merged and merged[-1][0] + merged[-1][1] == d:
            merged[-1][1] += n
        else:
            merged.append([d, n])
    assert arr.tolist() == merged and e.value == ext.value and s.value == size.value


def new(M, fn, *args):
    h = C.c_int()
    assert fn(*args, C.byref(h)) == 0
    return h.value


def test_tile_types_of_the_reference(M):
    # parsec_matrix_define_contiguous / _rectangle (matrixtypes.c:40-130): a full tile, and mb x nb out of ld = 40
    mb, nb, ld = 24, 17, 40
    d = prim(MPI_DOUBLE)
    check(M, new(M, M.MPI_Type_contiguous, mb * nb, MPI_DOUBLE), m_contig(mb * nb, d))
    rect = new(M, M.MPI_Type_vector, nb, mb, ld, MPI_DOUBLE)
    check(M, rect, m_vector(nb, mb, ld, d))
    # resized so that consecutive tiles of a column start mb rows apart (the LAPACK-storage datatype)
    check(M, new(M, M.MPI_Type_create_resized, rect, 0, mb * 8), m_resized(m_vector(nb, mb, ld, d), 0, mb * 8), count=2)
    # the same rectangle by bytes
    check(M, new(M, M.MPI_Type_create_hvector, nb, mb, ld * 8, MPI_DOUBLE), m_hvector(nb, mb, ld * 8, d))
    # lower triangle with the diagonal (matrixtypes.c:132-195): column j holds rows j .. n-1
    n = 9
    bls = (C.c_int * n)(*[n - j for j in range(n)])
    disps = (C.c_int * n)(*[j * ld + j for j in range(n)])
    check(M, new(M, M.MPI_Type_indexed, n, bls, disps, MPI_DOUBLE), m_indexed(list(bls), list(disps), d))
    ib = (C.c_int * 4)(0, 5, 11, 30)
    check(M, new(M, M.MPI_Type_create_indexed_block, 4, 3, ib, MPI_DOUBLE), m_indexed([3] * 4, list(ib), d))


def test_nested_and_struct_types(M):
    d, i4 = prim(MPI_DOUBLE), prim(MPI_INT)
    inner = new(M, M.MPI_Type_vector, 3, 2, 5, MPI_DOUBLE)
    m_inner = m_vector(3, 2, 5, d)
    check(M, new(M, M.MPI_Type_contiguous, 4, inner), m_contig(4, m_inner))                # spaced by the inner extent
    check(M, new(M, M.MPI_Type_vector, 2, 2, 3, inner), m_vector(2, 2, 3, m_inner))
    # struct {int a[3]; double b[2];} with natural padding, then an array of it
    bl = (C.c_int * 2)(3, 2)
    dp = (C.c_long * 2)(0, 16)
    tp = (C.c_int * 2)(MPI_INT, MPI_DOUBLE)
    s = new(M, M.MPI_Type_create_struct, 2, bl, dp, tp)
    out = ([], [], [])
    place(out, i4, 0, 3)
    place(out, d, 16, 2)
    check(M, s, finish(out))
    # MPI-1 resize through MPI_LB / MPI_UB markers (datatype_mpi.c:139-143 without MPI-2)
    bl3 = (C.c_int * 3)(1, 1, 1)
    dp3 = (C.c_long * 3)(0, 0, 64)
    tp3 = (C.c_int * 3)(MPI_LB, inner, MPI_UB)
    check(M, new(M, M.MPI_Type_struct, 3, bl3, dp3, tp3), m_resized(m_inner, 0, 64), count=2)


def test_copy_between_layouts_is_a_send_to_self(M):
    """PB2MPI_Copy: what the reference's reshape does with MPI_Sendrecv to itself (parsec_mpi_funnelled.c:384-407)."""
    mb, nb, ld = 6, 5, 11
    rect = new(M, M.MPI_Type_vector, nb, mb, ld, MPI_DOUBLE)
    full = new(M, M.MPI_Type_contiguous, mb * nb, MPI_DOUBLE)
    src = np.arange(ld * nb, dtype=np.float64)
    dst = np.zeros(mb * nb)
    assert M.PB2MPI_Copy(src.ctypes.data, 1, rect, dst.ctypes.data, 1, full) == 0
    assert np.array_equal(dst.reshape(nb, mb), src.reshape(nb, ld)[:, :mb])
    back = np.zeros(ld * nb)
    assert M.PB2MPI_Copy(dst.ctypes.data, mb * nb, MPI_DOUBLE, back.ctypes.data, 1, rect) == 0
    want = np.zeros((nb, ld))
    want[:, :mb] = src.reshape(nb, ld)[:, :mb]
    assert np.array_equal(back.reshape(nb, ld), want)
    assert M.PB2MPI_Copy(src.ctypes.data, 2, rect, dst.ctypes.data, 1, full) != 0          # would truncate


def test_a_world_of_one_and_handles(M):
    for k in ("PB2_CE_RANK", "PB2_CE_WORLD", "RANK", "WORLD_SIZE"):
        os.environ.pop(k, None)
    prov, flag, n, r = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    assert M.MPI_Init_thread(None, None, 2, C.byref(prov)) == 0 and prov.value == 2
    assert M.MPI_Initialized(C.byref(flag)) == 0 and flag.value == 1
    assert M.MPI_Comm_size(1, C.byref(n)) == 0 and M.MPI_Comm_rank(1, C.byref(r)) == 0 and (n.value, r.value) == (1, 0)
    dup, res = C.c_int(), C.c_int()
    assert M.MPI_Comm_dup(1, C.byref(dup)) == 0 and dup.value not in (0, 1)
    assert M.MPI_Comm_compare(1, dup, C.byref(res)) == 0 and res.value == 1               # MPI_CONGRUENT
    v = C.c_int(41)
    assert M.MPI_Allreduce(C.c_void_p(-1), C.byref(v), 1, MPI_INT, 1, dup) == 0 and v.value == 41
    assert M.MPI_Barrier(dup) == 0
    name, ln = C.create_string_buffer(64), C.c_int()
    assert M.MPI_Type_get_name(MPI_DOUBLE, name, C.byref(ln)) == 0 and name.value == b"MPI_DOUBLE"
    t = new(M, M.MPI_Type_contiguous, 4, MPI_DOUBLE)
    assert M.MPI_Type_set_name(t, b"CONT MPI_DOUBLE*   4") == 0
    assert M.MPI_Type_get_name(t, name, C.byref(ln)) == 0 and name.value == b"CONT MPI_DOUBLE*   4"
    comb = C.c_int()
    assert M.MPI_Type_get_envelope(t, C.byref(n), C.byref(r), C.byref(flag), C.byref(comb)) == 0 and comb.value == 3
    h = C.c_int(t)
    assert M.MPI_Type_free(C.byref(h)) == 0 and h.value == 0
    assert M.MPI_Type_size(t, C.byref(n)) != 0                                             # the handle is gone


def test_random_nested_types_against_the_model(M):
    """Randomly nested constructors (what a user of parsec_type_create_* can build), 60 types: size, bounds, packing,
    unpacking and the merged layout must all follow the typemap rules of the model above."""
    rng = np.random.default_rng(20260)

    def build(depth):
        if depth == 0 or rng.random() < 0.25:
            t = int(rng.choice([MPI_BYTE, MPI_INT, MPI_DOUBLE]))
            return t, prim(t)
        h, m = build(depth - 1)
        ext = m[2] - m[1]
        kind = int(rng.integers(0, 5))
        if kind == 0:
            n = int(rng.integers(1, 5))
            return new(M, M.MPI_Type_contiguous, n, h), m_contig(n, m)
        if kind == 1:
            c, bl = int(rng.integers(1, 4)), int(rng.integers(1, 4))
            st = bl + int(rng.integers(0, 4))
            return new(M, M.MPI_Type_vector, c, bl, st, h), m_vector(c, bl, st, m)
        if kind == 2:
            c, bl = int(rng.integers(1, 4)), int(rng.integers(1, 3))
            st = bl * ext + 8 * int(rng.integers(0, 5))
            return new(M, M.MPI_Type_create_hvector, c, bl, st, h), m_hvector(c, bl, st, m)
        if kind == 3:
            c = int(rng.integers(1, 4))
            bls = [int(rng.integers(1, 3)) for _ in range(c)]
            disps, at = [], 0
            for b in bls:                                   # increasing, non-overlapping displacements
                at += int(rng.integers(0, 3))
                disps.append(at)
                at += b
            return (new(M, M.MPI_Type_indexed, c, (C.c_int * c)(*bls), (C.c_int * c)(*disps), h), m_indexed(bls, disps, m))
        pad = 8 * int(rng.integers(0, 4))
        return new(M, M.MPI_Type_create_resized, h, m[1], ext + pad), m_resized(m, m[1], ext + pad)

    for _ in range(60):
        handle, model = build(int(rng.integers(1, 4)))
        if handle < 64:
            continue
        check(M, handle, model, count=int(rng.integers(1, 4)))
