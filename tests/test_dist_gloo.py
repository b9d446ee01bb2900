"""Host-side logic of the P x Q data plane on the CPU: the broadcast plan (pb2_comm_plan, through the C ABI) is
checked for consistency across ranks, and then executed by two gloo processes with torch.distributed.broadcast
standing in for ncclBroadcast; the tile arithmetic is the oracle's.  No GPU involved."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

H = load_hostapi()


# (kp, kq, a_kq, b_kp, ip, jq): k-cyclic super-tiles of C's rows / columns, of the k dimension in A / in B, grid origin
PLAIN = (1, 1, 1, 1, 0, 0)


@pytest.mark.parametrize("P,Q,MT,NT,KT,dist", [
    (1, 2, 4, 4, 4, PLAIN), (2, 2, 5, 3, 4, PLAIN), (2, 4, 8, 8, 8, PLAIN), (3, 2, 7, 5, 3, PLAIN), (1, 1, 3, 3, 3, PLAIN),
    (2, 2, 9, 7, 6, (2, 3, 1, 1, 0, 0)), (2, 3, 11, 10, 9, (2, 2, 3, 2, 0, 0)), (3, 2, 10, 9, 8, (3, 1, 2, 2, 1, 1)),
    (2, 4, 8, 13, 7, (1, 2, 1, 3, 1, 3))])
def test_plan_is_consistent_across_ranks(P, Q, MT, NT, KT, dist):
    kp, kq, a_kq, b_kp, ip, jq = dist
    plans = {r: H.comm_plan(P, Q, r, MT, NT, KT, *dist) for r in range(P * Q)}
    dA = O.bc(0, 1, 1, MT, KT, 0, 0, MT, KT, P, Q, kp, a_kq, ip, jq)
    dB = O.bc(0, 1, 1, KT, NT, 0, 0, KT, NT, P, Q, b_kp, kq, ip, jq)
    dC = O.bc(0, 1, 1, MT, NT, 0, 0, MT, NT, P, Q, kp, kq, ip, jq)
    rank_of = lambda d, m, n: O.lib().orc_bc_rank_of(C.byref(d), m, n)
    if P * Q == 1:
        assert plans[0] == []
        return
    # every member of a communicator issues the same sequence on it, exactly one member roots each op, the root is the
    # tile's owner under the reference's rank_of, and a process row moves exactly the A rows of the C tiles it owns
    for row in range(P):
        members = [row * Q + c for c in range(Q)]
        seqs = [[op[:4] for op in plans[r] if op[0] == 0] for r in members]
        assert all(s == seqs[0] for s in seqs)
        if Q > 1:
            mine = [m for m in range(MT) if rank_of(dC, m, 0) // Q == row]
            assert sorted((k, m) for (_, k, m, _) in seqs[0]) == [(k, m) for k in range(KT) for m in mine]
        for i, (_, k, m, root) in enumerate(seqs[0]):
            roots = [r for r in members if [op for op in plans[r] if op[0] == 0][i][4]]
            assert roots == [row * Q + root] == [rank_of(dA, m, k)]
    for col in range(Q):
        members = [r * Q + col for r in range(P)]
        seqs = [[op[:4] for op in plans[r] if op[0] == 1] for r in members]
        assert all(s == seqs[0] for s in seqs)
        if P > 1:
            mine = [n for n in range(NT) if rank_of(dC, 0, n) % Q == col]
            assert sorted((k, n) for (_, k, n, _) in seqs[0]) == [(k, n) for k in range(KT) for n in mine]
        for i, (_, k, n, root) in enumerate(seqs[0]):
            roots = [r for r in members if [op for op in plans[r] if op[0] == 1][i][4]]
            assert roots == [root * Q + col] == [rank_of(dB, k, n)]
    # canonical order inside a rank: k ascending, rows before columns inside a step
    for r, ops in plans.items():
        assert [(k, which) for (which, k, _, _, _) in ops] == sorted((k, which) for (which, k, _, _, _) in ops)
    # every operand of every GEMM a rank owns is either local or received
    for r, ops in plans.items():
        got = {(w, k, i) for (w, k, i, _, _) in ops}
        for m in range(MT):
            for n in range(NT):
                if rank_of(dC, m, n) != r:
                    continue
                for k in range(KT):
                    assert rank_of(dA, m, k) == r or (0, k, m) in got
                    assert rank_of(dB, k, n) == r or (1, k, n) in got


def test_plan_rejects_bad_arguments():
    with pytest.raises(ValueError):
        H.comm_plan(2, 2, 4, 4, 4, 4)
    with pytest.raises(ValueError):
        H.comm_plan(0, 2, 0, 4, 4, 4)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, P, Q, N, NB, out_dir, dist_spec):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        MT = N // NB
        kp, kq, a_kq, b_kp, ip, jq = dist_spec
        descs = [O.bc(rank, NB, NB, N, N, 0, 0, N, N, P, Q, r, c, ip, jq) for (r, c) in ((kp, a_kq), (b_kp, kq), (kp, kq))]
        local = [O.fill(d, s) for d, s in zip(descs, (1789, 1805, 1901))]
        myrow, mycol = rank // Q, rank % Q
        row_group = dist.new_group([myrow * Q + c for c in range(Q)]) if Q > 1 else None
        for r in range(P):  # new_group is collective over the world: every rank creates every group
            if r != myrow and Q > 1:
                dist.new_group([r * Q + c for c in range(Q)])
        col_groups = {}
        for c in range(Q):
            if P > 1:
                g = dist.new_group([r * Q + c for r in range(P)])
                if c == mycol:
                    col_groups[c] = g

        def tile_of(which, a, b):
            d, mat = (descs[0], local[0]) if which == 0 else (descs[1], local[1])
            pos = O.lib().orc_bc_position(C.byref(d), a, b)
            return mat[pos * NB * NB:(pos + 1) * NB * NB]

        got = {}
        for (which, k, idx, root, i_am_root) in H.comm_plan(P, Q, rank, MT, MT, MT, *dist_spec):
            buf = torch.from_numpy(tile_of(which, idx, k).copy() if which == 0 and i_am_root else
                                   tile_of(which, k, idx).copy() if i_am_root else np.zeros(NB * NB))
            if which == 0:
                dist.broadcast(buf, src=myrow * Q + root, group=row_group)
            else:
                dist.broadcast(buf, src=root * Q + mycol, group=col_groups[mycol])
            got[(which, k, idx)] = buf.numpy()

        def operand(which, a, b):  # A(a, b) or B(a, b)
            key = (0, b, a) if which == 0 else (1, a, b)
            if key in got:
                return got[key]
            return tile_of(which, a, b)

        result = {}
        for n in range(MT):
            for m in range(MT):
                if O.lib().orc_bc_rank_of(C.byref(descs[2]), m, n) != rank:
                    continue
                pos = O.lib().orc_bc_position(C.byref(descs[2]), m, n)
                c = local[2][pos * NB * NB:(pos + 1) * NB * NB].reshape((NB, NB), order="F").copy(order="F")
                for k in range(MT):
                    a = operand(0, m, k).reshape((NB, NB), order="F")
                    b = operand(1, k, n).reshape((NB, NB), order="F")
                    O.dgemm_tile(1.0, np.asfortranarray(a), np.asfortranarray(b), 1.0, c)
                result[(m, n)] = c
        np.save(os.path.join(out_dir, "rank%d.npy" % rank), result, allow_pickle=True)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("P,Q,dist_spec", [(1, 2, PLAIN), (2, 1, PLAIN), (1, 2, (1, 2, 2, 1, 0, 1)), (2, 1, (2, 1, 1, 2, 1, 0))])
def test_two_gloo_ranks_reproduce_the_single_rank_result(tmp_path, P, Q, dist_spec):
    import torch.multiprocessing as mp
    N, NB = 384, 64
    port = _free_port()
    mp.spawn(_worker, args=(2, port, P, Q, N, NB, str(tmp_path), dist_spec), nprocs=2, join=True)
    d = O.square(N, NB)
    A, B, Cm = O.fill(d, 1789), O.fill(d, 1805), O.fill(d, 1901)
    O.tiled_dgemm(N, N, N, NB, NB, NB, A, B, Cm)
    seen = 0
    for rank in range(2):
        res = np.load(os.path.join(str(tmp_path), "rank%d.npy" % rank), allow_pickle=True).item()
        for (m, n), tile in res.items():
            pos = O.lib().orc_bc_position(C.byref(d), m, n)
            want = Cm[pos * NB * NB:(pos + 1) * NB * NB].reshape((NB, NB), order="F")
            assert np.array_equal(tile, want)  # same tile arithmetic in the same k order: bit-identical
            seen += 1
    assert seen == (N // NB) ** 2


def test_plan_invariants_on_random_grids():
    """Property test over random grids, k-cyclicities and origins: on every communicator all members issue the same
    sequence with exactly one root per operation (the tile's owner), and every operand of every GEMM a rank owns is
    local or received exactly once."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=40, deadline=None)
    @given(st.integers(1, 3), st.integers(1, 4), st.integers(1, 9), st.integers(1, 9), st.integers(1, 7), st.integers(1, 3),
           st.integers(1, 3), st.integers(1, 3), st.integers(1, 3), st.integers(0, 2), st.integers(0, 3))
    def check(P, Q, MT, NT, KT, kp, kq, a_kq, b_kp, ip, jq):
        ip, jq = ip % P, jq % Q
        spec = (kp, kq, a_kq, b_kp, ip, jq)
        plans = {r: H.comm_plan(P, Q, r, MT, NT, KT, *spec) for r in range(P * Q)}
        dA = O.bc(0, 1, 1, MT, KT, 0, 0, MT, KT, P, Q, kp, a_kq, ip, jq)
        dB = O.bc(0, 1, 1, KT, NT, 0, 0, KT, NT, P, Q, b_kp, kq, ip, jq)
        dC = O.bc(0, 1, 1, MT, NT, 0, 0, MT, NT, P, Q, kp, kq, ip, jq)
        rank_of = lambda d, m, n: O.lib().orc_bc_rank_of(C.byref(d), m, n)
        for which, members_of in ((0, lambda r: [(r // Q) * Q + c for c in range(Q)]),
                                  (1, lambda r: [x * Q + (r % Q) for x in range(P)])):
            for r in range(P * Q):
                members = members_of(r)
                seqs = [[op[:4] for op in plans[x] if op[0] == which] for x in members]
                assert all(s == seqs[0] for s in seqs)
                for i, (_, k, idx, root) in enumerate(seqs[0]):
                    roots = [x for x in members if [op for op in plans[x] if op[0] == which][i][4]]
                    owner = rank_of(dA, idx, k) if which == 0 else rank_of(dB, k, idx)
                    assert roots == [owner] and members[root] == owner
        for r, ops in plans.items():
            keys = [(w, k, i) for (w, k, i, _, _) in ops]
            assert len(keys) == len(set(keys))
            got = set(keys)
            for m in range(MT):
                for n in range(NT):
                    if rank_of(dC, m, n) == r:
                        for k in range(KT):
                            assert rank_of(dA, m, k) == r or (0, k, m) in got
                            assert rank_of(dB, k, n) == r or (1, k, n) in got

    check()
