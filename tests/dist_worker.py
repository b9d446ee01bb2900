"""One rank of the multi-GPU parity check (launched by tests/test_gpu_dist.py through torch.distributed.run)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import oracle_lib as O  # noqa: E402
from conftest import load_hostapi  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    H = load_hostapi()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    N, NB = int(sys.argv[1]), int(sys.argv[2])
    # optional: kp kq a_kq b_kp ip jq (k-cyclic super-tiles of C's rows / columns, of k in A / in B, grid origin)
    kp, kq, a_kq, b_kp, ip, jq = [int(x) for x in sys.argv[3:9]] if len(sys.argv) >= 9 else (1, 1, 1, 1, 0, 0)
    layouts = ((kp, a_kq), (b_kp, kq), (kp, kq))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P, Q = H.grid_for(world)
    comm = H.Comm(local, P, Q, rank)
    dev = H.Device(local, memory_number_of_blocks=4096)
    mats, bufs = [], []
    for seed, (lr, lc) in zip((1789, 1805, 1901), layouts):
        m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB, P, Q, rank, kp=lr, kq=lc, ip=ip % P, jq=jq % Q)
        b = torch.zeros(max(m.nbytes, 8), dtype=torch.uint8).pin_memory()
        m.set_buffer(b)
        m.fill_lcg(seed, dev)
        mats.append(m)
        bufs.append(b)
    for step in range(2):   # twice: the second run finds the owned tiles cached and re-broadcasts the rest
        st = H.tiled_gemm(mats[0], mats[1], mats[2], dev, comm=comm)
    MT = -(-N // NB)
    dl = O.bc(rank, NB, NB, N, N, 0, 0, N, N, P, Q, kp, kq, ip % P, jq % Q)
    is_mine = lambda m, n: O.lib().orc_bc_rank_of(C.byref(dl), m, n) == rank
    owned = sum(1 for n in range(MT) for m in range(MT) if is_mine(m, n))
    assert st["nb_gemm"] == owned * MT, (st, owned)
    # oracle: the whole product on one rank, run twice as well; compare my tiles
    d1 = O.square(N, NB)
    A, B, Cm = O.fill(d1, 1789), O.fill(d1, 1805), O.fill(d1, 1901)
    for step in range(2):
        O.tiled_dgemm(N, N, N, NB, NB, NB, A, B, Cm)
    mine = bufs[2].numpy().view(np.float64)
    worst = 0.0
    Ad, Bd = O.tiled_to_dense(d1, A), O.tiled_to_dense(d1, B)
    scale = np.abs(Ad).sum(axis=1).max() * np.abs(Bd).sum(axis=1).max() * N * 2.0 ** -53
    for n in range(MT):
        for m in range(MT):
            if not is_mine(m, n):
                continue
            pl = O.lib().orc_bc_position(C.byref(dl), m, n)
            p1 = O.lib().orc_bc_position(C.byref(d1), m, n)
            got = mine[pl * NB * NB:(pl + 1) * NB * NB].reshape((NB, NB), order="F")
            want = Cm[p1 * NB * NB:(p1 + 1) * NB * NB].reshape((NB, NB), order="F")
            worst = max(worst, np.abs(got - want).sum(axis=1).max() / scale)
            assert mats[2].version(m, n) == 2 * MT
    assert worst < 10, worst
    s = dev.stats()
    assert s["data_in_from_peers"] > 0
    print("rank %d ok: norm=%.3g peers_bytes=%d gemm=%d" % (rank, worst, s["data_in_from_peers"], st["nb_gemm"]), flush=True)
    for m in mats:
        m.close()
    dev.close()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
