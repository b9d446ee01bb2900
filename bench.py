#!/usr/bin/env python
"""Benchmark of the hot path: tiled DGEMM C += A*B over a parsec_matrix_block_cyclic_t through the CUDA device
pipeline (stage-in -> hand-written sm_100a tile kernel -> stage-out, HBM tile cache).

    python bench.py --gpus 1 --steps 3 --warmup 3            # our arm
    python bench.py --impl reference --steps 1 --warmup 0    # CPU arm: the reference runtime with CPU bodies on the host cores

One JSON line on stdout (rank 0).  A "step" is one full tiled DGEMM over the workload's matrices.
  value : GFLOP/s with every tile already resident in the HBM cache when the timed region starts
  e2e   : GFLOP/s through the same C-ABI call with the HBM cache flushed first, i.e. every A/B/C tile is staged
          from pinned host memory and C is written back to the host inside the timed region
"""
import argparse
import ctypes as C
import importlib.util
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def load_hostapi():
    spec = importlib.util.spec_from_file_location("parsec1_b200_hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["parsec1_b200_hostapi"] = mod
    spec.loader.exec_module(mod)
    return mod


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu_index, self.rows, self.proc = gpu_index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, reasons, mx, power = [], set(), 0.0, 0.0
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                power = max(power, float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        sm.sort()
        busy = [x for x in sm if x > 0.5 * mx] or sm
        return {"sm_mhz": busy[len(busy) // 2] if busy else None, "sm_max_mhz": mx or None, "power_w_max": power or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_gflops(n, nb, threads):
    """The oracle port (oracle/pb2_oracle.c: orc_tiled_dgemm) on the host cores -- CPU baseline only."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as O
    d = O.square(n, nb)
    A, B, Cm = O.fill(d, 1789, threads), O.fill(d, 1805, threads), O.fill(d, 1901, threads)
    t0 = time.perf_counter()
    O.tiled_dgemm(n, n, n, nb, nb, nb, A, B, Cm, threads)
    dt = time.perf_counter() - t0
    assert np.isfinite(Cm[:16]).all()
    return 2.0 * n ** 3 / dt * 1e-9, dt


def reference_runtime_gflops(n, nb, threads):
    """The reference's own CPU path for this workload: the unmodified reference runtime (baseline/_ref/stock/libparsec.so,
    built from /root/reference by overlay/build_overlay.sh) executing the dgemm_NN JDF (parsec-ptgpp) with CPU task
    bodies on `threads` cores, --mca device_cuda_enabled 0.  The reference ships no tile GEMM (its CPU body calls an
    un-vendored cblas_dgemm), so the body's tile routine is the oracle's blocked plain-C one, handed in with --cpu-tile.
    Returns (GFLOP/s, seconds) or None when the binaries are not there."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    app = os.path.join(ref, "bin", "dgemm_NN_app")
    if not (os.path.exists(app) and os.path.exists(os.path.join(ref, "stock", "libparsec.so.4"))):
        return None
    env = dict(os.environ, LD_LIBRARY_PATH=os.pathsep.join([os.path.join(ref, "stock"), os.environ.get("LD_LIBRARY_PATH", "")]))
    cmd = [app, "-N", str(n), "-t", str(nb), "-c", str(threads), "--cpu-tile",
           os.path.join(ROOT, "oracle", "liboracle.so") + ":orc_dgemm_tile", "--", "--mca", "device_cuda_enabled", "0"]
    try:
        p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=1800)
        line = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    except Exception:
        return None
    if p.returncode != 0 or line.get("gemm_cpu") != (n // nb) ** 3:
        return None
    return line["gflops_best"], line["seconds_best"]


def cpu_baseline(n, nb, threads):
    """cpu_baseline object: the reference runtime with CPU bodies when it is built, else the oracle port."""
    r = reference_runtime_gflops(n, nb, threads)
    if r is not None:
        return {"value": r[0], "unit": "GFLOP/s", "cores": threads, "kind": "reference",
                "sample": "tiled DGEMM N=%d NB=%d (%.1f s): the reference runtime (unmodified libparsec, dgemm_NN JDF through "
                          "parsec-ptgpp) with CPU task bodies; tile routine = the oracle's blocked plain C" % (n, nb, r[1])}, r[1]
    g, dt = cpu_port_gflops(n, nb, threads)
    return {"value": g, "unit": "GFLOP/s", "cores": threads, "kind": "port",
            "sample": "tiled DGEMM N=%d NB=%d (%.1f s), oracle port with OpenMP (baseline/_ref not built)" % (n, nb, dt)}, dt


def run_reference(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n, nb = args.cpu_n, args.nb
    for _ in range(args.warmup):
        cpu_baseline(min(n, 2 * nb), nb, threads)
    vals, secs, base = [], [], None
    for _ in range(max(1, args.steps)):
        base, dt = cpu_baseline(n, nb, threads)
        vals.append(base["value"])
        secs.append(dt)
    v = sum(vals) / len(vals)
    base = dict(base, value=v)
    print(json.dumps({
        "impl": "reference", "metric": "tiled GEMM GFLOP/s", "value": v, "unit": "GFLOP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "tiled DGEMM N=%d NB=%d fp64 via CPU task bodies" % (args.n, nb), "sample": base["sample"]},
        "cpu_baseline": base,
        "e2e": {"value": v, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def fp64_peak_tflops(torch):
    """FP64 tensor peak measured the way MEASURED_PEAKS.json measures bf16: cuBLAS GEMM 8192^3, best of 10."""
    n = 8192
    a = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
    b = torch.rand(n, n, dtype=torch.float64, device="cuda") - 0.5
    c = torch.empty(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b, c
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / (best * 1e-3) * 1e-12


def parity_check(H, torch, dist, dev, bufs, mtypes, N, NB, P, Q, grid_rank, rank, world, knobs, bf16, vectors=4):
    """One extra step from freshly generated inputs, checked without the oracle and without our kernels:
      * Freivalds: for `vectors` random +-1 vectors X, C1 X must equal C0 X + A (B X); every rank multiplies the tiles it
        owns (torch on the GPU, fp64), partial products are summed over the ranks;
      * two C tiles per rank recomputed as C0 + sum_k A(m,k) B(k,n) with torch.matmul in fp64 (cuBLAS), operand tiles
        fetched from their owners with torch.distributed broadcasts;
      * every owned C tile's host copy must end at version KT.
    Both residuals are scaled like the north-star norm, ||.||inf / (||A||inf ||B||inf N eps), eps = 2^-53 (fp64) or
    2^-24 (fp32 accumulation of bf16 products); the worst over ranks is reported."""
    L = H.lib()
    T = N // NB
    assert T * NB == N
    cuda = torch.device("cuda", torch.cuda.current_device())
    tdt = {H.MATRIX_DOUBLE: torch.float64, H.MATRIX_FLOAT: torch.float32, H.MATRIX_BF16: torch.bfloat16}
    mats = []
    for seed, mtype, b in zip((1789, 1805, 1901), mtypes, bufs):
        m = H.Matrix(mtype, N, N, NB, NB, P, Q, grid_rank)     # fresh descriptors: versions start at 0
        m.set_buffer(b)
        m.fill_lcg(seed, dev)
        mats.append(m)
    host = [b[:m.nbytes].view(tdt[mt]) for b, m, mt in zip(bufs, mats, mtypes)]
    bs = NB * NB

    def owned_columns(mat):
        """{n: [(m, pos), ...] sorted by position}: the tiles of tile column n this rank stores."""
        cols = {}
        for n in range(T):
            for m in range(T):
                if L.pb2_matrix_rank_of(mat.h, m, n) == grid_rank:
                    cols.setdefault(n, []).append((m, int(L.pb2_matrix_position(mat.h, m, n))))
        for n in cols:
            cols[n].sort(key=lambda t: t[1])
        return cols

    cols = [owned_columns(m) for m in mats]

    def allsum(t):
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def matvec(i, X):
        """(M X, row sums of |M|) over the tiles of matrix i this rank owns, summed over the ranks."""
        Y = torch.zeros(N, X.shape[1], dtype=torch.float64, device=cuda)
        rows = torch.zeros(N, dtype=torch.float64, device=cuda)
        for n, tiles in cols[i].items():
            Xn = X[n * NB:(n + 1) * NB]
            for c0 in range(0, len(tiles), 32):
                chunk = tiles[c0:c0 + 32]
                if all(chunk[j + 1][1] == chunk[j][1] + 1 for j in range(len(chunk) - 1)):
                    slab = host[i][chunk[0][1] * bs:(chunk[-1][1] + 1) * bs].to(cuda, non_blocking=True)
                else:
                    slab = torch.cat([host[i][p * bs:(p + 1) * bs] for _, p in chunk]).to(cuda, non_blocking=True)
                slab = slab.view(len(chunk), NB, NB).to(torch.float64)       # [tile][column j][row i]
                part = torch.einsum("cji,jr->cir", slab, Xn)
                rs = slab.abs().sum(dim=1)
                for j, (m, _) in enumerate(chunk):
                    Y[m * NB:(m + 1) * NB] += part[j]
                    rows[m * NB:(m + 1) * NB] += rs[j]
        return allsum(Y), allsum(rows)

    def tile(i, pos):
        return host[i][pos * bs:(pos + 1) * bs].to(cuda).view(NB, NB).to(torch.float64).t()   # [row i][column j]

    g = torch.Generator(device="cpu")
    g.manual_seed(11)
    X = (torch.randint(0, 2, (N, vectors), generator=g).to(torch.float64) * 2.0 - 1.0).to(cuda)
    c0x, _ = matvec(2, X)
    # sampled tiles: the first and the last C tile each rank stores
    my_c = sorted((pos, m, n) for n, tiles in cols[2].items() for m, pos in tiles)
    picks_mine = [my_c[0], my_c[-1]] if len(my_c) > 1 else my_c
    c0_tiles = {(m, n): tile(2, pos) for pos, m, n in picks_mine}
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, [(m, n) for _, m, n in picks_mine])
    else:
        gathered = [[(m, n) for _, m, n in picks_mine]]

    st = H.tiled_gemm(mats[0], mats[1], mats[2], dev, **knobs)

    eps = 2.0 ** -24 if bf16 else 2.0 ** -53
    bx, rows_b = matvec(1, X)
    abx, rows_a = matvec(0, bx)
    c1x, _ = matvec(2, X)
    scale = float(rows_a.max()) * float(rows_b.max()) * N * eps
    freivalds = float((c1x - c0x - abx).abs().max()) / scale
    changed = float((c1x - c0x).abs().max()) > 1.0
    pos_of = lambda i, m, n: int(L.pb2_matrix_position(mats[i].h, m, n))
    tile_norm, nb_tiles = 0.0, 0
    for owner, picks in enumerate(gathered):
        for (m, n) in picks:
            acc = c0_tiles[(m, n)].clone() if owner == rank else None
            a_rows = torch.zeros(NB, dtype=torch.float64, device=cuda)
            b_cols = torch.zeros(NB, dtype=torch.float64, device=cuda)
            for k in range(T):
                ops = []
                for i, (tm, tn) in ((0, (m, k)), (1, (k, n))):
                    src = int(L.pb2_matrix_rank_of(mats[i].h, tm, tn))
                    t = tile(i, pos_of(i, tm, tn)).contiguous() if src == grid_rank else \
                        torch.empty(NB, NB, dtype=torch.float64, device=cuda)
                    if world > 1:
                        dist.broadcast(t, src=src)
                    ops.append(t)
                if owner == rank:
                    acc += ops[0] @ ops[1]
                    a_rows += ops[0].abs().sum(dim=1)
                    b_cols += ops[1].abs().sum(dim=0)
            if owner == rank:
                got = tile(2, pos_of(2, m, n))
                err = float((got - acc).abs().sum(dim=1).max())
                tile_norm = max(tile_norm, err / (float(a_rows.max()) * float(b_cols.max()) * N * eps))
                nb_tiles += 1
    versions_ok = all(mats[2].version(m, n) == T for n, tiles in cols[2].items() for m, _ in tiles)
    owned_c = sum(len(t) for t in cols[2].values())
    counts_ok = st["nb_gemm"] == owned_c * T
    res = torch.tensor([freivalds, tile_norm, 0.0 if (versions_ok and counts_ok and changed) else 1.0, float(nb_tiles)],
                       dtype=torch.float64, device=cuda)
    if world > 1:
        worst = res.clone()
        dist.all_reduce(worst, op=dist.ReduceOp.MAX)
        total = res.clone()
        dist.all_reduce(total, op=dist.ReduceOp.SUM)
        res = torch.stack([worst[0], worst[1], worst[2], total[3]])
    for m in mats:
        m.close()
    fr, tn, bad, nt = [float(x) for x in res.tolist()]
    return {"norm": max(fr, tn), "freivalds_norm": fr, "tile_norm": tn, "tiles": int(nt), "vectors": vectors,
            "versions_ok": bad == 0.0, "eps": "2^-24" if bf16 else "2^-53", "bound": 10.0,
            "how": "one extra step from fresh inputs: Freivalds with +-1 vectors and sampled C tiles recomputed with "
                   "torch fp64 (cuBLAS), scaled as ||.||inf / (||A||inf ||B||inf N eps); task counts and C versions == KT"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--order", dest="n", type=int, default=32768,
                    help="matrix order (BASELINE configs[1]: 32768); --order is the spelling to use under torchrun, whose "
                         "own parser claims every option that starts with --n")
    ap.add_argument("--nb", type=int, default=1024, help="tile size")
    ap.add_argument("--cpu-n", type=int, default=16384,
                    help="matrix order of the bounded CPU sample (16384: about 15 s of CPU work on the GPU box's 16 cores)")
    ap.add_argument("--window", type=int, default=0, help="C-tile chains in flight (0 = the taskpool default: 128, or 16 for tensor-core bodies)")
    ap.add_argument("--inflight", type=int, default=256)
    ap.add_argument("--streams", type=int, default=18, help="h2d + d2h + exec streams (the module spreads a task over as "
                    "many exec streams as its tile size needs: 5 at NB=1024, 16 at NB=512)")
    ap.add_argument("--batch", type=int, default=-1,
                    help="tasks per batched submission (one grouped tcgen05 launch per batch); -1 = 16 for --dtype bf16, "
                         "32 for f64 tiles of 512 and below, off for larger f64 tiles")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the extra checked step after the timed regions")
    ap.add_argument("--hbm-gib", type=float, default=0.0,
                    help="cap the HBM tile heap (GiB) to run out of core: the BASELINE configs[3] mechanism (LRU evict + "
                         "write-back) at a size the box's host memory allows; 0 = 95%% of free HBM")
    ap.add_argument("--strong", action="store_true",
                    help="N>1: keep --n fixed as GPUs are added (BASELINE configs[2] as written: --n 65536 on 2/4/8 GPUs) "
                         "instead of the default weak scaling N_g = n * g^(1/3)")
    ap.add_argument("--replicas", action="store_true",
                    help="diagnostic for N>1: independent 1x1 problems per GPU (isolates box effects from the data plane)")
    ap.add_argument("--dtype", default="f64", choices=["f64", "bf16"],
                    help="f64: the headline DGEMM; bf16: BF16-in/FP32-acc tiles on tcgen05 (BASELINE configs[4], use --nb 2048)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    # stdout carries exactly one JSON line.  Libraries write there too (NCCL prints its version banner with
    # NCCL_DEBUG=VERSION), so file descriptor 1 points at stderr until the line is printed.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    H = load_hostapi()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    NB = args.nb
    # weak scaling: flops per GPU constant, N_g = N_1 * g^(1/3) rounded to a whole number of tiles
    N = args.n if (world == 1 or args.strong) else int(round(args.n * world ** (1.0 / 3.0) / NB)) * NB
    P, Q = H.grid_for(world)
    if args.replicas:   # diagnostic: every rank runs its own 1x1 problem, no data-plane traffic at all
        N, P, Q = args.n, 1, 1
    grid_rank = 0 if args.replicas else rank
    bf16 = args.dtype == "bf16"
    # CTA budget of the NCCL communicators (comm_nccl.cu): an FP64 step needs ~10 GB/s of tile broadcasts per GPU and the
    # DMMA kernel wants every SM, so 2 CTAs; a bf16 step computes 25x faster on the same tiles and needs ~80 GB/s
    os.environ.setdefault("PB2_NCCL_MAX_CTAS", "8" if bf16 else "2")
    comm = H.Comm(local_rank, P, Q, rank) if (world > 1 and not args.replicas) else None
    mtypes = (H.MATRIX_BF16, H.MATRIX_BF16, H.MATRIX_FLOAT) if bf16 else (H.MATRIX_DOUBLE,) * 3
    if bf16:
        try:   # driver-written cuBLAS bf16 peaks; a kernel timed inside a seconds-long step compares to the sustained one
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak_tf, peak_src = float(peaks["bf16_tflops_sustained"]), "MEASURED_PEAKS.json bf16_tflops_sustained (of measured)"
        except Exception:
            peak_tf, peak_src = 1590.0, "B200_PROFILING.md fallback 1.59 PFLOP/s (of fallback)"
    else:
        # MEASURED_PEAKS.json has no FP64 entry.  profiles/fp64_peak_r2.json is the same measurement for FP64 taken on
        # this pool's B200 (parsec-1_b200/tools/fp64_peak.py: cuBLAS 8192^3, best of 10 and ~4 s sustained, clocks
        # recorded); the tile kernel is timed inside a seconds-long step, so it is compared with the sustained figure.
        # The live number of this very run is reported next to it.
        live_tf = fp64_peak_tflops(torch)
        try:
            pk = json.load(open(os.path.join(ROOT, "profiles", "fp64_peak_r2.json")))
            peak_tf = float(pk["fp64_tflops_sustained"])
            peak_src = ("profiles/fp64_peak_r2.json fp64_tflops_sustained (cuBLAS DGEMM 8192^3, ~4 s of back-to-back launches, "
                        "of measured); burst %.2f; live in this run %.2f" % (float(pk["fp64_tflops_burst"]), live_tf))
        except Exception:
            peak_tf = live_tf
            peak_src = "measured live: torch.matmul fp64 8192^3 (cuBLAS), best of 10 (profiles/fp64_peak_r2.json missing)"
    hbm = {"memory_number_of_blocks": int(args.hbm_gib * 2048)} if args.hbm_gib > 0 else {}   # 512 KiB blocks
    # tensor-core tile tasks share grouped launches; FP64 tiles of 512 and below are launch bound one by one (a 256^3
    # task is 4 CTAs), larger ones fill the GPU through stream concurrency
    batch = args.batch if args.batch >= 0 else (16 if bf16 else (32 if NB <= 512 else 0))
    if batch > 1:
        hbm["batch_bodies"] = batch
        if bf16 and not args.window:
            args.window = 3 * batch   # one batch on the GPU, one being staged, one retiring
    dev = H.Device(local_rank, max_inflight_tasks=args.inflight, max_streams=args.streams, profile_launches=0, **hbm)
    mats, bufs = [], []
    for seed, mtype in zip((1789, 1805, 1901), mtypes):
        m = H.Matrix(mtype, N, N, NB, NB, P, Q, grid_rank)
        b = torch.empty(max(m.nbytes, 8), dtype=torch.uint8, pin_memory=True)   # cudaHostAlloc: the collection's host memory
        m.set_buffer(b)
        m.fill_lcg(seed, dev)
        mats.append(m)
        bufs.append(b)
    A, B, Cm = mats
    flops = 2.0 * N ** 3
    knobs = {"window": args.window} if args.window else {}
    if comm is not None:
        knobs["comm"] = comm

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(nsteps, flush):
        """K steps between barrier+synchronize, device time from CUDA events; returns ms per step (max over ranks)."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(nsteps):
            if flush and H.lib().pb2_device_memory_release(dev.h) != H.SUCCESS:
                raise SystemExit("pb2_device_memory_release refused: the e2e steps would run from a warm cache")
            H.tiled_gemm(A, B, Cm, dev, **knobs)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / max(1, nsteps)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # warm-up: W steps (the first stages every tile into HBM; the rest run from the cache)
    for _ in range(max(3, args.warmup)):
        H.tiled_gemm(A, B, Cm, dev, **knobs)
    # only rank 0 polls nvidia-smi: every query takes driver locks, and eight pollers at 5 Hz measurably slowed the
    # submission path of all ranks (profiles/bench_r1_g8_notes.md)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev.stats(reset=True)
    ms_resident = timed(args.steps, flush=False)
    st_resident = dev.stats(reset=True)
    ms_e2e = timed(args.steps, flush=True)
    st_e2e = dev.stats(reset=True)
    clocks = sampler.stop()

    # roofline of the dominant kernel (dgemm_nn_tma): per-launch durations from CUDA events on the exec streams, taken
    # in one extra resident step so that the event records do not sit inside the timed region above
    dev2_params = dict(max_inflight_tasks=args.inflight, max_streams=args.streams, profile_launches=1, **hbm)
    prof = None
    if rank == 0:
        H.lib().pb2_device_memory_release(dev.h)
    barrier()
    # re-create the device with launch profiling on (same HBM budget)
    for m in mats:
        m.close()
    dev.close()
    dev = H.Device(local_rank, **dev2_params)
    mats = []
    for b, mtype in zip(bufs, mtypes):
        m = H.Matrix(mtype, N, N, NB, NB, P, Q, grid_rank)
        m.set_buffer(b)
        mats.append(m)
    A, B, Cm = mats
    H.tiled_gemm(A, B, Cm, dev, **knobs)          # stage everything
    dev.launch_ms()                                # drop the staging step's samples
    H.tiled_gemm(A, B, Cm, dev, **knobs)          # resident step, profiled
    lms, nstreams = dev.launch_ms()
    per_launch_flops = 2.0 * NB ** 3
    if len(lms):
        import numpy as np
        avg_ms = float(np.mean(lms))
        # `nstreams` launches are resident at a time (one per exec stream) and share the SMs
        achieved_tf = per_launch_flops * nstreams / (avg_ms * 1e-3) * 1e-12
        prof = {"avg_launch_ms": avg_ms, "samples": int(len(lms)), "resident_launches": nstreams}

    parity = None
    if not args.no_parity and not args.replicas:
        for m in mats:
            m.close()
        mats = []
        parity = parity_check(H, torch, dist if world > 1 else None, dev, bufs, mtypes, N, NB, P, Q, grid_rank, rank, world,
                              knobs, bf16)
    if args.replicas:
        flops *= world
    value = flops / (ms_resident * 1e-3) * 1e-9          # whole job: one N x N x N product spread over all ranks
    e2e = flops / (ms_e2e * 1e-3) * 1e-9
    tile_bytes = NB * NB * 8
    out = {
        "metric": "tiled GEMM GFLOP/s", "value": value, "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_resident, "higher_is_better": True,
        "scaling": "strong" if (args.strong and world > 1) else "weak",
        "vs_baseline": None, "dtype": "bf16 in / f32 accumulate" if bf16 else "f64", "data": "synthetic",
        "config": {"workload": ("mixed-precision tiled GEMM BF16-in/FP32-acc N=%d NB=%d via tcgen05 on %d B200 "
                                "(BASELINE configs[4] shape)" % (N, NB, world)) if bf16 else
                               ("tiled DGEMM N=%d NB=%d fp64 on 1 B200 via CUDA BODY + device stage-in/out "
                                "(BASELINE configs[1])" % (N, NB)) if world == 1 else
                               ("tiled DGEMM N=%d NB=%d fp64, 2D block-cyclic %dx%d over %d B200 with NCCL tile bcast "
                                "(BASELINE configs[2] shape; %s)" % (N, NB, P, Q, world,
                                                                    "strong scaling, N fixed" if args.strong else
                                                                    "weak scaling from N=%d" % args.n)),
                   "tasks_per_step": (N // NB) ** 3, "batch_bodies": batch, "window": args.window, "host_cores": os.cpu_count(), "replicas_diagnostic": bool(args.replicas),
                   "l2": "inputs (%.1f GiB) far larger than L2" % (N * N * (8 if bf16 else 24) / 2 ** 30),
                   "parallelism": "1 process per GPU, %dx%d grid; %s exec streams in use per GPU" % (
                       P, Q, prof["resident_launches"] if prof else "?"),
                   "pct_of_tensor_peak": 100.0 * value / world / (peak_tf * 1e3)},
        "e2e": {"value": e2e, "unit": "GFLOP/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": st_e2e["data_in_from_host"] // max(1, args.steps),
                "d2h_bytes_per_step": st_e2e["data_out_to_host"] // max(1, args.steps),
                "nvlink_bytes_in_per_step_rank0": st_e2e["data_in_from_peers"] // max(1, args.steps)},
        "gpu_launches": st_resident["kernel_launches"],
        "hbm_cache": {"heap_gib": st_resident["hbm_bytes_total"] / 2 ** 30,
                      "evictions_per_step": st_resident["nb_evictions"] // max(1, args.steps),
                      "writeback_tiles_per_step": st_resident["nb_w2r_tiles"] // max(1, args.steps),
                      "stalls_per_step": st_resident["nb_stalls"] // max(1, args.steps),
                      "h2d_bytes_per_resident_step": st_resident["data_in_from_host"] // max(1, args.steps)},
        "host_us_per_task": {"submit": st_resident["host_submit_ns"] / 1e3 / max(1, st_resident["kernel_launches"]),
                             "retire": st_resident["host_retire_ns"] / 1e3 / max(1, st_resident["kernel_launches"])},
        "clocks": clocks,
        "parity": parity,
        "roofline": {"bound": "tensor", "achieved": achieved_tf if prof else None, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": (achieved_tf / peak_tf) if prof else None,
                     # FP64 only: against the DMMA issue rate, 148 SMs x 128 flop per clock (one DMMA.8x8x4 per 16
                     # cycles per sub-partition), at the SM clock sampled during the timed region
                     "frac_of_dmma_issue_peak": (achieved_tf / (148 * 128 * clocks["sm_mhz"] * 1e-6))
                     if (prof and not bf16 and clocks and clocks.get("sm_mhz")) else None,
                     # dram bytes per launch from one ncu --set full capture (profiles/dgemm_tma_r1_ncu.md)
                     "traffic": 25206528 if (not bf16 and NB == 1024) else None,
                     "kernel": "pb2::gemm_umma_pair" if bf16 else "pb2::dgemm_nn_tma", "peak_source": peak_src,
                     "detail": prof, "algorithmic_flops_per_launch": per_launch_flops},
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:   # the CPU baseline is a single-GPU-run figure (N = 1 only)
        out["cpu_baseline"], _ = cpu_baseline(args.cpu_n, NB, os.cpu_count() or 1)
    for m in mats:
        m.close()
    dev.close()
    if rank == 0:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(out), flush=True)
        os.dup2(2, 1)
    if comm is not None:
        comm.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
