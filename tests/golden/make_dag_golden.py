"""Records what the REFERENCE RUNTIME does for the tiled GEMM DAG, with CPU task bodies, in this container.

Runs baseline/_ref/bin/dgemm_NN_app (overlay/apps/dgemm_NN.jdf through the reference's own parsec-ptgpp, linked with
the unmodified libparsec.so built from /root/reference by overlay/build_overlay.sh) with --mca device_cuda_enabled 0
and writes tests/golden/dag_reference_runtime.json: per-class task counts, the final version of the host
parsec_data_copy_t of every A, B and C tile, and the SHA-256 of C.  The oracle must reproduce all three
(tests/test_oracle.py::test_oracle_matches_the_reference_runtime_golden): this is what pins the oracle's task counts,
data-version results and tile-GEMM arithmetic on the reference itself instead of on remembered numbers.

    python tests/golden/make_dag_golden.py        # needs baseline/_ref (python -c 'import __graft_entry__ as g; g.build()')
"""
import hashlib
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "baseline", "_ref")

CASES = [(2048, 2048, 2048, 256),      # BASELINE configs[0]
         (1000, 1000, 1000, 192),      # ragged edge tiles
         (768, 512, 1280, 128),        # rectangular: MT, NT, KT all different
         (256, 256, 256, 256)]         # one tile


def run(M, N, K, NB, runs=1):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(REF, "stock"))
    with tempfile.TemporaryDirectory() as tmp:
        dump, ver = os.path.join(tmp, "C.bin"), os.path.join(tmp, "v.json")
        out = subprocess.run([os.path.join(REF, "bin", "dgemm_NN_app"), "-M", str(M), "-n", str(N), "-K", str(K), "-t", str(NB),
                              "-r", str(runs), "--dump", dump, "--versions", ver, "--", "--mca", "device_cuda_enabled", "0"],
                             env=env, capture_output=True, text=True, check=True).stdout
        line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
        versions = json.load(open(ver))
        sha = hashlib.sha256(open(dump, "rb").read()).hexdigest()
    return {"M": M, "N": N, "K": K, "NB": NB, "runs": runs,
            "counts": {"read_a": line["read_a"], "read_b": line["read_b"], "gemm": line["gemm_cpu"]},
            "versions": versions, "c_sha256": sha}


if __name__ == "__main__":
    golden = [run(*c) for c in CASES] + [run(1024, 1024, 1024, 256, runs=2)]
    with open(os.path.join(HERE, "dag_reference_runtime.json"), "w") as f:
        json.dump(golden, f, indent=0)
    for g in golden:
        print(g["M"], g["N"], g["K"], g["NB"], g["runs"], g["counts"], sorted(set(g["versions"]["C"])), g["c_sha256"][:16])
