"""The reference's OWN multi-rank tests, unmodified sources, over the new comm engine: every program its ctest files run
under mpiexec (tests/*/Testings.cmake, `MPI_TEST_CMD_LIST`), built by overlay/build_dist.sh against the distributed tree
(overlay/comm_engine_b200.c + libpb2mpi in place of the MPI engine and MPI) and started by overlay/pb2run with the
reference's rank counts and arguments.  CPU only, runs every round.

What they cover that the GEMM / Cholesky runs do not: DTD over several ranks (insertion, WAR, interleaved actions,
all-reduce, new tiles; two of them ask for MPI_THREAD_MULTIPLE, i.e. worker threads send through the engine themselves),
reshaping with triangular datatypes across ranks (the registrations that have no block-stride description and travel
through the packed bounce buffer), redistribution between two block-cyclic layouts on 8 ranks, band and stencil
collections, taskpool composition, and the ping-pong / bandwidth applications.  Pass criterion: the reference's own
(exit code 0, or the expected fatal message for the two "a NULL cannot be forwarded" tests, Testings.cmake:43-60)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
D = os.path.join(REF, "reftests", "dist")

CASES = [  # (ranks, program, arguments, text that must appear, text that must not)
    (2, "dtd_test_pingpong", [], "Time for each pingpong", None),
    (4, "dtd_test_task_inserting_task", [], None, None),
    (4, "dtd_test_task_insertion", [], None, None),
    (4, "dtd_test_war", [], None, None),
    (4, "dtd_test_interleave_actions", [], None, None),
    (4, "dtd_test_allreduce", [], None, None),
    (2, "dtd_test_new_tile", [], "which is the expected value", "failure"),
    (2, "jdf_forward_RW_NULL", [], "A NULL is forwarded", None),
    (2, "jdf_forward_READ_NULL", [], "A NULL is forwarded", None),
    (4, "write_check", [], "TEST SUCCESS", None),
    (4, "compose", [], None, None),
    (4, "project", ["-x"], None, None),
    (4, "merge_sort", [], None, None),
    (4, "reshape", ["-N", "120", "-t", "9", "-c", "10"], None, None),
    (4, "reshape", ["-N", "120", "-t", "9", "-c", "10", "-m", "1"], None, None),
    (4, "input_dep_reshape_single_copy", ["-N", "12", "-t", "2", "-c", "2"], None, None),
    (4, "remote_multiple_outs_same_pred_flow", ["-N", "12", "-t", "2", "-c", "2"], None, None),
    (8, "testing_redistribute", "-M 2400 -N 2400 -a 2400 -A 2400 -t 300 -T 300 -b 200 -B 200 -m 2000 -n 2000 -I 30 -J 40 -i 100 "
                                "-j 121 -v -z -x -P 2 -Q 4 -p 4 -q 2".split(), None, None),
    (8, "testing_redistribute_random", "-M 2400 -N 2400 -a 2400 -A 2400 -t 300 -T 300 -b 200 -B 200 -m 2000 -n 2000 -I 30 -J 40 "
                                       "-i 100 -j 121 -v -z -x -P 2 -Q 4 -p 4 -q 2".split(), None, None),
    (8, "testing_band", "-N 3200 -T 160 -P 4 -s 5 -S 10 -p 2 -f 2 -F 10 -b 2".split(), None, None),
    (8, "testing_stencil_1D", "-t 100 -T 100 -N 1000 -M 1000 -I 10 -R 2 -m 1".split(), "Stencil", None),
    (2, "bw_test", ["-n", "10", "-f", "10", "-l", "2097152"], None, None),
    (2, "schedmicro", ["-t", "10", "-l", "8", "-n", "512"], None, None),
    (2, "rtt", [], None, None),
]


@pytest.mark.parametrize("ranks,prog,args,must,must_not", CASES, ids=["%s-np%d%s" % (c[1], c[0], "-mt" if "-m" in c[2] and c[1] == "reshape" else "") for c in CASES])
def test_reference_multi_rank_test(ranks, prog, args, must, must_not):
    exe = os.path.join(D, prog)
    if not (os.path.exists(exe) and os.path.exists(os.path.join(REF, "dist", "libparsec.so.4"))):
        pytest.skip("baseline/_ref/reftests/dist is not built (overlay/build_dist.sh needs the reference tree)")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(REF, "dist"), os.path.join(ROOT, "parsec-1_b200"),
                                              env.get("LD_LIBRARY_PATH", "")])
    env["PARSEC_MCA_device_cuda_enabled"] = "0"
    p = subprocess.run([os.path.join(ROOT, "overlay", "pb2run"), "-n", str(ranks), "--timeout", "150", exe] + list(args),
                       env=env, capture_output=True, text=True, timeout=200)
    text = p.stdout + p.stderr
    if prog.startswith("jdf_forward_"):
        assert must in text, text[-2000:]                       # the reference's PASS_REGULAR_EXPRESSION
        return
    assert p.returncode == 0, (p.returncode, text[-2500:])
    assert "not in a loaded object" not in text and "SIGSEGV" not in text
    if must:
        assert must in text, text[-2000:]
    if must_not:
        assert must_not not in text, text[-2000:]
