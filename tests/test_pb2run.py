"""overlay/pb2run, the launcher that stands in for mpirun on one box: environment of the ranks, core sets, exit codes,
time limit, and that no shared-memory segment of the job is left behind.  No GPU, no reference tree needed."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PB2RUN = os.path.join(ROOT, "overlay", "pb2run")


def run(args, timeout=60):
    return subprocess.run([PB2RUN] + args, capture_output=True, text=True, timeout=timeout)


def segments():
    return {f for f in os.listdir("/dev/shm") if f.startswith("pb2ce.")}


def test_ranks_get_their_identity_and_their_own_cores():
    prog = ("import os; os.write(1, ('|'.join(str(x) for x in (os.environ['PB2_CE_RANK'], os.environ['PB2_CE_WORLD'], "
            "os.environ['PB2_CE_JOB'], sorted(os.sched_getaffinity(0)), os.environ.get('PARSEC_MCA_runtime_num_cores'), "
            "os.environ.get('PARSEC_MCA_runtime_bind_threads'), os.environ.get('PB2_CE_CORES'))) + '\\n').encode())")
    p = run(["-n", "2", sys.executable, "-c", prog])
    assert p.returncode == 0, p.stderr
    rows = sorted(l.split("|") for l in p.stdout.splitlines())
    assert [r[0] for r in rows] == ["0", "1"] and {r[1] for r in rows} == {"2"} and rows[0][2] == rows[1][2]
    ncpu = os.cpu_count()
    if ncpu >= 4:
        per = ncpu // 2
        assert eval(rows[0][3]) == list(range(0, per)) and eval(rows[1][3]) == list(range(per, 2 * per))   # disjoint halves
        assert rows[0][4] == str(per - 1) and rows[0][5] == "0"        # one core left for the communication thread
        assert rows[0][6] == "0-%d" % (per - 1) and rows[1][6] == "%d-%d" % (per, 2 * per - 1)
    q = run(["-n", "2", "--no-bind", sys.executable, "-c", "import os; os.write(1, ('%d %s\\n' % (len(os.sched_getaffinity(0)), os.environ.get('PB2_CE_CORES'))).encode())"])
    assert q.returncode == 0 and set(q.stdout.split()) == {str(len(os.sched_getaffinity(0))), "None"}


def test_gpu_per_rank_sets_the_visible_device():
    p = run(["-n", "3", "--gpu-per-rank", sys.executable, "-c", "import os; os.write(1, (os.environ['PB2_CE_RANK'] + ' ' + os.environ['CUDA_VISIBLE_DEVICES'] + '\\n').encode())"])
    assert p.returncode == 0 and sorted(p.stdout.splitlines()) == ["0 0", "1 1", "2 2"]


def test_a_failing_rank_fails_the_job_and_stops_the_others():
    before = segments()
    prog = "import os, sys, time; r = int(os.environ['PB2_CE_RANK']); sys.exit(7) if r == 1 else time.sleep(30)"
    p = run(["-n", "3", sys.executable, "-c", prog], timeout=25)
    assert p.returncode == 7 and "rank 1 exited with 7" in p.stderr
    assert segments() <= before


def test_time_limit():
    p = run(["-n", "2", "--timeout", "2", sys.executable, "-c", "import time; time.sleep(30)"], timeout=25)
    assert p.returncode != 0 and "time limit of 2 s reached" in p.stderr


def test_usage():
    p = run([])
    assert p.returncode == 2 and "usage" in p.stderr
