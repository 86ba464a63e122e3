"""CPU: host logic of bench.py that does not need a GPU."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpu_arm_uses_all_host_threads_under_torchrun_env():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the reference / cpu_baseline arm must still time the oracle on
    every host thread the process may use (bench.cpu_sample sets the OpenMP thread count itself)."""
    code = ("import sys; sys.argv=['bench.py']; import bench; w=bench.workload_cfg('tiny'); "
            "r=bench.cpu_sample(w, 2, 2.0, 0.05); print('CORES', r['cores'], r['value'] > 0)")
    env = dict(os.environ, OMP_NUM_THREADS="1", PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-800:]
    line = [l for l in out.stdout.splitlines() if l.startswith("CORES")][-1].split()
    assert int(line[1]) == len(os.sched_getaffinity(0)) and line[2] == "True"


def test_window_count_of_the_hot_path_frames_metric():
    # bench.py derives hot_path_frames_per_s from the reference's window schedule (mvtracker.py:522): 3 / 10 / 16 / 24 windows
    S = 12
    assert [len(range(0, max(T - S // 2, 1), S // 2)) for T in (24, 64, 100, 150)] == [3, 10, 16, 24]
