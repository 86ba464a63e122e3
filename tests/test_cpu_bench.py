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


def test_reference_arm_runs_the_references_own_code():
    """bench.py --impl reference times the reference's OWN init_pointcloud_from_rgbd / corr_sample (-> _knn_torch) when
    /root/reference or the staged oracle/_ref copy is present (kind "reference"), the C port otherwise (kind "port")."""
    import json
    sys.path.insert(0, ROOT)
    from oracle import ref_import
    env = dict(os.environ, OMP_NUM_THREADS="1", PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--workload", "kubric", "--steps", "1", "--warmup", "0",
                          "--cpu-budget", "2"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-800:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "queries/s"
    assert line["cpu_baseline"]["kind"] == ("reference" if ref_import.available() else "port")
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert line["config"]["parallelism"].startswith("CPU")
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    # the step this arm reports is the sample it really timed
    assert line["ms_per_step"] >= 1e3 * line["cpu_baseline"]["sample_seconds"] * 0.5


def test_staged_reference_is_unmodified():
    """oracle/_ref (what travels to the GPU box) is byte-identical to the files make_ref.py copied from the reference."""
    sys.path.insert(0, ROOT)
    from oracle import make_ref
    if not os.path.isfile(os.path.join(make_ref.DEST, "MANIFEST.json")):
        import pytest
        pytest.skip("oracle/_ref not staged")
    assert make_ref.verify()
    if os.path.isdir(os.path.join(make_ref.REF_ROOT, "mvtracker")):
        import filecmp
        for rel in make_ref.FILES:
            assert filecmp.cmp(os.path.join(make_ref.REF_ROOT, rel), os.path.join(make_ref.DEST, rel), shallow=False), rel
