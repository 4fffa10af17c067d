"""CPU: bench.py's host-side pieces — thread choice for kernel19, the --impl reference line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_pick_threads_gives_even_bands():
    for M in (2048, 4096, 8192, 1000, 6):
        for cores in (1, 3, 7, 8, 16, 96):
            n = bench.pick_threads(M, cores)
            assert 1 <= n <= cores and (M % (2 * n) == 0 or n == 1)
    assert bench.pick_threads(8192, 16) == 16 and bench.pick_threads(8192, 7) == 4


def test_reference_arm_prints_contract_line():
    env = dict(os.environ, BENCH_REF_SAMPLE="512")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, env=env, check=True).stdout
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "TFLOP/s" and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["steps"] == 2 and line["higher_is_better"] is True


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", BENCH_REF_SAMPLE="512")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
