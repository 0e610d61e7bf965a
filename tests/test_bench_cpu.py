"""bench.py's CPU legs run on tiny samples: the baselines of both metrics are part of the default bench line, and a crash in
one of them takes the whole line down (it did once: the rollout baseline passed the oracle a sub-tree of the parameters)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_cpu_baseline_of_the_rollout_metric_runs():
    import bench
    v, threads, steps = bench.cpu_rollout_env_steps_per_sec(42, 8, 16, 5, term_bias=-50.0)
    assert v > 0 and threads >= 1 and steps == 5   # bias -50: nothing terminates, every lockstep step runs


def test_cpu_baseline_of_the_update_metric_runs():
    import bench
    ups, threads = bench.cpu_oracle_updates_per_sec(28, 5, 32, 2, warm=1)[:2]
    assert ups > 0 and threads >= 1


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--batch", "32"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "fql_member_updates_per_sec" and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "port"
