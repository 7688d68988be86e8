"""`bench.py --impl reference` must always print its JSON line, quickly, whatever the host looks like: it timed out on the
driver's box in every run of round 1. Runs the arm with the driver's arguments pinned to ONE core with OMP_NUM_THREADS=1 in the
environment (the worst host it has to survive) and with an impossible budget (the line must still come, from a `finally`)."""
import json
import os
import subprocess
import sys
import time

import oracles as O

BENCH = os.path.join(O.ROOT, "bench.py")


def _run(extra_env, prefix=(), args=("--steps", "20", "--warmup", "5"), timeout=120):
    env = dict(os.environ)
    env.update(extra_env)
    t0 = time.time()
    p = subprocess.run(list(prefix) + [sys.executable, BENCH, "--impl", "reference", "--gpus", "1"] + list(args),
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout, env=env)
    dt = time.time() - t0
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, p.stdout  # exactly one JSON line on stdout
    return json.loads(lines[0]), p.stderr, dt


def test_reference_arm_finishes_pinned_to_one_core():
    taskset = ["taskset", "-c", str(sorted(os.sched_getaffinity(0))[0])]
    line, err, dt = _run({"OMP_NUM_THREADS": "1"}, prefix=taskset)
    assert dt < 120
    assert line["impl"] == "reference" and line["unit"] == "Mrays/s" and line["higher_is_better"] is True
    assert line["steps"] >= 1 and line["value"] > 0.5
    assert line["cpu_baseline"]["cores"] == 1          # thread count from the affinity mask, not os.cpu_count()
    assert line["cpu_baseline"]["kind"] in ("reference", "port")
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    assert "1 of the 256 samples per pixel" in line["cpu_baseline"]["sample"]
    assert "step 0:" in err                            # one progress line per step on stderr


def test_reference_arm_emits_its_line_even_when_nothing_completes():
    line, err, dt = _run({"DRT_REF_BUDGET_S": "0.05"}, args=("--steps", "3", "--warmup", "1"))
    assert dt < 60
    assert line["impl"] == "reference" and line["steps"] == 0 and line["value"] == 0.0
    assert "budget" in line["cpu_baseline"]["sample"]
    assert len(line["attempts"]) >= 1


def test_other_ranks_exit_without_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1"],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""
