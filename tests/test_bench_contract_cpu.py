"""The parts of the bench contract that run without a GPU: the reference arm answers with one JSON line and exit
code 0 (the reference is Lua / Torch7 over MPI and cannot be installed in this image), and the clock summary keeps
only the samples of the right GPU that fall inside the timed region and names the throttle reasons it saw."""
import datetime
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_module", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_reference_arm_prints_one_json_line_and_exits_zero():
    for n in (1, 8):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", str(n),
                            "--steps", "5", "--warmup", "3"], capture_output=True, text=True, timeout=120, cwd=ROOT)
        assert r.returncode == 0, r.stderr[-2000:]
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        assert len(lines) == 1
        d = json.loads(lines[0])
        assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and "\n" not in d["unavailable"]


def test_clock_summary_uses_the_samples_of_the_timed_region(tmp_path):
    b = _bench()
    t0 = datetime.datetime(2026, 1, 1, 12, 0, 0)
    stamp = lambda ms: (t0 + datetime.timedelta(milliseconds=ms)).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
    rows = [
        f"0, {stamp(0)}, 300, 1965, 150.0, 0x0, Not Active, Not Active, Not Active, Not Active",      # idle, before
        f"0, {stamp(100)}, 1965, 1965, 700.0, 0x4, Not Active, Not Active, Not Active, Active",       # inside, power cap
        f"1, {stamp(110)}, 1200, 1965, 400.0, 0x0, Not Active, Active, Not Active, Not Active",       # another GPU
        f"0, {stamp(120)}, 1950, 1965, 710.0, 0x0, Not Active, Not Active, Not Active, Not Active",   # inside
        f"0, {stamp(140)}, 1935, 1965, 705.0, 0x0, Not Active, Not Active, Not Active, Not Active",   # inside
        f"0, {stamp(900)}, 300, 1965, 150.0, 0x0, Active, Not Active, Not Active, Not Active",        # after
        "garbage line",
    ]
    p = tmp_path / "clocks.csv"
    p.write_text("\n".join(rows) + "\n")
    lo, hi = t0 + datetime.timedelta(milliseconds=50), t0 + datetime.timedelta(milliseconds=500)
    s = b._clock_summary(str(p), 0, lo, hi)
    assert s["samples_inside_timed_region"] == 3 and s["samples"] == 3
    assert s["sm_mhz"] == 1950 and s["sm_max_mhz"] == 1965
    assert s["reasons"] == ["sw_power_cap"]                                   # hw_slowdown was outside, GPU 1 ignored
    # a region too short to catch a sample falls back to everything recorded for that GPU
    s2 = b._clock_summary(str(p), 0, t0 + datetime.timedelta(milliseconds=60), t0 + datetime.timedelta(milliseconds=61))
    assert s2["samples_inside_timed_region"] == 0 and s2["samples"] == 5 and "hw_slowdown" in s2["reasons"]
    assert b._clock_summary(str(tmp_path / "missing.csv"), 0)["sm_mhz"] is None
