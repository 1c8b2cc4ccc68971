"""bench.py on the CPU: what can be checked without a GPU -- the reference arm's JSON line (the driver's contract for
`--impl reference`) on a tiny frame, the clock sampler's windowing to the timed region, and the argument surface."""
import json
import os
import subprocess
import sys
import time

from conftest import ROOT

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_clock_sampler_keeps_the_samples_of_the_timed_region():
    cs = bench.ClockSampler(0)
    now = time.monotonic()
    row = lambda mhz, cap, t: [str(mhz), "1965", "900", "Not Active", "Not Active", "Not Active", "Active" if cap else "Not Active", t]  # noqa: E731
    cs.rows = [row(1965, False, now - 1.0),          # warm-up
               row(1900, True, now + 0.05), row(1950, False, now + 0.15),
               ["garbage"], row(1200, False, now + 2.0)]     # after the region
    cs.t_begin, cs.t_end = now, now + 0.2
    r = cs.stop()
    assert r == {"sm_mhz": 1950, "sm_max_mhz": 1965, "reasons": ["sw_power_cap"], "samples": 2}
    # no sample landed inside the window: fall back to whatever came after its start instead of reporting nothing
    cs2 = bench.ClockSampler(0)
    cs2.rows = [row(1965, False, now - 1.0), row(1800, False, now + 5.0)]
    cs2.t_begin, cs2.t_end = now, now + 0.2
    assert cs2.stop()["sm_mhz"] == 1800
    # nvidia-smi absent (this container): keys still present
    assert bench.ClockSampler(0).stop() == {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}


def test_reference_arm_line_has_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--width", "96", "--height", "54"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "Msamples/s" and line["unit"] == "Msamples/s"
    assert line["higher_is_better"] is True and line["vs_baseline"] is None and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert "w5 mario" in line["config"]["workload"]


def test_reference_arm_runs_on_rank_zero_only():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_shard_modes_on_the_command_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--shard", "nonsense"], capture_output=True, text=True)
    assert r.returncode != 0 and "tiles_p2p" in r.stderr


def test_ab_identity_compare_reports_differences(tmp_path):
    """tools/ab_identity.py compare (the gate for adopting an experiment build): identical only if every array is byte-equal."""
    import numpy as np
    a, b = str(tmp_path / "a.npz"), str(tmp_path / "b.npz")
    x = np.arange(12, dtype=np.float32).reshape(3, 4)
    np.savez(a, acc=x, hits=np.zeros(3, np.uint32))
    y = x.copy(); y[1, 2] = np.nextafter(y[1, 2], np.float32(100))      # one ulp
    np.savez(b, acc=y, hits=np.zeros(3, np.uint32))
    for p in (a, b):
        json.dump({"w5": {"ms": [1.0, 1.0], "gsamples_per_s": 10.0 if p == a else 11.0}}, open(p + ".times.json", "w"))
    tool = os.path.join(ROOT, "tools", "ab_identity.py")
    r = subprocess.run([sys.executable, tool, "compare", a, a], capture_output=True, text=True)
    assert r.returncode == 0 and json.loads(r.stdout)["identical"] is True
    r = subprocess.run([sys.executable, tool, "compare", a, b], capture_output=True, text=True)
    out = json.loads(r.stdout)
    assert r.returncode == 1 and out["identical"] is False and out["differing"] == ["acc"]
    assert abs(out["speedup"]["w5"] - 1.1) < 1e-9
