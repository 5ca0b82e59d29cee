"""bench.py's line contract, checked on the CPU through the reference arm (one bounded step of the oracle port) and the
helpers the GPU arm uses: identical unit / config in both arms, required keys, profile-traffic reader, work figures."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_reference_arm_line():
    import bench
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == bench.UNIT
    assert d["config"] == bench.bench_config(1)                    # the dict the GPU arm prints as well
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["dtype"] == "f32" and d["n_gpus"] == 1
    assert d["value"] > 0 and d["e2e"] == dict(value=d["value"], unit=bench.UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0)
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "sample" in cb


def test_weak_scaled_config_and_unit_note():
    import bench
    c1, c8 = bench.bench_config(1), bench.bench_config(8)
    assert c1["particles_total"] == 20 and c8["particles_total"] == 160 and c1["particles_per_gpu"] == c8["particles_per_gpu"] == 20
    assert set(c1) == set(c8) and "workload" in c1 and "l2" in c1


def test_traffic_is_read_from_the_committed_profiles():
    import bench
    t = bench.traffic_from_profile("profiles/r02_ncu_c5_pair_gemms.txt", "gemm3xtf32_pair_kernel")
    assert t is not None and 5e10 < t < 2e11                       # two GEMM launches, tens of GB each
    assert bench.traffic_from_profile("profiles/r02_ncu_c5_pair_gemms.txt", "no_such_kernel") is None
    assert bench.traffic_from_profile("profiles/does_not_exist.txt", "x") is None


def test_work_figures():
    import bench
    pk = dict(hbm=6500.0, bf16=1600.0, bf16_sus=1360.0, src="test")
    p = bench.stein_point(16384, 10000, 38.0, 1, pk)
    assert p["flop_per_step"] == 4.0 * 16384 * 16384 * 10000       # W = 4 m^2 d
    r = p["roofline"]
    assert abs(r["peak"] - 1100.0 * 1600.0 / 2250.0 / 3.0) < 1e-9 and r["frac"] == r["achieved"] / r["peak"]
    # executed MMA work: 3 MMAs per term, symmetric Gram computes the 256 x 256 tiles touching j >= i
    T = 64
    want = 3.0 * (T * (T + 1) // 2 * 256 * 256 * 2.0 * 10016 + T * 40 * 256 * 256 * 2.0 * 16384)
    assert bench.executed_mma_flop(16384, 10000) == want
