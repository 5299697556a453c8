"""bench.py's reference arm (CPU, no GPU needed): exactly one JSON line on stdout from rank 0, nothing from the other
ranks, the keys the contract names."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(env_extra, *args):
    env = dict(os.environ, **env_extra)
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tvl1",
                           "--steps", "1", "--warmup", "0", *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)


def test_reference_arm_prints_one_json_line():
    r = _run({"OMP_NUM_THREADS": "1"})              # what torchrun exports; the arm must still use every core
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "fields/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] == (os.cpu_count() or 1)
    assert d["e2e"] == {"value": d["value"], "unit": "fields/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("configs[1]")


def test_reference_arm_other_ranks_stay_silent():
    r = _run({"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2"}, "--gpus", "2")
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_clock_sampler_parses_nvidia_smi_lines():
    from frame2frame_b200.benchutil import ClockSampler

    class FakeProc:
        def terminate(self):
            pass

    s = ClockSampler.__new__(ClockSampler)
    s.proc = FakeProc()
    s.samples = ["1965, 1965, Not Active, Not Active, Not Active, Not Active",
                 "1530, 1965, Not Active, Not Active, Not Active, Active",
                 "1500, 1965, Not Active, Not Active, Not Active, Active",
                 "garbage", "[N/A], 1965, Not Active, Not Active, Not Active, Not Active"]
    c = s.stop()
    assert c == {"sm_mhz": 1530.0, "sm_max_mhz": 1965.0, "reasons": ["sw_power_cap"], "samples": 3}


def test_clock_sampler_without_nvidia_smi_says_so():
    import shutil
    import pytest
    from frame2frame_b200.benchutil import ClockSampler
    if shutil.which("nvidia-smi"):
        pytest.skip("nvidia-smi present")
    assert ClockSampler(0).stop()["reasons"] == ["nvidia-smi unavailable"]


def test_peaks_fall_back_when_the_driver_file_is_absent(tmp_path, monkeypatch):
    from frame2frame_b200 import benchutil
    monkeypatch.setattr(benchutil, "ROOT", str(tmp_path))
    pk, kind = benchutil.peaks()
    assert kind == "fallback" and pk["hbm_gbs"] > 0 and pk["bf16_tflops_sustained"] <= pk["bf16_tflops"]
    (tmp_path / "MEASURED_PEAKS.json").write_text('{"hbm_gbs": 6551.7, "bf16_tflops": 1630.0, "bf16_tflops_sustained": 1385.1}')
    pk, kind = benchutil.peaks()
    assert kind == "measured" and pk["hbm_gbs"] == 6551.7


def test_f2f_reference_arm_reports_the_measured_sample():
    """The headline workload's reference arm: `ms_per_step` is the time of the bounded sample it really ran (so that
    steps x ms_per_step fits the run), `value` the frames/s extrapolated to iter = 20 with the extrapolation spelled out,
    and `config` is the very dict our own arm prints (the driver compares them)."""
    import time
    import bench_f2f
    t0 = time.time()
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--no-config0"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    wall = time.time() - t0
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["metric"] == bench_f2f.METRIC
    assert d["config"] == bench_f2f.workload_config()
    assert d["steps"] * d["ms_per_step"] * 1e-3 <= wall                      # the sample really is what was timed
    ex = d["extrapolation"]
    assert abs(ex["frame_ms"] - 1e3 * (ex["iter"] * ex["t_iteration_s"] + ex["t_eval_s"] + ex["t_flow_s"])) < 1.0
    assert abs(d["value"] - 1e3 / ex["frame_ms"]) < 1e-9 and d["ms_per_step"] < ex["frame_ms"]
    assert d["e2e"]["value"] == d["value"] and d["cpu_baseline"]["cores"] == (os.cpu_count() or 1)
