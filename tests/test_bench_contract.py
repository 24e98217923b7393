"""The bench line recorded on the B200 (profiles/) carries every key the measurement contract names, and the constants
bench.py multiplies the exact counts with are the ones DESIGN.md / SURVEY.md 8d state.  CPU only: reads committed files."""
import glob
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _latest(pattern):
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", pattern)))
    assert files, pattern
    return json.load(open(files[-1]))


def test_recorded_bench_line_has_the_contract_keys():
    d = _latest("r02_bench_mixed10k_?.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "roofline", "cpu_baseline", "clocks"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["dtype"] == "f64" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("mixed-cell") and "model" not in d["config"]
    assert abs(d["value"] - 1000.0 / d["ms_per_step"]) <= 1e-9 * d["value"]
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0
    assert 0.0 < e["value"] <= d["value"] * 1.001          # host buffers in the timed region: never faster than resident
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert abs(r["frac"] - r["achieved"] / r["peak"]) <= 1e-12
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["unit"] == d["unit"] and c["sample"]
    # the CPU baseline is a measurement with its parts reported separately; the EEQ part says whether it was extrapolated
    p = c["parts"]
    assert p["lapack"] is True and isinstance(p["eeq_extrapolated"], bool)
    assert abs(p["pair_atm_s"] + p["cn_c6_s"] + p["eeq_s"] - p["total_s"]) <= 1e-9 * p["total_s"]
    assert abs(c["value"] - 1.0 / p["total_s"]) <= 1e-9 * c["value"]
    assert d["gpu_launches"] > 0
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


def test_recorded_reference_arm_line():
    d = _latest("r02_bench_reference_?.json")
    assert d["impl"] == "reference" and d["unit"] == "evals/s"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"]
    assert d["cpu_baseline"]["cores"] > 1   # the arm sets its thread count itself (torchrun exports OMP_NUM_THREADS=1)


def test_work_constants_match_the_survey():
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert bench.FLOP_ATM_TRIPLE == 252.0 and bench.FLOP_PAIR_IMAGE == 88.0
    assert bench.METRIC == "D4 energy+gradient+virial evaluations/sec" and bench.UNIT == "evals/s"
