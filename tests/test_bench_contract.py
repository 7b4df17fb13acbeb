"""The bench lines committed under profiles/ carry every key of the measurement contract (bench.py's docstring / DESIGN.md
section 8): a guard against the JSON line losing a field the driver or the reviewer reads.  No GPU, no bench run."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        return json.loads(f.read())


@pytest.mark.parametrize("name,n", [("r02n_bench_n1.json", 1), ("r02g_bench_n2.json", 2), ("r02g_bench_n8.json", 8)])
def test_product_arm_line(name, n):
    d = _line(name)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in d, k
    assert d["n_gpus"] == n and d["metric"] == "Mpaths/s" and d["unit"] == "Mpaths/s" and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["dtype"] == "f32" and d["warmup"] >= 3
    assert "workload" in d["config"] and "configs[1]" in d["config"]["workload"] and "image_texture" in d["config"]
    assert "model" not in d["config"]
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["d2h_bytes_per_step"] >= 1920 * 1080 * 12 and e["h2d_bytes_per_step"] > 0
    assert 0.9 * d["value"] < e["value"] <= 1.01 * d["value"]          # through the host-buffer entry point, not a copy of `value`
    assert e["value"] != d["value"]
    assert d["gpu_launches"] > 0
    assert d["clocks"]["sm_mhz"] > 0 and not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "kernel", "kernel_ms", "units_per_launch", "ops_per_unit"):
        assert k in r, k
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    # achieved = counted work per launch / the kernel's measured launch duration
    assert abs(r["achieved"] - r["units_per_launch"] * r["ops_per_unit"] / (r["kernel_ms"] * 1e-3) / 1e12) < 1e-6 * r["achieved"]
    assert d["value"] * d["ms_per_step"] * 1e3 == pytest.approx(1920 * 1080 * 1024, rel=1e-6)      # whole-job paths / step time
    if n == 1:
        c = d["cpu_baseline"]
        for k in ("value", "unit", "cores", "kind", "sample"):
            assert k in c, k
        assert c["kind"] == "reference" and c["cores"] >= 1 and c["unit"] == d["unit"]
        assert r["traffic"] is not None and r["traffic"] > 0
        b = d["bvh_hit"]
        for cls in ("coherent", "coherent_near", "incoherent"):
            assert b[f"{cls}_fast_mrays_per_s"] > 0 and b[f"{cls}_fast_e2e_mrays_per_s"] > 0
            assert b["cpu_reference"][f"{cls}_gpu_matches_reference"] is True
        assert set(d["other_configs"]) == {"configs[2]_textured", "configs[3]_cornell_smoke"}


def test_reference_arm_line():
    d = _line("r02g_bench_reference.json")
    assert d["impl"] == "reference" and d["metric"] == "Mpaths/s" and d["higher_is_better"] is True
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    c = d["cpu_baseline"]
    assert c["kind"] == "reference" and c["value"] == d["value"] and c["cores"] >= 1 and "sample" in c
    assert "configs[1]" in d["config"]["workload"]
