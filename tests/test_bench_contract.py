"""The JSON line bench.py printed on a B200 (committed under profiles/) carries every key of the driver's contract, and the
roofline / e2e / cpu_baseline objects are internally consistent (frac = achieved / peak, the launch count is positive, the
end-to-end figure really moved bytes both ways)."""
import json
import os

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    txt = open(os.path.join(REPO, "profiles", name)).read().strip().splitlines()[-1]
    return json.loads(txt)


def test_single_gpu_line_has_the_contract_keys():
    d = _line("r2_bench_go2_n1_v4.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["data"] == "synthetic" and "workload" in d["config"] and "model" not in d["config"]
    assert d["gpu_launches"] > 0 and d["value"] > 0 and d["warmup"] >= 3
    r = d["roofline"]
    assert r["bound"] in ("hbm", "tensor") and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["traffic"] > 0
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] <= d["value"] * 1.001
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert set(d["configs"]) == {"h1_4096", "g1_4096", "h1_2_16384"}


def test_reference_arm_line():
    d = _line("r2_bench_go2_reference_arm.json")
    assert d["impl"] == "reference" and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["e2e"]["value"] == d["value"]


def test_multi_gpu_lines_are_whole_job_aggregates():
    one = _line("r2_bench_go2_n1_v4.json")["value"]
    for name, n in (("r2_bench_go2_n2_peer_fused.json", 2), ("r2_bench_go2_n8_nccl.json", 8)):
        d = _line(name)
        assert d["n_gpus"] == n and d["scaling"] == "weak"
        assert 0.9 * n * one < d["value"] < 1.1 * n * one          # whole-job env-steps/s, not per GPU
        assert "gradient_exchange" in d["run"]
