"""bench.py's reference arm (CPU oracle, no GPU needed) prints one JSON line with the keys the driver reads;
the workload table is self-consistent."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["unit"] == "examples/s" and j["higher_is_better"] is True
    assert j["value"] > 0 and j["ms_per_step"] > 0 and j["steps"] == 1 and j["warmup"] == 0 and j["n_gpus"] == 1
    assert j["scaling"] == "weak" and j["vs_baseline"] is None and j["data"] == "synthetic"
    assert "configs[1]" in j["config"]["workload"]
    cb = j["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == j["value"] and cb["sample"]
    e = j["e2e"]
    assert e["value"] == j["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert j["gpu_launches"] == 0


def test_workload_table():
    sys.path.insert(0, ROOT)
    import bench
    import hivemall_b200 as hm
    assert set(bench.WORKLOADS) == {"ftrl", "adagrad", "fm16", "adam26", "fm64"}
    for name, wl in bench.WORKLOADS.items():
        cfg = hm.parse_options(wl["fn"], wl["options"])          # every option string parses
        assert cfg.dims == wl["dims"] and cfg.mini_batch == bench.MINI_BATCH, name
        rows = bench.make_rows(wl, 64, 0)
        rp, f, v, y = rows.numpy()
        assert len(y) == 64 and rp[-1] == 64 * wl["nnz"] and f.max() < wl["dims"]
        if wl["gen"] == "kdd12":
            assert (v == 1.0).all()
