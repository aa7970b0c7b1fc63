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
    # the same config object as the GPU arm prints for these arguments (vs_reference.same_config)
    sys.path.insert(0, ROOT)
    import bench
    wl = bench.WORKLOADS["ftrl"]
    assert j["config"] == bench.config_of(wl, bench.MINI_BATCH, bench.ROWS_PER_STEP, 1, 1, 0, wl["total_rows"])
    assert j["config"]["rows_per_step"] == bench.ROWS_PER_STEP


def test_reference_arm_does_not_load_the_product():
    """The CPU arm's process maps oracle/ only: importing bench and building its oracle config must not
    import hivemall_b200 (synth.py is loaded by path)."""
    code = ("import sys; sys.argv=['bench.py']; import bench; sys.path.insert(0, 'tests'); import _oracle; "
            "bench.make_rows(bench.WORKLOADS['ftrl'], 8, 0); "
            "_oracle.config_from_options('train_classifier', bench.options_of(bench.WORKLOADS['ftrl'])); "
            "assert 'hivemall_b200' not in sys.modules, sorted(m for m in sys.modules if 'hivemall' in m); "
            "print(open('/proc/self/maps').read().count('libhivemall_b200'))")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip() == "0"


def test_workload_table():
    sys.path.insert(0, ROOT)
    import bench
    import hivemall_b200 as hm
    assert set(bench.WORKLOADS) == {"ftrl", "adagrad", "fm16", "adam26", "fm64", "pa1", "arow"}
    for name, wl in bench.WORKLOADS.items():
        cfg = hm.parse_options(wl["fn"], bench.options_of(wl))   # every option string parses
        assert cfg.dims == wl["dims"] and cfg.mini_batch == bench.MINI_BATCH, name
        assert hm.parse_options(wl["fn"], bench.options_of(wl, 128)).mini_batch == 128
        bpe = bench.bytes_per_example(wl)
        assert bpe["bytes"] == bpe["ingest"] + bpe["forward"] + bpe["update"]
        rows = bench.make_rows(wl, 64, 0)
        rp, f, v, y = rows.numpy()
        assert len(y) == 64 and rp[-1] == 64 * wl["nnz"] and f.max() < wl["dims"]
        if wl["gen"] == "kdd12":
            assert (v == 1.0).all()
    assert bench.bytes_per_example(bench.WORKLOADS["ftrl"])["bytes"] == 1412
