"""CPU test of bench.py's reference arm (`--impl reference`): it runs the CPU port of the reference path
(oracle/, the one place outside tests/ and smoke() that may execute it) and must print ONE JSON line that
carries the driver's contract keys with this project's metric, unit and workload."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--edges", "60000", "--ref-sample-edges", "20000", "--numba-sample-edges", "2000"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "tppr_streaming_edges_per_sec" and d["unit"] == "edges/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["dtype"] == "f64"
    assert d["gpu_launches"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    nb = cb["reference_numba"]          # the reference's own numba finder, timed beside the port
    assert ("unavailable" in nb) or (nb["value"] > 0 and nb["kind"] == "reference" and "1 of" in nb["cores"])
    if os.path.exists(os.path.join(ROOT, "baseline", "_ref", "utils", "util.py")):
        assert "unavailable" not in nb, nb
        assert nb["value"] < d["value"], "the C port is faster than the numba original it restates"
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"] and "model" not in d["config"]


def test_our_arm_refuses_to_run_without_a_gpu():
    """The product arm has no CPU fallback: on a box without CUDA it must fail loudly, not print a number."""
    import torch
    if torch.cuda.is_available():
        return
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1", "--edges", "2000"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode != 0
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{") and '"value"' in ln]
