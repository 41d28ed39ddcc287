"""bench.py's reference arm (the CPU port on the host cores; no GPU code on that path) keeps the JSON contract the driver
reads: one line on stdout with metric / value / unit / config and the cpu_baseline / e2e objects."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_reference_arm_emits_one_contract_line():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--n", "20000", "--batch", "1000",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"].startswith("batched Search QPS")
    assert line["unit"] == "queries/s" and line["value"] > 0 and line["higher_is_better"] is True
    assert line["n_gpus"] == 1 and line["steps"] == 2 and line["warmup"] == 1 and line["gpu_launches"] == 0
    assert line["config"]["rows_per_gpu"] == 20000 and line["config"]["batch"] == 1000
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"]
    e = line["e2e"]
    assert e["value"] == line["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0
