"""bench.py contract on CPU: the reference arm (the oracle port timed on host cores) prints exactly
ONE JSON line on stdout with the keys the driver reads; the B200 arm refuses to run without CUDA."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=600):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True,
                          text=True, timeout=timeout, cwd=ROOT)


def test_reference_arm_prints_one_json_line():
    r = _run("--impl", "reference", "--steps", "1", "--warmup", "0", "--batch", "4", "--size", "64")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["metric"] == "unet_train_images_per_sec_256x256"
    assert rec["unit"] == "images/s" and rec["higher_is_better"] is True and rec["value"] > 0
    assert rec["steps"] == 1 and rec["n_gpus"] == 1 and rec["dtype"] == "f32"
    assert rec["cpu_baseline"]["kind"] in ("port", "reference") and rec["cpu_baseline"]["cores"] >= 1
    assert rec["cpu_baseline"]["value"] == rec["value"]
    assert rec["e2e"] == {"value": rec["value"], "unit": "images/s", "h2d_bytes_per_step": 0,
                          "d2h_bytes_per_step": 0}
    assert "workload" in rec["config"]


def test_b200_arm_needs_cuda():
    import torch
    if torch.cuda.is_available():
        return  # covered by the -m gpu run and the driver's bench
    r = _run("--steps", "1", timeout=300)
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stderr + r.stdout)
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
