"""CPU: the bench.py contract the driver relies on -- the reference arm's JSON line, and that the product arm
refuses to run without a CUDA device (no CPU fallback)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                          timeout=600, cwd=ROOT, env=e)


def test_reference_arm_prints_one_contract_line():
    out = _run("--impl", "reference", "--steps", "1", "--warmup", "1", "--ref-frames", "1")
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines                                  # stdout carries the JSON line and nothing else
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["higher_is_better"] is True
    assert d["metric"] == "frames/s 5-cam Panoptic jln64" and d["dtype"] == "f32" and d["vs_baseline"] is None
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 1 and d["gpu_launches"] == 0
    cb = d["cpu_baseline"]
    # "reference" = the reference's own modules (oracle/_ref, made by oracle/make_ref.py where /root/reference exists),
    # "port" = the bit-identical torch restatement when they are not there
    have_ref = os.path.isfile(os.path.join(ROOT, "oracle", "_ref", "lib", "models", "human_detection_net.py")) \
        or os.path.isdir("/root/reference/lib")
    assert cb["kind"] == ("reference" if have_ref else "port")
    assert cb["cores"] >= 1 and cb["value"] == d["value"] and "frame" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_port_variant():
    out = _run("--impl", "reference", "--steps", "1", "--warmup", "1", "--ref-frames", "1", "--ref-port")
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads([l for l in out.stdout.splitlines() if l.strip()][0])
    assert d["cpu_baseline"]["kind"] == "port" and d["value"] > 0


def test_reference_arm_other_ranks_exit_quietly():
    out = _run("--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1", env={"RANK": "1", "WORLD_SIZE": "2"})
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_product_arm_needs_cuda():
    import torch
    if torch.cuda.is_available():
        return
    out = _run("--steps", "1", "--warmup", "1", "--no-e2e", "--no-cpu-baseline")
    assert out.returncode != 0 and out.stdout.strip() == ""
    assert "no CPU fallback" in out.stderr
