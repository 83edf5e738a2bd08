"""bench.py contract pieces that need no GPU: the reference arm answers with one JSON line and exit code 0, and the
argument surface the driver uses exists."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _json_line(out):
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout[-500:] + out.stderr[-1000:]
    return json.loads(lines[0])


def test_reference_arm_without_gpu_says_so():
    """No CUDA device: one JSON line, exit code 0, ``unavailable`` names the reason."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-1000:]
    j = _json_line(out)
    assert j["impl"] == "reference" and isinstance(j["unavailable"], str) and j["unavailable"]


def test_reference_arm_runs_the_torch_only_restatement():
    """The restatement (HF Qwen3 + LoRA modules + torch AdamW) steps on a tiny shape on the CPU and prints the bench line
    with the labels the judge asked for; it must not import this repository's package."""
    env = dict(os.environ, REF_DEVICE="cpu")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--model", "qwen3-tiny",
                          "--steps", "1", "--warmup", "1", "--seq-len", "32"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-1000:]
    j = _json_line(out)
    assert j["impl"] == "reference" and j["reference_class"] == "torch-only restatement" and j["same_config"] is False
    assert j["value"] > 0 and j["e2e"]["value"] > 0 and j["e2e"]["h2d_bytes_per_step"] == 2 * 32 * 8 * 2
    assert j["metric"] == "Qwen3-14B QLoRA ZeRO-3 train tok/s"
    src = open(os.path.join(ROOT, "baseline", "ref_headline.py")).read()
    assert "import llm_in_practise_b200" not in src and "from llm_in_practise_b200" not in src


def test_bench_help_lists_the_driver_flags():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in out.stdout
