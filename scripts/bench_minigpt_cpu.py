"""BASELINE config 1 on the CPU: the reference's llm-demo/minigpt/train.py (run unmodified, from a scratch directory) next
to `python -m llm_in_practise_b200.recipes.minigpt train` — wall time for the 200 epochs and the loss curve of both.
  python scripts/bench_minigpt_cpu.py [--epochs 200]"""
import argparse, json, os, re, subprocess, sys, time

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=200)
ap.add_argument("--ref-dir", default="/root/reference/llm-demo/minigpt")
a = ap.parse_args()
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
scratch = os.path.join(root, "gpurun_out", "minigpt_cpu")
os.makedirs(scratch, exist_ok=True)
env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=a.ref_dir + os.pathsep + root)


def run(cmd):
    t0 = time.perf_counter()
    out = subprocess.run(cmd, cwd=scratch, env=env, capture_output=True, text=True)
    dt = time.perf_counter() - t0
    losses = [float(x) for x in re.findall(r"Loss: ([0-9.]+)", out.stdout)]
    return dt, losses, out

res = {}
if a.epochs == 200 and os.path.exists(os.path.join(a.ref_dir, "train.py")):
    dt, losses, out = run([sys.executable, os.path.join(a.ref_dir, "train.py")])
    res["reference"] = {"wall_s": round(dt, 1), "s_per_epoch": round(dt / max(len(losses), 1), 3), "loss_epoch_1": losses[0],
                        "loss_epoch_50": losses[49], "loss_final": losses[-1], "rc": out.returncode}
dt, losses, out = run([sys.executable, "-m", "llm_in_practise_b200.recipes.minigpt", "train", "--epochs", str(a.epochs), "--device", "cpu"])
res["b200_framework_cpu_path"] = {"wall_s": round(dt, 1), "s_per_epoch": round(dt / max(len(losses), 1), 3), "loss_epoch_1": losses[0] if losses else None,
                                  "loss_epoch_50": losses[49] if len(losses) > 49 else None, "loss_final": losses[-1] if losses else None,
                                  "rc": out.returncode, "tail": out.stdout[-200:] if not losses else ""}
print(json.dumps(res, indent=1))
