"""bench.py's output contract on the CPU arm (no GPU needed): `--impl reference` times the CPU restatement of the reference
integrator and prints exactly ONE JSON line on stdout — library chatter (e.g. NCCL's version banner) goes to stderr — carrying the
base keys plus cpu_baseline / e2e as the reference arm defines them; non-zero ranks print nothing."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(extra_env=None, *args):
    env = dict(os.environ, **(extra_env or {}))
    return subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--config", "fire", "--volume-scale", "0.25",
                           "--steps", "1", "--warmup", "0", *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)


def test_reference_arm_prints_one_json_line(built):
    r = _run()
    assert r.returncode == 0, r.stderr
    lines = [ln for ln in r.stdout.split("\n") if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "pixel_samples_per_sec" and d["unit"] == "pixel-samples/s"
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None and d["dtype"] == "f32"
    assert d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 0 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["gpu_launches"] == 0


def test_reference_arm_other_ranks_do_no_work(built):
    r = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}, "--gpus", "2")
    assert r.returncode == 0 and r.stdout.strip() == ""
