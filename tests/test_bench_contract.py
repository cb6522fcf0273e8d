"""The bench.py JSON contract (CPU part): the reference arm prints one line with the keys the driver reads, on the
same metric / unit / config vocabulary as the GPU arm."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, env=None):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         timeout=600, env={**os.environ, **(env or {})}, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line_has_the_contract_keys():
    d = _run("--impl", "reference", "--workload", "gpt2-small", "--steps", "2", "--warmup", "1")
    assert d["impl"] == "reference" and d["metric"] == "decode_tokens_per_sec" and d["unit"] == "tokens/s"
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] >= 1
    assert d["config"]["workload"].startswith("gpt2-small decode, batch 1")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "oracle" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, env={**os.environ, "RANK": "1"}, cwd=ROOT)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_workload_names_match_between_arms():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.workload_name("m", 1, 512, 1, 22) == "m decode, batch 1, 512-token prompt, 1 stage"
    assert "32 layers over 4 stages, 4 micro-batches of 1" in bench.workload_name("m", 1, 512, 4, 32)
    for name, (model, batch, prompt) in bench.WORKLOADS.items():
        assert batch >= 1 and prompt >= 1 and isinstance(model, str)
