"""bench.py's JSON contract, checked on the CPU arm (the GPU arm needs a B200).

`bench.py --impl reference` times the reference's own CPU path for this hot path -- which, for a
reference whose generator is a remote service, is the library it pins (transformers
GPT2LMHeadModel.generate) -- and must print ONE JSON line with the same
metric / unit / config as the GPU arm plus `impl`, `cpu_baseline` and a zero-copy `e2e` block.
"""
import json
import os
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    env = dict(os.environ, HF_HUB_OFFLINE="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload",
                        "smart_reply_b1", "--steps", "1", "--warmup", "0"], capture_output=True, text=True,
                       env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "tokens/s" and d["higher_is_better"] is True
    assert d["metric"].startswith("generated tokens/sec") and d["n_gpus"] == 1 and d["steps"] == 1
    assert "smart_reply_b1" in d["config"]["workload"] and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and "sample" in d["cpu_baseline"]
    assert d["e2e"] == {"value": d["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0


def test_workload_table_names_every_baseline_config():
    sys.path.insert(0, ROOT)
    import bench

    assert {"smart_reply_b1", "smart_reply_b64", "summarize_medium", "ask_large_b32", "ask_large_b256"} <= set(bench.WORKLOADS)
    assert bench.WORKLOADS["smart_reply_b64"] == ("small", 64, 128, 32)          # BASELINE.json configs[1]
    assert bench.WORKLOADS["summarize_medium"] == ("medium", 16, 1024, 128)      # configs[2]
    assert bench.WORKLOADS["ask_large_b256"] == ("large", 256, 64, 256)          # configs[3]
    # decode bytes of one step, SURVEY.md 8d: weights 247.3 MB + 64 rows x (ctx + 1) x 36,864 B
    import importlib
    E = importlib.import_module("distributed-real-time-chat-and-collaboration-tool_b200.engine")
    b = bench.decode_step_bytes(E.GPT2_SMALL, 64, 144)
    assert abs(b - (247.3e6 + 64 * 145 * 36864)) < 0.2e6
