"""bench.py's reference arm runs on the host cores (no GPU needed): one JSON line with the contract's keys."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--height", "256", "--width", "320"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "impl"):
        assert k in d, k
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "pairs/s"
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["value"] > 0 and "workload" in d["config"]


def test_bench_defaults_and_configs_parse():
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    old = sys.argv
    try:
        sys.argv = ["bench.py"]
        a = bench.parse_args()
        assert a.gpus == 1 and a.steps >= 1 and a.warmup >= 3 and a.pairs == 32 and (a.height, a.width) == (512, 640)
        assert a.with_fine_transformer is True
        for cfg, hw in (("1280x1024", (1024, 1280)), ("seq320", (320, 320)), ("seq10k", (320, 320))):
            sys.argv = ["bench.py", "--config", cfg]
            a = bench.parse_args()
            assert (a.height, a.width) == hw and a.job == (cfg == "seq10k")
    finally:
        sys.argv = old
