"""tools/make_gguf.py writes files the REFERENCE's loader accepts: the reference's own llama-bench (built unmodified into oracle/_ref by
`make -f oracle/Makefile.ref llama`) loads a tiny synthetic model and runs it on its CPU backend.  No GPU needed; skipped where the
binary has not been built.  (The 7B / 8B presets are the same writer with the public Llama shapes; bench.py and tools/gpu_llama_bench.sh use them.)"""
import json
import os
import subprocess
import sys

import pytest

import pyref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench_exe():
    d = pyref.ref_dir()
    exe = os.path.join(d or "", "llama-bench")
    if not d or not os.path.exists(exe):
        pytest.skip("oracle/_ref/<variant>/llama-bench not built")
    return exe


@pytest.mark.parametrize("typ,extra", [("q4_0", []), ("q8_0", []), ("q4_K", ["--out-type", "q6_K"]), ("q4_0", ["--rope-freqs"]), ("q4_K", ["--kmix"])])
def test_tiny_gguf_loads_and_runs_on_the_reference_cpu_backend(tmp_path, typ, extra):
    exe = _bench_exe()
    gguf = str(tmp_path / "tiny.gguf")
    # K-quants need rows of 256: the tiny preset has n_embd = 256, n_ff = 512
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_gguf.py"), gguf, "--preset", "tiny", "--type", typ] + extra, check=True, capture_output=True, timeout=120)
    env = dict(os.environ)
    env.pop("GGML_BACKEND_PATH", None)
    r = subprocess.run([exe, "-m", gguf, "-p", "16", "-n", "4", "-r", "1", "-ngl", "0", "-t", "2", "-o", "jsonl"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    rows = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(rows) == 2 and all(x["avg_ts"] > 0 for x in rows), r.stdout[-500:]
    assert rows[0]["model_type"].startswith("llama") and rows[0]["model_n_params"] > 1_000_000
    assert {(x["n_prompt"], x["n_gen"]) for x in rows} == {(16, 0), (0, 4)}
