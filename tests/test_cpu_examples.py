"""Run a few of the example scripts end to end with small settings (the reference converts notebooks to scripts and
runs them with fewer epochs: apps/run-app-tests.sh)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [
    ("recommendation/ncf_explicit_feedback.py", ["--epochs", "1", "--ratings", "20000", "--batch-size", "2000"], "validation"),
    ("recommendation/wide_n_deep.py", ["--epochs", "1", "--rows", "4096", "--batch-size", "512"], "Top1Accuracy"),
    ("keras/functional_api.py", [], "mse after"),
    ("serving/quick_start.py", [], "results dequeued"),
    ("inference/inference_model.py", [], "all equal: True"),
    ("qaranker/qa_ranker.py", [], "NDCG@3"),
]


@pytest.mark.parametrize("script,args,expect", CASES)
def test_example_runs(script, args, expect):
    env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script), *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and expect in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]
