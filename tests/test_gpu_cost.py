"""cost_estimator.py / cost_calculator.py drop-ins against the reference's outputs."""
import json
import os
import shutil
import subprocess
import sys

import pytest

from conftest import FAKE_BIN, GOLDEN, REPO, diff_keys, load_json

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,kube", [("cfg1", "kube"), ("repro100", None)])
def test_cost_estimator_cli_drop_in(name, kube, tmp_path):
    d = os.path.join(GOLDEN, name)
    run = tmp_path / "run"
    run.mkdir()
    # state after the reference's analyze.py: results.json + requests_classified.csv
    shutil.copy(os.path.join(d, "requests.csv"), run / "requests.csv")
    shutil.copy(os.path.join(d, "expected_classified.csv"), run / "requests_classified.csv")
    shutil.copy(os.path.join(d, "expected_results_analyze.json"), run / "results.json")
    env = dict(os.environ)
    env["PATH"] = FAKE_BIN + os.pathsep + env["PATH"]
    if kube:
        env["KVM_FAKE_KUBECTL_DIR"] = os.path.join(d, kube)
    r = subprocess.run([sys.executable, os.path.join(REPO, "cost_estimator.py"), "--run-dir", str(run),
                        "--namespace", "ml-prod", "--service", "demo-llm", "--cost-file",
                        os.path.join(d, "cost.yaml")], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = load_json(run / "results.json")
    exp = load_json(os.path.join(d, "expected_results_cost.json"))
    assert diff_keys(got, exp) == []
    if name == "cfg1":
        assert json.loads(r.stdout) == json.loads(open(os.path.join(d, "expected_cost_stdout.json")).read())


def test_cost_estimator_without_kubectl_raises_like_reference(tmp_path):
    shutil.copy(os.path.join(GOLDEN, "ci5", "requests.csv"), tmp_path / "requests.csv")
    env = dict(os.environ)
    env["PATH"] = "/nonexistent"
    r = subprocess.run([sys.executable, os.path.join(REPO, "cost_estimator.py"), "--run-dir", str(tmp_path),
                        "--namespace", "n", "--service", "s", "--cost-file",
                        os.path.join(GOLDEN, "cfg1", "cost.yaml")], env=env, capture_output=True, text=True)
    assert r.returncode != 0 and "FileNotFoundError" in r.stderr


def test_cost_calculator_cli_drop_in():
    d = os.path.join(GOLDEN, "costcalc")
    r = subprocess.run([sys.executable, os.path.join(REPO, "cost_calculator.py"), "lat.txt", "1.50",
                        "--requests-per-1k-tokens", "10"], cwd=d, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout == open(os.path.join(d, "expected_stdout.txt")).read()
    exp = load_json(os.path.join(d, "expected_none.json"))
    r = subprocess.run([sys.executable, os.path.join(REPO, "cost_calculator.py"), "none.txt", "1.50"],
                       cwd=d, capture_output=True, text=True)
    assert r.returncode == exp["returncode"] and r.stdout == exp["stdout"]
    r = subprocess.run([sys.executable, os.path.join(REPO, "cost_calculator.py"), "missing.txt", "1.50"],
                       cwd=d, capture_output=True, text=True)
    assert r.returncode == 1 and "not found" in r.stderr
