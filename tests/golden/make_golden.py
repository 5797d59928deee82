#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(/root/reference) on seeded inputs.  Run in the build container only - the reference does
not travel to the GPU box; the fixtures this script writes are what the tests read.

  python tests/golden/make_golden.py            # rewrites tests/golden/*

Fixtures
  known_answers.json   function-level vectors: analyze.percentile, analyze.compute_histograms,
                       energy.collector.trapezoidal_wh / run_window_bounds (incl. the
                       reference's own tests/test_energy.py and lint-test.yml:119-135 cases)
  ci5/                 the 5-row CSV of lint-test.yml:92-99 -> results.json (no kubectl)
  cfg1/                config 1 (500 requests, concurrency 20) + fake-kubectl pod + power.json
                       -> analyze.py, cost_estimator.py, energy/collector.py outputs
  repro100/            repro-smoke.yml:43-88 seeded 100-row CSV -> analyze + cost outputs
  edge5k/              5000 adversarial rows (ties, nan cells, zero tokens, negative generation, unsorted,
                       cold starts) -> analyze.py outputs
  costcalc/            cost_calculator.py input text + stdout
"""

import csv
import importlib.util
import json
import math
import os
import random
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
FAKE_BIN = os.path.join(REPO, "tests", "fake_kubectl")

sys.path.insert(0, REPO)
sys.dont_write_bytecode = True
os.environ["PYTHONDONTWRITEBYTECODE"] = "1"

from kvmini_b200 import synth  # noqa: E402


def ref_module(name, relpath):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def jnum(x):
    """JSON-safe float (NaN/inf as strings so plain json can round-trip them)."""
    if isinstance(x, float) and (math.isnan(x) or math.isinf(x)):
        return repr(x)
    return x


def run_ref(script, args, env_extra=None, cwd=None):
    env = dict(os.environ)
    env["PATH"] = FAKE_BIN + os.pathsep + env["PATH"]
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(REF, script)] + args, env=env, cwd=cwd,
                          capture_output=True, text=True)


def known_answers():
    an = ref_module("ref_analyze", "analyze.py")
    en = ref_module("ref_collector", "energy/collector.py")
    rng = random.Random(20240611)
    out = {"percentile": [], "histogram": [], "trapezoid": [], "window_bounds": []}

    # the reference CI's own vectors (lint-test.yml:119-135)
    ci = [10, 20, 30, 40, 50, 60, 70, 80, 90, 100]
    cases = [(ci, 0.5), (ci, 0.95), ([], 0.5), ([7.5], 0.99), ([float("nan")], 0.5),
             ([3.0, float("nan"), 1.0], 0.5), ([1.0, 2.0], 1.0), ([1.0, 2.0], 0.0), ([1.0, 2.0], -3.0),
             ([1.0, 2.0], 7.0), ([5.0] * 17, 0.95)]
    for _ in range(140):
        n = rng.choice([1, 2, 3, 5, 10, 33, 100, 257])
        kind = rng.choice(["lognormal3", "uniform", "dups", "neg", "tiny"])
        if kind == "lognormal3":
            vals = [round(rng.lognormvariate(5.5, 0.6), 3) for _ in range(n)]
        elif kind == "uniform":
            vals = [rng.uniform(0, 1e6) for _ in range(n)]
        elif kind == "dups":
            vals = [float(rng.randint(0, 5)) for _ in range(n)]
        elif kind == "neg":
            vals = [rng.gauss(0, 50) for _ in range(n)]
        else:
            vals = [rng.uniform(0, 1) * 1e-300 for _ in range(n)]
        if rng.random() < 0.3:
            for _k in range(max(1, n // 10)):
                vals[rng.randrange(n)] = float("nan")
        cases.append((vals, rng.choice([0.0, 0.01, 0.25, 0.5, 0.9, 0.95, 0.99, 0.999, 1.0])))
    for vals, p in cases:
        out["percentile"].append({"values": [jnum(float(v)) for v in vals], "p": p,
                                  "expected": jnum(an.percentile(list(map(float, vals)), p))})

    hist_cases = [([1, 2, 3, 4, 5] * 10, 5), ([], 20), ([2.0] * 9, 20), ([float("nan")] * 3, 20)]
    for _ in range(40):
        n = rng.choice([2, 7, 50, 400])
        vals = [round(rng.lognormvariate(4.0, 0.8), 3) for _ in range(n)]
        if rng.random() < 0.3:
            vals[rng.randrange(n)] = float("nan")
        hist_cases.append((vals, rng.choice([20, 20, 20, 5, 64])))
    for vals, nb in hist_cases:
        h = an.compute_histograms([float(v) for v in vals], nb)
        out["histogram"].append({"values": [jnum(float(v)) for v in vals], "num_bins": nb,
                                 "expected": {k: ([jnum(x) for x in v] if isinstance(v, list) else jnum(v))
                                              for k, v in h.items()}})

    # tests/test_energy.py:4-34
    trap_cases = [
        ([0.0, 5.0, 10.0], [100.0, 100.0, 100.0], 0.0, 10.0),
        ([0.0, 5.0, 10.0, 12.0], [None, 100.0, 100.0, None], 0.0, 12.0),
        ([], [], 0.0, 1.0), ([1.0], [5.0], 0.0, 2.0), ([1.0, 2.0], [5.0, 7.0], 3.0, 2.0),
        ([1.0, 1.0, 2.0], [5.0, 9.0, 7.0], 0.0, 5.0),
    ]
    for _ in range(40):
        m = rng.choice([2, 3, 10, 100, 300])
        ts = sorted(rng.uniform(1000.0, 1100.0) for _ in range(m))
        if rng.random() < 0.4:
            rng.shuffle(ts)
        if rng.random() < 0.3 and m > 3:
            ts[1] = ts[2]
        w = [None if rng.random() < 0.1 else rng.uniform(60.0, 1000.0) for _ in range(m)]
        t0 = rng.uniform(990.0, 1040.0)
        t1 = rng.uniform(1050.0, 1110.0)
        trap_cases.append((ts, w, t0, t1))
    for ts, w, t0, t1 in trap_cases:
        samples = [en.PowerSample(ts=t, watts=x) for t, x in zip(ts, w)]
        out["trapezoid"].append({"ts": ts, "watts": w, "t0": t0, "t1": t1,
                                 "expected": en.trapezoidal_wh(samples, t0, t1)})

    out["window_bounds"].append({"rows": [{"start_ms": 1000.0, "latency_ms": 500.0},
                                          {"start_ms": 4000.0, "latency_ms": 1000.0}],
                                 "expected": list(en.run_window_bounds(
                                     [{"start_ms": 1000.0, "latency_ms": 500.0},
                                      {"start_ms": 4000.0, "latency_ms": 1000.0}]))})
    out["window_bounds"].append({"rows": [], "expected": list(en.run_window_bounds([]))})
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump(out, f)


CI5_CSV = """id,scheduled_ms,start_ms,ttfb_ms,tllt_ms,latency_ms,status,prompt_tokens,completion_tokens,total_tokens,error
1,1000.0,1001.0,1050.0,1080.0,150.0,200,10,15,25,
2,1100.0,1101.0,1140.0,1190.0,180.0,200,12,18,30,
3,1200.0,1201.0,1260.0,1310.0,220.0,200,8,20,28,
4,1300.0,1301.0,,450.0,500,10,9,12,21,timeout
5,1400.0,1401.0,1420.0,1450.0,95.0,200,11,8,19,
"""


def fresh(d):
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)


def load_json_text(path):
    """results.json may hold bare NaN; keep the raw text so tests can parse it the same way."""
    with open(path) as f:
        return f.read()


def ci5():
    d = os.path.join(HERE, "ci5")
    fresh(d)
    with open(os.path.join(d, "requests.csv"), "w") as f:
        f.write(CI5_CSV)
    r = run_ref("analyze.py", ["--run-dir", d, "--namespace", "test-ns", "--service", "test-svc"])
    assert r.returncode == 0, r.stderr
    os.rename(os.path.join(d, "results.json"), os.path.join(d, "expected_results.json"))
    os.rename(os.path.join(d, "requests_classified.csv"), os.path.join(d, "expected_classified.csv"))


COST_YAML = """gpu:
  default: 1.50
  NVIDIA-H100-80GB: 6.50
cpu:
  hourly_per_core: 0.04
memory:
  hourly_per_gib: 0.005
overhead:
  fraction: 0.10
calculation:
  use_requests: true
  include_sidecars: false
"""


def cfg1():
    d = os.path.join(HERE, "cfg1")
    fresh(d)
    rows = synth.cfg1_rows()
    synth.write_requests_csv(os.path.join(d, "requests.csv"), rows)
    synth.write_json(os.path.join(d, "power.json"), synth.cfg1_power(rows))
    kube = os.path.join(d, "kube")
    os.makedirs(kube)
    synth.write_json(os.path.join(kube, "pods.json"), synth.cfg1_pods(rows))
    synth.write_json(os.path.join(kube, "node.json"),
                     {"metadata": {"name": "gpu-node-0", "labels": {"nvidia.com/gpu.product": "NVIDIA-H100-80GB"}}})
    with open(os.path.join(d, "cost.yaml"), "w") as f:
        f.write(COST_YAML)
    env = {"KVM_FAKE_KUBECTL_DIR": kube}
    common = ["--run-dir", d, "--namespace", "ml-prod", "--service", "demo-llm"]

    r = run_ref("analyze.py", common, env)
    assert r.returncode == 0, r.stderr
    shutil.copy(os.path.join(d, "results.json"), os.path.join(d, "expected_results_analyze.json"))
    r = run_ref("cost_estimator.py", common + ["--cost-file", os.path.join(d, "cost.yaml")], env)
    assert r.returncode == 0, r.stderr
    with open(os.path.join(d, "expected_cost_stdout.json"), "w") as f:
        f.write(r.stdout)
    shutil.copy(os.path.join(d, "results.json"), os.path.join(d, "expected_results_cost.json"))
    for mode, extra in (("series", ["--idle-tax", "series", "--merge-results"]),
                        ("baseline", ["--idle-tax", "baseline"]),
                        ("warmup", ["--include-warmup"]),
                        ("plain", [])):
        r = run_ref("energy/collector.py", ["integrate", "--run-dir", d] + extra, env)
        assert r.returncode == 0, r.stderr
        shutil.copy(os.path.join(d, "energy.json"), os.path.join(d, f"expected_energy_{mode}.json"))
        if mode == "series":
            shutil.copy(os.path.join(d, "results.json"), os.path.join(d, "expected_results_final.json"))
    # gate.py on the final outputs (consumer acceptance)
    r = run_ref("tools/gate.py", ["--results", os.path.join(d, "expected_results_final.json"),
                                  "--energy", os.path.join(d, "expected_energy_series.json"),
                                  "--slo", os.path.join(REF, "slo.json")])
    with open(os.path.join(d, "expected_gate.json"), "w") as f:
        json.dump({"returncode": r.returncode, "stdout": r.stdout}, f, indent=1)
    os.rename(os.path.join(d, "requests_classified.csv"), os.path.join(d, "expected_classified.csv"))
    for name in ("results.json", "energy.json"):
        os.remove(os.path.join(d, name))


def repro100():
    """repro-smoke.yml:43-88 generator, verbatim parameters (seed 42)."""
    d = os.path.join(HERE, "repro100")
    fresh(d)
    random.seed(42)
    base_time = 1000000
    with open(os.path.join(d, "requests.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "scheduled_ms", "start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status",
                    "prompt_tokens", "completion_tokens", "total_tokens", "error", "is_cold_start"])
        for i in range(100):
            base = 200 + i * 2
            status = 200 if i < 95 else 500
            w.writerow([i + 1, base_time + i * 100, base_time + i * 100 + random.randint(0, 5),
                        base * 0.3 if status == 200 else "", base * 0.8 if status == 200 else "",
                        base + random.randint(-10, 10), status, 10 + random.randint(0, 5),
                        15 + random.randint(0, 10) if status == 200 else "",
                        25 + random.randint(0, 15) if status == 200 else "",
                        "" if status == 200 else "timeout", i < 10])
    with open(os.path.join(d, "cost.yaml"), "w") as f:
        f.write(COST_YAML)
    common = ["--run-dir", d, "--namespace", "test-ns", "--service", "test-svc"]
    r = run_ref("analyze.py", common)
    assert r.returncode == 0, r.stderr
    shutil.copy(os.path.join(d, "results.json"), os.path.join(d, "expected_results_analyze.json"))
    r = run_ref("cost_estimator.py", common + ["--cost-file", os.path.join(d, "cost.yaml")])
    assert r.returncode == 0, r.stderr
    os.rename(os.path.join(d, "results.json"), os.path.join(d, "expected_results_cost.json"))
    os.rename(os.path.join(d, "requests_classified.csv"), os.path.join(d, "expected_classified.csv"))


def edge5k():
    """5000 rows built to stress what the two small fixtures do not: heavy ties (latencies rounded to
    whole milliseconds), literal `nan` cells, zero completion tokens, tllt < ttfb (negative
    generation time), several status codes, rows in file order that is NOT time order, three pods
    (two started inside the window -> cold starts).  Outputs come from the reference's analyze.py."""
    d = os.path.join(HERE, "edge5k")
    fresh(d)
    rnd = random.Random(20260101)
    t0 = 1_700_000_000_000.0
    rows = []
    for i in range(5000):
        status = rnd.choices([200, 500, 503, 404], weights=[90, 5, 3, 2])[0]
        lat = float(round(rnd.lognormvariate(5.5, 0.7)))
        start = t0 + rnd.uniform(0.0, 600_000.0)
        ok = status == 200
        ttfb = round(0.2 * lat + rnd.uniform(0.0, 3.0), 3) if ok else ""
        tllt = round(0.9 * lat, 3) if ok else ""
        if ok and i % 97 == 0:
            ttfb = "nan"
        if ok and i % 89 == 0:
            tllt = round(0.1 * lat, 3)              # last token "before" the first one
        ctok = rnd.randint(0, 40) if ok else ""     # zeros included
        rows.append([f"r{i}", "%.3f" % start, "%.3f" % start, ttfb, tllt, "%.3f" % lat, status, 12, ctok,
                     (12 + ctok) if ok else "", "" if ok else "upstream error, code %d" % status, "%032x" % i])
    with open(os.path.join(d, "requests.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "scheduled_ms", "start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status",
                    "prompt_tokens", "completion_tokens", "total_tokens", "error", "trace_id"])
        w.writerows(rows)
    kube = os.path.join(d, "kube")
    os.makedirs(kube)
    import datetime as dt

    def iso(ms):
        return dt.datetime.fromtimestamp(ms / 1000.0, dt.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")
    pods = {"items": [{"metadata": {"name": f"demo-llm-predictor-{k}"},
                       "status": {"containerStatuses": [{"state": {"running": {"startedAt": iso(ms)}}}]}}
                      for k, ms in enumerate((t0 - 3_600_000.0, t0 + 20_000.0, t0 + 300_000.0))]}
    synth.write_json(os.path.join(kube, "pods.json"), pods)
    with open(os.path.join(d, "cost.yaml"), "w") as f:
        f.write(COST_YAML)
    env = {"KVM_FAKE_KUBECTL_DIR": kube}
    common = ["--run-dir", d, "--namespace", "ml-prod", "--service", "demo-llm"]
    r = run_ref("analyze.py", common, env)
    assert r.returncode == 0, r.stderr
    shutil.copy(os.path.join(d, "results.json"), os.path.join(d, "expected_results_analyze.json"))
    os.remove(os.path.join(d, "results.json"))
    os.rename(os.path.join(d, "requests_classified.csv"), os.path.join(d, "expected_classified.csv"))


def costcalc():
    d = os.path.join(HERE, "costcalc")
    fresh(d)
    rng = random.Random(5)
    with open(os.path.join(d, "lat.txt"), "w") as f:
        for _ in range(300):
            f.write(f"{rng.lognormvariate(5.5, 0.5):.3f} {200 if rng.random() > 0.05 else 500}\n")
    r = run_ref("cost_calculator.py", [os.path.join(d, "lat.txt"), "1.50", "--requests-per-1k-tokens", "10"])
    assert r.returncode == 0, r.stderr
    with open(os.path.join(d, "expected_stdout.txt"), "w") as f:
        f.write(r.stdout.replace(d + os.sep, ""))
    with open(os.path.join(d, "none.txt"), "w") as f:
        f.write("100.0 500\n")
    r = run_ref("cost_calculator.py", [os.path.join(d, "none.txt"), "1.50"])
    with open(os.path.join(d, "expected_none.json"), "w") as f:
        json.dump({"returncode": r.returncode, "stdout": r.stdout}, f)


if __name__ == "__main__":
    assert os.path.isdir(REF), "the reference is only present in the build container"
    known_answers()
    ci5()
    cfg1()
    repro100()
    edge5k()
    costcalc()
    print("golden fixtures written under", HERE)
