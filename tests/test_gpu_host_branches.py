"""The host-side branches that splice into results.json (SURVEY.md A10 and its neighbours):
`--prom-url` utilisation + inline energy (analyze.py:529-552), the io_probe.json passthrough
(:554-562,:591-593) and cache_hit_ratio (:565-571).

Method: a stub Prometheus on 127.0.0.1 answers /api/v1/query and /api/v1/query_range; the
UNMODIFIED reference (oracle/_ref/analyze.py, installed by __graft_entry__.build()) and the
drop-in analyze.py are run as subprocesses on copies of the same run directory against the same
stub, with the same fake kubectl on PATH.  results.json must agree key by key in the same order
(mean TTFT <= 1e-12, everything else bit-exact), the drop-in's extra `timeseries` key aside.
Also pins two documented deviations against what the reference does.
"""
import http.server
import json
import os
import shutil
import subprocess
import sys
import threading
import urllib.parse

import pytest

from conftest import FAKE_BIN, GOLDEN, REPO, diff_keys, load_json

pytestmark = pytest.mark.gpu
REF = os.path.join(REPO, "oracle", "_ref", "analyze.py")


class _Prom(http.server.BaseHTTPRequestHandler):
    """Answers by the metric named in the query; unknown metrics -> empty result."""
    table = {}
    seen = []

    def do_GET(self):
        u = urllib.parse.urlparse(self.path)
        q = urllib.parse.parse_qs(u.query).get("query", [""])[0]
        type(self).seen.append((u.path, q))
        value = None
        for key, v in type(self).table.items():
            if key in q:
                value = v
                break
        if value is None:
            result = []
        elif u.path.endswith("query_range"):
            result = [{"metric": {}, "values": [[1.0, str(value)], [2.0, str(value * 1.5)]]}]
        else:
            result = [{"metric": {}, "value": [1.0, str(value)]}, {"metric": {}, "value": [1.0, str(value * 3.0)]}]
        body = json.dumps({"status": "success", "data": {"resultType": "vector", "result": result}}).encode()
        self.send_response(200)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def log_message(self, *a):
        pass


@pytest.fixture()
def prom():
    _Prom.seen = []
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), _Prom)
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    yield f"http://127.0.0.1:{srv.server_address[1]}"
    srv.shutdown()


def _run(script, run_dir, extra, kube_dir=None):
    env = dict(os.environ)
    env["PATH"] = FAKE_BIN + os.pathsep + env["PATH"]
    env["PYTHONDONTWRITEBYTECODE"] = "1"
    if kube_dir:
        env["KVM_FAKE_KUBECTL_DIR"] = kube_dir
    return subprocess.run([sys.executable, script, "--run-dir", str(run_dir), "--namespace", "ml-prod",
                           "--service", "demo-llm"] + extra, env=env, capture_output=True, text=True)


def _both(tmp_path, fixture, extra, files=None, kube=None):
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref not installed (run __graft_entry__.build() where /root/reference exists)")
    d = os.path.join(GOLDEN, fixture)
    out = {}
    for who, script in (("ref", REF), ("ours", os.path.join(REPO, "analyze.py"))):
        run = tmp_path / who
        run.mkdir()
        shutil.copy(os.path.join(d, "requests.csv"), run / "requests.csv")
        for name, text in (files or {}).items():
            (run / name).write_text(text)
        r = _run(script, run, extra, os.path.join(d, kube) if kube else None)
        assert r.returncode == 0, (who, r.stderr[-2000:])
        out[who] = (load_json(run / "results.json"), r)
    return out


def _compare(out):
    ours, ref = dict(out["ours"][0]), out["ref"][0]
    ours.pop("timeseries", None)
    assert diff_keys(ours, ref, rel={"time_to_first_token_ms": 1e-12}) == []
    return ours


def test_prom_url_utilisation_energy_and_cache_hit(tmp_path, prom):
    _Prom.table = {"DCGM_FI_DEV_GPU_UTIL": 71.5, "DCGM_FI_DEV_FB_USED": 40321.0, "DCGM_FI_DEV_POWER_USAGE": 412.25,
                   "container_cpu_usage_seconds_total": 3.25, "container_memory_working_set_bytes": 7.5e9,
                   "vllm_prompt_cache_hits_total": 0.125}
    got = _compare(_both(tmp_path, "cfg1", ["--prom-url", prom], kube="kube"))
    assert got["gpu_power_watts_avg"] == (412.25 + 3 * 412.25) / 2
    assert got["energy_wh"] == got["gpu_power_watts_avg"] * (got["window"]["seconds"] / 3600.0)
    assert got["energy_wh_per_request"] is not None and got["energy_wh_per_1k_tokens"] is not None
    assert got["cache_hit_ratio"] == 0.25          # mean of the stub's two series
    keys = list(got)
    assert keys.index("energy_wh") == keys.index("mem_used_avg") + 1      # spliced right after the util block


def test_prom_fallback_metric_names_and_missing_power(tmp_path, prom):
    """First metric name unknown -> the next candidate is asked (analyze.py:262-285); no power metric at
    all -> no energy keys; cache metrics absent -> kubectl-log fallback -> None."""
    _Prom.table = {"nvidia_dcgm_gpu_utilization": 55.0, "nvidia_dcgm_fb_used_bytes": 1.0e9}
    got = _compare(_both(tmp_path, "repro100", ["--prom-url", prom]))
    assert got["gpu_util_avg"] == 110.0 and got["gpu_power_watts_avg"] is None and "energy_wh" not in got
    assert got["cache_hit_ratio"] is None


def test_prom_unreachable_is_swallowed(tmp_path):
    out = _both(tmp_path, "ci5", ["--prom-url", "http://127.0.0.1:9/"])
    got = _compare(out)
    assert got["gpu_util_avg"] is None
    assert "Prometheus query failed" in out["ours"][1].stderr and "Prometheus query failed" in out["ref"][1].stderr


def test_io_probe_passthrough_and_merge_on_write(tmp_path):
    probe = json.dumps({"rtt": {"p50_ms": 1.25, "p95_ms": 3.5}, "s3": {"avg_MBps": 812.0}})
    prev = json.dumps({"cost_per_1k_tokens": 0.0123, "p95_ms": -1.0, "custom": "kept"})
    got = _compare(_both(tmp_path, "ci5", [], files={"io_probe.json": probe, "results.json": prev}))
    assert (got["network_rtt_p50_ms"], got["network_rtt_p95_ms"], got["s3_avg_MBps"]) == (1.25, 3.5, 812.0)
    assert list(got)[:3] == ["cost_per_1k_tokens", "p95_ms", "custom"] and got["custom"] == "kept" and got["p95_ms"] != -1.0


def test_io_probe_malformed_or_partial(tmp_path):
    a = tmp_path / "a"
    a.mkdir()
    got = _compare(_both(a, "ci5", [], files={"io_probe.json": "{not json"}))
    assert got["network_rtt_p50_ms"] is None
    b = tmp_path / "b"
    b.mkdir()
    got = _compare(_both(b, "ci5", [], files={"io_probe.json": json.dumps({"rtt": None, "s3": {}})}))
    assert got["network_rtt_p95_ms"] is None and got["s3_avg_MBps"] is None


# ---- documented deviations, pinned against the reference's behaviour --------------------------
HDR = "id,scheduled_ms,start_ms,ttfb_ms,tllt_ms,latency_ms,status,prompt_tokens,completion_tokens,total_tokens,error,trace_id\n"


def _ref_and_ours_on(tmp_path, text):
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref not installed")
    res = {}
    for who, script in (("ref", REF), ("ours", os.path.join(REPO, "analyze.py"))):
        run = tmp_path / who
        run.mkdir()
        (run / "requests.csv").write_text(text)
        res[who] = (_run(script, run, []), run)
    return res


def test_float_token_cells_reference_sums_floats_we_refuse(tmp_path):
    """the columnar log keeps token counts as int64.  The reference float()s the
    cell and sums it (analyze.py:502-511); a non-integer token count in a success row is not a
    value any producer writes (loadtest.py:598 writes ints) - the drop-in refuses it loudly
    instead of rounding silently."""
    rows = "".join(f"{i},,{1000.0 + 10 * i:.3f},20.000,90.000,100.000,200,12,{7.5 if i == 2 else 8},20,,\n" for i in range(5))
    res = _ref_and_ours_on(tmp_path, HDR + rows)
    assert res["ref"][0].returncode == 0
    ref = load_json(res["ref"][1] / "results.json")
    assert ref["tokens_per_sec"] == (8 * 4 + 7.5) / ref["window"]["seconds"]
    assert res["ours"][0].returncode != 0 and "ValueError" in res["ours"][0].stderr
    # integer-valued float spellings are the same number in both
    rows = "".join(f"{i},,{1000.0 + 10 * i:.3f},20.000,90.000,100.000,200,12,{'8.0' if i == 2 else 8},20,,\n" for i in range(5))
    both = tmp_path / "int"
    both.mkdir()
    res = _ref_and_ours_on(both, HDR + rows)
    assert res["ours"][0].returncode == 0, res["ours"][0].stderr
    ours = load_json(res["ours"][1] / "results.json")
    ours.pop("timeseries", None)
    assert diff_keys(ours, load_json(res["ref"][1] / "results.json"), rel={"time_to_first_token_ms": 1e-12}) == []


def test_nan_start_ms_window_is_order_dependent_in_the_reference(tmp_path):
    """Python's min()/max() with a NaN operand depend on where the NaN sits
    (analyze.py:185-187).  With the NaN row anywhere but first the reference's window skips it -
    that case is reproduced bit for bit; with the NaN row first the reference's window IS NaN and
    it dies with a traceback in get_cold_start_times (datetime.fromtimestamp(nan), analyze.py:380),
    where the device reports the window of the finite rows."""
    def rows(nan_at):
        out = ""
        for i in range(6):
            start = "nan" if i == nan_at else f"{1000.0 + 100 * i:.3f}"
            out += f"{i},,{start},20.000,90.000,100.000,200,12,8,20,,\n"
        return HDR + out
    mid = tmp_path / "mid"
    mid.mkdir()
    res = _ref_and_ours_on(mid, rows(3))
    assert res["ref"][0].returncode == 0 and res["ours"][0].returncode == 0, res["ours"][0].stderr
    ours = load_json(res["ours"][1] / "results.json")
    ours.pop("timeseries", None)
    assert diff_keys(ours, load_json(res["ref"][1] / "results.json"), rel={"time_to_first_token_ms": 1e-12}) == []
    first = tmp_path / "first"
    first.mkdir()
    res = _ref_and_ours_on(first, rows(0))
    assert res["ref"][0].returncode != 0 and "ValueError: Invalid value NaN" in res["ref"][0].stderr
    assert res["ours"][0].returncode == 0, res["ours"][0].stderr
    ours = load_json(res["ours"][1] / "results.json")
    assert ours["window"] == {"start": 1.1, "end": 1.6, "seconds": 0.5}
    assert ours["p50_ms"] == 100.0 and ours["requests"] == {"total": 6, "success": 6}
