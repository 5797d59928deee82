"""Columnar ingest of power samples (csrc/kvm_samples.cu) against the reference's
own loader: `load_power_samples` of the unmodified energy/collector.py (oracle/_ref) or, where
that is not installed, the json-module restatement in energy_api.  CPU only (host code)."""
import json
import math
import os
import random

import numpy as np
import pytest

from conftest import GOLDEN, REPO


def _ref_loader():
    from oracle import ref_run
    if ref_run.available():
        import importlib.util
        import sys
        path = os.path.join(ref_run.REF_DST, "energy", "collector.py")
        spec = importlib.util.spec_from_file_location("_kvm_reference_collector", path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules["_kvm_reference_collector"] = mod     # dataclasses look the module up
        old = sys.dont_write_bytecode
        sys.dont_write_bytecode = True
        try:
            spec.loader.exec_module(mod)
        finally:
            sys.dont_write_bytecode = old
        return mod.load_power_samples
    from kvmini_b200 import energy_api
    return energy_api.load_power_samples


def _same_columns(path):
    from kvmini_b200 import energy_api
    ts, w = energy_api.load_power_columns(path)
    ref = _ref_loader()(path)
    assert ts.size == len(ref)
    assert ts.tolist() == [s.ts for s in ref]
    assert [None if math.isnan(x) else x for x in w.tolist()] == [s.watts for s in ref]
    return ts, w


def test_golden_power_json_equals_reference_loader():
    _same_columns(os.path.join(GOLDEN, "cfg1", "power.json"))


@pytest.mark.parametrize("style", ["dump", "compact", "indent", "bare_list", "ts_s", "string_numbers", "extra_keys"])
def test_generated_power_json_variants(style, tmp_path):
    rnd = random.Random(7)
    n = 50_000
    samples = []
    for i in range(n):
        t = 1.7e9 + i * 0.001 + rnd.random() * 5e-5
        w = None if rnd.random() < 0.01 else round(400 + 300 * rnd.random(), rnd.choice([0, 1, 3, 6]))
        if style == "ts_s":
            s = {"ts_s": t, "ts": -1.0, "watts": w}
        elif style == "string_numbers":
            s = {"ts": repr(t), "watts": (None if w is None else repr(w))}
        elif style == "extra_keys":
            s = {"gpu": "GPU-0", "ok": True, "watts": w, "note": None, "ts": t}
        else:
            s = {"ts": t, "watts": w}
        samples.append(s)
    samples[5]["watts"] = 1e-300
    samples[6]["watts"] = 6.02e23
    samples[7]["ts"] = 12 if style not in ("ts_s", "string_numbers") else samples[7].get("ts")
    obj = samples if style == "bare_list" else {"interval_s": 0.001, "samples": samples, "source": "dcgm"}
    path = tmp_path / "power.json"
    with open(path, "w") as f:
        if style == "compact":
            json.dump(obj, f, separators=(",", ":"))
        elif style == "indent":
            json.dump(obj, f, indent=2)
        else:
            json.dump(obj, f)
    from kvmini_b200 import _native as N
    import ctypes as C
    h = C.c_void_p()
    assert N.lib().kvm_power_json_parse(os.fsencode(str(path)), 8, C.byref(h)) == 0     # the native path took it
    N.lib().kvm_samples_free(h)
    ts, w = _same_columns(str(path))
    assert ts.size == n


@pytest.mark.parametrize("text", [
    '{"samples": [{"ts": 1.0, "watts": NaN}]}',                 # Python-only literal
    '{"samples": [{"ts": 1.0, "watts": {"v": 2}}]}',           # nested value
    '{"samples": [{"ts": 1.0, "wat\\u0074s": 3.0}]}',           # escape in a key
    '{"samples": [{"ts": 1.0}]}',                               # KeyError in the reference
    '{"samples": [{"ts": null, "watts": 1.0}]}',                # TypeError in the reference
    '[1, 2, 3]',
])
def test_irregular_power_json_is_left_to_the_json_module(text, tmp_path):
    """The scanner declines; the fallback then behaves like the reference (same values or same exception)."""
    from kvmini_b200 import _native as N
    from kvmini_b200 import energy_api
    import ctypes as C
    path = tmp_path / "power.json"
    path.write_text(text)
    h = C.c_void_p()
    assert N.lib().kvm_power_json_parse(os.fsencode(str(path)), 4, C.byref(h)) == N.KVM_ERR_UNSUPPORTED
    ref = _ref_loader()
    try:
        exp = ref(str(path))
    except Exception as e:
        with pytest.raises(type(e)):
            energy_api.load_power_columns(str(path))
        return
    ts, w = energy_api.load_power_columns(str(path))
    assert ts.tolist() == [s.ts for s in exp]


def test_empty_and_foreign_documents(tmp_path):
    from kvmini_b200 import energy_api
    for text in ('{"samples": []}', '[]', '{"other": 1}'):
        p = tmp_path / "p.json"
        p.write_text(text)
        ts, w = energy_api.load_power_columns(str(p))
        assert ts.size == 0 and w.size == 0 and _ref_loader()(str(p)) == []


def test_prometheus_matrix_and_vector_bodies():
    from kvmini_b200 import cluster, energy_api
    rnd = random.Random(3)
    series = []
    for g in range(3):
        vals = [[1.7e9 + i * 0.5, repr(round(300 + 200 * rnd.random(), 3))] for i in range(40_000)]
        vals[10][1] = "NaN"
        vals[11][1] = "+Inf"
        series.append({"metric": {"gpu": str(g), "Hostname": "node-[0]"}, "values": vals})
    body = json.dumps({"status": "success", "data": {"resultType": "matrix", "result": series}}).encode()
    cols = energy_api.prom_result_columns(body)
    assert cols is not None and len(cols) == 3
    for (ts, v), s in zip(cols, series):
        assert ts.tolist() == [float(a) for a, _ in s["values"]]
        exp = [float(b) for _, b in s["values"]]
        assert all((math.isnan(x) and math.isnan(y)) or x == y for x, y in zip(v.tolist(), exp))
    # the time average of analyze.py:233-247, finite values only in this body
    for s in series:
        s["values"][10][1] = "1.5"
        s["values"][11][1] = "2.5"
    doc = {"status": "success", "data": {"resultType": "matrix", "result": series}}
    body = json.dumps(doc).encode()
    assert cluster.prom_matrix_timeavg_body(body) == cluster.prom_matrix_timeavg(doc)
    vec = {"status": "success", "data": {"resultType": "vector", "result": [
        {"metric": {}, "value": [1.7e9, "412.25"]}, {"metric": {"a": "b"}, "value": [1.7e9, "7"]}]}}
    cols = energy_api.prom_result_columns(json.dumps(vec).encode())
    assert [c[1].tolist() for c in cols] == [[412.25], [7.0]]
    assert cluster.prom_matrix_timeavg_body(b"not json at all") is None
