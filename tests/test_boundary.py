"""CPU checks of the drop-in boundary: the C-ABI library builds, loads and exports every
symbol include/kvmini_b200.h declares; the product never imports the oracle; the drop-in
scripts keep the reference's importable names; without a GPU the product fails loudly."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

from conftest import REPO

PKG = os.path.join(REPO, "kserve-vllm-mini_b200")
HEADER = os.path.join(REPO, "include", "kvmini_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kvm_[a-z0-9_]+)\s*\(", text)))


def test_build_and_exports():
    import __graft_entry__ as g
    g.build()
    lib = ctypes.CDLL(os.path.join(PKG, "csrc", "libkvmini_b200.so"))
    syms = declared_symbols()
    assert len(syms) >= 25
    missing = [s for s in syms if not hasattr(lib, s)]
    assert missing == []
    # and the ctypes layer declares an argtypes signature for each of them
    from kvmini_b200 import _native
    undeclared = [s for s in syms if s not in _native.lib()._kvm_declared]
    assert undeclared == []


def test_struct_layouts_match_header():
    """ctypes mirrors vs the sizes the C compiler sees."""
    from kvmini_b200 import _native as N
    src = r'''
    #include <stdio.h>
    #include "kvmini_b200.h"
    int main(void){ printf("%zu %zu %zu %zu %zu\n", sizeof(kvm_request_log), sizeof(kvm_set_info),
        sizeof(kvm_log_stats), sizeof(kvm_query), sizeof(kvm_query_result)); return 0; }'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.run(["gcc", "-I", os.path.join(REPO, "include"), os.path.join(d, "t.c"), "-o",
                        os.path.join(d, "t")], check=True)
        sizes = list(map(int, subprocess.check_output([os.path.join(d, "t")]).split()))
    assert sizes == [ctypes.sizeof(N.RequestLogC), ctypes.sizeof(N.SetInfoC), ctypes.sizeof(N.LogStatsC),
                     ctypes.sizeof(N.QueryC), ctypes.sizeof(N.QueryResultC)]


def test_product_never_imports_oracle():
    offenders = []
    roots = [PKG] + [os.path.join(REPO, f) for f in ("analyze.py", "cost_estimator.py", "cost_calculator.py",
                                                      "energy", "kvmini_b200.py")]
    for root in roots:
        files = [root] if os.path.isfile(root) else [os.path.join(dp, f) for dp, _, fs in os.walk(root) for f in fs]
        for path in files:
            if path.endswith(".py") and re.search(r"^\s*(from|import)\s+oracle\b", open(path).read(), flags=re.M):
                offenders.append(path)
    assert offenders == []


def test_reference_import_names():
    # lint-test.yml:117,143 and tests/test_energy.py:1 of the reference
    sys.path.insert(0, REPO)
    from analyze import compute_histograms, percentile, read_requests_csv  # noqa: F401


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from analyze import percentile
    from kvmini_b200._native import NativeError
    with pytest.raises(NativeError):
        percentile([1.0, 2.0, 3.0], 0.5)


def test_read_requests_csv_ci_fixture():
    # lint-test.yml:131-135: 5 rows, rows[0]['status'] == 200
    from analyze import read_requests_csv
    rows = read_requests_csv(os.path.join(REPO, "tests", "golden", "ci5", "requests.csv"))
    assert len(rows) == 5 and rows[0]["status"] == 200
    assert rows[3]["ttfb_ms"] == "" and rows[3]["status"] == 10.0   # the malformed 4th row


def test_ingest_matches_reference_errors():
    from kvmini_b200 import columns
    log = columns.columns_from_rows([{"start_ms": 1.0, "latency_ms": 2.0, "ttfb_ms": "", "tllt_ms": 3.0,
                                      "status": 200.0, "completion_tokens": 5.0, "total_tokens": 9.0}])
    with pytest.raises(TypeError, match="must be real number, not str"):
        columns.check_analyzable(log)
    with pytest.raises(ValueError):
        columns.columns_from_rows([{"status": ""}])
    ok = columns.columns_from_rows([{"start_ms": 1.0, "latency_ms": 2.0, "ttfb_ms": "", "tllt_ms": "",
                                     "status": 500.0, "completion_tokens": "", "total_tokens": ""}])
    columns.check_analyzable(ok)   # failed rows may have empty timing cells
    assert ok.cols["completion_tokens"][0] == 0
