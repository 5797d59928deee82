import json
import math
import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
GOLDEN = os.path.join(REPO, "tests", "golden")
FAKE_BIN = os.path.join(REPO, "tests", "fake_kubectl")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """GPU tests are skipped (not failed) where no device exists; `-m gpu` on the GPU box runs them."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def unj(x):
    """Inverse of make_golden.jnum: 'nan'/'inf' strings back to floats."""
    if isinstance(x, str) and x in ("nan", "inf", "-inf"):
        return float(x)
    if isinstance(x, list):
        return [unj(v) for v in x]
    if isinstance(x, dict):
        return {k: unj(v) for k, v in x.items()}
    return x


def same(a, b):
    """Bit-level equality for floats (NaN == NaN, 0.0 != -0.0 ignored), recursive."""
    if isinstance(a, float) and isinstance(b, float):
        return (math.isnan(a) and math.isnan(b)) or a == b
    if isinstance(a, (list, tuple)) and isinstance(b, (list, tuple)):
        return len(a) == len(b) and all(same(x, y) for x, y in zip(a, b))
    if isinstance(a, dict) and isinstance(b, dict):
        return list(a.keys()) == list(b.keys()) and all(same(a[k], b[k]) for k in a)
    if isinstance(a, bool) or isinstance(b, bool):
        return a is b
    if isinstance(a, (int, float)) and isinstance(b, (int, float)):
        return float(a) == float(b) and isinstance(a, int) == isinstance(b, int)
    return a == b


def diff_keys(got, exp, rel=None, prefix=""):
    """List of human-readable differences between two result dicts.
    rel: {key: tolerance} for fields compared relatively instead of bit-exactly."""
    rel = rel or {}
    out = []
    if list(got.keys()) != list(exp.keys()):
        out.append(f"{prefix}key order/sets differ: {list(got.keys())} vs {list(exp.keys())}")
    for k in exp:
        if k not in got:
            continue
        g, e = got[k], exp[k]
        if isinstance(e, dict) and isinstance(g, dict):
            out += diff_keys(g, e, rel, prefix + k + ".")
        elif k in rel and isinstance(e, float) and isinstance(g, float):
            if not (abs(g - e) <= rel[k] * abs(e)):
                out.append(f"{prefix}{k}: {g!r} vs {e!r} (rel tol {rel[k]})")
        elif not same(g, e):
            out.append(f"{prefix}{k}: {g!r} vs {e!r}")
    return out


@pytest.fixture(scope="session")
def known_answers():
    with open(os.path.join(GOLDEN, "known_answers.json")) as f:
        return json.load(f)


def load_json(path):
    with open(path) as f:
        return json.load(f)
