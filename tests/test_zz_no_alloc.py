"""kvm_column_select_static: the select for callers that cannot allocate device memory.

Some users of this library run next to an inference server that has already taken the
GPU's memory pool (weights + KV cache), or under a policy that forbids allocating once
serving has started; they cannot hand the library a workspace.  For them the library
carries a fixed scratch array of its own.  This test drives that entry point through the
C ABI and requires bit-identical results to kvm_column_select (the workspace-taking entry
point it mirrors) on two consecutive calls - the second call reuses the scratch the first
one left dirty.  Collected last (file name) so a `-x` run has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _queries(N, ps):
    arr = (N.QueryC * (2 * len(ps)))()
    for i, p in enumerate(ps):
        arr[2 * i] = N.QueryC(1, N.RANK_LERP_LO, p, 0)
        arr[2 * i + 1] = N.QueryC(1, N.RANK_LERP_HI, p, 0)
    return arr


def _run_both(values_np, status_np, ps):
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200.engine import RESULT_DTYPE
    lib = N.lib()
    dev = torch.device("cuda", 0)
    values = torch.from_numpy(values_np).to(dev)
    status = torch.from_numpy(status_np).to(dev)
    n = int(values.shape[0])
    q = _queries(N, ps)
    nq = len(q)
    stream = torch.cuda.current_stream(dev).cuda_stream
    out_static = torch.zeros(nq * C.sizeof(N.QueryResultC), dtype=torch.uint8, device=dev)
    out_ws = torch.zeros_like(out_static)

    # entry point under test: inputs and outputs are ours, the scratch is the library's
    N.check(lib.kvm_column_select_static(values.data_ptr(), n, N.MEMBER_STATUS, status.data_ptr(), q, nq,
                                         out_static.data_ptr(), stream), "kvm_column_select_static")
    # existing entry point with a caller-provided workspace
    cap = n + 1024
    ws = torch.empty(lib.kvm_select_workspace_bytes(cap), dtype=torch.uint8, device=dev)
    N.check(lib.kvm_column_select(values.data_ptr(), n, N.MEMBER_STATUS, status.data_ptr(), 1, q, nq,
                                  out_ws.data_ptr(), ws.data_ptr(), ws.numel(), cap, stream), "kvm_column_select")
    N.check(lib.kvm_select_status(ws.data_ptr(), stream), "kvm_select_status")
    a = np.frombuffer(out_static.cpu().numpy().tobytes(), dtype=RESULT_DTYPE)
    b = np.frombuffer(out_ws.cpu().numpy().tobytes(), dtype=RESULT_DTYPE)
    return a, b


def test_static_scratch_select_matches_workspace_select_twice():
    rng = np.random.default_rng(2024)
    ps = [0.5, 0.95, 0.99]
    # call 1: large, concentrated latencies; call 2: smaller, different range, with NaNs -
    # anything the first call left in the scratch must not leak into the second
    v1 = np.round(rng.lognormal(5.5, 0.6, 700_001), 3)
    s1 = np.where(rng.random(v1.size) < 0.98, 200, 500).astype(np.int64)
    v2 = rng.normal(0.0, 1e6, 33_333)
    v2[rng.integers(0, v2.size, 500)] = np.nan
    s2 = np.where(rng.random(v2.size) < 0.5, 200, 503).astype(np.int64)
    for values, status in ((v1, s1), (v2, s2)):
        a, b = _run_both(values, status, ps)
        assert np.array_equal(a["status"], b["status"])
        assert np.array_equal(a["n"], b["n"]) and np.array_equal(a["rank"], b["rank"])
        assert a["value"].tobytes() == b["value"].tobytes()          # bit-exact, like every order statistic here
        kept = np.sort(values[(status == 200) & ~np.isnan(values)])   # and both are right
        assert np.array_equal(a["value"], kept[a["rank"]])


def test_static_scratch_select_rejects_what_does_not_fit():
    import torch
    from kvmini_b200 import _native as N
    lib = N.lib()
    dev = torch.device("cuda", 0)
    limit = int(lib.kvm_column_select_static_max_n())
    assert limit == 1 << 20
    values = torch.zeros(limit + 1, dtype=torch.float64, device=dev)
    q = _queries(N, [0.5])
    out = torch.zeros(2 * C.sizeof(N.QueryResultC), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.kvm_column_select_static(values.data_ptr(), limit + 1, N.MEMBER_ALL, None, q, 2, out.data_ptr(), stream)
    assert rc == -1 and b"exceeds the built-in scratch" in lib.kvm_last_error()
    ids = torch.zeros(16, dtype=torch.uint8, device=dev)
    rc = lib.kvm_column_select_static(values.data_ptr(), 16, N.MEMBER_SET_IDS, ids.data_ptr(), q, 2,
                                      out.data_ptr(), stream)
    assert rc == -1 and b"not supported" in lib.kvm_last_error()


def test_static_scratch_select_with_mixed_set_masks_at_the_size_limit():
    """Queries whose set masks differ only in bits no set answers to (1, 3, 0xFFFFFFFF) all mean
    "set 0" here.  Without the entry point reducing them to bit 0 every distinct mask would get
    its own copy of a candidate bucket and, on concentrated data at the size limit, the copies
    would outgrow the built-in scratch; the caller - who has no workspace to read a status word
    from - must still get exact answers."""
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200.engine import RESULT_DTYPE
    lib = N.lib()
    dev = torch.device("cuda", 0)
    n = int(lib.kvm_column_select_static_max_n())
    rng = np.random.default_rng(11)
    # two distinct keys far apart + a narrow cluster: nearly every value falls into ONE fine bin
    v = np.full(n, 100.0)
    v[:1000] = rng.normal(100.0, 1e-9, 1000)
    v[-1] = 1e12
    v[0] = -1e12
    values = torch.from_numpy(v).to(dev)
    masks = (1, 3, 0xFFFFFFFF, 5, 9, 17)
    q = (N.QueryC * len(masks))(*[N.QueryC(m, N.RANK_LERP_LO, 0.5, 0) for m in masks])
    out = torch.zeros(len(masks) * C.sizeof(N.QueryResultC), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    for _ in range(2):
        N.check(lib.kvm_column_select_static(values.data_ptr(), n, N.MEMBER_ALL, None, q, len(masks),
                                             out.data_ptr(), stream), "kvm_column_select_static")
        r = np.frombuffer(out.cpu().numpy().tobytes(), dtype=RESULT_DTYPE)
        assert (r["status"] == N.Q_OK).all(), r["status"]
        assert (r["n"] == n).all()
        expect = np.sort(v)[(n - 1) // 2]
        assert (r["value"] == expect).all()
