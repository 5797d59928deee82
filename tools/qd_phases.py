#!/usr/bin/env python3
"""Phase time stamps of kvm_column_select_sampled (the 16 B/record percentile path): the kernels
write %globaltimer at their phase boundaries into the workspace; this prints the deltas.
  python tools/qd_phases.py [rows]"""
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import numpy as np
import torch


def main():
    from kvmini_b200 import _native as N
    from kvmini_b200 import synth
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
    dev = torch.device("cuda", 0)
    lib = N.lib()
    lib.kvm_column_select_sampled_dbg_offset.restype = __import__("ctypes").c_int64
    off = int(lib.kvm_column_select_sampled_dbg_offset())
    c = synth.request_columns_device(n, dev, seed=1234)
    lat, st = c["latency_ms"], c["status"]
    q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
    arr = (N.QueryC * len(q))(*[N.QueryC(*x) for x in q])
    res = torch.zeros(len(q) * 32, dtype=torch.uint8, device=dev)
    ws = torch.empty(lib.kvm_column_select_sampled_workspace_bytes(n), dtype=torch.uint8, device=dev)
    names = ["hist:entry", "hist:sample_done", "hist:cta0_done", "hist:last_cta", "hist:resolved", "collect:entry",
             "collect:cta0_streamed", "collect:cta0_done", "rank:entry"]
    rows = []
    for rep in range(6):
        N.check(lib.kvm_column_select_sampled(lat.data_ptr(), n, N.MEMBER_STATUS, st.data_ptr(), arr, len(q),
                                              res.data_ptr(), ws.data_ptr(), ws.numel(),
                                              torch.cuda.current_stream(dev).cuda_stream))
        torch.cuda.synchronize()
        d = ws[off:off + 128].view(torch.int64).cpu().numpy()
        rows.append(d.copy())
    d = rows[-1]
    t0 = int(d[0])
    names[7], names[8] = "collect:cta0_done", "rank:entry"
    idx = {0: 0, 1: 1, 2: 2, 3: 3, 4: 4, 5: 5, 6: 6, 7: 7, 8: 8}
    out = {names[i]: (int(d[j]) - t0) / 1000.0 for i, j in idx.items()}
    out["status"] = int(ws[0:4].view(torch.int32).item())
    print(json.dumps({"rows": n, "us_since_hist_entry": out}))


if __name__ == "__main__":
    main()
