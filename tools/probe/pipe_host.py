#!/usr/bin/env python3
"""Where does a pipelined step go: GPU or host?  Times the host pieces of bench.py's loop."""
import os, sys, time
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
import torch
from kvmini_b200 import engine, synth, report, cost_api

dev = torch.device("cuda", 0)
n = 10_000_000
dlog = engine.DeviceLog.from_tensors(synth.request_columns_device(n, dev, seed=1234))
cold = [1.7e9 + 5.0]
pair = (engine.LogAnalyzer(dev, stream=torch.cuda.Stream(dev)), engine.LogAnalyzer(dev, stream=torch.cuda.Stream(dev)))

for an in pair:
    for _ in range(3):
        an.analyze(dlog, cold, use_graph=True)
torch.cuda.synchronize()
K = 200
t_submit = t_wait = t_post = 0.0
pending = None
t0 = time.perf_counter()
for it in range(K):
    a = time.perf_counter()
    nxt = pair[it & 1].analyze_async(dlog, cold, use_graph=True)
    b = time.perf_counter()
    t_submit += b - a
    if pending is not None:
        raw = pending()
        c = time.perf_counter()
        t_wait += c - b
        core = report.core_results(raw, len(cold))
        t_post += time.perf_counter() - c
    pending = nxt
raw = pending()
torch.cuda.synchronize()
tot = time.perf_counter() - t0
print("per step us: total %.1f submit %.1f wait+collect %.1f post %.1f" % (tot / K * 1e6, t_submit / K * 1e6, t_wait / K * 1e6, t_post / K * 1e6))
