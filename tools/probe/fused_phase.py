#!/usr/bin/env python3
"""Phase stamps of the one-pass kernel (kvm_log_fused_phase_us):  python tools/probe/fused_phase.py [rows]"""
import ctypes
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
import torch
from kvmini_b200 import engine, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
dev = torch.device("cuda", 0)
dlog = engine.DeviceLog.from_tensors(synth.request_columns_device(n, dev, seed=1234))
an = engine.LogAnalyzer(dev)
out = (ctypes.c_double * 4)()
for _ in range(4):
    an.analyze(dlog, [1.7e9 + 5.0], series_window_s=1.0)
    torch.cuda.synchronize()
    an.lib.kvm_log_fused_phase_us(an._scan_ws.data_ptr(), out)
    print(n, "us since first CTA entry: set-up %.1f  streamed %.1f  flushed %.1f  stats %.1f" % tuple(out))
