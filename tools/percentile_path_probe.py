import sys, time
sys.path.insert(0, '/root/repo')
import torch, numpy as np
from kvmini_b200 import engine, synth, _native as N
cols = synth.request_columns(10_000_000, seed=1234)
dev = torch.device('cuda', 0)
lat = torch.from_numpy(cols['latency_ms']).to(dev); st = torch.from_numpy(cols['status']).to(dev)
ops = engine.ColumnOps(dev)
q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
for _ in range(3): r = ops.order_stats(lat, q, member_mode=N.MEMBER_STATUS, member=st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
K = 20
for _ in range(K): r = ops.order_stats(lat, q, member_mode=N.MEMBER_STATUS, member=st)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
print('percentile path ms', ms, 'eff GB/s', 16e7 / ms / 1e6, 'frac of 6551', 16e7 / ms / 1e6 / 6551.4, [float(x['value']) for x in r])
