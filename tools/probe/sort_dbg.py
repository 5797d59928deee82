import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from kvmini_b200 import _native as N
lib = N.lib(); dev = torch.device('cuda', 0)
for m in (4097, 4128, 4129, 6144, 20000, 300000, 2000003):
    rng = np.random.default_rng(m)
    keys = 1.7e9 + np.round(rng.uniform(0, 5000.0, m), 2)
    vals = np.arange(m, dtype=np.float64)
    k_in, v_in = torch.from_numpy(keys).to(dev), torch.from_numpy(vals).to(dev)
    k_out, v_out = torch.full((m,), -7.0, dtype=torch.float64, device=dev), torch.full((m,), -7.0, dtype=torch.float64, device=dev)
    ws = torch.zeros(lib.kvm_sort_pairs_workspace_bytes(m), dtype=torch.uint8, device=dev)
    N.check(lib.kvm_sort_pairs_f64(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(), m, ws.data_ptr(), ws.numel(), torch.cuda.current_stream(dev).cuda_stream))
    torch.cuda.synchronize()
    order = np.argsort(keys, kind='stable')
    got_k, got_v = k_out.cpu().numpy(), v_out.cpu().numpy()
    bad = np.flatnonzero(got_k != keys[order])
    plan = ws[:16640].cpu().numpy()
    run = plan[16384:16384+32].view(np.int32); ordr = plan[16416:16416+32].view(np.int32); ex = plan[16448:16452].view(np.int32)
    print(m, 'bad', bad.size, 'multiset_equal', np.array_equal(np.sort(got_k), np.sort(keys)), 'untouched', int((got_k == -7.0).sum()),
          'run', run.tolist(), 'executed', ex.tolist(), flush=True)
    if bad.size: print('   bad idx', bad[:8], 'got', got_k[bad[:4]], 'vals', got_v[bad[:4]])
