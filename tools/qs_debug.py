import sys, struct
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import kvmini_b200
from kvmini_b200 import _native as N, engine
lib = N.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
rng = np.random.default_rng(1)
lat = rng.lognormal(5.5, 0.6, n)
status = np.where(rng.random(n) < 0.97, 200, 500).astype(np.int64)
d, st = torch.from_numpy(lat).cuda(), torch.from_numpy(status).cuda()
q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
arr = (N.QueryC * len(q))(*[N.QueryC(*x) for x in q])
ws = torch.zeros(lib.kvm_column_select_sampled_workspace_bytes(n), dtype=torch.uint8, device='cuda')
res = torch.zeros(8 * 32, dtype=torch.uint8, device='cuda')
rc = lib.kvm_column_select_sampled(d.data_ptr(), n, N.MEMBER_STATUS, st.data_ptr(), arr, len(q), res.data_ptr(), ws.data_ptr(), ws.numel(), None)
torch.cuda.synchronize()
print('rc', rc, lib.kvm_last_error())
raw = ws[:4096].cpu().numpy().tobytes()
status_, nq, nb, grid, cap = struct.unpack_from('<iiiiq', raw, 0)
print('status', status_, 'nq', nq, 'nb', nb, 'grid', grid, 'cap', cap)
off = 24
def unkey(k):
    bits = k ^ (0x8000000000000000 if (k >> 63) else 0xFFFFFFFFFFFFFFFF)
    return struct.unpack('<d', struct.pack('<Q', bits))[0]
for b in range(4):
    lo, hi, base, sh, act = struct.unpack_from('<QQQii', raw, off + 32 * b)
    print('bracket', b, hex(lo), hex(hi), sh, act, unkey(lo) if 0 < lo < 2**64 - 2 else None, unkey(hi) if 0 < hi < 2**64 - 2 else None)
off += 128
for i in range(8):
    f = struct.unpack_from('<iidiiqqqii', raw, off + 56 * i)
    print('q', i, f)
off += 56 * 8
print('list_n', struct.unpack_from('<8Q', raw, off))
ok = np.sort(lat[status == 200]); print('n_ok', len(ok), 'p50', ok[int(0.5 * (len(ok) - 1))])
