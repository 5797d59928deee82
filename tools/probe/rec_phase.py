import sys, ctypes
sys.path.insert(0, '/root/repo')
import torch
from kvmini_b200 import engine, synth, _native as N
dev = torch.device("cuda", 0)
cols = synth.request_columns_device(10_000_000, dev, seed=1234)
dlog = engine.DeviceLog.from_tensors(cols)
an = engine.LogAnalyzer(dev)
lib = N.lib()
lib.kvm_rec_phase_times.argtypes = [ctypes.c_void_p, ctypes.c_int32]
ph = (ctypes.c_uint64 * 3)()
for _ in range(3): an.analyze(dlog, [1.7e9 + 5.0])
torch.cuda.synchronize()
lib.kvm_rec_phase_times(ph, 1)
an.analyze(dlog, [1.7e9 + 5.0])
torch.cuda.synchronize()
lib.kvm_rec_phase_times(ph, 1)
print("per-CTA mean ns: prologue, body+drain, epilogue:", [int(x) / 296 for x in ph])
