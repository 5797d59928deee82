"""The peer-memory exchange primitives (csrc/kvm_peer.cu) on one GPU: two "ranks" are two buffers
of the same device, which is all the kernels see (base pointers).  The multi-process form is
covered by tests/dist_gpu_check.py on boxes with several GPUs."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _table(bufs):
    return (C.c_void_p * len(bufs))(*[b.data_ptr() for b in bufs])


def test_put_add_barrier_between_two_buffers():
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    lib = N.lib()
    dev = engine.require_cuda()
    st = engine._stream_ptr(dev)
    flag_bytes = int(lib.kvm_peer_flag_bytes())
    assert flag_bytes == 8 * (N.KVM_PEER_MAX_WORLD + 2)
    nbytes = 256 + 8 * 4096
    bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]
    peers = _table(bufs)
    rng = np.random.default_rng(3)
    # put: rank r's record lands at the same offset of both buffers
    recs = [torch.from_numpy(rng.integers(0, 1 << 60, 37)).to(dev) for _ in range(2)]
    for r in range(2):
        N.check(lib.kvm_peer_put(recs[r].data_ptr(), 37 * 8, 256 + r * 37 * 8, peers, 2, r, st))
    # add: both ranks add a sparse u64 array into the accumulator at offset 256 + 1024
    adds = []
    for r in range(2):
        a = rng.integers(0, 1 << 40, 1000)
        a[rng.random(1000) < 0.7] = 0
        adds.append(a)
        N.check(lib.kvm_peer_add_u64(torch.from_numpy(a).to(dev).data_ptr(), 1000, 256 + 1024 * 8, peers, 2, r, st))
    # barrier: both ranks arrive (stream-ordered on one device: rank 0's kernel would spin for
    # rank 1's arrival, so rank 1 must be able to run - use a second stream)
    side = torch.cuda.Stream(dev)
    for rep in range(3):
        with torch.cuda.stream(side):
            N.check(lib.kvm_peer_barrier(0, peers, 2, 1, engine._stream_ptr(dev)))
        N.check(lib.kvm_peer_barrier(0, peers, 2, 0, st))
    torch.cuda.synchronize(dev)
    for b in bufs:
        words = b.view(torch.int64).cpu().numpy()
        for r in range(2):
            assert np.array_equal(words[32 + r * 37: 32 + (r + 1) * 37], recs[r].cpu().numpy())
        assert np.array_equal(words[32 + 1024: 32 + 2024], adds[0] + adds[1])
        flags = words[: N.KVM_PEER_MAX_WORLD + 2]
        assert flags[0] == 3 and flags[1] == 3            # three arrivals from each rank
        assert flags[N.KVM_PEER_MAX_WORLD] == 3           # local epoch
        assert flags[N.KVM_PEER_MAX_WORLD + 1] == 0       # no timeout


def test_barrier_times_out_instead_of_hanging():
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    lib = N.lib()
    dev = engine.require_cuda()
    bufs = [torch.zeros(256, dtype=torch.uint8, device=dev) for _ in range(2)]
    old = lib.kvm_peer_set_timeout_s(0.5)
    try:
        N.check(lib.kvm_peer_barrier(0, _table(bufs), 2, 0, engine._stream_ptr(dev)))   # rank 1 never arrives
        torch.cuda.synchronize(dev)
        # the waiting rank tells EVERY rank (the late one included) which peer was missing
        for b in bufs:
            flags = b.view(torch.int64).cpu().numpy()
            assert flags[N.KVM_PEER_MAX_WORLD + 1] == 2           # 1 + the rank that was missing
        # sticky: the late rank's own barrier returns at once instead of waiting for a full time-out
        lib.kvm_peer_set_timeout_s(30.0)
        t0 = __import__("time").perf_counter()
        N.check(lib.kvm_peer_barrier(0, _table(bufs), 2, 1, engine._stream_ptr(dev)))
        N.check(lib.kvm_peer_barrier(0, _table(bufs), 2, 1, engine._stream_ptr(dev)))   # rank 0 is one epoch behind now
        torch.cuda.synchronize(dev)
        assert __import__("time").perf_counter() - t0 < 5.0
    finally:
        lib.kvm_peer_set_timeout_s(old)


def test_bad_arguments_are_rejected():
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    lib = N.lib()
    dev = engine.require_cuda()
    buf = torch.zeros(512, dtype=torch.uint8, device=dev)
    peers = _table([buf])
    assert lib.kvm_peer_put(buf.data_ptr(), 12, 0, peers, 1, 0, None) == -1          # size not a multiple of 8
    assert lib.kvm_peer_barrier(0, peers, 1, 1, None) == -1                          # rank outside the world
    assert lib.kvm_peer_barrier(0, peers, N.KVM_PEER_MAX_WORLD + 1, 0, None) == -1
    assert b"peer table" in lib.kvm_last_error()
