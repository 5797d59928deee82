"""Multi-GPU run analysis: one process per GPU, the log sharded by time range.

SURVEY.md section 8(e).  Rank r holds a contiguous time slice of the log.  Nothing on the
data path waits on another rank inside a kernel; the host exchanges small tensors between
the stages with torch.distributed (NCCL on GPUs, gloo in the CPU tests):

  K1 partials   int64 sums  -> all_reduce(SUM)      key minima / maxima -> all_reduce(MIN/MAX)
                fp sums (double-double)             -> all_gather, added in RANK ORDER on the host
  quantiles     per-set digit histograms (7 x 2048 u64) -> all_reduce(SUM); every rank then
                resolves identically, compacts its own candidates, all_gathers the candidate
                arenas and finishes identically  => exact global order statistics
  histograms / series   int64 bin counts -> all_reduce(SUM)

All integer reductions are order-independent, the fp ones have a fixed order, so every rank
ends with bit-identical results.
"""

from __future__ import annotations

import ctypes as C
import math
from fractions import Fraction
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from . import _native as N
from . import engine

_SIGN = -(1 << 63)

# int64 fields of kvm_log_stats that add up across shards
_SUM_FIELDS = ("n_total", "n_success", "n_cold", "n_cold_success", "sum_completion_tokens",
               "sum_total_tokens", "sum_total_tokens_cold", "n_gen_rows")


def _key_to_signed(k: int) -> int:
    """u64 ordered key -> int64 with the same order (torch has no uint64 reductions)."""
    v = k ^ (1 << 63)
    return v - (1 << 64) if v >= (1 << 63) else v


def _signed_to_key(v: int) -> int:
    return (v + (1 << 64) if v < 0 else v) ^ (1 << 63)


def _exact_dd_total(pairs: Sequence[Sequence[float]]):
    """Sum of double-double partials in rank order, exactly (rational), back to (hi, lo)."""
    tot = Fraction(0)
    for hi, lo in pairs:
        tot += Fraction(hi) + Fraction(lo)
    hi = float(tot)
    lo = float(tot - Fraction(hi))
    return hi, lo


def combine_stats(local: N.LogStatsC, device: torch.device, group=None) -> N.LogStatsC:
    """Global kvm_log_stats from every rank's local one.  Works on any backend (the CPU tests
    drive it over gloo with world_size 2)."""
    world = dist.get_world_size(group)
    sums = torch.tensor([getattr(local, f) for f in _SUM_FIELDS] + list(local.sets.count),
                        dtype=torch.int64, device=device)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    kmin = torch.tensor([_key_to_signed(int(k)) for k in local.sets.kmin], dtype=torch.int64, device=device)
    kmax = torch.tensor([_key_to_signed(int(k)) for k in local.sets.kmax], dtype=torch.int64, device=device)
    dist.all_reduce(kmin, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=group)
    inf = math.inf
    lo = torch.tensor([inf if math.isnan(local.min_start_ms) else local.min_start_ms], dtype=torch.float64, device=device)
    hi = torch.tensor([-inf if math.isnan(local.max_end_ms) else local.max_end_ms], dtype=torch.float64, device=device)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    dd_local = torch.tensor([local.ttfb_sum_hi, local.ttfb_sum_lo, local.lat_sum_hi, local.lat_sum_lo],
                            dtype=torch.float64, device=device)
    dd_all = torch.empty(world * 4, dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(dd_all, dd_local, group=group) if device.type == "cuda" else \
        dist.all_gather(list(dd_all.view(world, 4).unbind(0)), dd_local, group=group)
    dd = dd_all.view(world, 4).cpu().tolist()

    out = N.LogStatsC.from_buffer_copy(bytes(local))
    s = sums.cpu().tolist()
    for i, f in enumerate(_SUM_FIELDS):
        setattr(out, f, s[i])
    for b in range(N.KVM_MAX_SETS):
        out.sets.count[b] = s[len(_SUM_FIELDS) + b]
    for c, (a, b) in enumerate(zip(kmin.cpu().tolist(), kmax.cpu().tolist())):
        out.sets.kmin[c] = _signed_to_key(a)
        out.sets.kmax[c] = _signed_to_key(b)
    lo_v, hi_v = float(lo.item()), float(hi.item())
    out.min_start_ms = math.nan if lo_v == inf else lo_v
    out.max_end_ms = math.nan if hi_v == -inf else hi_v
    out.ttfb_sum_hi, out.ttfb_sum_lo = _exact_dd_total([(r[0], r[1]) for r in dd])
    out.lat_sum_hi, out.lat_sum_lo = _exact_dd_total([(r[2], r[3]) for r in dd])
    return out


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class PeerFabric:
    """One exchange buffer per rank, mapped into every rank (torch symmetric memory does the
    allocation and the address exchange - plumbing); the data movement is libkvmini_b200's own
    kernels over NVLink (csrc/kvm_peer.cu).  Layout (bytes, 256-aligned regions):
      flags | stats[world] | hist | hist2 | series[3][cap] | post[world] | candidates[world][slot]
    hist..series are accumulators the peers add into: zeroed by the owner at the start of a step."""

    def __init__(self, device: torch.device, group, world: int, rank: int, stats_words: int, hist_words: int,
                 hist2_words: int, series_cap: int, post_words: int, slot_keys: int, sample_words: int = 16):
        import torch.distributed._symmetric_memory as symm
        lib = N.lib()
        self.world, self.rank = world, rank
        slot_keys = (slot_keys + 1) & ~1       # 16-byte aligned candidate slots
        self.series_cap, self.slot_keys = series_cap, slot_keys
        off = 0
        self.flags_off = off
        off = _align(off + int(lib.kvm_peer_flag_bytes()))
        self.stats_off, self.stats_stride = off, 8 * stats_words
        off = _align(off + world * self.stats_stride)
        self.sample_off, self.sample_stride = off, 8 * sample_words   # sample records of the one-pass form
        off = _align(off + world * self.sample_stride)
        self.hist_off = off
        off = _align(off + 8 * hist_words)
        self.hist2_off = off
        off = _align(off + 8 * hist2_words)
        self.series_off = off
        off = _align(off + 8 * 3 * series_cap)
        self.zero_end = off
        self.post_off, self.post_stride = off, 8 * post_words
        off = _align(off + world * self.post_stride)
        self.cand_off, self.cand_stride = off, 8 * slot_keys
        off = _align(off + world * self.cand_stride)
        self.nbytes = off
        self.buf = symm.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        pg = group if group is not None else dist.group.WORLD
        self.handle = symm.rendezvous(self.buf, pg)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        assert len(ptrs) == world and ptrs[rank] == self.buf.data_ptr()
        self.peers = (C.c_void_p * world)(*ptrs)
        dist.barrier(group=group)          # every buffer is zeroed before anybody signals into it
        torch.cuda.synchronize(device)
        self.base = self.buf.data_ptr()
        self.zero_view = self.buf[self.hist_off:self.zero_end]
        self.stats_all = self.buf[self.stats_off:self.stats_off + world * self.stats_stride].view(torch.int64)
        self.hist = self.buf[self.hist_off:self.hist_off + 8 * hist_words].view(torch.int64)
        self.hist2 = self.buf[self.hist2_off:self.hist2_off + 8 * hist2_words].view(torch.int64)
        self.series = self.buf[self.series_off:self.series_off + 8 * 3 * series_cap].view(torch.int64).view(3, series_cap)
        self.post_all = self.buf[self.post_off:self.post_off + world * self.post_stride].view(torch.int64)
        self.error_word = self.buf[self.flags_off + 8 * (N.KVM_PEER_MAX_WORLD + 1):
                                   self.flags_off + 8 * (N.KVM_PEER_MAX_WORLD + 2)].view(torch.int64)
        self.cand_ptrs = (C.c_void_p * world)(*[self.base + self.cand_off + r * self.cand_stride
                                                for r in range(world)])

    def put(self, src_ptr: int, nbytes: int, dst_off: int, st) -> None:
        N.check(N.lib().kvm_peer_put(src_ptr, nbytes, dst_off, self.peers, self.world, self.rank, st), "kvm_peer_put")

    def add(self, src_ptr: int, words: int, dst_off: int, st) -> None:
        N.check(N.lib().kvm_peer_add_u64(src_ptr, words, dst_off, self.peers, self.world, self.rank, st),
                "kvm_peer_add_u64")

    def barrier(self, st) -> None:
        N.check(N.lib().kvm_peer_barrier(self.flags_off, self.peers, self.world, self.rank, st), "kvm_peer_barrier")


class ShardedAnalyzer(engine.LogAnalyzer):
    """LogAnalyzer whose analyze() returns GLOBAL results for a log sharded over the ranks of
    the default process group (one rank per GPU).  Everything between the first launch and the
    final read-back is stream-ordered: collectives are enqueued between the kernels, the host
    synchronises once at the end."""

    GATHER_KEYS = 1 << 17   # candidate keys exchanged per rank (1 MB); more -> one exact retry

    def __init__(self, device: Optional[torch.device] = None, group=None, exchange: str = "auto",
                 stream: Optional[torch.cuda.Stream] = None):
        """exchange: "peer"  - the intermediate results travel by our own kernels over NVLink peer
        memory (PeerFabric); "nccl" - by torch.distributed collectives; "auto" - peer when the
        symmetric-memory mapping can be set up on this node, else nccl.
        stream: own stream, so that two analyzers can overlap (peer exchange only: collectives of
        one communicator must not be issued from two streams at once)."""
        super().__init__(device, stream)
        assert exchange in ("auto", "peer", "nccl")
        self.exchange = exchange
        self._fabric: Optional[PeerFabric] = None
        self._fabric_slot = self.GATHER_KEYS
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self._bins = int(self.lib.kvm_select_bins())
        dev = self.device
        words = int(self.lib.kvm_log_stats_pack_words())
        self._pack_local = torch.zeros(words, dtype=torch.int64, device=dev)
        self._pack_all = torch.zeros(self.world * words, dtype=torch.int64, device=dev)
        # after the compact pass: [bucket counts | histogram edge counters | dropped] per rank
        self._post_words = int(self.lib.kvm_select_post_words())
        self._views: dict = {}
        self._post_local = torch.zeros(self._post_words, dtype=torch.int64, device=dev)
        self._post_all = torch.zeros(self.world * self._post_words, dtype=torch.int64, device=dev)
        self._used_h = torch.zeros(1, dtype=torch.int64).pin_memory()
        self._counts_all = torch.zeros(self.world * N.KVM_MAX_QUERIES, dtype=torch.int64, device=dev)
        self._gather_n = self.GATHER_KEYS
        self._arena_all: Optional[torch.Tensor] = None
        self._hist_glob = torch.zeros(N.KVM_MAX_HISTS * N.KVM_HIST_MAX_BINS, dtype=torch.int64, device=dev)
        pw = int(self.lib.kvm_log_fused_preset_words())
        self._sample_local = torch.zeros(pw, dtype=torch.int64, device=dev)
        self._preset_d = torch.zeros(pw, dtype=torch.int64, device=dev)
        # a refine round costs about one more 9 B/row pass and shrinks the candidate exchange
        # (world x candidates x 8 B into every rank) and the finish to almost nothing; measured on
        # 2 x 1e7 rows it already pays at 64K candidates (profiles/r1_scaling.md)
        self.refine_min_keys = 1 << 16

    def _refine_rounds(self, n_rows: int) -> int:
        """Sharded runs always launch one refine round: the shard sizes of the other ranks are
        unknown to this host, the device decides from the global bucket sizes."""
        return 1 if self.refine_rounds is None else int(self.refine_rounds)

    def _wrap(self, ptr: int, count: int, dtype=torch.int64) -> torch.Tensor:
        """A tensor view over `count` 8-byte words inside the select workspace."""
        off = ptr - self._sel_ws.data_ptr()
        return self._sel_ws[off: off + 8 * count].view(dtype)

    # ---- peer-memory exchange -------------------------------------------------------------
    def _ensure_fabric(self, want_series: bool) -> None:
        """(Re)build the exchange buffers when the series capacity or the candidate slot changed.
        Collective: every rank takes the same decisions (they derive from global quantities)."""
        if self.exchange == "nccl":
            return
        cap = int(self._series_d.shape[1]) if (want_series and self._series_d is not None) else 0
        fb = self._fabric
        if fb is not None and fb.series_cap >= cap and fb.slot_keys >= self._fabric_slot:
            return
        if fb is not None:
            self.close()
            self._fabric = None
        try:
            self._fabric = PeerFabric(self.device, self.group, self.world, self.rank,
                                      int(self.lib.kvm_log_stats_pack_words()), N.KVM_MAX_SETS * self._bins,
                                      int(self.lib.kvm_select_hist2_words()), max(cap, 1), self._post_words,
                                      self._fabric_slot, int(self.lib.kvm_log_fused_preset_words()))
            self._peer_err_h = torch.zeros(1, dtype=torch.int64).pin_memory()
        except Exception as exc:   # no symmetric memory on this node / torch build
            if self.exchange == "peer":
                raise
            import warnings
            warnings.warn(f"peer-memory exchange unavailable ({exc!r}); using NCCL collectives")
            self.exchange = "nccl"
            self._fabric = None

    def _enqueue_peer(self, dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c):
        """One sharded analysis with the exchanges done by kvm_peer_* kernels (no collective call)."""
        lib, tm, dev, fb = self.lib, self.timer, self.device, self._fabric
        st = engine._stream_ptr(dev)
        world, rank = self.world, self.rank
        # accumulators the peers add into: cleared before this rank's first barrier of the step -
        # the peers' adds of this step start only after they have passed that barrier
        fb.zero_view.zero_()
        fused = bool(self.use_fused) and not getattr(self, "_fused_off", False)
        self.last_path = "fused" if fused else "two-pass"
        if not fused:
            self._launch_scan(dlog, cold_times)
            N.check(lib.kvm_log_stats_pack_words_into(self._stats_ptr, self._pack_local.data_ptr(), st))
            with tm.stage("exchange_stats"):
                fb.put(self._pack_local.data_ptr(), fb.stats_stride, fb.stats_off + rank * fb.stats_stride, st)
                fb.barrier(st)
            N.check(lib.kvm_log_stats_combine(self._stats_ptr, fb.base + fb.stats_off, world, st))
            self.launches += 4
        extras = N.LogExtrasC()
        if want_series:
            extras.series_window_ms = float(series_window_s) * 1000.0
            extras.series_max_bins = int(self._series_d.shape[1])
            extras.series = self._series_d.data_ptr()
            extras.series_dropped = self._dropped_ptr
        extras.digit_records = self._digits.data_ptr()
        ws, fl = self._sel_ws.data_ptr(), self._flags.data_ptr()
        vk = (ws, "peer")
        views = self._views.get(vk)
        if views is None:
            self._views.clear()
            views = (self._wrap(lib.kvm_select_hist_ptr(ws), N.KVM_MAX_SETS * self._bins), None,
                     self._wrap(lib.kvm_select_bucket_offsets_ptr(ws), N.KVM_MAX_QUERIES + 1),
                     self._wrap(lib.kvm_select_hist2_ptr(ws), int(lib.kvm_select_hist2_words())))
            self._views[vk] = views
        hist, _, offsets, hist2 = views
        rounds = self._refine_rounds(dlog.n)
        if fused:
            # one-pass form: the ranks first agree on the digit parameters (a 512-row sample per
            # shard, 88 bytes exchanged), then ONE pass per shard produces statistics AND histograms,
            # which travel together: two exchange points instead of the two-pass form's three
            cold_d, nc = self._cold_device(cold_times)
            N.check(lib.kvm_log_fused_sample(C.byref(log_c), self._sample_local.data_ptr(), st), "kvm_log_fused_sample")
            with tm.stage("exchange_sample"):
                fb.put(self._sample_local.data_ptr(), fb.sample_stride, fb.sample_off + rank * fb.sample_stride, st)
                fb.barrier(st)
            N.check(lib.kvm_log_fused_preset_combine(fb.base + fb.sample_off, world, fb.sample_stride // 8,
                                                     self._preset_d.data_ptr(), st), "kvm_log_fused_preset_combine")
            with tm.stage("log_fused"):
                N.check(lib.kvm_log_fused_pass_preset(C.byref(log_c), cold_d.data_ptr(), nc, 30.0, fl, self._stats_ptr,
                                                      C.byref(extras), self._scan_ws.data_ptr(), self._scan_ws.numel(),
                                                      ws, self._sel_ws.numel(), arena, self._preset_d.data_ptr(), st),
                        "kvm_log_fused_pass_preset")
            N.check(lib.kvm_log_stats_pack_words_into(self._stats_ptr, self._pack_local.data_ptr(), st))
            with tm.stage("exchange_hist"):
                fb.put(self._pack_local.data_ptr(), fb.stats_stride, fb.stats_off + rank * fb.stats_stride, st)
                fb.add(hist.data_ptr(), hist.numel(), fb.hist_off, st)
                fb.barrier(st)
                hist.copy_(fb.hist, non_blocking=True)
            N.check(lib.kvm_log_stats_combine(self._stats_ptr, fb.base + fb.stats_off, world, st))
            with tm.stage("select_plan"):
                N.check(lib.kvm_select_plan_preset(self._stats_ptr + N.LogStatsC.sets.offset, queries, nq,
                                                   C.byref(hspec), ws, self._sel_ws.numel(), arena, rounds,
                                                   int(self.refine_min_keys), st), "kvm_select_plan_preset")
            self.launches += 9
        else:
            with tm.stage("select_plan"):
                N.check(lib.kvm_select_plan_refine(self._stats_ptr + N.LogStatsC.sets.offset, queries, nq,
                                                   C.byref(hspec), ws, self._sel_ws.numel(), arena, rounds,
                                                   int(self.refine_min_keys), st), "kvm_select_plan")
            with tm.stage("select_hist"):
                N.check(lib.kvm_log_select_hist(C.byref(log_c), fl, self._stats_ptr, C.byref(extras), ws, st),
                        "kvm_log_select_hist")
            with tm.stage("exchange_hist"):
                fb.add(hist.data_ptr(), hist.numel(), fb.hist_off, st)
                fb.barrier(st)
                hist.copy_(fb.hist, non_blocking=True)
        with tm.stage("select_resolve"):
            N.check(lib.kvm_select_resolve(ws, st), "kvm_select_resolve")
        self.launches += 2
        assert rounds <= 1, "the peer exchange clears its hist2 accumulator once per step"
        for _ in range(rounds):
            with tm.stage("select_refine_hist"):
                N.check(lib.kvm_log_select_refine_hist(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                        "kvm_log_select_refine_hist")
            with tm.stage("exchange_hist2"):
                fb.add(hist2.data_ptr(), hist2.numel(), fb.hist2_off, st)
                fb.barrier(st)
                hist2.copy_(fb.hist2, non_blocking=True)
            with tm.stage("select_refine"):
                N.check(lib.kvm_select_refine(ws, st), "kvm_select_refine")
            self.launches += 4
        with tm.stage("select_compact"):
            N.check(lib.kvm_log_select_compact(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                    "kvm_log_select_compact")
        N.check(lib.kvm_select_post_pack(ws, self._dropped_ptr if want_series else None,
                                         self._post_local.data_ptr(), st))
        with tm.stage("exchange_candidates"):
            fb.put(self._post_local.data_ptr(), fb.post_stride, fb.post_off + rank * fb.post_stride, st)
            N.check(lib.kvm_peer_put_candidates(ws, fb.cand_off + rank * fb.cand_stride, fb.slot_keys,
                                                fb.peers, world, rank, st), "kvm_peer_put_candidates")
            if want_series:
                fb.add(self._series_d.data_ptr(), 3 * int(self._series_d.shape[1]), fb.series_off, st)
                self.launches += 1
            fb.barrier(st)
        N.check(lib.kvm_select_post_unpack(ws, fb.base + fb.post_off, world, self._counts_all.data_ptr(),
                                           self._dropped_ptr if want_series else None, st))
        with tm.stage("select_finish"):
            N.check(lib.kvm_select_finish(ws, fb.cand_ptrs, world, self._counts_all.data_ptr(),
                                          self._results_ptr, self._hist_ptr, st), "kvm_select_finish")
        self.launches += 5 + 5
        self._out_d[self._OFF_STATUS:self._OFF_STATUS + 4].copy_(self._sel_ws[0:4], non_blocking=True)
        self._out_h.copy_(self._out_d, non_blocking=True)
        self._used_h.copy_(offsets[N.KVM_MAX_QUERIES:N.KVM_MAX_QUERIES + 1], non_blocking=True)
        self._peer_err_h.copy_(fb.error_word, non_blocking=True)
        if want_series:
            p = self._series_h.shape[1]
            self._series_h.copy_(fb.series[:, :p], non_blocking=True)

    def _enqueue_sharded(self, dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c):
        """One sharded analysis, stream-ordered (kernels + exchanges + read-back copies)."""
        if self._fabric is not None:
            return self._enqueue_peer(dlog, cold_times, queries, nq, hspec, want_series, series_window_s,
                                      arena, log_c)
        lib, tm, dev, grp = self.lib, self.timer, self.device, self.group
        st = engine._stream_ptr(dev)
        # ---- K1 on the shard, then the shard statistics become global ones on the device
        self._launch_scan(dlog, cold_times)
        N.check(lib.kvm_log_stats_pack_words_into(self._stats_ptr, self._pack_local.data_ptr(), st))
        with tm.stage("allgather_stats"):
            dist.all_gather_into_tensor(self._pack_all, self._pack_local, group=grp)
        N.check(lib.kvm_log_stats_combine(self._stats_ptr, self._pack_all.data_ptr(), self.world, st))
        self.launches += 2
        # ---- select with global digit parameters
        extras = N.LogExtrasC()
        if want_series:
            extras.series_window_ms = float(series_window_s) * 1000.0
            extras.series_max_bins = int(self._series_d.shape[1])
            extras.series = self._series_d.data_ptr()
            extras.series_dropped = self._dropped_ptr
        extras.digit_records = self._digits.data_ptr()
        ws, fl = self._sel_ws.data_ptr(), self._flags.data_ptr()
        vk = (ws, self._gather_n)
        views = self._views.get(vk)
        if views is None:
            self._views.clear()
            views = (self._wrap(lib.kvm_select_hist_ptr(ws), N.KVM_MAX_SETS * self._bins),
                     self._wrap(lib.kvm_select_arena_ptr(ws), self._gather_n),
                     self._wrap(lib.kvm_select_bucket_offsets_ptr(ws), N.KVM_MAX_QUERIES + 1),
                     self._wrap(lib.kvm_select_hist2_ptr(ws), int(lib.kvm_select_hist2_words())))
            self._views[vk] = views
        hist, arena_local, _, hist2 = views
        rounds = self._refine_rounds(dlog.n)
        with tm.stage("select_plan"):
            N.check(lib.kvm_select_plan_refine(self._stats_ptr + N.LogStatsC.sets.offset, queries, nq,
                                               C.byref(hspec), ws, self._sel_ws.numel(), arena, rounds,
                                               int(self.refine_min_keys), st), "kvm_select_plan")
        with tm.stage("select_hist"):
            N.check(lib.kvm_log_select_hist(C.byref(log_c), fl, self._stats_ptr, C.byref(extras), ws, st),
                    "kvm_log_select_hist")
        with tm.stage("allreduce_hist"):
            dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=grp)
        with tm.stage("select_resolve"):
            N.check(lib.kvm_select_resolve(ws, st), "kvm_select_resolve")
        # ---- refine: whether a round does work is decided on the device from the GLOBAL bucket
        # sizes, so every rank takes the same branch and the collective below always matches
        for _ in range(rounds):
            with tm.stage("select_refine_hist"):
                N.check(lib.kvm_log_select_refine_hist(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                        "kvm_log_select_refine_hist")
            with tm.stage("allreduce_hist2"):
                dist.all_reduce(hist2, op=dist.ReduceOp.SUM, group=grp)
            with tm.stage("select_refine"):
                N.check(lib.kvm_select_refine(ws, st), "kvm_select_refine")
            self.launches += 2
        with tm.stage("select_compact"):
            N.check(lib.kvm_log_select_compact(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                    "kvm_log_select_compact")
        # ---- candidates of every rank -> every rank (fixed-size exchange, no host round trip)
        N.check(lib.kvm_select_post_pack(ws, self._dropped_ptr if want_series else None,
                                         self._post_local.data_ptr(), st))
        with tm.stage("allgather_candidates"):
            dist.all_gather_into_tensor(self._post_all, self._post_local, group=grp)
            dist.all_gather_into_tensor(self._arena_all[: self.world * self._gather_n], arena_local, group=grp)
            if want_series:
                dist.all_reduce(self._series_d, op=dist.ReduceOp.SUM, group=grp)
        N.check(lib.kvm_select_post_unpack(ws, self._post_all.data_ptr(), self.world,
                                           self._counts_all.data_ptr(),
                                           self._dropped_ptr if want_series else None, st))
        N.check(lib.kvm_select_require_capacity(ws, self._gather_n, st))   # finish stays inside the slots
        self.launches += 3
        ptrs = (C.c_void_p * self.world)(*[self._arena_all.data_ptr() + 8 * r * self._gather_n
                                           for r in range(self.world)])
        with tm.stage("select_finish"):
            # global fine histogram + global edge counters -> global value histograms on every rank
            N.check(lib.kvm_select_finish(ws, ptrs, self.world, self._counts_all.data_ptr(),
                                          self._results_ptr, self._hist_ptr, st), "kvm_select_finish")
        self.launches += 5
        # ---- one read-back: outbox (+status), gathered fp partials, arena fill, series prefix
        offsets = views[2]
        self._out_d[self._OFF_STATUS:self._OFF_STATUS + 4].copy_(self._sel_ws[0:4], non_blocking=True)
        self._out_h.copy_(self._out_d, non_blocking=True)
        self._used_h.copy_(offsets[N.KVM_MAX_QUERIES:N.KVM_MAX_QUERIES + 1], non_blocking=True)
        if want_series:
            p = self._series_h.shape[1]
            self._series_h.copy_(self._series_d[:, :p], non_blocking=True)

    def _run(self, dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c, use_graph):
        """Eager launch, or replay of a captured CUDA graph (kernels AND NCCL collectives) when the
        same buffers are analysed again."""
        if not (use_graph and isinstance(self.timer, engine._NoTimer)):
            self._enqueue_sharded(dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c)
            return
        key = (tuple(dlog.cols[k].data_ptr() for k in engine.F64_COLS + engine.I64_COLS), dlog.n,
               tuple(sorted(float(c) for c in cold_times)), hspec.num_bins, series_window_s, id(queries), arena,
               self._sel_ws.data_ptr(), self._flags.data_ptr(), self._gather_n,
               self._refine_rounds(dlog.n), self.refine_min_keys,
               bool(self.use_fused) and not getattr(self, "_fused_off", False),
               self._fabric.base if self._fabric is not None else 0,
               self._series_d.data_ptr() if want_series else 0,
               self._arena_all.data_ptr() if self._arena_all is not None else 0)
        entry = self._graphs.get(key)
        if entry is None:
            self._enqueue_sharded(dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c)
            torch.cuda.current_stream(self.device).synchronize()
            g = torch.cuda.CUDAGraph()
            self.launches = 0
            with torch.cuda.graph(g, capture_error_mode=getattr(self, "capture_error_mode", "global")):
                self._enqueue_sharded(dlog, cold_times, queries, nq, hspec, want_series, series_window_s, arena, log_c)
            if len(self._graphs) >= 4:
                self._graphs.pop(next(iter(self._graphs)))
            entry = (g, self.launches, self.last_path)
            self._graphs[key] = entry
        entry[0].replay()
        self.launches = entry[1]
        self.last_path = entry[2]

    def analyze_async(self, dlog: engine.DeviceLog, cold_times: Sequence[float] = (), num_bins: int = 20,
                      series_window_s: Optional[float] = 1.0, max_series_bins: int = 1 << 24,
                      quantiles: Sequence[engine.Quantile] = engine.ANALYZE_QUANTILES,
                      use_graph: bool = False):
        """Submit one sharded analysis (on this analyzer's stream when it has one) and return the
        callable that waits for it and builds the global RawAnalysis.  Every rank must submit and
        collect in the same order; retries (arena / series capacity) are decided from global
        quantities, so all ranks take them together."""
        dev = self.device
        assert num_bins <= N.KVM_HIST_MAX_BINS, "sharded analysis folds the histograms into the select"
        queries = self._queries(quantiles)
        nq = len(queries)
        hspec = self.hist_spec(num_bins)
        want_series = series_window_s is not None
        self._ensure(dlog.n)
        state = {"arena": self._sel_cap}
        log_c = dlog.c_struct()

        def submit():
            self.launches = 0
            if want_series and (self._series_d is None or self._series_d.shape[1] < self._series_cap):
                self._series_d = torch.empty((3, self._series_cap), dtype=torch.int64, device=dev)
                self._series_h = torch.empty((3, min(self._series_cap, self.SERIES_PREFIX)),
                                             dtype=torch.int64).pin_memory()
            self._ensure_fabric(want_series)
            if self._fabric is None and (self._arena_all is None
                                         or self._arena_all.shape[0] < self.world * self._gather_n):
                self._arena_all = torch.empty(self.world * self._gather_n, dtype=torch.int64, device=dev)
            args = (dlog, cold_times, queries, nq, hspec, want_series, series_window_s, state["arena"], log_c,
                    use_graph)
            if self.stream is None:
                self._run(*args)
                self._done.record(torch.cuda.current_stream(dev))
            else:
                self.stream.wait_stream(torch.cuda.current_stream(dev))   # inputs were produced there
                with torch.cuda.stream(self.stream):
                    self._run(*args)
                    self._done.record(self.stream)

        submit()

        def result() -> engine.RawAnalysis:
            while True:
                self._done.synchronize()       # the one synchronisation
                out = self._out_h.numpy().tobytes()
                rc = int(np.frombuffer(out, dtype=np.int32, count=1, offset=self._OFF_STATUS)[0])
                stats = N.LogStatsC.from_buffer_copy(out[:C.sizeof(N.LogStatsC)])
                n_global = int(stats.n_total)
                used = int(self._used_h[0])   # total of the GLOBAL bucket sizes (valid on overflow too)
                if self._fabric is not None and int(self._peer_err_h[0]) != 0:
                    raise N.NativeError(f"peer barrier timed out waiting for rank {int(self._peer_err_h[0]) - 1}")
                if rc == N.KVM_ERR_ARENA:       # the local arena is laid out with the global bucket sizes
                    grown = False
                    if used > state["arena"]:
                        state["arena"] = max(2 * state["arena"], used + 1024)
                        self._ensure(dlog.n, state["arena"])
                        grown = True
                    if self._fabric is not None and used > self._fabric.slot_keys:
                        self._fabric_slot = used + 1024
                        grown = True
                    if self._fabric is None and used > self._gather_n:
                        self._gather_n = used + 1024   # (`used` sums GLOBAL bucket sizes: same on every rank)
                        if state["arena"] < self._gather_n:
                            state["arena"] = self._gather_n
                            self._ensure(dlog.n, state["arena"])
                        grown = True
                    if not grown:
                        raise N.NativeError("select reported an arena overflow that no buffer explains")
                    submit()
                    continue
                if rc != N.KVM_OK:
                    raise N.NativeError(f"select failed with status {rc}")
                if self._fabric is None and used > self._gather_n:
                    # more candidates than the fixed exchange carries: redo exactly
                    self._gather_n = used + 1024   # (`used` sums GLOBAL bucket sizes: same on every rank)
                    if state["arena"] < self._gather_n:
                        state["arena"] = self._gather_n
                        self._ensure(dlog.n, state["arena"])
                    submit()
                    continue
                series = None
                if (want_series and self.last_path == "fused" and self._fabric is not None and n_global > 0
                        and not math.isnan(stats.min_start_ms)
                        and np.frombuffer(out, dtype=np.float64, count=1, offset=self._OFF_DROPPED + 16)[0]
                        != stats.min_start_ms):
                    # the one-pass kernels binned the series relative to the minimum of the ranks' samples,
                    # which is not the log's minimum (rows out of time order): every rank sees the same
                    # two global numbers and redoes the analysis with the two-pass form
                    self._fused_off = True
                    try:
                        submit()
                    finally:
                        self._fused_off = False
                    continue
                if want_series and n_global > 0 and not math.isnan(stats.min_start_ms):
                    t0, t1 = stats.min_start_ms / 1000.0, stats.max_end_ms / 1000.0
                    nb = max(1, int(math.ceil((t1 - t0) / series_window_s)) + 1)
                    dropped = int(np.frombuffer(out, dtype=np.int64, count=1, offset=self._OFF_DROPPED)[0])
                    if dropped != 0:
                        if nb > max_series_bins:
                            raise N.NativeError("window series needs more than max_series_bins windows")
                        self._series_cap = nb + 1
                        submit()
                        continue
                    nb = min(nb, self._series_d.shape[1])
                    series_glob = self._fabric.series if self._fabric is not None else self._series_d
                    series = (self._series_h[:, :nb].numpy().copy() if nb <= self._series_h.shape[1]
                              else series_glob[:, :nb].cpu().numpy())
                break
            # (fp sums were combined on the device in rank order - identical on every rank)
            res = np.frombuffer(out, dtype=engine.RESULT_DTYPE, count=nq, offset=self._OFF_RESULTS)
            hist_counts = np.frombuffer(out, dtype=np.int64, count=4 * num_bins,
                                        offset=self._OFF_HIST).reshape(4, num_bins).copy()
            return engine.RawAnalysis(stats=stats, order=engine.results_to_orderstats(res, quantiles),
                                      hist_counts=hist_counts, num_bins=num_bins, series=series,
                                      series_window_s=series_window_s or 1.0, flags=self._flags[:dlog.n])

        return result

    def analyze(self, dlog: engine.DeviceLog, cold_times: Sequence[float] = (), num_bins: int = 20,
                series_window_s: Optional[float] = 1.0, max_series_bins: int = 1 << 24,
                quantiles: Sequence[engine.Quantile] = engine.ANALYZE_QUANTILES,
                use_graph: bool = False) -> engine.RawAnalysis:
        return self.analyze_async(dlog, cold_times, num_bins, series_window_s, max_series_bins, quantiles,
                                  use_graph)()
