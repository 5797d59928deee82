"""Multi-GPU energy path: one process per GPU (SURVEY.md section 8e(4), configs[3]).

The reference integrates ONE window over ONE sample list (energy/collector.py:133-149) and lets
Prometheus add the GPUs of the served model (`sum(DCGM_FI_DEV_POWER_USAGE)`, collector.py:45).
Two ways a node's GPUs share that work here:

  trace-sharded   rank r owns the power trace of GPU r (8 traces -> 8 ranks).  Every rank
                  integrates its trace alone - no exchange on the data path; what is exchanged
                  is RESULTS, and they are added in GPU-index order on every rank:
                    scalar Wh per window   all-gather of (hi, lo) pairs, rank-ordered exact sum
                    per-interval energies  the interval index space is cut into `world` slices;
                                           every rank sends its energies of slice q to rank q
                                           (all-to-all), rank q adds the `world` contributions
                                           in rank order (kvm_ordered_sum_f64) and the summed
                                           slices are all-gathered back
                  so every value is ((E_0 + E_1) + E_2) + ... exactly like the single-GPU loop
                  over the traces; no floating-point all-reduce is used anywhere.
  time-sharded    one long trace cut into `world` contiguous time shards.  The segment that
                  crosses a cut belongs to the later shard, which gets the previous shard's last
                  non-null sample as a one-sample halo (all-gather of one (ts, watts) pair per
                  rank); per-window Wh of the shards are added in rank order.

The host logic below works on any backend (the CPU tests drive it over gloo with numpy
stand-ins for the kernels); the kernels are reached through energy_api.PowerTrace.
"""

from __future__ import annotations

import math
from fractions import Fraction
from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def _all_gather_f64(local: torch.Tensor, group=None) -> torch.Tensor:
    """[k] float64 per rank -> [world, k] on every rank (works on gloo and nccl)."""
    world = dist.get_world_size(group)
    out = torch.empty((world,) + tuple(local.shape), dtype=local.dtype, device=local.device)
    if local.device.type == "cuda":
        dist.all_gather_into_tensor(out.view(-1), local.contiguous().view(-1), group=group)
    else:
        dist.all_gather(list(out.unbind(0)), local.contiguous(), group=group)
    return out


def rank_ordered_total(pairs: Sequence[Tuple[float, float]]) -> float:
    """Sum of per-rank double-double values in rank order, rounded once (exact rational sum)."""
    tot = Fraction(0)
    for hi, lo in pairs:
        tot += Fraction(hi) + Fraction(lo)
    return float(tot)


def window_wh_across_ranks(local_dd: Sequence[Tuple[float, float]], device: torch.device, group=None) -> List[float]:
    """Per-window Wh of this rank's trace (or time shard) as (hi, lo) pairs -> per-window Wh of all
    ranks together, added in rank order; identical on every rank."""
    k = len(local_dd)
    loc = torch.tensor([x for p in local_dd for x in p], dtype=torch.float64, device=device)
    allv = _all_gather_f64(loc, group).cpu().tolist()
    return [rank_ordered_total([(allv[r][2 * w], allv[r][2 * w + 1]) for r in range(len(allv))]) for w in range(k)]


def ordered_sum(slices: torch.Tensor) -> torch.Tensor:
    """[world, n] -> [n]: ((s0 + s1) + s2) + ... elementwise.  On CUDA this is kvm_ordered_sum_f64
    (one pass, one write); on CPU tensors (tests) the same additions in the same order."""
    if slices.device.type == "cuda":
        from . import _native as N
        world, n = int(slices.shape[0]), int(slices.shape[1])
        out = torch.empty(n, dtype=torch.float64, device=slices.device)
        N.check(N.lib().kvm_ordered_sum_f64(slices.data_ptr(), world, n, n, out.data_ptr(),
                                            torch.cuda.current_stream(slices.device).cuda_stream),
                "kvm_ordered_sum_f64")
        return out
    out = slices[0].clone()
    for r in range(1, slices.shape[0]):
        out = out + slices[r]
    return out


def interval_energy_across_ranks(local: torch.Tensor, group=None) -> torch.Tensor:
    """local[n] = this rank's per-interval energies (its own trace) -> the rank-ordered sum over
    all ranks, complete on every rank."""
    world = dist.get_world_size(group)
    n = int(local.shape[0])
    if world == 1:
        return local
    per = (n + world - 1) // world
    padded = torch.zeros(world * per, dtype=local.dtype, device=local.device)
    padded[:n] = local
    recv = torch.empty_like(padded)
    if local.device.type == "cuda":
        dist.all_to_all_single(recv, padded, group=group)       # recv[r*per:(r+1)*per] = rank r's slice for me
    else:   # gloo has no all_to_all_single: gather every rank's full vector, keep my slice of each
        rank = dist.get_rank(group)
        allv = _all_gather_f64(padded, group)
        recv = allv[:, rank * per:(rank + 1) * per].contiguous().view(-1)
    mine = ordered_sum(recv.view(world, per))
    return _all_gather_f64(mine, group).view(-1)[:n]


def halo_for_rank(last_samples: Sequence[Tuple[float, float]], rank: int) -> Optional[Tuple[float, float]]:
    """last_samples[r] = (ts, watts) of rank r's last non-null sample (watts NaN: the shard has none).
    The halo of `rank` is the nearest earlier shard's last non-null sample."""
    for r in range(rank - 1, -1, -1):
        ts, w = last_samples[r]
        if not math.isnan(w):
            return ts, w
    return None


class TraceShardedEnergy:
    """One power trace per rank.  `trace` is this rank's energy_api.PowerTrace (or, in the CPU
    tests, any object with trapezoid_dd(windows) and interval_energy(start_ms, latency_ms))."""

    def __init__(self, trace, device: torch.device, group=None):
        self.trace, self.device, self.group = trace, device, group

    def window_wh(self, windows: Sequence[Tuple[float, float]]) -> List[float]:
        return window_wh_across_ranks(self.trace.trapezoid_dd(windows), self.device, self.group)

    def interval_energy(self, start_ms: torch.Tensor, latency_ms: torch.Tensor) -> torch.Tensor:
        return interval_energy_across_ranks(self.trace.interval_energy(start_ms, latency_ms), self.group)


class TimeShardedEnergy:
    """One long trace cut into contiguous time shards (rank order = time order).  make_trace(ts,
    watts) builds the per-shard trace object once the halo sample has been prepended."""

    def __init__(self, ts, watts, device: torch.device, make_trace: Callable, last_non_null: Tuple[float, float],
                 group=None):
        import numpy as np
        self.device, self.group = device, group
        rank = dist.get_rank(group)
        loc = torch.tensor(list(last_non_null), dtype=torch.float64, device=device)
        allv = _all_gather_f64(loc, group).cpu().tolist()
        halo = halo_for_rank([(a[0], a[1]) for a in allv], rank)
        if halo is not None:
            ts = np.concatenate([[halo[0]], ts])
            watts = np.concatenate([[halo[1]], watts])
        self.halo = halo
        self.trace = make_trace(ts, watts)

    def window_wh(self, windows: Sequence[Tuple[float, float]]) -> List[float]:
        """Windows whose edges fall inside a shard are clipped by sample membership exactly like the
        reference: a segment counts iff both its samples lie in the window - the halo sample makes
        the cross-shard segment available to (only) the later shard."""
        return window_wh_across_ranks(self.trace.trapezoid_dd(windows), self.device, self.group)
