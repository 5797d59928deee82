#!/usr/bin/env python3
"""Run under torchrun (one rank per GPU): sharded analysis == analysis of the whole log.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tests/dist_gpu_check.py [rows]
Every rank builds the same seeded log, keeps its time slice, runs ShardedAnalyzer; the global
results must equal the numpy oracle on the full log bit for bit (mean TTFT <= 1e-12)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from conftest import diff_keys, same
    from kvmini_b200 import dist_api, engine, report, synth
    from oracle import np_port

    failures = []
    # (seed, uneven shards, refine_min_keys, candidate keys per rank in the exchange, exchange mode)
    #   refine_min_keys = 0 forces the refine round (it is launched in every sharded run; the device
    #   skips it for small logs); a tiny candidate slot forces the "redo with an exact size" path
    cases = ((5, False, None, None, "nccl"), (6, True, None, None, "nccl"), (7, True, 0, None, "nccl"),
             (8, False, None, 64, "nccl"), (9, True, 0, 16, "nccl"),
             (5, False, None, None, "peer"), (6, True, None, None, "peer"), (7, True, 0, None, "peer"),
             (8, False, None, 64, "peer"), (9, True, 0, 16, "peer"))
    if os.environ.get("KVM_DIST_EXCHANGE"):
        cases = tuple(c for c in cases if c[4] == os.environ["KVM_DIST_EXCHANGE"])
    for seed, uneven, refine_min, gather, exchange in cases:
        full = synth.request_columns(rows, seed=seed, span_ms=600_000.0)
        rng = np.random.default_rng(seed)
        full["ttfb_ms"][rng.integers(0, rows, rows // 100)] = np.nan
        full["completion_tokens"][rng.integers(0, rows, rows // 80)] = 0
        t0 = float(full["start_ms"].min()) / 1000.0
        cold = [t0 + 3.0, t0 + 290.0]
        # uneven shards (odd boundaries -> unaligned columns on some ranks)
        cuts = [0] + [(r * rows // world) + (7 * r + 1 if uneven else 0) for r in range(1, world)] + [rows]
        lo, hi = cuts[rank], cuts[rank + 1]
        shard = {k: torch.from_numpy(np.ascontiguousarray(v[lo:hi])).to(dev) for k, v in full.items()}
        print(f"[rank {rank}] case seed={seed} uneven={uneven} refine_min={refine_min} slot={gather} {exchange}",
              file=sys.stderr, flush=True)
        an = dist_api.ShardedAnalyzer(dev, exchange=exchange)
        if refine_min is not None:
            an.refine_min_keys = refine_min
        if gather is not None:
            an._gather_n = gather
            an._fabric_slot = gather
        raw = an.analyze(engine.DeviceLog.from_tensors(shard), cold)
        got = report.core_results(raw, len(cold))
        exp = np_port.analyze(full, cold)
        d = diff_keys(got, {k: exp[k] for k in got}, rel={"time_to_first_token_ms": 1e-12})
        if not same(report.trim_series(raw), np_port.window_series(full)):
            d.append("timeseries differs")
        failures += [f"{exchange}/{seed}: {k}" for k in d]
        if seed in (5, 7):   # the same analysis replayed from a CUDA graph (kernels + collectives)
            for _ in range(3):
                rep = report.core_results(an.analyze(engine.DeviceLog.from_tensors(shard), cold, use_graph=True),
                                          len(cold))
            failures += ["graph: " + k for k in diff_keys(rep, got)]
        an.close()           # captured collectives must be released before the group is destroyed
    flag = torch.tensor([len(failures)], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("DIST_CHECK", "OK" if int(flag.item()) == 0 else "FAIL", f"world={world} rows={rows}")
    if failures:
        print(f"rank {rank}:", failures[:6])
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
