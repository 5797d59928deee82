#!/usr/bin/env python3
"""Probe: can this stack capture NCCL collectives in a CUDA graph, and does ShardedAnalyzer's
captured sequence replay?  Run under torchrun with 2+ ranks and a short `timeout`; progress goes
to stderr line by line, a hang dumps the Python stack after 40 s (faulthandler)."""
import faulthandler
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def say(rank, *a):
    print(f"[rank {rank}]", *a, file=sys.stderr, flush=True)


def main():
    faulthandler.dump_traceback_later(45, exit=True)
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    world = int(os.environ["WORLD_SIZE"])
    mode = sys.argv[1] if len(sys.argv) > 1 else "global"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    x = torch.ones(1 << 14, dtype=torch.int64, device=dev)
    y = torch.zeros(world << 14, dtype=torch.int64, device=dev)
    dist.all_reduce(x)
    dist.all_gather_into_tensor(y, x)
    torch.cuda.synchronize()
    say(rank, "eager collectives ok")
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode=mode):
        x.add_(1)
        dist.all_reduce(x)
        dist.all_gather_into_tensor(y, x)
        y.mul_(2)
    say(rank, "capture ok")
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    say(rank, "replay ok", int(y[0]))

    from kvmini_b200 import dist_api, engine, synth
    cols = synth.request_columns_device(1_000_000, dev, seed=77 + rank, t_epoch_ms=1.7e12, t_offset_ms=rank * 3.6e6)
    dlog = engine.DeviceLog.from_tensors(cols)
    an = dist_api.ShardedAnalyzer(dev)
    an.capture_error_mode = mode
    cold = [1.7e9 + 5.0]
    a = an.analyze(dlog, cold)
    say(rank, "eager analyze ok")
    b = an.analyze(dlog, cold, use_graph=True)
    say(rank, "graph analyze (capture + first replay) ok")
    for _ in range(5):
        b = an.analyze(dlog, cold, use_graph=True)
    say(rank, "graph replays ok")
    same = all(p.v_lo == q.v_lo and p.v_hi == q.v_hi and p.n == q.n for p, q in zip(a.order.values(), b.order.values()))
    c = an.analyze(dlog, cold)      # eager again after replays
    say(rank, "eager after graph ok", same)
    dist.barrier()
    if rank == 0:
        print("NCCL_GRAPH_PROBE OK", mode)
    an.close()
    del g
    torch.cuda.synchronize()
    dist.barrier()
    say(rank, "graphs released")
    dist.destroy_process_group()
    say(rank, "process group destroyed")
    faulthandler.cancel_dump_traceback_later()


if __name__ == "__main__":
    main()
