#!/usr/bin/env python3
"""Bare pinned-copy probe at N = 1/2/4/8 ranks: what does the PLATFORM give the
host-resident path when N processes each move a tick's bytes over PCIe at the same time, with no
kernel and none of this library's slicing in the way?

    python tools/pcie_probe.py                                   # one rank
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_probe.py

Every rank copies the bytes its shard of BASELINE configs[3] moves per tick (strong: 2^20 / N sessions,
weak: 2^20 each; 3840 B in + 640 B out per session), H2D and D2H concurrently, `--reps` times; ranks
start together (gloo barrier).  Rank 0 prints one JSON line with per-rank and aggregate GB/s."""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "voice-enabled-browser-automation_b200"))
import bench  # noqa: E402  (Dist: gloo barrier / gather)
import pcm_frontend  # noqa: E402
from pcm_frontend._capi import check  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sessions", type=int, default=1 << 20)
    ap.add_argument("--scaling", choices=["strong", "weak"], default="strong")
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    rank, local_rank, world = bench.dist_env()
    ctx = pcm_frontend.Context(local_rank)
    dist = bench.Dist(world, True, local_rank)
    S = args.sessions if args.scaling == "weak" else pcm_frontend.shard_sessions(args.sessions, world, rank)[1]
    h2d, d2h = S * 960 * 4, S * 320 * 2
    ms = C.c_float(0)
    out = {}
    for name, a, b in (("both", h2d, d2h), ("h2d_only", h2d, 0), ("d2h_only", 0, d2h)):
        dist.barrier()
        check(pcm_frontend.load().pcmb200_pcie_probe(ctx.handle, a, b, args.reps, C.byref(ms)), ctx.handle)
        per_rank = dist.gather(float(ms.value))
        if rank == 0:
            t = max(per_rank) * 1e-3
            out[name] = {"per_rank_ms": per_rank,
                         "h2d_gbs_per_rank": [a * args.reps / (m * 1e-3) / 1e9 for m in per_rank] if a else None,
                         "d2h_gbs_per_rank": [b * args.reps / (m * 1e-3) / 1e9 for m in per_rank] if b else None,
                         "aggregate_h2d_gbs": a * world * args.reps / t / 1e9 if args.scaling == "weak" else None,
                         "job_h2d_gbs": dist_total(a, world, args, per_rank) if a else None}
    if rank == 0:
        print(json.dumps({"probe": "bare pinned cudaMemcpyAsync, H2D and D2H in flight together, no kernels", "n_gpus": world,
                          "scaling": args.scaling, "sessions_this_rank": S, "h2d_bytes_per_round": h2d, "d2h_bytes_per_round": d2h,
                          "reps": args.reps, "results": out,
                          "samples_per_s_ceiling": args.sessions * (world if args.scaling == "weak" else 1) * 960 * args.reps /
                          (max(out["both"]["per_rank_ms"]) * 1e-3)}), flush=True)
    dist.close()
    ctx.close()


def dist_total(bytes_this_rank, world, args, per_rank):
    """Whole-job H2D GB/s: every rank's bytes over the slowest rank's time (rank 0's bytes stand for all: shards are equal)."""
    return bytes_this_rank * world * args.reps / (max(per_rank) * 1e-3) / 1e9


if __name__ == "__main__":
    main()
