#!/usr/bin/env python3
"""Tuning sweep of the device-resident frame kernels (one process, buffers allocated once).

    python tools/sweep.py [--sessions N] [--out file.json]

For every (rate, frame) workload and every tuning point it times K launches with CUDA events on
the context's stream and reports algorithmic GB/s (4*N_in + 2*N_out bytes per frame) and the
fraction of the measured HBM peak.  Measurement only; parity is covered by tests/.
"""
import argparse
import itertools
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "voice-enabled-browser-automation_b200"))
import pcm_frontend  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sessions", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    root = os.path.dirname(HERE)
    try:
        peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    ctx = pcm_frontend.Context(0)
    S = args.sessions
    workloads = [(48000, 20), (16000, 20), (44100, 20), (32000, 20), (96000, 20), (48000, 60)]
    if args.quick:
        workloads = workloads[:3]
    results = []
    for rate, ms in workloads:
        n_in = rate * ms // 1000
        n_out = pcm_frontend.resample_len(n_in, rate)
        s = S if n_in <= 1000 else S // (n_in // 960 + 1)
        bufs = [(ctx.device_buffer(s * n_in * 4), ctx.device_buffer(s * n_out * 2)) for _ in range(2)]
        for b, (di, _) in enumerate(bufs):
            ctx.synth_fill(di.ptr, n_in, s, n_in, rate, 0x5EED + b, 0)
        ctx.sync()
        algo = s * (n_in * 4 + n_out * 2)
        fast = rate % 16000 == 0 and rate // 16000 in (1, 2, 3, 4, 6)
        grid = []
        if fast:
            grid += [dict(path=0, coalesced=c, block=b, unroll=u, blocks_per_sm=p, store_hint=h)
                     for c, b, u, p, h in itertools.product((0, 1) if rate != 16000 else (0,), (128, 256, 512), (1, 2, 4), (0,), (0, 1, 2))]
            grid += [dict(path=0, coalesced=1, block=256, unroll=u, blocks_per_sm=p, store_hint=1)
                     for u, p in itertools.product((1, 2), (1, 2, 3, 4, 6))]
        grid += [dict(path=1, coalesced=0, block=b, unroll=u, blocks_per_sm=0, store_hint=0)
                 for b, u in itertools.product((128, 256, 512), (1, 2, 4))]
        grid = [dict(tile=0, tile_kb=28, stages=2, **t) for t in grid]
        if n_out % 8 == 0:
            grid += [dict(path=2, coalesced=0, block=256, unroll=0, blocks_per_sm=p, store_hint=1, tile=1, tile_kb=kb, stages=st)
                     for kb, st, p in itertools.product((8, 14, 20, 28, 40, 56), (1, 2, 3, 4), (0,))]
            grid += [dict(path=2, coalesced=0, block=256, unroll=0, blocks_per_sm=p, store_hint=1, tile=1, tile_kb=kb, stages=2)
                     for kb, p in itertools.product((14, 28), (1, 2, 3, 4))]
        for t in grid:
            ctx.tune(**t)
            for i in range(3):
                ctx.frames_encode_ptr(bufs[i % 2][0].ptr, n_in, bufs[i % 2][1].ptr, n_out, s, n_in, rate)
            ctx.sync()
            ctx.timer_start()
            for i in range(args.steps):
                ctx.frames_encode_ptr(bufs[i % 2][0].ptr, n_in, bufs[i % 2][1].ptr, n_out, s, n_in, rate)
            ms_total = ctx.timer_stop()
            gbs = algo * args.steps / (ms_total * 1e-3) / 1e9
            r = dict(rate=rate, frame_ms=ms, sessions=s, n_in=n_in, n_out=n_out, **t,
                     us_per_launch=ms_total * 1e3 / args.steps, gbs=gbs, frac=gbs / peak)
            results.append(r)
            print(json.dumps(r), flush=True)
        for di, do in bufs:
            di.free(); do.free()
    ctx.tune(path=0, coalesced=1, block=256, unroll=0, blocks_per_sm=0, store_hint=1, tile=1, tile_kb=28, stages=2)
    if args.out:
        json.dump(results, open(args.out, "w"), indent=0)
    best = {}
    for r in results:
        k = (r["rate"], r["frame_ms"], r["path"], r["coalesced"], r["tile"])
        if k not in best or r["gbs"] > best[k]["gbs"]:
            best[k] = r
    print("---- best per (rate, frame_ms, path, coalesced) ----")
    for k, r in sorted(best.items()):
        print(k, f"{r['gbs']:.0f} GB/s frac {r['frac']:.3f}", {x: r[x] for x in ("block", "unroll", "blocks_per_sm", "store_hint", "tile_kb", "stages")})


if __name__ == "__main__":
    main()
