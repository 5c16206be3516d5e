"""Throughput floor: BASELINE.json's target is >= 70 % of the B200 HBM peak device-resident on the
batched kernels.  Measured values are 1.0-1.1 of the measured copy bandwidth; this
test keeps a regression from slipping below the stated target.  CUDA-event timing, batches far
larger than L2, two alternating buffers."""
import json
import os

import pytest

import pcm_frontend
from pcm_frontend import MEM_DEVICE

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0          # B200_PROFILING.md fallback


@pytest.mark.parametrize("rate,n_in,floor", [(48000, 960, 0.70), (44100, 882, 0.70), (16000, 320, 0.70), (96000, 1920, 0.70),
                                             (80000, 1600, 0.70), (128000, 2560, 0.70), (192000, 3840, 0.70)])
def test_batched_kernels_stay_above_70_percent_of_hbm_peak(ctx, rate, n_in, floor):
    rows = (1 << 19) if n_in <= 1000 else (1 << 18)
    n_out = pcm_frontend.resample_len(n_in, rate)
    bufs = [(ctx.device_buffer(rows * n_in * 4), ctx.device_buffer(rows * n_out * 2)) for _ in range(2)]
    for t, (di, _) in enumerate(bufs):
        ctx.synth_fill(di.ptr, n_in, rows, n_in, rate, 0x5EED + t, 0)

    def step(i):
        di, do = bufs[i % 2]
        ctx.frames_encode_ptr(di.ptr, n_in, do.ptr, n_out, rows, n_in, rate, MEM_DEVICE)
    for i in range(4):
        step(i)
    ctx.sync()
    best = 0.0
    for _ in range(3):                      # best of three timed bursts: robust against a noisy neighbour
        ctx.timer_start()
        for i in range(20):
            step(i)
        ms = ctx.timer_stop() / 20
        best = max(best, rows * (n_in * 4 + n_out * 2) / (ms * 1e-3) / 1e9)
    for di, do in bufs:
        di.free(); do.free()
    print(f"PERF_FLOOR rate={rate} n_in={n_in} gbs={best:.0f} frac={best / _peak():.3f}")
    assert best / _peak() >= floor, f"{rate} Hz: {best:.0f} GB/s is {best / _peak():.2f} of the HBM peak"


@pytest.mark.parametrize("n_in,rows", [(960, 1 << 19), (2944, 1 << 17)])
def test_fused_meter_stays_above_its_floor(ctx, n_in, rows):
    """Frames + level meter (kernel 2d), 48 kHz: the stated goal is >= 0.90 of the HBM peak; the floor here
    is set below the measured value so that only a real regression trips it."""
    n_out = n_in // 3
    bufs = [(ctx.device_buffer(rows * n_in * 4), ctx.device_buffer(rows * n_out * 2)) for _ in range(2)]
    for t, (di, _) in enumerate(bufs):
        ctx.synth_fill(di.ptr, n_in, rows, n_in, 48000, 0x5EED + t, 0)
    st = ctx.device_buffer(rows * 8)
    st.memset(0)

    def step(i):
        di, do = bufs[i % 2]
        ctx.frames_encode_meter_ptr(di.ptr, n_in, do.ptr, n_out, rows, n_in, 48000, st.ptr, None)
    for i in range(4):
        step(i)
    ctx.sync()
    best = 0.0
    for _ in range(3):
        ctx.timer_start()
        for i in range(20):
            step(i)
        ms = ctx.timer_stop() / 20
        best = max(best, rows * (n_in * 4 + n_out * 2 + 16) / (ms * 1e-3) / 1e9)
    for di, do in bufs:
        di.free(); do.free()
    st.free()
    print(f"PERF_FLOOR meter n_in={n_in} gbs={best:.0f} frac={best / _peak():.3f}")
    assert best / _peak() >= 0.75, f"meter, {n_in}-sample frames: {best:.0f} GB/s is {best / _peak():.2f} of the HBM peak"


def test_aggregator_push_stays_below_its_ceiling(ctx):
    """Device-side aggregator, 65 536 sessions x 23 quanta per push, everything device-resident: 0.25-0.31 ms measured
    (round 1: 3.2 ms).  The ceiling is far above the measured value: only a real regression -- a
    kernel of the chain falling off its fast path -- trips it."""
    import time

    import numpy as np
    S, nq, rate = 65536, 23, 48000
    n = S * nq
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=n, max_out_samples=S * 984)
    agg.reset_range(0, S, 0.0)
    d = {k: ctx.device_buffer(v) for k, v in dict(s=n * 4, t=n * 8, x=n * 512, o=S * 984 * 2, fs=n * 4, fl=n * 4, fo=n * 8).items()}
    d["s"].upload(np.tile(np.arange(S, dtype=np.uint32), nq))
    ctx.synth_fill(d["x"].ptr, 128, n, 128, rate)
    base = np.repeat(np.arange(1, nq + 1) * (128e3 / rate), S)
    best = 1e9
    for p in range(5):
        d["t"].upload(base + p * nq * (128e3 / rate))
        ctx.sync()
        t0 = time.perf_counter()
        agg.push_device(n, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984, d["fs"].ptr, d["fl"].ptr, d["fo"].ptr, n)
        nf, _ = agg.wait()
        dt = time.perf_counter() - t0
        assert nf == S
        if p:
            best = min(best, dt)
    agg.close()
    for b in d.values():
        b.free()
    print(f"PERF_FLOOR aggregator push_ms={best * 1e3:.3f}")
    assert best * 1e3 <= 0.6, f"aggregator push took {best * 1e3:.3f} ms (measured 0.25-0.31 ms)"
