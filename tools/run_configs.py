#!/usr/bin/env python3
"""Measures BASELINE.json configs 1, 2, 3 and 5 on one B200 (config 4 is bench.py itself).

    python tools/run_configs.py --out gpurun_out/configs.json

Every GPU figure is timed with CUDA events on the context's stream after warm-up; CPU figures are
the oracle port (oracle/pcm_oracle.c) on this box's host cores, on the same input bits.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "voice-enabled-browser-automation_b200"))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402  (measurement baseline + checker only)
import pcm_frontend  # noqa: E402
from pcm_frontend import MEM_DEVICE, MEM_HOST  # noqa: E402

try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0


def timed(ctx, fn, reps, warm=3):
    for _ in range(warm):
        fn()
    ctx.sync()
    ctx.timer_start()
    for _ in range(reps):
        fn()
    return ctx.timer_stop() / reps


def config1(ctx):
    """One 1 s 48 kHz mono buffer: resample -> encode (the reference's own call sequence)."""
    n = 48000
    d = ctx.device_buffer(n * 4)
    ctx.synth_fill(d.ptr, n, 1, n, 48000)
    ctx.sync()
    x = d.download(np.float32, n)
    L = oracle.lib()
    out = np.empty(16000, np.int16)
    ts = []
    for _ in range(1000):
        t0 = time.perf_counter_ns()
        L.oracle_frame(x.ctypes.data_as(C.POINTER(C.c_float)), n, 48000.0, out.ctypes.data_as(C.POINTER(C.c_int16)))
        ts.append(time.perf_counter_ns() - t0)
    cpu_ns = statistics.median(ts)
    got = ctx.frames_encode(x.reshape(1, n), 48000)[0]
    assert np.array_equal(got, out)
    two = ctx.floatTo16BitPCM(ctx.resampleTo16k(x, 48000))
    assert np.array_equal(two, out)
    ts = []
    for _ in range(300):
        t0 = time.perf_counter_ns()
        ctx.frames_encode(x.reshape(1, n), 48000)
        ts.append(time.perf_counter_ns() - t0)
    gpu_fused_ns = statistics.median(ts)
    ts = []
    for _ in range(300):
        t0 = time.perf_counter_ns()
        ctx.floatTo16BitPCM(ctx.resampleTo16k(x, 48000))
        ts.append(time.perf_counter_ns() - t0)
    gpu_two_ns = statistics.median(ts)
    return {"config": 1, "n_in": n, "n_out": 16000, "node": "Node unavailable on this box",
            "cpu_oracle_1thread_ns_per_in_sample": cpu_ns / n, "cpu_oracle_us_per_call": cpu_ns / 1e3,
            "gpu_host_call_fused_us": gpu_fused_ns / 1e3, "gpu_host_call_two_functions_us": gpu_two_ns / 1e3,
            "note": "single stream: fixed costs (pageable H2D staging, launch, D2H, sync) dominate; on a par with one CPU core, no speed-up to be had"}


def config2(ctx):
    """1,024 x 16 kHz x 20 ms encode-only: steady state over 2,048 ticks per launch + single-tick latency."""
    S, n, T = 1024, 320, 2048
    rows = S * T
    d_in, d_out = ctx.device_buffer(rows * n * 4), ctx.device_buffer(rows * n * 2)
    ctx.synth_fill(d_in.ptr, n, rows, n, 16000)
    ctx.sync()
    ms = timed(ctx, lambda: ctx.frames_encode_ptr(d_in.ptr, n, d_out.ptr, n, rows, n, 16000, MEM_DEVICE), 20)
    algo = rows * n * 6
    tick = timed(ctx, lambda: ctx.frames_encode_ptr(d_in.ptr, n, d_out.ptr, n, S, n, 16000, MEM_DEVICE), 2000, warm=20)
    g = ctx.tick_graph(d_in.ptr, n, d_out.ptr, n, S, n, 16000)
    tick_g = timed(ctx, g.launch, 2000, warm=20)
    # host-observed latency of one tick (launch + sync)
    ts = []
    for _ in range(500):
        t0 = time.perf_counter_ns()
        g.launch()
        ctx.sync()
        ts.append(time.perf_counter_ns() - t0)
    g.destroy()
    x = d_in.download(np.float32, S * n).reshape(S, n)
    assert np.array_equal(d_out.download(np.int16, S * n).reshape(S, n), oracle.frames(x, n, 16000))
    d_in.free(); d_out.free()
    return {"config": 2, "sessions": S, "n": n, "ticks_per_launch": T, "working_set_gb": algo / 1e9,
            "steady_ms_per_launch": ms, "steady_samples_per_s": rows * n / (ms * 1e-3),
            "steady_gbs": algo / (ms * 1e-3) / 1e9, "steady_frac_of_measured_peak": algo / (ms * 1e-3) / 1e9 / PEAK,
            "single_tick_us_plain_launch_back_to_back": tick * 1e3, "single_tick_us_graph_replay_back_to_back": tick_g * 1e3,
            "single_tick_us_graph_launch_plus_sync_host_clock": statistics.median(ts) / 1e3,
            "note": "one tick is 1.97 MB (L2 resident, 0.3 us of HBM time): only the steady-state line is compared with the HBM roofline"}


def config3(ctx, ms_frame):
    """65,536 x 48 kHz frames, 16 distinct tick buffers (4.7 GB at 20 ms) rotated to defeat L2."""
    S, n_in = 65536, 48 * ms_frame
    n_out = n_in // 3
    ctx.tune(overlap_ticks=1)        # 16 / 8 distinct tick buffers: consecutive ticks are independent
    ticks = 16 if ms_frame == 20 else 8
    bufs = [(ctx.device_buffer(S * n_in * 4), ctx.device_buffer(S * n_out * 2)) for _ in range(ticks)]
    for t, (di, _) in enumerate(bufs):
        ctx.synth_fill(di.ptr, n_in, S, n_in, 48000, 0x5EED + t, 0)
    ctx.sync()
    state = {"i": 0}

    def one():
        di, do = bufs[state["i"] % ticks]
        state["i"] += 1
        ctx.frames_encode_ptr(di.ptr, n_in, do.ptr, n_out, S, n_in, 48000, MEM_DEVICE)
    ms = timed(ctx, one, 10 * ticks, warm=ticks)
    algo = S * (n_in * 4 + n_out * 2)
    x = bufs[0][0].download(np.float32, 4096 * n_in).reshape(4096, n_in)
    assert np.array_equal(bufs[0][1].download(np.int16, 4096 * n_out).reshape(4096, n_out), oracle.frames(x, n_in, 48000, threads=8))
    for di, do in bufs:
        di.free(); do.free()
    ctx.tune(overlap_ticks=0)
    return {"config": 3, "sessions": S, "frame_ms": ms_frame, "n_in": n_in, "n_out": n_out, "distinct_ticks": ticks,
            "bytes_per_tick": algo, "us_per_tick": ms * 1e3, "in_samples_per_s": S * n_in / (ms * 1e-3),
            "gbs": algo / (ms * 1e-3) / 1e9, "frac_of_measured_peak": algo / (ms * 1e-3) / 1e9 / PEAK}


def config5(ctx):
    """Host-resident end to end through pinned buffers; frame sweep 10-100 ms; vs raw PCIe and vs the CPU."""
    cap = 1 << 30
    pin_in = ctx.pinned(cap // 4, np.float32)
    pin_out = ctx.pinned(cap // 8, np.int16)
    dev = ctx.device_buffer(cap)
    h2d = timed(ctx, lambda: ctx.memcpy_h2d(dev.ptr, pin_in.ptr, cap), 5, warm=1)
    d2h = timed(ctx, lambda: ctx.memcpy_d2h(pin_in.ptr, dev.ptr, cap), 5, warm=1)
    res = {"config": 5, "pinned_h2d_gbs": cap / (h2d * 1e-3) / 1e9, "pinned_d2h_gbs": cap / (d2h * 1e-3) / 1e9, "points": []}
    cores = os.cpu_count() or 1
    for ms_frame in (10, 20, 40, 60, 100):
        n_in = 48 * ms_frame
        n_out = n_in // 3
        for batch_bytes in (1 << 20, 1 << 24, 1 << 27, 1 << 30):
            S = max(1, batch_bytes // (n_in * 4))
            ctx.synth_fill(dev.ptr, n_in, S, n_in, 48000)
            ctx.sync()
            ctx.memcpy_d2h(pin_in.ptr, dev.ptr, S * n_in * 4)
            reps = 3 if batch_bytes >= (1 << 27) else 30
            ms = timed(ctx, lambda: ctx.frames_encode_ptr(pin_in.ptr, n_in, pin_out.ptr, n_out, S, n_in, 48000, MEM_HOST), reps, warm=2)
            x = pin_in.array[: S * n_in].reshape(S, n_in)
            rows_cpu = min(S, 1 << 15)
            t0 = time.perf_counter()
            want = oracle.frames(x[:rows_cpu], n_in, 48000, threads=min(cores, rows_cpu))
            cpu_dt = time.perf_counter() - t0
            assert np.array_equal(pin_out.array[: rows_cpu * n_out].reshape(rows_cpu, n_out), want)
            res["points"].append({"frame_ms": ms_frame, "sessions": S, "batch_mb": S * n_in * 4 / 2**20, "ms_per_batch": ms,
                                  "in_samples_per_s": S * n_in / (ms * 1e-3), "h2d_gbs": S * n_in * 4 / (ms * 1e-3) / 1e9,
                                  "frac_of_pinned_h2d": S * n_in * 4 / (ms * 1e-3) / 1e9 / res["pinned_h2d_gbs"],
                                  "cpu_all_cores_samples_per_s": rows_cpu * n_in / cpu_dt, "cpu_cores": cores})
    pin_in.free(); pin_out.free(); dev.free()
    return res


def meter(ctx, n_in, S, reps=20):
    """8f next-1: frames + fused level meter vs frames alone on the same dense batch."""
    n_out = n_in // 3
    nq = n_in // 128
    bufs = [(ctx.device_buffer(S * n_in * 4), ctx.device_buffer(S * n_out * 2)) for _ in range(2)]
    for t, (di, _) in enumerate(bufs):
        ctx.synth_fill(di.ptr, n_in, S, n_in, 48000, 0x5EED + t, 0)
    st = ctx.device_buffer(S * 8)
    st.memset(0)
    ctx.sync()
    k = {"i": 0}

    def plain():
        di, do = bufs[k["i"] % 2]; k["i"] += 1
        ctx.frames_encode_ptr(di.ptr, n_in, do.ptr, n_out, S, n_in, 48000, MEM_DEVICE)

    def fused():
        di, do = bufs[k["i"] % 2]; k["i"] += 1
        ctx.frames_encode_meter_ptr(di.ptr, n_in, do.ptr, n_out, S, n_in, 48000, st.ptr, None)
    ms_plain = timed(ctx, plain, reps)
    ms_fused = timed(ctx, fused, reps)
    algo = S * (n_in * 4 + n_out * 2)
    algo_m = algo + S * 16          # + level state read and written
    for di, do in bufs:
        di.free(); do.free()
    st.free()
    return {"config": "next-1 fused level meter", "sessions": S, "n_in": n_in, "quanta_per_frame": nq,
            "frames_only_ms": ms_plain, "frames_only_gbs": algo / (ms_plain * 1e-3) / 1e9,
            "frames_plus_meter_ms": ms_fused, "frames_plus_meter_gbs": algo_m / (ms_fused * 1e-3) / 1e9,
            "frac_of_measured_peak": algo_m / (ms_fused * 1e-3) / 1e9 / PEAK,
            "in_samples_per_s": S * n_in / (ms_fused * 1e-3)}


def aggregator(ctx, S=65536, rate=48000, meter=False, order="quantum"):
    """8f next-2: S sessions, every push carries 23 quanta per session (61.3 ms of 48 kHz audio), so
    every session flushes one 2944-sample frame per push.  Everything device-resident
    (pcmb200_aggregator_push_device): `enqueue_ms` is the host time of the call, `total_ms` runs from
    the call to the end of pcmb200_aggregator_wait (plan + convert on the GPU, 24 bytes read back).
    Also timed: the same push through pcmb200_aggregator_push with HOST session/arrival arrays."""
    import math
    nq = math.ceil(60.0 / (128e3 / rate) - 1e-9)                           # quanta until elapsed >= 60 ms: 23 at 48 kHz, 21 at 44.1 kHz
    n = S * nq
    n_out = pcm_frontend.resample_len(nq * 128, rate)
    pad = (n_out + 7) // 8 * 8
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=max(32, nq + 9), max_push_quanta=n, max_out_samples=S * pad)
    if meter:
        agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    if order == "quantum":                                                 # quantum-major arrival order (what time order gives)
        sessions = np.tile(np.arange(S, dtype=np.uint32), nq)
        base = np.repeat(np.arange(1, nq + 1) * (128e3 / rate), S)
    else:                                                                  # diagnostic: every session's quanta adjacent in the push
        sessions = np.repeat(np.arange(S, dtype=np.uint32), nq)
        base = np.tile(np.arange(1, nq + 1) * (128e3 / rate), S)
    d_samples = ctx.device_buffer(n * 128 * 4)
    ctx.synth_fill(d_samples.ptr, 128, n, 128, rate)
    d_out = ctx.device_buffer(S * pad * 2)
    d_sess, d_arr = ctx.device_buffer(n * 4), ctx.device_buffer(n * 8)
    d_fs, d_fl, d_fo = ctx.device_buffer(n * 4), ctx.device_buffer(n * 4), ctx.device_buffer(n * 8)
    d_sess.upload(sessions)
    enq, total = [], []
    for p in range(6):
        d_arr.upload(base + p * nq * (128e3 / rate))
        ctx.sync()
        t0 = time.perf_counter()
        agg.push_device(n, d_sess.ptr, d_arr.ptr, d_samples.ptr, d_out.ptr, S * pad, d_fs.ptr, d_fl.ptr, d_fo.ptr, n)
        t1 = time.perf_counter()
        nf, ns = agg.wait()
        t2 = time.perf_counter()
        enq.append(t1 - t0); total.append(t2 - t0)
        assert nf == S and ns == S * pad, (nf, ns)
    fl, fo, fs = d_fl.download(np.uint32, S), d_fo.download(np.uint64, S), d_fs.download(np.uint32, S)
    assert (fl == n_out).all() and (fs == np.arange(S)).all()
    # every frame against the oracle: session s holds quanta s, s+S, ... of the push
    x = d_samples.download(np.float32, n * 128)
    x = x.reshape(nq, S, 128).transpose(1, 0, 2).reshape(S, nq * 128) if order == "quantum" else x.reshape(S, nq * 128)
    want = oracle.frames(np.ascontiguousarray(x), nq * 128, rate, threads=os.cpu_count() or 1)
    got = d_out.download(np.int16, S * pad).reshape(S, pad)[:, :n_out]
    assert np.array_equal(got, want)
    # the same push with host metadata arrays (pageable numpy): pcmb200_aggregator_push
    L = pcm_frontend.load()
    from pcm_frontend._capi import check
    fs_h, fl_h, fo_h = np.empty(n, np.uint32), np.empty(n, np.uint32), np.empty(n, np.uint64)
    nf_c = C.c_uint32(0)
    host_meta = []
    for p in range(6, 9):
        arr = base + p * nq * (128e3 / rate)
        ctx.sync()
        t0 = time.perf_counter()
        check(L.pcmb200_aggregator_push(ctx.handle, agg.handle, n, sessions.ctypes.data, arr.ctypes.data, d_samples.ptr, MEM_DEVICE,
                                        d_out.ptr, S * pad, MEM_DEVICE, fs_h.ctypes.data, fl_h.ctypes.data, fo_h.ctypes.data, n,
                                        C.byref(nf_c)), ctx.handle)
        host_meta.append(time.perf_counter() - t0)
        assert nf_c.value == S
    agg.close()
    for d in (d_samples, d_out, d_sess, d_arr, d_fs, d_fl, d_fo):
        d.free()
    t = min(total[1:])
    # bytes the push has to move: every quantum read once (512 B), every frame written once
    algo = n * 512 + S * n_out * 2
    return {"config": "next-2 device aggregator (device-resident quanta, ids, times and frames)", "meter": meter, "sessions": S, "rate": rate, "quanta_per_frame": nq, "push_order": order,
            "quanta_per_push": n, "enqueue_ms": min(enq[1:]) * 1e3, "total_ms": t * 1e3, "in_samples_per_s": n * 128 / t,
            "realtime_sessions_at_48k": n * 128 / t / rate, "algorithmic_gbs": algo / t / 1e9, "frac_of_measured_peak": algo / t / 1e9 / PEAK,
            "host_metadata_push_ms": min(host_meta) * 1e3, "frames_checked_against_oracle": S,
            "note": "flush plan, frame list and conversion all on the GPU (agg_kernels.cuh); total_ms is host wall clock from "
                    "the call to the end of pcmb200_aggregator_wait"}


def aggregator_streaming(ctx, S=65536, rate=48000, meter=False, pushes=92):
    """8f next-2 in the regime a live server runs in: every push carries ONE quantum per session (2.67 ms
    of 48 kHz audio, sessions in random order), so 22 pushes out of 23 only append to the sessions'
    buffers and the 23rd flushes every session.  Device-resident; per-push wall time from the call to the
    end of pcmb200_aggregator_wait, averaged over `pushes` pushes (4 flush cycles)."""
    rng = np.random.RandomState(1)
    q_ms = 128e3 / rate
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=S, max_out_samples=16)
    if meter:
        agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    d_samples = ctx.device_buffer(S * 128 * 4)
    ctx.synth_fill(d_samples.ptr, 128, S, 128, rate)
    d_out = ctx.device_buffer(S * 984 * 2)
    d_sess, d_arr = ctx.device_buffer(S * 4), ctx.device_buffer(S * 8)
    d_fs, d_fl, d_fo = ctx.device_buffer(S * 4), ctx.device_buffer(S * 4), ctx.device_buffer(S * 8)
    d_sess.upload(rng.permutation(S).astype(np.uint32))
    times, frames = [], 0
    for p in range(pushes):
        d_arr.upload(np.full(S, (p + 1) * q_ms))
        ctx.sync()
        t0 = time.perf_counter()
        agg.push_device(S, d_sess.ptr, d_arr.ptr, d_samples.ptr, d_out.ptr, S * 984, d_fs.ptr, d_fl.ptr, d_fo.ptr, S)
        nf, _ = agg.wait()
        times.append(time.perf_counter() - t0)
        frames += nf
    assert frames == S * (pushes // 23), frames
    agg.close()
    for d in (d_samples, d_out, d_sess, d_arr, d_fs, d_fl, d_fo):
        d.free()
    t = np.array(times[23:])                      # skip the first cycle (warm-up)
    flush = t[22::23]
    return {"config": "next-2 aggregator, streaming regime (one quantum per session per push)", "meter": meter, "sessions": S,
            "ms_per_push_mean": float(t.mean() * 1e3), "ms_per_append_only_push": float(np.median(np.delete(t, np.s_[22::23])) * 1e3),
            "ms_per_flushing_push": float(np.median(flush) * 1e3), "audio_ms_per_push": q_ms,
            "realtime_sessions_at_48k": S * (q_ms * 1e-3) / float(t.mean()),
            "note": "real-time capacity = sessions x (audio time per push / wall time per push)"}


def quanta_ingest(ctx, S=65536, rate=48000, tick_ms=20.0, ticks=12):
    """next-2 + next-4 end to end from host memory: S connections with "framing": "quanta"; every tick the quanta that
    arrived in tick_ms (pinned staging, as GpuQuantumAggregator fills it) go through ONE aggregator push -- H2D of ids,
    times and samples, the planned push on the GPU, D2H of the frames that flushed -- and come back as views.  Timed:
    push() alone (staging is the socket layer's copy).  Frames of the last tick are checked against the oracle."""
    from pcm_frontend.ingest import GpuQuantumAggregator
    q_ms = 128e3 / rate
    per_tick = int(round(tick_ms / q_ms))                                   # quanta per connection per tick
    n = S * per_tick
    g = GpuQuantumAggregator(ctx, float(rate), max_sessions=S, max_push_quanta=n, max_frame_quanta=32)
    for _ in range(S):
        g.open(0.0)
    rng = np.random.RandomState(7)
    x = rng.uniform(-1, 1, size=(ticks * per_tick, 1024, 128)).astype(np.float32)    # audio: 1024 patterns, tiled over S
    sess = np.tile(np.arange(S, dtype=np.uint32), per_tick)
    times, calls, frames_last, hist = [], [], None, {}
    for t in range(ticks):
        q0 = t * per_tick
        g._sess.array[:n] = sess
        g._time.array[:n] = np.repeat((np.arange(q0, q0 + per_tick) + 1) * q_ms, S)
        for k in range(per_tick):
            g._x.array[k * S:(k + 1) * S] = np.tile(x[q0 + k], (S // 1024, 1))
        g.n = n
        ctx.sync()
        t0 = time.perf_counter()
        frames_last = g.push()
        times.append(time.perf_counter() - t0)
        calls.append(g.last_call_s)
        hist[t] = len(frames_last)
    # the last flushing tick against the oracle: every connection s holds pattern s % 1024
    flush, _ = oracle.aggregator_plan((np.arange(ticks * per_tick) + 1) * q_ms, 60.0, 0.0)
    ends = np.nonzero(flush)[0]
    checked = 0
    if frames_last:
        last_end = ends[ends < ticks * per_tick][-1]
        start = ends[ends < last_end][-1] + 1 if (ends < last_end).any() else 0
        for slot, b in frames_last[:4096]:
            buf = x[start:last_end + 1, slot % 1024].reshape(1, -1)
            assert np.array_equal(np.frombuffer(bytes(b), dtype="<i2"), oracle.frames(buf, buf.shape[1], rate)[0]), slot
            checked += 1
    flushing = [c for c, k in zip(calls, hist.values()) if k][1:] or [0.0]
    quiet = [c for c, k in zip(calls, hist.values()) if not k][1:] or [0.0]
    med = (sum(sorted(quiet)[len(quiet) // 2:len(quiet) // 2 + 1]) * 2 + sorted(flushing)[len(flushing) // 2]) / 3   # 2 quiet ticks per flushing one
    h2d_bytes = n * (128 * 4 + 4 + 8)
    dev = ctx.device_buffer(n * 512)
    bare = timed(ctx, lambda: ctx.memcpy_h2d(dev.ptr, g._x.ptr, n * 512), 5, warm=1)     # the same pinned block, copy alone
    dev.free()
    g.close()
    return {"config": "next-2 + next-4: quanta-framed ingest from pinned host memory, one aggregator push per tick",
            "sessions": S, "tick_ms": tick_ms, "quanta_per_tick": n, "h2d_mb_per_tick": h2d_bytes / 2**20,
            "ms_per_quiet_tick": sorted(quiet)[len(quiet) // 2] * 1e3, "ms_per_flushing_tick": sorted(flushing)[len(flushing) // 2] * 1e3,
            "ms_per_tick_mean": med * 1e3, "bare_h2d_of_the_samples_ms": bare, "h2d_gbs": h2d_bytes / med / 1e9,
            "in_samples_per_s": n * 128 / med, "realtime_sessions_at_48k": n * 128 / med / rate,
            "fraction_of_a_tick": med * 1e3 / tick_ms, "python_push_ms_median_incl_views": sorted(times[2:])[len(times[2:]) // 2] * 1e3,
            "frames_per_tick": hist, "frames_checked_against_oracle": checked,
            "note": "times are the library call alone (pcmb200_aggregator_push with pinned host arrays: three H2D copies, the "
                    "planned push, D2H of the frame arrays and frames); building 65 536 Python memoryviews on a flushing tick "
                    "costs more than the call and is reported separately"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ctx = pcm_frontend.Context(0)
    results = {"device": ctx.device_info(), "host_cores": os.cpu_count(), "hbm_peak_gbs": PEAK, "configs": []}
    ap.add_argument("--only-meter", action="store_true")
    ap.add_argument("--meter-once", type=int, default=0, help="n_in: a short frames+meter run of that frame length (for ncu)")
    ap.add_argument("--tune", default="", help="comma list key=value for pcmb200_tune")
    ap.add_argument("--agg-once", action="store_true", help="the aggregator benchmark alone (for ncu)")
    ap.add_argument("--sweep-agg", action="store_true", help="sweep the aggregator convert kernel's CTA size and step size")
    ap.add_argument("--sweep-meter", action="store_true")
    ap.add_argument("--quanta-ingest", action="store_true", help="the quanta-framed host-ingest benchmark alone")
    ap.add_argument("--agg-stream", type=int, default=0, help="the streaming-regime aggregator benchmark alone with this many sessions (for ncu)")
    ap.add_argument("--agg-rate", type=int, default=48000, help="with --agg-order: capture rate")
    ap.add_argument("--agg-meter", action="store_true", help="with --agg-order: the meter on")
    ap.add_argument("--agg-order", default="", help="diagnostic: run the big-push aggregator benchmark alone with this push order (quantum | session)")
    args = ap.parse_args()
    if args.tune:
        ctx.tune(**{k: int(v) for k, v in (kv.split("=") for kv in args.tune.split(","))})
    if args.sweep_agg:
        for blk in (64, 96, 128, 192, 256):
            for kb in (17, 34, 51):
                ctx.tune(agg_block=blk, agg_stage_kb=kb)
                for m in (False, True):
                    r = aggregator(ctx, meter=m)
                    print(json.dumps(dict(agg_block=blk, agg_stage_kb=kb, meter=m, total_ms=round(r["total_ms"], 4))), flush=True)
        return
    if args.quanta_ingest:
        print(json.dumps(quanta_ingest(ctx)), flush=True)
        return
    if args.agg_stream:
        print(json.dumps(aggregator_streaming(ctx, S=args.agg_stream, pushes=46)), flush=True)
        return
    if args.agg_order:
        print(json.dumps(aggregator(ctx, order=args.agg_order, meter=args.agg_meter, rate=args.agg_rate)), flush=True)
        return
    if args.agg_once:
        print(json.dumps(aggregator(ctx)), flush=True)
        print(json.dumps(aggregator(ctx, meter=True)), flush=True)
        print(json.dumps(aggregator_streaming(ctx)), flush=True)
        print(json.dumps(aggregator_streaming(ctx, meter=True)), flush=True)
        print(json.dumps(aggregator_streaming(ctx, S=1 << 20, pushes=69)), flush=True)
        return
    if args.meter_once:
        print(json.dumps(meter(ctx, args.meter_once, (1 << 19) if args.meter_once < 2000 else (1 << 17), reps=3)), flush=True)
        return
    if args.sweep_meter:
        # kernel 2c: tile height (via meter_tile_kb: 32/64/96/128 KB = 8/16/24/32 rows of 960 samples),
        # stages and grid cap; meter_kernel=0 is the old kernel 2b for comparison
        pts = [dict(meter_kernel=0, meter_tile_kb=64, meter_stages=1, blocks_per_sm=0),
               dict(meter_kernel=1, meter_tile_kb=64, meter_stages=1, blocks_per_sm=0)]
        # kernel 2d: stage size, ring size, CTAs per SM (2 CTAs need the ring to be <= ~110 KB each)
        for pairs in (4, 6, 8, 12, 16):
            for depth in (1, 2, 3):
                for chunk in (4, 8):
                    pts.append(dict(meter_kernel=2, ws_smem_kb=212, ws_pairs=pairs, ws_depth=depth, ws_chunk=chunk, blocks_per_sm=0))
        for pt in pts:
            try:
                ctx.tune(**pt)
                for n_in, S in ((960, 1 << 19), (2944, 1 << 17)):
                    r = meter(ctx, n_in, S)
                    print(json.dumps(dict(pt, n_in=n_in, ms=r["frames_plus_meter_ms"], frac=round(r["frac_of_measured_peak"], 4))), flush=True)
            except Exception as e:
                print("skip", pt, e, flush=True)
        return
    if args.only_meter:
        for fn in (lambda: meter(ctx, 960, 1 << 20), lambda: meter(ctx, 2944, 1 << 18), lambda: aggregator(ctx), lambda: aggregator(ctx, meter=True)):
            print(json.dumps(fn()), flush=True)
        return
    for fn in (lambda: meter(ctx, 960, 1 << 20), lambda: meter(ctx, 2944, 1 << 18), lambda: aggregator(ctx), lambda: aggregator(ctx, meter=True), lambda: aggregator(ctx, rate=44100), lambda: aggregator_streaming(ctx), lambda: aggregator_streaming(ctx, meter=True), lambda: aggregator_streaming(ctx, S=1 << 20, pushes=69), lambda: config1(ctx), lambda: config2(ctx), lambda: config3(ctx, 20), lambda: config3(ctx, 60), lambda: config5(ctx), lambda: quanta_ingest(ctx)):
        r = fn()
        results["configs"].append(r)
        print(json.dumps(r), flush=True)
    if args.out:
        json.dump(results, open(args.out, "w"), indent=1)
    ctx.close()


if __name__ == "__main__":
    main()
