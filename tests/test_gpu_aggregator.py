"""GPU parity of the device-side capture aggregator (App.tsx:273-288) against
the oracle: same flush decisions (injected clock), same variable-length frames (multiples of 128
samples), bit-exact PCM."""
import numpy as np
import pytest

import oracle
import pcm_frontend

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rate", [48000, 44100, 16000, 96000])
def test_aggregator_matches_oracle(ctx, rate):
    rng = np.random.RandomState(int(rate) % 1000)
    n_sessions, seconds = 37, 0.5
    q_ms = 128e3 / rate
    # build the event stream (kept here so the audio of every session is addressable)
    audio, times, want = {}, {}, {}
    events = []
    for s in range(n_sessions):
        nq = int(seconds * 1e3 / q_ms)
        t = 5.0 * rng.rand() + np.cumsum(q_ms + rng.uniform(-0.4, 0.4, nq))
        x = (rng.uniform(-1.1, 1.1, size=(nq, 128)) * rng.uniform(0.05, 1)).astype(np.float32)
        if s == 0:
            x[0, :3] = [np.nan, np.inf, -0.0]       # poisons session 0's level for good (Math.max NaN rule)
        elif s == 1:
            x[2, 5] = np.float32(np.inf)
        elif s == 2:
            x[:4] = 0.0                              # starts with silence
        audio[s], times[s] = x, t
        flush, _ = oracle.aggregator_plan(t, 60.0, 0.0)
        frames, start = [], 0
        for q in np.nonzero(flush)[0]:
            buf = x[start:q + 1].reshape(1, -1)
            frames.append(oracle.frames(buf, buf.shape[1], rate)[0])
            start = q + 1
        want[s] = frames
        events += [(t[q], s, q) for q in range(nq)]
    events.sort()
    agg = pcm_frontend.Aggregator(ctx, n_sessions, rate, max_frame_quanta=96, max_push_quanta=4096, max_out_samples=1 << 20)
    agg.set_meter(True)
    for s in range(n_sessions):
        agg.reset_session(s, 0.0)
    got = {s: [] for s in range(n_sessions)}
    # pushes of varying size: small (a few ms of wall time) up to very large (several flushes of the
    # same session inside one push, which forces the scheduler to split it into rounds)
    i, sizes = 0, [7, 64, 301, 1500, 33, 4096, 900]
    k = 0
    while i < len(events):
        n = min(sizes[k % len(sizes)], len(events) - i)
        k += 1
        chunk = events[i:i + n]
        i += n
        frames = agg.push(np.array([e[1] for e in chunk]), np.array([e[0] for e in chunk]),
                          np.stack([audio[e[1]][e[2]] for e in chunk]))
        for s, pcm in frames:
            got[s].append(pcm)
    # level meter (App.tsx:267-271): binary64 level after the last FLUSHED quantum of every session
    levels = agg.levels(0, n_sessions)
    for s in range(n_sessions):
        flush, _ = oracle.aggregator_plan(times[s], 60.0, 0.0)
        last = np.nonzero(flush)[0][-1] + 1
        want_lv = oracle.level_meter(audio[s][:last].reshape(-1))[-1]
        assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], (s, levels[s], want_lv)
    agg.close()
    n_frames = 0
    for s in range(n_sessions):
        assert len(got[s]) == len(want[s]), (s, len(got[s]), len(want[s]))
        for a, b in zip(got[s], want[s]):
            assert a.size == b.size and np.array_equal(a, b), s
            n_frames += 1
    assert n_frames >= n_sessions * 6
    if rate == 48000:   # 23 quanta = 2944 samples -> 981 outputs is the typical frame
        sizes_seen = {f.size for s in got for f in got[s]}
        assert 981 in sizes_seen


def test_aggregator_rejects_overlong_buffer_and_rolls_back(ctx):
    agg = pcm_frontend.Aggregator(ctx, 2, 48000, max_frame_quanta=4, max_push_quanta=64, max_out_samples=4096)
    agg.reset_session(0, 0.0)
    x = np.ones((6, 128), np.float32) * 0.25
    with pytest.raises(pcm_frontend.PcmError):        # 6 quanta, no flush (clock never reaches 60 ms), cap 4
        agg.push(np.zeros(6, np.uint32), np.arange(6) * 1.0, x)
    # the rejected push left the session untouched: 4 quanta still fit, and the 4th flushes
    frames = agg.push(np.zeros(4, np.uint32), np.array([1.0, 2.0, 3.0, 61.0]), x[:4])
    assert len(frames) == 1 and frames[0][0] == 0 and frames[0][1].size == 170 and (frames[0][1] == 8191).all()
    agg.close()


def test_aggregator_rejects_out_of_range_session(ctx):
    agg = pcm_frontend.Aggregator(ctx, 37, 48000, max_frame_quanta=8, max_push_quanta=64, max_out_samples=4096)
    x = np.zeros((3, 128), np.float32)
    with pytest.raises(pcm_frontend.PcmError):
        agg.push(np.array([1, 40, 2], np.uint32), np.array([1.0, 2.0, 3.0]), x)
    # the valid sessions of the rejected push were rolled back: session 1 still flushes a 1-quantum frame
    frames = agg.push(np.array([1], np.uint32), np.array([70.0]), x[:1])
    assert len(frames) == 1 and frames[0][1].size == 42
    agg.close()


def _session_frames(x, t, rate, last0=0.0, desired_ms=60.0):
    """Oracle frames of one session: x [nq, 128] quanta, t [nq] arrival times."""
    flush, last = oracle.aggregator_plan(t, desired_ms, last0)
    frames, start = [], 0
    for q in np.nonzero(flush)[0]:
        buf = x[start:q + 1].reshape(1, -1)
        frames.append(oracle.frames(buf, buf.shape[1], rate)[0])
        start = q + 1
    return frames, start, last


@pytest.mark.parametrize("rate", [48000, 44100])
def test_aggregator_65536_sessions_vs_oracle(ctx, rate):
    """The device-side planner at scale: 65 536 sessions, three pushes of jittered quanta (sessions
    interleaved at random inside a push, some sessions skipping a push, buffers carried from push to
    push), every flushed frame and every level compared with the oracle."""
    rng = np.random.RandomState(int(rate) // 100)
    S, q_ms = 65536, 128e3 / rate
    per_push = (9, 11, 8)                                  # quanta per session per push (~60 ms flushes span pushes)
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=40, max_push_quanta=S * 11, max_out_samples=S * 1100)
    agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    start_t = 5.0 * rng.rand(S)
    period = q_ms + rng.uniform(-0.3, 0.3, S)             # each session's own quantum period
    audio, times, got = [], [], {}
    done = np.zeros(S, np.int64)                           # quanta sent so far per session
    for p, m in enumerate(per_push):
        active = rng.rand(S) > (0.02 if p else 0.0)       # a few sessions skip later pushes
        sess = np.repeat(np.nonzero(active)[0].astype(np.uint32), m)
        k = np.tile(np.arange(m), int(active.sum())) + done[sess]
        t = start_t[sess] + (k + 1) * period[sess]
        order = np.lexsort((sess, t))                      # arrival order: by time, sessions interleaved
        sess, t, k = sess[order], t[order], k[order]
        x = (rng.uniform(-1.1, 1.1, size=(sess.size, 128)) * rng.uniform(0.05, 1, size=(sess.size, 1))).astype(np.float32)
        audio.append((sess, k, x)); times.append(t)
        done[active] += m
        for s, pcm in agg.push(sess, t, x):
            got.setdefault(s, []).append(pcm)
    levels = agg.levels(0, S)
    agg.close()
    # regroup the audio per session in arrival order
    all_s = np.concatenate([a[0] for a in audio]); all_k = np.concatenate([a[1] for a in audio])
    all_x = np.concatenate([a[2] for a in audio]); all_t = np.concatenate(times)
    order = np.lexsort((all_k, all_s))
    all_s, all_x, all_t = all_s[order], all_x[order], all_t[order]
    bounds = np.searchsorted(all_s, np.arange(S + 1))
    n_frames = 0
    for s in range(S):
        a, b = bounds[s], bounds[s + 1]
        want, flushed_q, _ = _session_frames(all_x[a:b], all_t[a:b], rate)
        have = got.get(s, [])
        assert len(have) == len(want), (s, len(have), len(want))
        for u, v in zip(have, want):
            assert u.size == v.size and np.array_equal(u, v), s
        n_frames += len(want)
        if s % 16 == 0:                                    # levels: every 16th session (the oracle loop is Python)
            want_lv = oracle.level_meter(all_x[a:a + flushed_q].reshape(-1))[-1] if flushed_q else 0.0
            assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], (s, levels[s], want_lv)
    assert n_frames >= 0.9 * S                            # ~75 ms of audio per session: one flush each, a few skipped


def test_aggregator_capacity_is_checked_on_the_padded_total(ctx):
    """frames start on 8-sample boundaries, so the output must hold every frame rounded up to
    8 samples.  A capacity equal to the exact, unpadded total is refused cleanly (nothing is written
    past it, the session is untouched); the padded total is accepted."""
    import ctypes as C
    from pcm_frontend._capi import check, MEM_HOST
    L = pcm_frontend.load()
    agg = pcm_frontend.Aggregator(ctx, 4, 48000, max_frame_quanta=32, max_push_quanta=64, max_out_samples=4096)
    agg.reset_range(0, 4, 0.0)
    x = (np.arange(23 * 128, dtype=np.float32).reshape(23, 128) % 97) / 97.0
    t = (np.arange(23) + 1) * (128e3 / 48000)             # the 23rd quantum arrives at 61.3 ms -> 981 outputs
    sess = np.zeros(23, np.uint32)
    out = np.full(984 + 64, 0x1111, np.int16)
    fs, fl, fo = np.empty(8, np.uint32), np.empty(8, np.uint32), np.empty(8, np.uint64)
    nf = C.c_uint32(0)

    def push(cap):
        return L.pcmb200_aggregator_push(ctx.handle, agg.handle, 23, sess.ctypes.data, t.ctypes.data, x.ctypes.data, MEM_HOST,
                                         out.ctypes.data, cap, MEM_HOST, fs.ctypes.data, fl.ctypes.data, fo.ctypes.data, 8, C.byref(nf))
    assert push(981) == pcm_frontend._capi.ERR_CAPACITY
    assert (out == 0x1111).all()
    check(push(984), ctx.handle)
    assert nf.value == 1 and fl[0] == 981 and fo[0] == 0
    assert np.array_equal(out[:981], oracle.frames(x.reshape(1, -1), 23 * 128, 48000)[0]) and (out[984:] == 0x1111).all()
    agg.close()


def test_aggregator_device_resident_push_and_wait(ctx):
    """pcmb200_aggregator_push_device: ids, times, quanta, frames and frame arrays all in device memory."""
    rng = np.random.RandomState(77)
    S, nq, rate = 1000, 25, 48000
    n = S * nq
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=n, max_out_samples=16)
    agg.reset_range(0, S, 0.0)
    sess = np.tile(np.arange(S, dtype=np.uint32), nq)
    t = np.repeat((np.arange(nq) + 1) * (128e3 / rate), S)
    x = rng.uniform(-1, 1, size=(n, 128)).astype(np.float32)
    d = {k: ctx.device_buffer(v) for k, v in dict(s=n * 4, t=n * 8, x=n * 512, o=S * 984 * 2, fs=n * 4, fl=n * 4, fo=n * 8).items()}
    d["s"].upload(sess); d["t"].upload(t); d["x"].upload(x)
    agg.push_device(n, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984, d["fs"].ptr, d["fl"].ptr, d["fo"].ptr, n)
    with pytest.raises(pcm_frontend.PcmError):            # one push may be pending
        agg.push_device(n, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984)
    nf, ns = agg.wait()
    assert nf == S and ns == S * 984                       # 23 quanta flush (61.3 ms), quanta 24 and 25 stay buffered
    fs, fl, fo = d["fs"].download(np.uint32, S), d["fl"].download(np.uint32, S), d["fo"].download(np.uint64, S)
    assert (fs == np.arange(S)).all() and (fl == 981).all() and (fo == np.arange(S) * 984).all()
    got = d["o"].download(np.int16, S * 984).reshape(S, 984)[:, :981]
    xs = x.reshape(nq, S, 128).transpose(1, 0, 2)
    want = oracle.frames(np.ascontiguousarray(xs[:, :23].reshape(S, 23 * 128)), 23 * 128, rate, threads=4)
    assert np.array_equal(got, want)
    # second push: 21 more quanta per session -> the 2 buffered quanta + 21 new ones flush as 23 again
    t2 = np.repeat((np.arange(nq, nq + 21) + 1) * (128e3 / rate), S)
    x2 = rng.uniform(-1, 1, size=(S * 21, 128)).astype(np.float32)
    d["t"].upload(t2); d["x"].upload(x2)
    agg.push_device(S * 21, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984, None, None, None, n)
    nf, ns = agg.wait()
    assert nf == S
    got = d["o"].download(np.int16, S * 984).reshape(S, 984)[:, :981]
    x2s = x2.reshape(21, S, 128).transpose(1, 0, 2)
    frame = np.concatenate([xs[:, 23:25], x2s], axis=1).reshape(S, 23 * 128)
    assert np.array_equal(got, oracle.frames(np.ascontiguousarray(frame), 23 * 128, rate, threads=4))
    # too small an output: rejected, and the state is unchanged (the same push then succeeds)
    t3 = np.repeat((np.arange(nq + 21, nq + 44) + 1) * (128e3 / rate), S)
    d["t"].upload(t3)
    x3 = rng.uniform(-1, 1, size=(S * 23, 128)).astype(np.float32)
    d["x"].upload(x3)
    agg.push_device(S * 23, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984 - 1)
    with pytest.raises(pcm_frontend.PcmError) as ei:
        agg.wait()
    assert ei.value.status == pcm_frontend._capi.ERR_CAPACITY
    agg.push_device(S * 23, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr, S * 984)
    assert agg.wait() == (S, S * 984)
    got = d["o"].download(np.int16, S * 984).reshape(S, 984)[:, :981]
    x3s = x3.reshape(23, S, 128).transpose(1, 0, 2).reshape(S, 23 * 128)
    assert np.array_equal(got, oracle.frames(np.ascontiguousarray(x3s), 23 * 128, rate, threads=4))
    agg.close()
    for b in d.values():
        b.free()


@pytest.mark.parametrize("rate, knobs", [(48000, {}), (44100, {}), (48000, dict(agg_walk_regs=32)), (48000, dict(agg_warp_frames=0))])
def test_aggregator_streaming_regime_at_scale(ctx, rate, knobs):
    """The regime a live server runs in, at a size that takes the large-push kernels: 60 000 sessions, every push
    carries ONE quantum of most sessions (a few send two, a few skip), the flushes (16 ms here, ~6 quanta) fall in
    different pushes for different sessions.  After the first push the library takes the kernels built for short
    lists (4-entry flush-rule kernel, one tile per append trip, one frame per warp in the convert kernel); every
    flushed frame is mostly quanta converted on arrival in earlier pushes plus one or two new ones.  With
    agg_walk_regs = 32 the other builds run on the same pushes, with agg_warp_frames = 0 the CTA-per-frame convert.  Every frame against the oracle, levels of every 16th session."""
    rng = np.random.RandomState(int(rate) // 100 + len(knobs))
    S, q_ms, desired, n_push = 60000, 128e3 / rate, 16.0, 12
    try:
        ctx.tune(**knobs)
        agg = pcm_frontend.Aggregator(ctx, S, rate, desired_ms=desired, max_frame_quanta=16, max_push_quanta=2 * S, max_out_samples=S * 360)
        agg.set_meter(True)
        agg.reset_range(0, S, 0.0)
        start_t = desired * rng.rand(S)                    # so that the sessions flush in different pushes
        period = q_ms + rng.uniform(-0.05, 0.05, S)
        done = np.zeros(S, np.int64)
        audio, times, got = [], [], {}
        for p in range(n_push):
            u = rng.rand(S)
            m = np.where(u < 0.03, 0, np.where(u < 0.06, 2, 1)) if p else np.ones(S, np.int64)
            sess = np.repeat(np.arange(S, dtype=np.uint32), m)
            k = done[sess] + (np.arange(sess.size) - np.repeat(np.cumsum(m) - m, m))
            t = start_t[sess] + (k + 1) * period[sess]
            order = np.lexsort((sess, t))
            sess, t, k = sess[order], t[order], k[order]
            x = (rng.uniform(-1.1, 1.1, size=(sess.size, 128)) * rng.uniform(0.05, 1, size=(sess.size, 1))).astype(np.float32)
            audio.append((sess, k, x)); times.append(t)
            done += m
            for s_, pcm in agg.push(sess, t, x):
                got.setdefault(s_, []).append(pcm)
        levels = agg.levels(0, S)
        agg.close()
    finally:
        ctx.tune(agg_walk_regs=0, agg_warp_frames=1)
    all_s = np.concatenate([a[0] for a in audio]); all_k = np.concatenate([a[1] for a in audio])
    all_x = np.concatenate([a[2] for a in audio]); all_t = np.concatenate(times)
    order = np.lexsort((all_k, all_s))
    all_s, all_x, all_t = all_s[order], all_x[order], all_t[order]
    bounds = np.searchsorted(all_s, np.arange(S + 1))
    n_frames = 0
    for s_ in range(S):
        a, b = bounds[s_], bounds[s_ + 1]
        want, flushed_q, _ = _session_frames(all_x[a:b], all_t[a:b], rate, desired_ms=desired)
        have = got.get(s_, [])
        assert len(have) == len(want), (s_, len(have), len(want))
        for u_, v_ in zip(have, want):
            assert u_.size == v_.size and np.array_equal(u_, v_), s_
        n_frames += len(want)
        if s_ % 16 == 0:
            want_lv = oracle.level_meter(all_x[a:a + flushed_q].reshape(-1))[-1] if flushed_q else 0.0
            assert levels[s_:s_ + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], (s_, levels[s_], want_lv)
    assert n_frames >= S                                   # ~32 ms of audio per session against 16 ms flushes


@pytest.mark.parametrize("rate, max_q", [(96000, 64), (96000, 96), (88200, 64), (16000, 32), (8000, 32), (24000, 32), (192000, 96),
                                         (22050, 32), (11025, 32), (32000, 32), (47999, 32), (48001, 32), (12345.678, 32), (176400, 96)])
def test_aggregator_long_frames(ctx, rate, max_q):
    """60 ms at 96 kHz is 45-46 quanta per frame: more than one warp's worth of push quanta per frame, and
    -- with pushes cut at uneven places -- more than 32 quanta carried over in a session's buffer.  max_q = 64
    takes the one-frame-per-step convert kernel (descriptors in registers, two index registers per lane),
    max_q = 96 the general one; 88.2 kHz adds the pick table.  The low rates run the one-frame kernel at ratio 1
    (16 kHz: encode only), 1.5 (24 kHz, table) and 0.5 (8 kHz: twice as many outputs as inputs, table).
    The second line of rates: picks by table at ratios whose per-quantum output count varies from quantum to quantum
    (the staging of waiting quanta cuts the output sequence at exactly those places).
    Frames and levels against the oracle."""
    rng = np.random.RandomState(int(rate) // 100 + max_q)
    S, q_ms = 23, 128e3 / rate
    nq = int(400 / q_ms)
    audio = (rng.uniform(-1.05, 1.05, size=(S, nq, 128)) * rng.uniform(0.1, 1, size=(S, 1, 1))).astype(np.float32)
    times = 3.0 * rng.rand(S, 1) + np.cumsum(q_ms + rng.uniform(-0.2, 0.2, size=(S, nq)), axis=1)
    events = sorted((times[s, q], s, q) for s in range(S) for q in range(nq))
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=max_q, max_push_quanta=2048, max_out_samples=1 << 18)
    agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    got = {s: [] for s in range(S)}
    i, k, sizes = 0, 0, [S * 40, 7, S * 3, S * 37, 100, S * 50, 1]      # cuts that leave 0 ... 44 quanta waiting
    while i < len(events):
        chunk = events[i:i + sizes[k % len(sizes)]]
        i += len(chunk); k += 1
        for s, pcm in agg.push(np.array([e[1] for e in chunk]), np.array([e[0] for e in chunk]),
                               np.stack([audio[e[1], e[2]] for e in chunk])):
            got[s].append(pcm)
    levels = agg.levels(0, S)
    agg.close()
    for s in range(S):
        want, start, _ = _session_frames(audio[s], times[s], rate)
        assert len(got[s]) == len(want) >= 5, (s, len(got[s]), len(want))
        if rate >= 88200:
            assert max(f.size for f in want) > 32 * 128 * 16000 // rate      # frames longer than 32 quanta
        for a, b in zip(got[s], want):
            assert a.size == b.size and np.array_equal(a, b), s
        want_lv = oracle.level_meter(audio[s, :start].reshape(-1))[-1]
        assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], s


@pytest.mark.parametrize("max_q", [32, 96])
@pytest.mark.parametrize("rate", [48000, 44100])
def test_aggregator_unaligned_device_output(ctx, rate, max_q):
    """An output pointer that is only 2-byte aligned (frames still start every 8 samples from it): both convert
    kernels fall back to 16-bit stores; nothing outside the frames is written."""
    rng = np.random.RandomState(5)
    S, nq = 300, 24
    n = S * nq
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=max_q, max_push_quanta=n, max_out_samples=16)
    agg.reset_range(0, S, 0.0)
    sess = np.tile(np.arange(S, dtype=np.uint32), nq)
    t = np.repeat((np.arange(nq) + 1) * (128e3 / rate), S)
    x = rng.uniform(-1, 1, size=(n, 128)).astype(np.float32)
    flush, _ = oracle.aggregator_plan(t[::S], 60.0, 0.0)
    fq = int(np.nonzero(flush)[0][0]) + 1                           # quanta in the one frame every session flushes
    n_out = oracle.resample_len(fq * 128, rate)
    pad = (n_out + 7) // 8 * 8
    d = {k: ctx.device_buffer(v) for k, v in dict(s=n * 4, t=n * 8, x=n * 512, o=(S * pad + 16) * 2, fl=n * 4, fo=n * 8).items()}
    d["s"].upload(sess); d["t"].upload(t); d["x"].upload(x)
    d["o"].upload(np.full(S * pad + 16, 0x2222, np.int16))
    agg.push_device(n, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr + 2, S * pad, None, d["fl"].ptr, d["fo"].ptr, n)
    nf, ns = agg.wait()
    assert nf == S and ns == S * pad
    fl, fo = d["fl"].download(np.uint32, S), d["fo"].download(np.uint64, S)
    assert (fl == n_out).all() and (fo == np.arange(S) * pad).all()
    raw = d["o"].download(np.int16, S * pad + 16)
    assert raw[0] == 0x2222 and (raw[1 + S * pad:] == 0x2222).all()
    got = raw[1:1 + S * pad].reshape(S, pad)
    xs = x.reshape(nq, S, 128).transpose(1, 0, 2)[:, :fq].reshape(S, fq * 128)
    assert np.array_equal(got[:, :n_out], oracle.frames(np.ascontiguousarray(xs), fq * 128, rate, threads=4))
    assert (got[:, n_out:] == 0x2222).all()                          # the padding between frames is left alone
    agg.close()
    for b in d.values():
        b.free()


@pytest.mark.parametrize("shift", [0, 1])
@pytest.mark.parametrize("rate", [48000, 44100])
def test_aggregator_streaming_device_output_canaries(ctx, rate, shift):
    """One quantum per session per push, everything in device memory: after the first push the frames are converted
    one per warp (k_agg_convert_warp), mostly by copying the outputs staged on arrival.  The output pointer is 16-byte
    aligned (shift 0: 16-byte copies) or only 2-byte aligned (shift 1: 16-bit stores); the frames must be the
    oracle's and nothing around or between them may be written."""
    rng = np.random.RandomState(7 + shift)
    S, nq, q_ms = 500, 26, 128e3 / rate
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=S, max_out_samples=16)
    agg.reset_range(0, S, 0.0)
    x = rng.uniform(-1.05, 1.05, size=(nq, S, 128)).astype(np.float32)
    t_all = (np.arange(nq) + 1) * q_ms
    flush, _ = oracle.aggregator_plan(t_all, 60.0, 0.0)
    fq = int(np.nonzero(flush)[0][0]) + 1                           # the push on which every session flushes
    n_out = oracle.resample_len(fq * 128, rate)
    pad = (n_out + 7) // 8 * 8
    d = {k: ctx.device_buffer(v) for k, v in dict(s=S * 4, t=S * 8, x=S * 512, o=(S * pad + 16) * 2, fl=S * 4, fo=S * 8).items()}
    d["s"].upload(np.arange(S, dtype=np.uint32))
    try:
        for p in range(nq):
            d["t"].upload(np.full(S, t_all[p])); d["x"].upload(np.ascontiguousarray(x[p]))
            d["o"].upload(np.full(S * pad + 16, 0x2222, np.int16))
            agg.push_device(S, d["s"].ptr, d["t"].ptr, d["x"].ptr, d["o"].ptr + 2 * shift, S * pad, None, d["fl"].ptr, d["fo"].ptr, S)
            nf, ns = agg.wait()
            raw = d["o"].download(np.int16, S * pad + 16)
            if p + 1 != fq:
                assert nf == 0 and ns == 0 and (raw == 0x2222).all(), p
                continue
            assert nf == S and ns == S * pad
            fl, fo = d["fl"].download(np.uint32, S), d["fo"].download(np.uint64, S)
            assert (fl == n_out).all() and (fo == np.arange(S) * pad).all()
            assert (raw[:shift] == 0x2222).all() and (raw[shift + S * pad:] == 0x2222).all()
            got = raw[shift:shift + S * pad].reshape(S, pad)
            xs = x[:fq].transpose(1, 0, 2).reshape(S, fq * 128)
            assert np.array_equal(got[:, :n_out], oracle.frames(np.ascontiguousarray(xs), fq * 128, rate, threads=4))
            assert (got[:, n_out:] == 0x2222).all()                  # the padding between frames is left alone
    finally:
        agg.close()
        for b in d.values():
            b.free()


@pytest.mark.parametrize("knobs", [dict(agg_pdl=0), dict(agg_persistent=0), dict(agg_block=64), dict(agg_block=256),
                                   dict(agg_block=128, agg_stage_kb=68), dict(agg_block=256, agg_stage_kb=51, agg_pdl=0),
                                   dict(agg_walk_regs=4), dict(agg_walk_regs=32), dict(agg_walk_regs=4, agg_warp_frames=0)])
@pytest.mark.parametrize("rate", [48000, 44100])
def test_aggregator_every_kernel_form(ctx, knobs, rate):
    """The tuning knobs choose between kernel forms (one frame per step with descriptors in registers / several
    frames per step with shared-memory descriptors; persistent or oversized grid; plain or programmatic dependent
    launches; the flush-rule kernel sorting its lists of 9-12 quanta in registers or -- forced to the streaming
    build -- in global memory, and with it the warp-per-frame convert kernel taking frames of 9-12 new quanta through
    its four-slot window three times over).  Every form must give the oracle's frames and levels: 700 sessions, buffers carried across three
    pushes, sessions interleaved at random, the meter on."""
    defaults = dict(agg_pdl=1, agg_persistent=1, agg_block=128, agg_stage_kb=20, agg_walk_regs=0, agg_warp_frames=1)
    rng = np.random.RandomState(sum(knobs.values()) + rate // 100)
    S, q_ms, per_push = 700, 128e3 / rate, (10, 9, 12)
    try:
        ctx.tune(**knobs)                                          # read when the aggregator is created / on every push
        agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=S * 12, max_out_samples=S * 1000)
        agg.set_meter(True)
        agg.reset_range(0, S, 0.0)
        nq = sum(per_push)
        audio = rng.uniform(-1.02, 1.02, size=(S, nq, 128)).astype(np.float32)
        times = 2.0 * rng.rand(S, 1) + np.cumsum(q_ms + rng.uniform(-0.3, 0.3, size=(S, nq)), axis=1)
        got = {s: [] for s in range(S)}
        q0 = 0
        for pp in per_push:
            ev = sorted((times[s, q], s, q) for s in range(S) for q in range(q0, q0 + pp))
            q0 += pp
            for s, pcm in agg.push(np.array([e[1] for e in ev]), np.array([e[0] for e in ev]), np.stack([audio[e[1], e[2]] for e in ev])):
                got[s].append(pcm)
        levels = agg.levels(0, S)
        agg.close()
    finally:
        ctx.tune(**defaults)
    n_frames = 0
    for s in range(S):
        want, start, _ = _session_frames(audio[s], times[s], rate)
        assert len(got[s]) == len(want), (s, len(got[s]), len(want))
        for a, b in zip(got[s], want):
            assert a.size == b.size and np.array_equal(a, b), s
        n_frames += len(want)
        if start:
            want_lv = oracle.level_meter(audio[s, :start].reshape(-1))[-1]
            assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], s
    assert n_frames >= S


@pytest.mark.parametrize("quantum, max_q", [(64, 64), (64, 96), (256, 16), (32, 128), (512, 8)])
def test_aggregator_other_quantum_sizes(ctx, quantum, max_q):
    """The AudioWorklet render quantum is 128 today, but nothing in App.tsx:264-289 depends on it: the handler appends
    whatever e.data holds.  Quanta of 32 ... 512 samples take the generic copy loops of every aggregator kernel
    (the 128-sample case has its own); frames and levels against the oracle, the meter on."""
    rate = 48000
    rng = np.random.RandomState(quantum + max_q)
    S, q_ms = 61, quantum * 1e3 / rate
    nq = int(260 / q_ms)
    audio = (rng.uniform(-1.05, 1.05, size=(S, nq, quantum)) * rng.uniform(0.1, 1, size=(S, 1, 1))).astype(np.float32)
    times = 3.0 * rng.rand(S, 1) + np.cumsum(q_ms + rng.uniform(-0.1, 0.1, size=(S, nq)) * min(q_ms, 1.0), axis=1)
    events = sorted((times[s, q], s, q) for s in range(S) for q in range(nq))
    agg = pcm_frontend.Aggregator(ctx, S, rate, quantum=quantum, max_frame_quanta=max_q, max_push_quanta=4096, max_out_samples=1 << 18)
    agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    got = {s: [] for s in range(S)}
    i, k, sizes = 0, 0, [S * 11, 5, S * 30, 97, 4096]
    while i < len(events):
        chunk = events[i:i + sizes[k % len(sizes)]]
        i += len(chunk); k += 1
        for s, pcm in agg.push(np.array([e[1] for e in chunk]), np.array([e[0] for e in chunk]),
                               np.stack([audio[e[1], e[2]] for e in chunk])):
            got[s].append(pcm)
    levels = agg.levels(0, S)
    agg.close()
    for s in range(S):
        flush, _ = oracle.aggregator_plan(times[s], 60.0, 0.0)
        want, start = [], 0
        for q in np.nonzero(flush)[0]:
            buf = audio[s, start:q + 1].reshape(1, -1)
            want.append(oracle.frames(buf, buf.shape[1], rate)[0])
            start = q + 1
        assert len(got[s]) == len(want) >= 3, (s, len(got[s]), len(want))
        for a, b in zip(got[s], want):
            assert a.size == b.size and np.array_equal(a, b), s
        want_lv = oracle.level_meter(audio[s, :start].reshape(-1), quantum)[-1]
        assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], s


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_aggregator_fuzz_with_rejected_pushes(ctx, seed):
    """Random pushes, some of which the library must refuse -- an unknown session id in the middle of a push, more
    flushed frames than the output holds -- and then accept once the cause is removed (the bad id dropped, the push
    cut in two).  A refused push must leave no trace: in the end every session has exactly the oracle's frames and level."""
    rng = np.random.RandomState(1000 + seed)
    S, rate, q_ms = 97, 48000, 128e3 / 48000
    nq = 150
    audio = (rng.uniform(-1.05, 1.05, size=(S, nq, 128)) * rng.uniform(0.05, 1, size=(S, 1, 1))).astype(np.float32)
    times = 4.0 * rng.rand(S, 1) + np.cumsum(q_ms + rng.uniform(-0.3, 0.3, size=(S, nq)), axis=1)
    events = sorted((times[s, q], s, q) for s in range(S) for q in range(nq))
    agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=4096, max_out_samples=30 * 984)
    agg.set_meter(True)
    agg.reset_range(0, S, 0.0)
    got = {s: [] for s in range(S)}
    refused = {"bad_id": 0, "too_big": 0}

    def push(chunk):
        try:
            frames = agg.push(np.array([e[1] for e in chunk], dtype=np.uint32), np.array([e[0] for e in chunk]),
                              np.stack([audio[e[1] % S, e[2]] for e in chunk]))
        except pcm_frontend.PcmError as err:
            if err.status == pcm_frontend._capi.ERR_INVALID:             # the injected unknown session: drop it, push again
                refused["bad_id"] += 1
                return push([e for e in chunk if e[1] < S])
            assert err.status == pcm_frontend._capi.ERR_CAPACITY and len(chunk) > 1, err
            refused["too_big"] += 1                                       # more frames than the output holds: halves, in order
            return push(chunk[:len(chunk) // 2]), push(chunk[len(chunk) // 2:])
        for s, pcm in frames:
            got[s].append(pcm)

    i = 0
    while i < len(events):
        n = int(rng.choice([1, 17, 97, 400, 1500, 4000]))
        chunk = list(events[i:i + n])
        i += len(chunk)
        if rng.rand() < 0.3:                                              # an unknown session somewhere in the push
            k = int(rng.randint(0, len(chunk)))
            chunk.insert(k, (chunk[k][0], S + int(rng.randint(0, 5)), 0))
        push(chunk)
    levels = agg.levels(0, S)
    agg.close()
    assert refused["bad_id"] >= 3 and refused["too_big"] >= 3, refused
    for s in range(S):
        want, start, _ = _session_frames(audio[s], times[s], rate)
        assert len(got[s]) == len(want) >= 5, (s, len(got[s]), len(want))
        for a, b in zip(got[s], want):
            assert a.size == b.size and np.array_equal(a, b), s
        want_lv = oracle.level_meter(audio[s, :start].reshape(-1))[-1]
        assert levels[s:s + 1].view(np.uint64)[0] == np.array([want_lv]).view(np.uint64)[0], s


def test_aggregator_quantum_limit(ctx):
    """A quantum the append kernel could not stage 64 of is refused when the aggregator is created, with a message that
    says so (not by a CUDA error on the first push); the largest allowed one works."""
    with pytest.raises(pcm_frontend.PcmError) as ei:
        pcm_frontend.Aggregator(ctx, 4, 48000, quantum=1024, max_frame_quanta=4, max_push_quanta=64, max_out_samples=1 << 14)
    assert "quantum too long" in str(ei.value)
    agg = pcm_frontend.Aggregator(ctx, 4, 48000, quantum=876, max_frame_quanta=4, max_push_quanta=64, max_out_samples=1 << 14)
    agg.reset_range(0, 4, 0.0)
    x = np.linspace(-1, 1, 3 * 876, dtype=np.float32).reshape(3, 876)
    frames = agg.push(np.array([1, 1, 1]), np.array([10.0, 30.0, 61.0]), x)
    agg.close()
    assert len(frames) == 1 and frames[0][0] == 1
    assert np.array_equal(frames[0][1], oracle.frames(x.reshape(1, -1), 3 * 876, 48000)[0])
