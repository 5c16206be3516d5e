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
