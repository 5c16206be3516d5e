"""GPU tests of the ingestion glue with the real converter: Float32
frames of many connections at mixed rates go through ONE pcmb200_frames_encode call per (rate, length)
class, and every connection receives a zero-copy VIEW of its 2 * N_out bytes in the pinned output
block (pcmb200_egress_table) that equals the oracle's bytes; late frames, an over-full tick and the
keep-alive frames ride along."""
import json

import numpy as np
import pytest

import oracle
import pcm_frontend
from pcm_frontend import ingest

pytestmark = pytest.mark.gpu

SHAPES = [(48000, 960), (44100, 882), (16000, 320), (48000, 2944), (8000, 160), (44100, 2646), (96000, 1920)]


def test_tick_batching_mixed_rates_zero_copy_views(ctx):
    rng = np.random.RandomState(77)
    conv = ingest.GpuGroupConverter(ctx, max_rows=64, max_n_in=1024)      # small on purpose: it has to grow
    b = ingest.TickBatcher(conv, tick_ms=20.0)
    got, want, kinds = {}, {}, {}
    conns = []
    for sid in range(300):
        rate, n_in = SHAPES[sid % len(SHAPES)]

        def sink(view, sid=sid):
            kinds[sid] = type(view)
            got[sid] = bytes(view)                                        # what sendAudio would put on the wire
        c = ingest.Connection(sid, sink)
        fmt = json.dumps({"type": "audio_format", "encoding": "f32le", "sample_rate": rate}).encode()
        b.offer(c, *ingest.classify_message(fmt, False), now_ms=0.0)     # text frame, delivered as bytes like ws ^8 does
        x = rng.uniform(-1.2, 1.2, n_in).astype("<f4")
        x[:3] = [np.nan, np.inf, -0.0]
        b.offer(c, *ingest.classify_message(x.tobytes(), True), now_ms=float(sid) * 0.01)
        want[sid] = oracle.frames(x.reshape(1, -1), n_in, rate)[0].astype("<i2").tobytes()
        conns.append(c)
    before = ctx.launches
    assert b.due(20.0) and b.flush(20.0) == 300
    assert got == want
    assert all(k is memoryview for k in kinds.values())                   # views into the pinned block, not copies
    assert b.stats["conversions"] == len(SHAPES) and ctx.launches - before >= len(SHAPES)
    # a legacy PCM16 client on the same batcher is relayed untouched, without the GPU
    legacy_out = []
    legacy = ingest.Connection("legacy", lambda p: legacy_out.append(bytes(p)))
    pcm = np.arange(-100, 100, dtype="<i2").tobytes()
    before = ctx.launches
    b.offer(legacy, ingest.AUDIO, pcm, 21.0)
    assert legacy_out == [pcm] and ctx.launches == before
    conv.close()


def test_late_frames_and_over_full_tick_on_the_gpu(ctx):
    rng = np.random.RandomState(3)
    conv = ingest.GpuGroupConverter(ctx, max_rows=16, max_n_in=960)
    b = ingest.TickBatcher(conv, tick_ms=20.0, max_frames=8, overflow="drop_oldest")
    out = []
    conns = [ingest.Connection(s, lambda v, s=s: out.append((s, bytes(v)))) for s in range(12)]
    frames = {}
    for s, c in enumerate(conns):
        c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 48000})
        frames[s] = rng.uniform(-1, 1, 960).astype("<f4")
    late = rng.uniform(-1, 1, 960).astype("<f4")

    def sink0(v, _orig=conns[4].send_audio):
        _orig(v)
        b.offer(conns[0], ingest.AUDIO, late.tobytes(), 25.0)            # arrives while the flush is running
    conns[4].send_audio = sink0
    for s in range(12):                                                   # 12 frames into a queue of 8: 0..3 are dropped
        b.offer(conns[s], ingest.AUDIO, frames[s].tobytes(), float(s))
    assert b.stats["dropped_oldest"] == 4 and b.pending == 8 and b.due(12.0)
    assert b.flush(12.0) == 8
    assert [s for s, _ in out] == list(range(4, 12))
    for s, payload in out:
        assert payload == oracle.frames(frames[s].reshape(1, -1), 960, 48000)[0].astype("<i2").tobytes()
    assert b.pending == 1 and not b.due(30.0) and b.due(45.0)            # the late frame waits for its own tick
    out.clear()
    assert b.flush(45.0) == 1
    assert out == [(0, oracle.frames(late.reshape(1, -1), 960, 48000)[0].astype("<i2").tobytes())]
    conv.close()


def test_keepalive_frames_for_a_fleet(ctx):
    """Every 2 s each open session gets the reference's 1600-sample silence frame (App.tsx:291-296)."""
    sched = ingest.KeepAliveScheduler(1000)
    for s in range(1000):
        sched.start_session(s, now_ms=float(s))                          # sessions start 1 ms apart
    silence = pcm_frontend.keepalive_frame().astype("<i2").tobytes()
    assert silence == b"\x00" * 3200
    sent = np.zeros(1000, int)
    for now in np.arange(0.0, 6100.0, 20.0):                             # the server's 20 ms tick
        for s in sched.poll(now):
            sent[s] += 1
    # session s fires at s + 2000, s + 4000, s + 6000 (ms): three firings before 6100 only for s <= 80 (polls land on
    # multiples of 20 ms, the last one at 6080)
    want = np.array([sum(1 for k in (1, 2, 3) if s + 2000 * k <= 6080) for s in range(1000)])
    assert np.array_equal(sent, want)
