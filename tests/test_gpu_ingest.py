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


def test_quanta_framing_server_side_aggregation(ctx):
    """"framing": "quanta": 52 connections (48 kHz and 44.1 kHz) forward raw AudioWorklet quanta, 1-3 per message,
    with jittered arrival; the server aggregates on the GPU and every connection must receive exactly the frames the
    reference's aggregator (App.tsx:273-288, via the oracle) makes of its quanta with the message arrival times as
    performance.now().  One connection drops out half way and its slot is reused by a newcomer."""
    rng = np.random.RandomState(11)
    made = []

    def make(rate):
        made.append(pcm_frontend.GpuQuantumAggregator(ctx, rate, max_sessions=48, max_push_quanta=512))
        return made[-1]
    qb = pcm_frontend.QuantumBatcher(make, tick_ms=20.0)
    conns, sent, audio, stamps = [], {}, {}, {}
    for c in range(52):
        rate = 48000.0 if c < 40 else 44100.0
        sent[c] = []
        conn = pcm_frontend.Connection(c, (lambda b, c=c: sent[c].append(np.frombuffer(bytes(b), dtype="<i2").copy())))
        assert conn.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": rate, "framing": "quanta"}) == "audio_format"
        conns.append(conn)
        audio[c], stamps[c] = [], []
    assert conns[0].on_control({"type": "audio_format", "encoding": "pcm16", "sample_rate": 48000, "framing": "quanta"}) == "bad_audio_format"
    # the event stream: every connection sends a message of k quanta every k quantum periods (+ jitter)
    events = []
    for c, conn in enumerate(conns):
        q_ms, t = 128e3 / conn.sample_rate, 5.0 * rng.rand()
        while t < 400.0:
            k = int(rng.randint(1, 4))
            t += k * q_ms + rng.uniform(-0.3, 0.3)
            events.append((t, c, k))
    events.sort()
    newcomer = pcm_frontend.Connection(99, lambda b: sent.setdefault(99, []).append(np.frombuffer(bytes(b), dtype="<i2").copy()))
    newcomer.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 48000, "framing": "quanta"})
    audio[99], stamps[99] = [], []
    for t, c, k in events:
        if c == 7 and t > 200.0:
            if conns[7].slot >= 0:
                qb.disconnect(conns[7])
            c, conn = 99, newcomer                                # the newcomer takes over connection 7's cadence (and slot)
        else:
            conn = conns[c]
        x = (rng.uniform(-1.05, 1.05, size=(k, 128)) * rng.uniform(0.1, 1)).astype(np.float32)
        assert qb.offer(conn, x.tobytes(), t)
        audio[c].append(x)
        stamps[c] += [t] * k
        if qb.due(t):
            qb.flush()
    qb.flush()
    assert newcomer.slot >= 0 and conns[7].slot == -1 and qb.stats["protocol_errors"] == 0 and qb.stats["blocked"] == 0
    before = conns[0].protocol_errors
    assert qb.offer(conns[0], b"\x00" * 100, 500.0) and conns[0].protocol_errors == before + 1   # not a whole quantum: counted, dropped
    n_frames = 0
    for c in list(range(52)) + [99]:
        x, t = np.concatenate(audio[c]), np.asarray(stamps[c])
        rate = 44100.0 if 40 <= c < 52 else 48000.0
        flush, _ = oracle.aggregator_plan(t, 60.0, t[0])           # the slot is opened by the first message: lastSend = its time
        want, start = [], 0
        for q in np.nonzero(flush)[0]:
            buf = x[start:q + 1].reshape(1, -1)
            want.append(oracle.frames(buf, buf.shape[1], rate)[0])
            start = q + 1
        assert len(sent[c]) == len(want), (c, len(sent[c]), len(want))
        for a, b in zip(sent[c], want):
            assert a.size == b.size and np.array_equal(a, b), c
        n_frames += len(want)
    assert n_frames > 250 and qb.stats["frames"] == n_frames and len(made) == 2
    qb.close()


def test_quanta_tick_larger_than_the_output_block_is_pushed_in_halves(ctx):
    """A tick whose frames do not fit the pinned output block is rejected by the library with nothing changed and then
    pushed in halves, in order: same frames as the oracle's, none lost."""
    rng = np.random.RandomState(3)
    S, nq = 64, 24
    g = pcm_frontend.GpuQuantumAggregator(ctx, 48000.0, max_sessions=S, max_push_quanta=S * nq)
    g.max_out = 20 * 984                                             # room for 20 of the 64 frames
    slots = [g.open(0.0) for _ in range(S)]
    x = rng.uniform(-1, 1, size=(nq, S, 128)).astype(np.float32)
    for q in range(nq):
        for s in range(S):
            assert g.add(slots[s], x[q, s], (q + 1) * (128e3 / 48000))
    frames = g.push()
    assert len(frames) == S and sorted(f[0] for f in frames) == sorted(slots)
    for slot, b in frames:
        s = slots.index(slot)
        want = oracle.frames(np.ascontiguousarray(x[:23, s].reshape(1, -1)), 23 * 128, 48000)[0]
        assert np.array_equal(np.frombuffer(bytes(b), dtype="<i2"), want), s
    g.close()
