"""CPU tests of the ingestion glue: message classification with the
`ws` isBinary flag, the control frames of apps/voice/src/server.ts:267-280 plus the audio_format
announcement, the bounded queue and its three overflow policies, tick deadlines, late frames, the
egress table and the keep-alive schedule (against the oracle's event-by-event timer).  The converter
is a stand-in built on the oracle here (no GPU); tests/test_gpu_ingest.py runs the real one."""
import ctypes as C
import json

import numpy as np
import pytest

import oracle
import pcm_frontend
from pcm_frontend import ingest


class OracleConverter:
    """convert_group() stand-in for CPU tests: same contract as ingest.GpuGroupConverter."""

    def __init__(self):
        self.calls = []

    def convert_group(self, rate, n_in, frames):
        self.calls.append((rate, n_in, len(frames)))
        x = np.stack(frames) if n_in else np.zeros((len(frames), 0), np.float32)
        pcm = oracle.frames(np.ascontiguousarray(x, dtype=np.float32), n_in, rate)
        block = memoryview(pcm.tobytes())
        n = pcm.shape[1] * 2
        return block, [block[r * n:(r + 1) * n] for r in range(pcm.shape[0])]


def _conn(sid, sink):
    return ingest.Connection(sid, lambda b: sink.append((sid, bytes(b))))


def test_classify_uses_the_is_binary_flag_not_the_type():
    # with ws ^8 a text frame also arrives as a Buffer: only the flag tells them apart
    text = json.dumps({"type": "context_update", "payload": {"url": "x"}}).encode()
    assert ingest.classify_message(text, False) == (ingest.CONTROL, {"type": "context_update", "payload": {"url": "x"}})
    assert ingest.classify_message(text, True) == (ingest.AUDIO, text)          # same bytes, binary frame: audio
    assert ingest.classify_message(b"not json", False) == (ingest.IGNORED, None)   # server.ts:281-283
    assert ingest.classify_message(b"[1, 2]", False) == (ingest.IGNORED, None)
    assert ingest.classify_message(b"\xff\xfe", False) == (ingest.IGNORED, None)


def test_control_frames():
    closed = []
    c = ingest.Connection(1, lambda b: None, close_upstream=lambda: closed.append(1))
    assert c.encoding == "pcm16" and c.sample_rate == 16000.0                    # the reference's clients
    assert c.on_control({"type": "context_update", "payload": {"a": 1}}) == "context_update"
    assert c.on_control({"type": "context_update", "payload": {"b": 2}}) == "context_update" and c.context == {"a": 1, "b": 2}
    assert c.on_control({"type": "context_update", "payload": "nope"}) is None
    assert c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 44100}) == "audio_format"
    assert (c.encoding, c.sample_rate) == ("f32le", 44100.0)
    for bad in ({"encoding": "mp3", "sample_rate": 48000}, {"encoding": "f32le", "sample_rate": 0},
                {"encoding": "f32le", "sample_rate": "48k"}, {"encoding": "f32le", "sample_rate": True}, {}):
        assert c.on_control(dict(bad, type="audio_format")) == "bad_audio_format"
    assert c.protocol_errors == 5 and (c.encoding, c.sample_rate) == ("f32le", 44100.0)
    assert c.on_control({"type": "something_else"}) is None
    assert c.on_control({"type": "close"}) == "close" and c.closed and closed == [1]


def test_pcm16_clients_are_relayed_untouched_like_the_reference():
    sink = []
    b = ingest.TickBatcher(OracleConverter())
    c = _conn("legacy", sink)
    payload = np.arange(960, dtype="<i2").tobytes()
    assert b.offer(c, *ingest.classify_message(payload, True), now_ms=0.0)
    assert sink == [("legacy", payload)] and b.pending == 0 and b.stats["relayed_pcm16"] == 1


def test_mixed_rates_one_call_per_class_and_bit_exact_payloads():
    rng = np.random.RandomState(1)
    sink, conv = [], OracleConverter()
    b = ingest.TickBatcher(conv, tick_ms=20.0)
    conns, want = [], {}
    for sid, (rate, n_in) in enumerate([(48000, 2944), (44100, 2646), (48000, 2944), (16000, 960), (48000, 2816), (44100, 2646)]):
        c = _conn(sid, sink)
        b.offer(c, ingest.CONTROL, {"type": "audio_format", "encoding": "f32le", "sample_rate": rate}, 0.0)
        x = rng.uniform(-1.2, 1.2, n_in).astype("<f4")
        assert b.offer(c, *ingest.classify_message(x.tobytes(), True), now_ms=1.0 + sid)
        want[sid] = oracle.frames(x.reshape(1, -1), n_in, rate)[0].astype("<i2").tobytes()
        conns.append(c)
    assert not b.due(20.9) and b.due(21.0)                                      # the oldest frame has waited tick_ms
    assert b.flush(21.0) == 6
    assert sorted(conv.calls) == sorted([(48000.0, 2944, 2), (44100.0, 2646, 2), (16000.0, 960, 1), (48000.0, 2816, 1)])
    assert dict(sink) == want and b.pending == 0 and not b.due(1e9)


def test_malformed_and_oversized_float_frames_are_counted_not_queued():
    b = ingest.TickBatcher(OracleConverter(), max_frame_samples=1000)
    c = _conn(0, [])
    c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 48000})
    assert b.offer(c, ingest.AUDIO, b"\x00" * 10, 0.0) and b.offer(c, ingest.AUDIO, b"\x00" * 4004, 0.0)
    assert b.pending == 0 and c.protocol_errors == 2 and b.stats["protocol_errors"] == 2


@pytest.mark.parametrize("policy", ingest.OVERFLOW_POLICIES)
def test_over_full_tick(policy):
    sink = []
    b = ingest.TickBatcher(OracleConverter(), max_frames=4, overflow=policy)
    c = _conn(0, sink)
    c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 16000})
    frames = [np.full(8, 0.1 * k, "<f4") for k in range(6)]
    results = [b.offer(c, ingest.AUDIO, f.tobytes(), float(k)) for k, f in enumerate(frames)]
    assert b.pending == 4 and b.due(5.0)                                         # a full queue is due at once
    b.flush(5.0)
    got = [np.frombuffer(p, "<i2")[0] for _, p in sink]
    want = [int(oracle.float_to_16bit_pcm(f)[0]) for f in frames]
    if policy == "drop_oldest":
        assert results == [True] * 6 and got == want[2:] and b.stats["dropped_oldest"] == 2
    elif policy == "drop_newest":
        assert results == [True] * 6 and got == want[:4] and b.stats["dropped_newest"] == 2
    else:
        assert results == [True] * 4 + [False] * 2 and got == want[:4] and b.stats["blocked"] == 2


def test_late_frames_go_to_the_next_tick_and_closed_connections_get_nothing():
    sink = []
    b = ingest.TickBatcher(OracleConverter(), tick_ms=20.0)
    a, late, gone = _conn("a", sink), _conn("late", sink), _conn("gone", sink)
    for c in (a, late, gone):
        c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 16000})
    x = np.full(16, 0.5, "<f4").tobytes()

    def send_and_inject(payload, _orig=a.send_audio):
        _orig(payload)
        b.offer(late, ingest.AUDIO, x, 21.5)                                      # arrives while the flush is running
    a.send_audio = send_and_inject
    b.offer(a, ingest.AUDIO, x, 0.0)
    b.offer(gone, ingest.AUDIO, x, 0.5)
    gone.on_control({"type": "close"})
    assert b.flush(21.0) == 1 and [s for s, _ in sink] == ["a"] and b.pending == 1
    assert not b.due(30.0) and b.due(41.5)
    assert b.flush(41.5) == 1 and [s for s, _ in sink] == ["a", "late"]
    assert b.offer(gone, ingest.AUDIO, x, 50.0) and b.pending == 0                # closed: ignored


def test_egress_table_rows_and_aggregator_frames():
    L = pcm_frontend.load()
    off, ln = np.empty(5, np.uint64), np.empty(5, np.uint32)
    assert L.pcmb200_egress_table(5, 981, 984, off.ctypes.data, ln.ctypes.data) == 0
    assert off.tolist() == [r * 984 * 2 for r in range(5)] and ln.tolist() == [1962] * 5      # byteLength = 2 * N_out
    fo, fl = np.array([0, 984, 1968 + 8], np.uint64), np.array([981, 981, 42], np.uint32)
    assert L.pcmb200_egress_table_frames(3, fo.ctypes.data, fl.ctypes.data, off.ctypes.data, ln.ctypes.data) == 0
    assert off[:3].tolist() == [0, 1968, 3952] and ln[:3].tolist() == [1962, 1962, 84]
    assert L.pcmb200_egress_table(2, 10, 5, off.ctypes.data, ln.ctypes.data) == pcm_frontend._capi.ERR_INVALID


def test_keepalive_schedule_matches_the_timer_simulation():
    """pcmb200_keepalive_due (closed form over a fleet) against the oracle's event-by-event timer of
    App.tsx:291-296 per session: random start times, irregular polls (some > 2 s apart, so ticks
    coalesce), sockets that close and reopen, a session stopped half way."""
    rng = np.random.RandomState(7)
    S = 64
    polls = np.cumsum(rng.choice([20.0, 20.0, 20.0, 700.0, 2500.0, 5100.0], size=400))
    sched = ingest.KeepAliveScheduler(S)
    starts = rng.uniform(0, 3000, S)
    started = np.zeros(S, bool)
    open_at = rng.rand(len(polls), S) > 0.2
    stop_poll = {3: 150, 9: 10}
    sent = np.zeros((len(polls), S), np.uint8)
    for p, now in enumerate(polls):
        for s in range(S):
            if not started[s] and now >= starts[s]:
                sched.start_session(s, starts[s]); started[s] = True
            if stop_poll.get(s) == p:
                sched.stop_session(s)
            elif started[s] and stop_poll.get(s, 1 << 30) > p:
                sched.set_open(s, bool(open_at[p, s]))
        for s in sched.poll(now):
            sent[p, s] = 1
    for s in range(S):
        first = int(np.searchsorted(polls, starts[s]))                            # polls before the start do not see the timer
        last = stop_poll.get(s, len(polls))
        want = np.zeros(len(polls), np.uint8)
        if last > first:
            want[first:last] = oracle.keepalive_timer(starts[s], polls[first:last], open_at[first:last, s])
        assert np.array_equal(sent[:, s], want), s
    assert sent.sum() > S * 10
    assert np.array_equal(pcm_frontend.keepalive_frame(), oracle.keepalive_frame())       # the payload: 1600 zeros


def test_keepalive_due_capacity_leaves_state_untouched():
    L = pcm_frontend.load()
    start, done, op = np.zeros(4), np.zeros(4, np.uint64), np.ones(4, np.uint8)
    due, n = np.empty(4, np.uint32), C.c_uint32(0)
    assert L.pcmb200_keepalive_due(2000.0, 4, start.ctypes.data, done.ctypes.data, op.ctypes.data, 2000.0,
                                   due.ctypes.data, 2, C.byref(n)) == pcm_frontend._capi.ERR_CAPACITY
    assert done.tolist() == [0, 0, 0, 0]
    assert L.pcmb200_keepalive_due(2000.0, 4, start.ctypes.data, done.ctypes.data, op.ctypes.data, 2000.0,
                                   due.ctypes.data, 4, C.byref(n)) == 0 and n.value == 4 and done.tolist() == [1] * 4


def test_quanta_framing_is_announced_in_the_audio_format_frame():
    """"framing": "quanta" (raw AudioWorklet quanta, aggregated by the server) needs Float32; anything else is refused."""
    c = ingest.Connection("q", lambda b: None)
    assert c.framing == "frames" and c.slot == -1
    assert c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 48000}) == "audio_format" and c.framing == "frames"
    assert c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 44100, "framing": "quanta"}) == "audio_format"
    assert c.framing == "quanta" and c.sample_rate == 44100.0
    for bad in ({"encoding": "pcm16", "framing": "quanta"}, {"encoding": "f32le", "framing": "pages"}):
        assert c.on_control({"type": "audio_format", "sample_rate": 48000, **bad}) == "bad_audio_format"
    assert c.framing == "quanta" and c.protocol_errors == 2


def test_quantum_batcher_protocol_slots_and_back_pressure():
    """QuantumBatcher with a stand-in aggregator: malformed payloads are counted, slots are handed out on the first
    message and reused after disconnect, a full tick refuses the message (back-pressure), frames go to their connection."""
    class FakeAgg:
        def __init__(self):
            self.free, self.staged, self.cap = [1, 0], [], 4
        def open(self, now):
            return self.free.pop() if self.free else -1
        def release(self, slot):
            self.free.append(slot)
        def add(self, slot, samples, now):
            if len(self.staged) + samples.size // 128 > self.cap:
                return False
            self.staged += [(slot, now)] * (samples.size // 128)
            return True
        def push(self):
            out, self.staged = [(s, bytes([s])) for s, _ in self.staged], []
            return out
        def close(self):
            pass
    got = {k: [] for k in "abc"}
    conns = {k: ingest.Connection(k, got[k].append) for k in "abc"}
    for c in conns.values():
        c.on_control({"type": "audio_format", "encoding": "f32le", "sample_rate": 48000, "framing": "quanta"})
    qb = ingest.QuantumBatcher(lambda rate: FakeAgg(), tick_ms=20.0)
    q = np.zeros(128, np.float32).tobytes()
    assert qb.offer(conns["a"], q[:100], 0.0) and conns["a"].protocol_errors == 1 and not qb.due(100.0)
    assert qb.offer(conns["a"], q * 2, 1.0) and qb.offer(conns["b"], q, 2.0) and conns["a"].slot == 0 and conns["b"].slot == 1
    assert qb.offer(conns["c"], q, 3.0) and conns["c"].slot == -1 and qb.stats["no_slot"] == 1      # both slots taken
    assert not qb.offer(conns["b"], q * 2, 4.0) and qb.stats["blocked"] == 1                          # 3 staged + 2 > 4
    assert not qb.due(20.9) and qb.due(21.0)
    assert qb.flush() == 3 and got["a"] == [b"\x00", b"\x00"] and got["b"] == [b"\x01"] and not qb.due(1e9)
    qb.disconnect(conns["a"])
    assert qb.offer(conns["c"], q, 30.0) and conns["c"].slot == 0                                     # a's slot, reused
    conns["b"].closed = True
    assert qb.offer(conns["b"], q, 31.0) and qb.flush() == 1 and got["c"] == [b"\x00"]
    qb.close()
