"""Voice-server ingestion glue: what sits between the WebSocket handler of
apps/voice/src/server.ts:97-109, :260-284 and DeepgramClient.sendAudio (apps/voice/src/deepgram.ts:55-58)
when the PCM front-end runs on the server's GPU instead of in every browser.

The reference's handler relays every binary message, untouched, to sendAudio and treats everything
else as an optional JSON control frame ("close", "context_update"; non-JSON text is ignored).  This
module keeps that behaviour for clients that still send PCM16 and adds a Float32 sub-protocol:

    text frame   {"type": "audio_format", "encoding": "f32le", "sample_rate": 48000}
    binary frame raw little-endian Float32 samples of ONE flushed capture buffer
                 (what the browser would have passed to resampleTo16k, App.tsx:281)

and, with "framing": "quanta" in the announcement, a thinner client still: it forwards the AudioWorklet's
128-sample quanta as they come (one or more per message) and the SERVER runs the capture aggregator of
App.tsx:148-157, :273-288 for every connection on the GPU (QuantumBatcher below: the arrival time of a message
plays performance.now(), frames come out when `elapsed >= desiredMs`, exactly as in the browser).

Telling the two kinds of frame apart: with `ws` ^8 (apps/voice/package.json:19) a TEXT frame is also
delivered as a Buffer, so the reference's `Buffer.isBuffer(data)` test (server.ts:261) is true for both
and its control branch is unreachable; the glue must use the `isBinary` flag that `ws` passes as the
second argument of the "message" event.  classify_message() takes that flag.

Float32 frames of all connections are collected per tick and converted with ONE fused call per
(rate, length) class (pcmb200_frames_encode); every session then gets a VIEW of its 2 * N_out bytes in
the pinned output block (pcmb200_egress_table) for the unchanged sendAudio.  The queue is bounded.

node/ingest.ts is the line-for-line TypeScript twin of this file (it cannot run here: no Node).
"""
from __future__ import annotations

import ctypes as C
import json
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

import numpy as np

from ._capi import check, load

AUDIO, CONTROL, IGNORED = "audio", "control", "ignored"
OVERFLOW_POLICIES = ("drop_oldest", "drop_newest", "block")


def classify_message(data: bytes, is_binary: bool) -> Tuple[str, Any]:
    """("audio", bytes) for a binary frame; ("control", dict) for a JSON text frame; ("ignored", None)
    for any other text (the reference swallows non-JSON control messages, server.ts:281-283)."""
    if is_binary:
        return AUDIO, data
    try:
        msg = json.loads(bytes(data).decode("utf-8"))
    except (ValueError, UnicodeDecodeError):
        return IGNORED, None
    return (CONTROL, msg) if isinstance(msg, dict) else (IGNORED, None)


@dataclass
class Connection:
    """Per-connection state (the ClientState of server.ts:88-95 plus the audio format)."""
    session_id: Any
    send_audio: Callable[[Any], None]                 # DeepgramClient.sendAudio, unchanged
    close_upstream: Optional[Callable[[], None]] = None
    encoding: str = "pcm16"                           # the reference's clients send ready-made PCM16
    sample_rate: float = 16000.0
    context: Dict[str, Any] = field(default_factory=dict)
    closed: bool = False
    protocol_errors: int = 0
    framing: str = "frames"                           # "frames": one flushed capture buffer per message; "quanta": raw
    slot: int = -1                                    #   AudioWorklet quanta, aggregated on the GPU (QuantumBatcher slot)

    def on_control(self, msg: dict) -> Optional[str]:
        """server.ts:267-280 ("close", "context_update") + the audio_format announcement.
        Returns the action taken (for logs/tests) or None."""
        t = msg.get("type")
        if t == "close":
            self.closed = True
            if self.close_upstream:
                self.close_upstream()
            return "close"
        if t == "context_update" and isinstance(msg.get("payload"), dict):
            self.context = {**self.context, **msg["payload"]}
            return "context_update"
        if t == "audio_format":
            enc, rate = msg.get("encoding"), msg.get("sample_rate")
            if enc not in ("f32le", "pcm16") or not isinstance(rate, (int, float)) or isinstance(rate, bool) or not rate > 0:
                self.protocol_errors += 1
                return "bad_audio_format"
            framing = msg.get("framing", "frames")
            if framing not in ("frames", "quanta") or (framing == "quanta" and enc != "f32le"):
                self.protocol_errors += 1
                return "bad_audio_format"
            self.encoding, self.sample_rate, self.framing = enc, float(rate), framing
            return "audio_format"
        return None


@dataclass
class _Pending:
    conn: Connection
    samples: np.ndarray        # float32 view of the frame's payload (not copied)
    t_ms: float


class TickBatcher:
    """Bounded cross-connection queue + per-tick conversion.

    offer()  one WebSocket message of one connection (already classified);
    due()    is it time to flush: the oldest queued frame has waited tick_ms, or the queue is full;
    flush()  convert everything queued, one fused call per (rate, length) class, hand every
             connection a view of its PCM16 bytes.

    Overflow policy when max_frames / max_bytes would be exceeded:
      "drop_oldest"  make room by dropping the oldest queued frames (live audio: the newest matters);
      "drop_newest"  refuse the new frame;
      "block"        refuse the new frame and tell the caller to stop reading that socket
                     (offer() returns False; in Node: ws.pause() until the next flush).
    Frames that arrive while a flush is running belong to the next tick.
    """

    def __init__(self, converter, tick_ms: float = 20.0, max_frames: int = 1 << 16, max_bytes: int = 1 << 30,
                 overflow: str = "drop_oldest", max_frame_samples: int = 1 << 20):
        if overflow not in OVERFLOW_POLICIES:
            raise ValueError(f"overflow must be one of {OVERFLOW_POLICIES}")
        self.converter, self.tick_ms = converter, float(tick_ms)
        self.max_frames, self.max_bytes, self.overflow = int(max_frames), int(max_bytes), overflow
        self.max_frame_samples = int(max_frame_samples)
        self._q: List[_Pending] = []
        self._bytes = 0
        self._busy = False
        self._next: List[_Pending] = []              # arrivals during a flush
        self.stats = {"frames": 0, "dropped_oldest": 0, "dropped_newest": 0, "blocked": 0, "protocol_errors": 0,
                      "relayed_pcm16": 0, "flushes": 0, "conversions": 0}

    # -- ingestion -----------------------------------------------------------------------------
    def offer(self, conn: Connection, kind: str, payload: Any, now_ms: float) -> bool:
        """Returns False only under the "block" policy when the frame was refused (back-pressure)."""
        if conn.closed or kind == IGNORED:
            return True
        if kind == CONTROL:
            conn.on_control(payload)
            return True
        if conn.encoding == "pcm16":                 # the reference's path: relay the bytes untouched (server.ts:263)
            conn.send_audio(payload)
            self.stats["relayed_pcm16"] += 1
            return True
        n_bytes = len(payload)
        if n_bytes % 4 != 0 or n_bytes // 4 > self.max_frame_samples:
            conn.protocol_errors += 1
            self.stats["protocol_errors"] += 1
            return True
        item = _Pending(conn, np.frombuffer(payload, dtype="<f4"), float(now_ms))
        q = self._next if self._busy else self._q
        if len(self._q) + len(self._next) + 1 > self.max_frames or self._bytes + n_bytes > self.max_bytes:
            if self.overflow == "drop_newest":
                self.stats["dropped_newest"] += 1
                return True
            if self.overflow == "block":
                self.stats["blocked"] += 1
                return False
            while self._q and (len(self._q) + len(self._next) + 1 > self.max_frames or self._bytes + n_bytes > self.max_bytes):
                old = self._q.pop(0)
                self._bytes -= old.samples.nbytes
                self.stats["dropped_oldest"] += 1
            if len(self._q) + len(self._next) + 1 > self.max_frames or self._bytes + n_bytes > self.max_bytes:
                self.stats["dropped_newest"] += 1    # nothing left to drop (everything is in the running flush)
                return True
        q.append(item)
        self._bytes += n_bytes
        self.stats["frames"] += 1
        return True

    @property
    def pending(self) -> int:
        return len(self._q) + len(self._next)

    def due(self, now_ms: float) -> bool:
        if not self._q:
            return False
        return now_ms - self._q[0].t_ms >= self.tick_ms or len(self._q) >= self.max_frames or self._bytes >= self.max_bytes

    # -- conversion ----------------------------------------------------------------------------
    def flush(self, now_ms: float = 0.0) -> int:
        """Convert and deliver everything queued.  Returns the number of frames delivered.  The views
        handed to send_audio stay valid until the next flush (sendAudio hands the bytes to the
        socket layer before returning)."""
        if self._busy or not self._q:
            return 0
        self._busy = True
        batch, self._q = self._q, []
        delivered = 0
        try:
            groups: Dict[Tuple[float, int], List[_Pending]] = {}
            for it in batch:                                     # arrival order is kept inside a group
                groups.setdefault((it.conn.sample_rate, it.samples.size), []).append(it)
            for (rate, n_in), items in groups.items():
                live = [it for it in items if not it.conn.closed]
                if not live:
                    continue
                block, views = self.converter.convert_group(rate, n_in, [it.samples for it in live])
                self.stats["conversions"] += 1
                for it, view in zip(live, views):
                    it.conn.send_audio(view)                     # raw little-endian int16, 2 * N_out bytes, no header
                    delivered += 1
        finally:
            self._bytes -= sum(it.samples.nbytes for it in batch)
            self._q, self._next = self._next, []
            self._busy = False
            self.stats["flushes"] += 1
        return delivered


class GpuGroupConverter:
    """convert_group() on the CUDA library: rows are gathered into a pinned input block (the one host
    copy on the way in), one pcmb200_frames_encode call converts the group, and every row's payload
    is a memoryview into the pinned output block (pcmb200_egress_table: no copy on the way out)."""

    def __init__(self, ctx, max_rows: int = 4096, max_n_in: int = 4096):
        self.ctx = ctx
        self._in = self._out = None
        self._cap_in = self._cap_out = 0
        self._reserve(max_rows * max_n_in, max_rows * max_n_in)

    def _reserve(self, n_in_total: int, n_out_total: int):
        if n_in_total > self._cap_in:
            if self._in is not None:
                self._in.free()
            self._in, self._cap_in = self.ctx.pinned(n_in_total, np.float32), n_in_total
        if n_out_total > self._cap_out:
            if self._out is not None:
                self._out.free()
            self._out, self._cap_out = self.ctx.pinned(n_out_total, np.int16), n_out_total

    def convert_group(self, rate: float, n_in: int, frames: List[np.ndarray]):
        from .api import resample_len
        from ._capi import MEM_HOST
        rows = len(frames)
        n_out = resample_len(n_in, rate)
        self._reserve(rows * max(n_in, 1), rows * max(n_out, 1))
        x = self._in.array[: rows * n_in].reshape(rows, n_in)
        for r, f in enumerate(frames):
            x[r] = f
        if n_out:
            self.ctx.frames_encode_ptr(self._in.ptr, max(n_in, 1), self._out.ptr, n_out, rows, n_in, rate, MEM_HOST)
        off = np.empty(rows, dtype=np.uint64)
        ln = np.empty(rows, dtype=np.uint32)
        check(load().pcmb200_egress_table(rows, n_out, max(n_out, 1), off.ctypes.data, ln.ctypes.data), self.ctx.handle)
        block = memoryview((C.c_char * max(rows * n_out * 2, 1)).from_address(self._out.ptr)).cast("B")
        return block, [block[int(o): int(o) + int(n)] for o, n in zip(off, ln)]

    def close(self):
        for b in (self._in, self._out):
            if b is not None:
                b.free()
        self._in = self._out = None


class GpuQuantumAggregator:
    """One device-side capture aggregator (pcmb200_aggregator_*) for the connections of ONE capture rate, with
    pinned staging for a tick's quanta and a pinned output block: push() is one H2D of the tick's quanta, the
    planned push on the GPU, one D2H of the flushed frames; every frame is handed out as a view (no copy)."""

    def __init__(self, ctx, rate: float, max_sessions: int, max_push_quanta: int, quantum: int = 128, desired_ms: float = 60.0,
                 max_frame_quanta: int = 64):
        from .api import Aggregator, resample_len
        self.ctx, self.rate, self.quantum, self.max_push = ctx, float(rate), int(quantum), int(max_push_quanta)
        n_out_max = resample_len(max_frame_quanta * quantum, rate)
        self.max_out = int(max_push_quanta) * ((n_out_max + 7) // 8 * 8)         # worst case: every quantum flushes a frame
        self.max_out = min(self.max_out, max(1 << 16, 2 * max_push_quanta * quantum))
        self.agg = Aggregator(ctx, max_sessions, rate, quantum=quantum, desired_ms=desired_ms, max_frame_quanta=max_frame_quanta,
                              max_push_quanta=max_push_quanta, max_out_samples=self.max_out)
        self._sess, self._time = ctx.pinned(max_push_quanta, np.uint32), ctx.pinned(max_push_quanta, np.float64)
        self._x = ctx.pinned((max_push_quanta, quantum), np.float32)
        self._out = ctx.pinned(self.max_out, np.int16)
        self._fs, self._fl = np.empty(max_push_quanta, np.uint32), np.empty(max_push_quanta, np.uint32)
        self._fo = np.empty(max_push_quanta, np.uint64)
        self._free = list(range(max_sessions - 1, -1, -1))
        self.n = 0
        self.last_call_s = 0.0

    def open(self, now_ms: float) -> int:
        """A slot for a new connection: empty buffer, lastSend = now (App.tsx:148-157).  -1 if all are taken."""
        if not self._free:
            return -1
        slot = self._free.pop()
        self.agg.reset_session(slot, now_ms)
        return slot

    def release(self, slot: int):
        if slot >= 0:
            self._free.append(slot)

    def add(self, slot: int, samples: np.ndarray, now_ms: float) -> bool:
        """Stage the quanta of one message (samples.size is a multiple of the quantum).  False if the tick is full."""
        k = samples.size // self.quantum
        if self.n + k > self.max_push:
            return False
        self._sess.array[self.n:self.n + k] = slot
        self._time.array[self.n:self.n + k] = now_ms
        self._x.array[self.n:self.n + k] = samples.reshape(k, self.quantum)
        self.n += k
        return True

    def push(self):
        """Run the staged quanta through the aggregator.  Returns [(slot, frame bytes)] in flush order; normally the
        frames are views into the pinned output block, valid until the next push.  The block is sized for real-time
        traffic, not for the worst case (every quantum of the tick completing a full-length frame); a tick that
        overflows it is rejected by the library with nothing changed, and is then pushed in halves, in order."""
        n, self.n = self.n, 0
        return self._push_range(0, n, copy=False) if n else []

    def _push_range(self, lo: int, hi: int, copy: bool):
        from ._capi import ERR_CAPACITY, MEM_HOST, PcmError
        import time
        n = hi - lo
        nf = C.c_uint32(0)
        t0 = time.perf_counter()
        try:
            check(load().pcmb200_aggregator_push(self.ctx.handle, self.agg.handle, n, self._sess.ptr + 4 * lo, self._time.ptr + 8 * lo,
                                                 self._x.ptr + 4 * self.quantum * lo, MEM_HOST, self._out.ptr, self.max_out, MEM_HOST,
                                                 self._fs.ctypes.data, self._fl.ctypes.data, self._fo.ctypes.data, self._fs.size,
                                                 C.byref(nf)), self.ctx.handle)
        except PcmError as e:
            if e.status != ERR_CAPACITY or n < 2:
                raise
            mid = lo + n // 2
            return self._push_range(lo, mid, copy=True) + self._push_range(mid, hi, copy=True)
        self.last_call_s = time.perf_counter() - t0                   # the library call alone (copies + kernels), for benchmarks
        if nf.value == 0:
            return []
        off, ln = np.empty(nf.value, np.uint64), np.empty(nf.value, np.uint32)
        check(load().pcmb200_egress_table_frames(nf.value, self._fo.ctypes.data, self._fl.ctypes.data, off.ctypes.data, ln.ctypes.data),
              self.ctx.handle)
        block = memoryview((C.c_char * (self.max_out * 2)).from_address(self._out.ptr)).cast("B")
        views = [(s, block[o: o + k]) for s, o, k in zip(self._fs[:nf.value].tolist(), off.tolist(), ln.tolist())]
        return [(s, bytes(v)) for s, v in views] if copy else views

    def close(self):
        self.agg.close()
        for b in (self._sess, self._time, self._x, self._out):
            b.free()


class QuantumBatcher:
    """The "framing": "quanta" path: connections forward raw AudioWorklet quanta, the server aggregates.

    offer()  one binary message (k * quantum Float32 samples) of a quanta-framed connection; the connection gets an
             aggregator slot on its first message (lastSend = that moment, as the browser sets it at start);
    flush()  one aggregator push per capture rate; every frame that flushed goes to its connection's send_audio as a
             view of its raw little-endian int16 bytes (frames of one connection in flush order).
    Bounded by max_push_quanta per rate and tick; a message that does not fit is refused (offer() returns False:
    back-pressure, as TickBatcher's "block" policy).  make_aggregator(rate) builds a GpuQuantumAggregator."""

    def __init__(self, make_aggregator: Callable[[float], "GpuQuantumAggregator"], tick_ms: float = 20.0, quantum: int = 128):
        self.make, self.tick_ms, self.quantum = make_aggregator, float(tick_ms), int(quantum)
        self._aggs: Dict[float, GpuQuantumAggregator] = {}
        self._conns: Dict[Tuple[float, int], Connection] = {}
        self._oldest: Optional[float] = None
        self.stats = {"quanta": 0, "blocked": 0, "protocol_errors": 0, "no_slot": 0, "frames": 0, "flushes": 0}

    def offer(self, conn: Connection, payload: Any, now_ms: float) -> bool:
        if conn.closed:
            return True
        n_bytes = len(payload)
        if n_bytes == 0 or n_bytes % (4 * self.quantum) != 0:
            conn.protocol_errors += 1
            self.stats["protocol_errors"] += 1
            return True
        agg = self._aggs.get(conn.sample_rate)
        if agg is None:
            agg = self._aggs[conn.sample_rate] = self.make(conn.sample_rate)
        if conn.slot < 0:
            conn.slot = agg.open(now_ms)
            if conn.slot < 0:
                self.stats["no_slot"] += 1
                return True
            self._conns[(conn.sample_rate, conn.slot)] = conn
        if not agg.add(conn.slot, np.frombuffer(payload, dtype="<f4"), float(now_ms)):
            self.stats["blocked"] += 1
            return False
        self.stats["quanta"] += n_bytes // (4 * self.quantum)
        if self._oldest is None:
            self._oldest = float(now_ms)
        return True

    def due(self, now_ms: float) -> bool:
        return self._oldest is not None and now_ms - self._oldest >= self.tick_ms

    def flush(self) -> int:
        delivered = 0
        for rate, agg in self._aggs.items():
            for slot, view in agg.push():
                conn = self._conns.get((rate, slot))
                if conn is not None and not conn.closed:
                    conn.send_audio(view)
                    delivered += 1
        self._oldest = None
        self.stats["frames"] += delivered
        self.stats["flushes"] += 1
        return delivered

    def disconnect(self, conn: Connection):
        """The socket closed: its slot can serve another connection (whatever it had buffered is dropped, as the
        browser drops agg.buf when the page goes away)."""
        if conn.slot >= 0:
            self._conns.pop((conn.sample_rate, conn.slot), None)
            self._aggs[conn.sample_rate].release(conn.slot)
            conn.slot = -1

    def close(self):
        for agg in self._aggs.values():
            agg.close()
        self._aggs.clear()


class KeepAliveScheduler:
    """The reference's keep-alive timers (App.tsx:291-296, cleared at :309-312) for a fleet, on an
    injected clock (pcmb200_keepalive_due): poll(now) returns the sessions that must be sent the
    1600-sample silence frame now."""

    def __init__(self, max_sessions: int, period_ms: float = 2000.0):
        self.n, self.period = int(max_sessions), float(period_ms)
        self.start = np.full(self.n, np.nan)
        self.done = np.zeros(self.n, dtype=np.uint64)
        self.open = np.zeros(self.n, dtype=np.uint8)
        self._due = np.empty(self.n, dtype=np.uint32)

    def start_session(self, s: int, now_ms: float, socket_open: bool = True):   # start(): setInterval armed
        self.start[s], self.done[s], self.open[s] = now_ms, 0, int(socket_open)

    def stop_session(self, s: int):                                              # stop(): clearInterval
        self.start[s], self.open[s] = np.nan, 0

    def set_open(self, s: int, socket_open: bool):
        self.open[s] = int(socket_open)

    def poll(self, now_ms: float) -> np.ndarray:
        n = C.c_uint32(0)
        check(load().pcmb200_keepalive_due(float(now_ms), self.n, self.start.ctypes.data, self.done.ctypes.data,
                                           self.open.ctypes.data, self.period, self._due.ctypes.data, self.n, C.byref(n)))
        return self._due[: n.value].copy()
