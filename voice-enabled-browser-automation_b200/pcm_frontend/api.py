"""Host-side objects over the C ABI: Context, device / pinned buffers and the mirrored functions."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _capi
from ._capi import MEM_DEVICE, MEM_HOST, check, load


def resample_len(n_in: int, rate: float) -> int:
    """Output length of resampleTo16k (App.tsx:22-24).  PcmRangeError where JS throws RangeError."""
    out = C.c_size_t(0)
    check(load().pcmb200_resample_len(int(n_in), float(rate), C.byref(out)))
    return int(out.value)


def index_table(n_in: int, rate: float, n_out: Optional[int] = None) -> np.ndarray:
    """Pick indices of resampleTo16k's running sum (App.tsx:26-30); 0xFFFFFFFF = outside input."""
    if n_out is None:
        n_out = resample_len(n_in, rate)
    idx = np.empty(n_out, dtype=np.uint32)
    check(load().pcmb200_index_table(int(n_in), float(rate), int(n_out), idx.ctypes.data))
    return idx


def keepalive_frame() -> np.ndarray:
    """The reference's keep-alive payload (App.tsx:291-296): 1600 zero samples, sent every 2 s."""
    out = np.empty(1600, dtype=np.int16)
    n = C.c_size_t(0)
    check(load().pcmb200_keepalive_frame(out.ctypes.data, out.size, C.byref(n)))
    return out[: n.value]


def shard_sessions(n_sessions: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Session-affine sharding: contiguous blocks, the first n % world ranks get one
    extra session.  Returns (first_session, count).  Sessions never migrate and never exchange
    data, so there is no collective anywhere on the data path."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    base, extra = divmod(int(n_sessions), world_size)
    count = base + (1 if rank < extra else 0)
    first = rank * base + min(rank, extra)
    return first, count


class DeviceBuffer:
    """A device allocation owned through the C ABI (pcmb200_device_alloc)."""

    def __init__(self, ctx: "Context", nbytes: int):
        self.ctx, self.nbytes = ctx, int(nbytes)
        p = C.c_void_p()
        check(load().pcmb200_device_alloc(ctx.handle, self.nbytes, C.byref(p)), ctx.handle)
        self.ptr = p.value

    def offset(self, nbytes: int) -> int:
        return int(self.ptr + int(nbytes))

    def upload(self, host: np.ndarray, byte_offset: int = 0):
        host = np.ascontiguousarray(host)
        byte_offset = int(byte_offset)
        assert byte_offset + host.nbytes <= self.nbytes
        check(load().pcmb200_memcpy_h2d(self.ctx.handle, self.ptr + byte_offset, host.ctypes.data, host.nbytes), self.ctx.handle)

    def download(self, dtype, count: int, byte_offset: int = 0) -> np.ndarray:
        out = np.empty(int(count), dtype=dtype)
        byte_offset = int(byte_offset)
        assert byte_offset + out.nbytes <= self.nbytes
        check(load().pcmb200_memcpy_d2h(self.ctx.handle, out.ctypes.data, self.ptr + byte_offset, out.nbytes), self.ctx.handle)
        return out

    def memset(self, value: int = 0):
        check(load().pcmb200_memset_device(self.ctx.handle, self.ptr, value, self.nbytes), self.ctx.handle)

    def free(self):
        if self.ptr:
            check(load().pcmb200_device_free(self.ctx.handle, self.ptr), self.ctx.handle)
            self.ptr = None

    def __del__(self):
        try:
            if getattr(self, "ptr", None) and self.ctx.handle:
                self.free()
        except Exception:
            pass


class PinnedBuffer:
    """Page-locked host memory (pcmb200_pinned_alloc) exposed as a numpy array."""

    def __init__(self, ctx: "Context", shape, dtype, write_combined: bool = False):
        self.ctx = ctx
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) if np.ndim(shape) else int(shape)
        self.nbytes = n * dtype.itemsize
        p = C.c_void_p()
        alloc = load().pcmb200_pinned_alloc_write_combined if write_combined else load().pcmb200_pinned_alloc
        check(alloc(ctx.handle, self.nbytes, C.byref(p)), ctx.handle)
        self.ptr = p.value
        raw = (C.c_char * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(raw, dtype=dtype, count=n).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            check(load().pcmb200_pinned_free(self.ctx.handle, self.ptr), self.ctx.handle)
            self.ptr = None

    def __del__(self):
        try:
            if getattr(self, "ptr", None) and self.ctx.handle:
                self.free()
        except Exception:
            pass


class Context:
    """One CUDA device, one stream, one staging pool.  Not thread-safe (one per device per thread)."""

    def __init__(self, device: int = 0):
        self.handle = None
        L = load()
        h = C.c_void_p()
        check(L.pcmb200_ctx_create(int(device), C.byref(h)))
        self.handle = h.value
        self.device = int(device)
        self._L = L

    # -- lifecycle -------------------------------------------------------------------------
    def close(self):
        if self.handle:
            self._L.pcmb200_ctx_destroy(self.handle)
            self.handle = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(self._L.pcmb200_sync(self.handle), self.handle)

    @property
    def launches(self) -> int:
        n = C.c_uint64(0)
        check(self._L.pcmb200_launch_count(self.handle, C.byref(n)), self.handle)
        return int(n.value)

    def reserve(self, max_rows: int, max_n_in: int, max_n_out: Optional[int] = None, max_rate_classes: int = 4):
        """Capacity up front (pcmb200_ctx_reserve): after this, host-resident calls within the limits
        neither allocate nor free device memory."""
        if max_n_out is None:
            max_n_out = max_n_in
        check(self._L.pcmb200_ctx_reserve(self.handle, int(max_rows), int(max_n_in), int(max_n_out), int(max_rate_classes)),
              self.handle)

    @property
    def alloc_stats(self) -> dict:
        m, f, s = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        check(self._L.pcmb200_alloc_stats(self.handle, C.byref(m), C.byref(f), C.byref(s)), self.handle)
        return {"device_mallocs": int(m.value), "device_frees": int(f.value), "stream_syncs": int(s.value)}

    def device_info(self) -> dict:
        name = C.create_string_buffer(256)
        sm, mem = C.c_int(0), C.c_size_t(0)
        check(self._L.pcmb200_device_info(self.handle, name, 256, C.byref(sm), C.byref(mem)), self.handle)
        return {"name": name.value.decode(), "sm_count": sm.value, "total_mem": mem.value}

    def tune(self, **kv):
        for k, v in kv.items():
            check(self._L.pcmb200_tune(self.handle, k.encode(), int(v)), self.handle)

    # -- memory ----------------------------------------------------------------------------
    def device_buffer(self, nbytes: int) -> DeviceBuffer:
        return DeviceBuffer(self, nbytes)

    def pinned(self, shape, dtype, write_combined: bool = False) -> PinnedBuffer:
        return PinnedBuffer(self, shape, dtype, write_combined)

    def memcpy_h2d(self, dst_device: int, src_host: int, nbytes: int):
        check(self._L.pcmb200_memcpy_h2d(self.handle, dst_device, src_host, int(nbytes)), self.handle)

    def memcpy_d2h(self, dst_host: int, src_device: int, nbytes: int):
        check(self._L.pcmb200_memcpy_d2h(self.handle, dst_host, src_device, int(nbytes)), self.handle)

    # -- timing ----------------------------------------------------------------------------
    def timer_start(self):
        check(self._L.pcmb200_timer_start(self.handle), self.handle)

    def timer_stop(self) -> float:
        ms = C.c_float(0)
        check(self._L.pcmb200_timer_stop(self.handle, C.byref(ms)), self.handle)
        return float(ms.value)

    # -- the reference's functions, host arrays in / host arrays out ------------------------------
    def floatTo16BitPCM(self, float32: np.ndarray) -> np.ndarray:
        """App.tsx:7-16.  Returns a fresh int16 array of the same length; input is not modified."""
        x = _as_f32_1d(float32)
        out = np.empty(x.size, dtype=np.int16)
        check(self._L.pcmb200_encode_f32_s16(self.handle, x.ctypes.data, out.ctypes.data, x.size, MEM_HOST), self.handle)
        return out

    def resampleTo16k(self, input: np.ndarray, inputSampleRate: float) -> np.ndarray:  # noqa: A002
        """App.tsx:18-32.  At 16 kHz the reference returns the SAME array object; so does this."""
        if inputSampleRate == 16000:
            return input
        x = _as_f32_1d(input)
        n_out = resample_len(x.size, inputSampleRate)
        out = np.empty(n_out, dtype=np.float32)
        got = C.c_size_t(0)
        check(self._L.pcmb200_resample_to_16k(self.handle, x.ctypes.data, x.size, float(inputSampleRate),
                                              out.ctypes.data, out.size, C.byref(got), MEM_HOST), self.handle)
        assert got.value == n_out
        return out

    def frames_encode(self, batch: np.ndarray, rate: float, n_in: Optional[int] = None,
                      out: Optional[np.ndarray] = None) -> np.ndarray:
        """floatTo16BitPCM(resampleTo16k(row, rate)) for every row of a [rows, pitch] float32 batch
        (App.tsx:281-282), fused on the device.  Returns [rows, n_out] int16."""
        if batch.ndim != 2 or batch.dtype != np.float32:
            raise TypeError("expected a 2-D float32 batch [rows, samples]")
        assert batch.size == 0 or batch.strides[1] == 4, "samples of a row must be contiguous"
        rows = batch.shape[0]
        pitch_in = batch.strides[0] // 4 if rows > 1 and batch.size else max(batch.shape[1], 1)
        n_in = batch.shape[1] if n_in is None else int(n_in)
        n_out = resample_len(n_in, rate)
        if out is None:
            out = np.empty((rows, n_out), dtype=np.int16)
        assert out.dtype == np.int16 and out.shape[0] == rows and out.shape[1] >= n_out
        pitch_out = out.strides[0] // 2 if rows > 1 and out.size else max(out.shape[1], 1)
        got = C.c_size_t(0)
        check(self._L.pcmb200_frames_encode(self.handle, batch.ctypes.data, pitch_in, out.ctypes.data, pitch_out,
                                            rows, n_in, float(rate), MEM_HOST, C.byref(got)), self.handle)
        return out[:, :n_out]

    # -- raw pointer variants (device- or host-resident, caller-owned) -----------------------------
    def frames_encode_ptr(self, in_ptr: int, pitch_in: int, out_ptr: int, pitch_out: int, rows: int,
                          n_in: int, rate: float, mem: int = MEM_DEVICE) -> int:
        got = C.c_size_t(0)
        check(self._L.pcmb200_frames_encode(self.handle, int(in_ptr), int(pitch_in), int(out_ptr), int(pitch_out),
                                            int(rows), int(n_in), float(rate), int(mem), C.byref(got)), self.handle)
        return int(got.value)

    def encode_ptr(self, in_ptr: int, out_ptr: int, n: int, mem: int = MEM_DEVICE):
        check(self._L.pcmb200_encode_f32_s16(self.handle, in_ptr, out_ptr, n, mem), self.handle)

    def resample_ptr(self, in_ptr: int, n_in: int, rate: float, out_ptr: int, out_cap: int, mem: int = MEM_DEVICE) -> int:
        got = C.c_size_t(0)
        check(self._L.pcmb200_resample_to_16k(self.handle, in_ptr, n_in, float(rate), out_ptr, out_cap,
                                              C.byref(got), mem), self.handle)
        return int(got.value)

    def frames_encode_meter_ptr(self, in_ptr: int, pitch_in: int, out_ptr: int, pitch_out: int, rows: int, n_in: int,
                                rate: float, state_ptr: int, levels_ptr: Optional[int] = None, quantum: int = 128) -> int:
        """Fused frames + level meter (App.tsx:267-271, :281-282) on device buffers."""
        got = C.c_size_t(0)
        check(self._L.pcmb200_frames_encode_meter(self.handle, int(in_ptr), int(pitch_in), int(out_ptr), int(pitch_out),
                                                  int(rows), int(n_in), float(rate), int(quantum), int(state_ptr),
                                                  int(levels_ptr) if levels_ptr else None, C.byref(got)), self.handle)
        return int(got.value)

    def frames_encode_meter(self, batch: np.ndarray, rate: float, level_state: np.ndarray, quantum: int = 128):
        """Host-array convenience over frames_encode_meter_ptr: returns (pcm16 [rows, n_out],
        levels [rows, n_in // quantum], new level_state [rows]).  Stages through device buffers."""
        if batch.ndim != 2 or batch.dtype != np.float32:
            raise TypeError("expected a 2-D float32 batch [rows, samples]")
        batch = np.ascontiguousarray(batch)
        rows, n_in = batch.shape
        n_out, nq = resample_len(n_in, rate), n_in // quantum
        state = np.ascontiguousarray(level_state, dtype=np.float64)
        assert state.size == rows
        d_in, d_out = self.device_buffer(max(batch.nbytes, 16)), self.device_buffer(max(rows * n_out * 2, 16))
        d_state, d_lv = self.device_buffer(max(rows * 8, 16)), self.device_buffer(max(rows * nq * 8, 16))
        try:
            if batch.size:
                d_in.upload(batch)
            if rows:
                d_state.upload(state)
            self.frames_encode_meter_ptr(d_in.ptr, max(n_in, 1), d_out.ptr, max(n_out, 1), rows, n_in, rate, d_state.ptr,
                                         d_lv.ptr, quantum)
            self.sync()
            pcm = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
            levels = d_lv.download(np.float64, rows * nq).reshape(rows, nq)
            new_state = d_state.download(np.float64, rows)
        finally:
            for d in (d_in, d_out, d_state, d_lv):
                d.free()
        return pcm, levels, new_state

    def tick_graph(self, in_ptr: int, pitch_in: int, out_ptr: int, pitch_out: int, rows: int, n_in: int,
                   rate: float) -> "TickGraph":
        """Record one device-resident tick as a CUDA graph (pcmb200_tick_graph_create)."""
        return TickGraph(self, in_ptr, pitch_in, out_ptr, pitch_out, rows, n_in, rate)

    def synth_fill(self, out_ptr: int, pitch: int, rows: int, n: int, rate: float, seed: int = 0x5EED, row0: int = 0):
        check(self._L.pcmb200_synth_fill(self.handle, out_ptr, pitch, rows, n, float(rate), seed, row0), self.handle)


class TickGraph:
    """A recorded tick; launch() replays it on the context's stream."""

    def __init__(self, ctx: Context, in_ptr, pitch_in, out_ptr, pitch_out, rows, n_in, rate):
        self.ctx = ctx
        h = C.c_void_p()
        check(ctx._L.pcmb200_tick_graph_create(ctx.handle, int(in_ptr), int(pitch_in), int(out_ptr), int(pitch_out),
                                               int(rows), int(n_in), float(rate), C.byref(h)), ctx.handle)
        self.handle = h.value

    def launch(self):
        check(self.ctx._L.pcmb200_tick_graph_launch(self.ctx.handle, self.handle), self.ctx.handle)

    def destroy(self):
        if self.handle and self.ctx.handle:
            check(self.ctx._L.pcmb200_tick_graph_destroy(self.ctx.handle, self.handle), self.ctx.handle)
        self.handle = None


class Aggregator:
    """The reference's capture aggregator (App.tsx:148-157, :273-288) for many sessions on the device:
    push() takes 128-sample quanta with their arrival times and returns the frames that flushed."""

    def __init__(self, ctx: Context, max_sessions: int, rate: float, quantum: int = 128, desired_ms: float = 60.0,
                 max_frame_quanta: int = 64, max_push_quanta: int = 1 << 16, max_out_samples: int = 1 << 24):
        self.ctx, self.quantum, self.max_push, self.max_out = ctx, quantum, max_push_quanta, max_out_samples
        h = C.c_void_p()
        check(ctx._L.pcmb200_aggregator_create(ctx.handle, int(max_sessions), float(rate), int(quantum), float(desired_ms),
                                               int(max_frame_quanta), int(max_push_quanta), int(max_out_samples),
                                               C.byref(h)), ctx.handle)
        self.handle = h.value
        self._out = np.empty(max_out_samples, dtype=np.int16)
        self._fs = np.empty(max_push_quanta, dtype=np.uint32)
        self._fl = np.empty(max_push_quanta, dtype=np.uint32)
        self._fo = np.empty(max_push_quanta, dtype=np.uint64)

    def set_meter(self, enabled: bool = True):
        check(self.ctx._L.pcmb200_aggregator_set_meter(self.ctx.handle, self.handle, int(enabled)), self.ctx.handle)

    def levels(self, first_session: int, count: int) -> np.ndarray:
        """Current meter level (App.tsx:267-271) of `count` sessions, as of their last flushed frame."""
        out = np.empty(count, dtype=np.float64)
        check(self.ctx._L.pcmb200_aggregator_get_levels(self.ctx.handle, self.handle, int(first_session), int(count),
                                                        out.ctypes.data), self.ctx.handle)
        return out

    def reset_session(self, session: int, now_ms: float):
        check(self.ctx._L.pcmb200_aggregator_reset_session(self.ctx.handle, self.handle, int(session), float(now_ms)),
              self.ctx.handle)

    def reset_range(self, first_session: int, count: int, now_ms: float):
        """Initial state (App.tsx:148-157) for `count` sessions at once."""
        check(self.ctx._L.pcmb200_aggregator_reset_range(self.ctx.handle, self.handle, int(first_session), int(count),
                                                         float(now_ms)), self.ctx.handle)

    def push_device(self, n: int, sessions_ptr: int, arrival_ptr: int, samples_ptr: int, out_ptr: int, out_capacity: int,
                    frame_session_ptr: Optional[int] = None, frame_len_ptr: Optional[int] = None,
                    frame_offset_ptr: Optional[int] = None, max_frames: Optional[int] = None):
        """Everything device-resident, asynchronous (pcmb200_aggregator_push_device); collect with wait()."""
        check(self.ctx._L.pcmb200_aggregator_push_device(self.ctx.handle, self.handle, int(n), int(sessions_ptr), int(arrival_ptr),
                                                         int(samples_ptr), int(out_ptr), int(out_capacity),
                                                         frame_session_ptr, frame_len_ptr, frame_offset_ptr,
                                                         int(n if max_frames is None else max_frames)), self.ctx.handle)

    def wait(self) -> Tuple[int, int]:
        """(n_frames, output samples incl. per-frame padding) of the pending push; raises if it was rejected."""
        nf, ns = C.c_uint32(0), C.c_size_t(0)
        check(self.ctx._L.pcmb200_aggregator_wait(self.ctx.handle, self.handle, C.byref(nf), C.byref(ns)), self.ctx.handle)
        return int(nf.value), int(ns.value)

    def push(self, sessions: np.ndarray, arrival_ms: np.ndarray, samples: np.ndarray):
        """sessions[n] uint32, arrival_ms[n] float64, samples[n, quantum] float32 (arrival order).
        Returns [(session, int16 frame), ...] in flush order."""
        sessions = np.ascontiguousarray(sessions, dtype=np.uint32)
        arrival_ms = np.ascontiguousarray(arrival_ms, dtype=np.float64)
        samples = np.ascontiguousarray(samples, dtype=np.float32)
        n = sessions.size
        assert arrival_ms.size == n and samples.size == n * self.quantum
        nf = C.c_uint32(0)
        check(self.ctx._L.pcmb200_aggregator_push(self.ctx.handle, self.handle, n, sessions.ctypes.data, arrival_ms.ctypes.data,
                                                  samples.ctypes.data, MEM_HOST, self._out.ctypes.data, self._out.size, MEM_HOST,
                                                  self._fs.ctypes.data, self._fl.ctypes.data, self._fo.ctypes.data,
                                                  self._fs.size, C.byref(nf)), self.ctx.handle)
        return [(int(self._fs[f]), self._out[int(self._fo[f]): int(self._fo[f]) + int(self._fl[f])].copy())
                for f in range(nf.value)]

    def close(self):
        if self.handle and self.ctx.handle:
            check(self.ctx._L.pcmb200_aggregator_destroy(self.ctx.handle, self.handle), self.ctx.handle)
        self.handle = None


def _as_f32_1d(a) -> np.ndarray:
    x = np.asarray(a)
    if x.dtype != np.float32:
        raise TypeError("expected a float32 array (the reference takes a Float32Array)")
    return np.ascontiguousarray(x).reshape(-1)


_default: Optional[Context] = None


def default_context() -> Context:
    global _default
    if _default is None or _default.handle is None:
        _default = Context(0)
    return _default


def floatTo16BitPCM(float32: np.ndarray) -> np.ndarray:
    """Module-level mirror of App.tsx:7-16 on the default context (cuda:0)."""
    return default_context().floatTo16BitPCM(float32)


def resampleTo16k(input: np.ndarray, inputSampleRate: float) -> np.ndarray:  # noqa: A002
    """Module-level mirror of App.tsx:18-32 on the default context (cuda:0)."""
    if inputSampleRate == 16000:
        return input
    return default_context().resampleTo16k(input, inputSampleRate)


def frames_encode(batch: np.ndarray, rate: float) -> np.ndarray:
    return default_context().frames_encode(batch, rate)
