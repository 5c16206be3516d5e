"""ctypes binding of the CPU oracle (oracle/pcm_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference``
legs may import this package; the product path never does.  Parity: pinned to the reference's
source text run by tests/golden/mini_js.py (tests/test_ref_exec.py), not yet by a production
JavaScript engine -- see the header of pcm_oracle.c.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpcm_oracle.so")
_lib = None


def build() -> str:
    """Compile libpcm_oracle.so with gcc (oracle/Makefile) if it is missing or stale."""
    src = os.path.join(_HERE, "pcm_oracle.c")
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _SO


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        f32p, i16p, u32p, f64p, u8p = (C.POINTER(t) for t in (C.c_float, C.c_int16, C.c_uint32, C.c_double, C.c_uint8))
        L.oracle_to_int16.restype = C.c_int16
        L.oracle_to_int16.argtypes = [C.c_double]
        L.oracle_scale_sample.restype = C.c_double
        L.oracle_scale_sample.argtypes = [C.c_float]
        L.oracle_float_to_16bit_pcm.restype = None
        L.oracle_float_to_16bit_pcm.argtypes = [f32p, C.c_size_t, i16p]
        L.oracle_resample_len.restype = C.c_size_t
        L.oracle_resample_len.argtypes = [C.c_size_t, C.c_double, C.POINTER(C.c_int)]
        L.oracle_index_table.restype = None
        L.oracle_index_table.argtypes = [C.c_size_t, C.c_double, C.c_size_t, u32p]
        L.oracle_resample_to_16k.restype = C.c_size_t
        L.oracle_resample_to_16k.argtypes = [f32p, C.c_size_t, C.c_double, f32p]
        L.oracle_frame.restype = C.c_size_t
        L.oracle_frame.argtypes = [f32p, C.c_size_t, C.c_double, i16p]
        L.oracle_frames.restype = C.c_size_t
        L.oracle_frames.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double, i16p, C.c_size_t]
        L.oracle_frames_mt.restype = C.c_size_t
        L.oracle_frames_mt.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double, i16p, C.c_size_t, C.c_int]
        L.oracle_level_meter.restype = C.c_double
        L.oracle_level_meter.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_double, f64p]
        L.oracle_aggregator_plan.restype = C.c_size_t
        L.oracle_aggregator_plan.argtypes = [f64p, C.c_size_t, C.c_double, f64p, u8p]
        L.oracle_keepalive_frame.restype = C.c_size_t
        L.oracle_keepalive_frame.argtypes = [i16p]
        L.oracle_keepalive_timer.restype = C.c_size_t
        L.oracle_keepalive_timer.argtypes = [C.c_double, C.c_double, f64p, C.c_size_t, u8p, u8p]
        L.oracle_synth_fill.restype = None
        L.oracle_synth_fill.argtypes = [f32p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double, C.c_uint64, C.c_uint64, C.c_int]
        _lib = L
    return _lib


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(C.POINTER(t))


def float_to_16bit_pcm(x: np.ndarray) -> np.ndarray:
    """App.tsx:7-16."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.size, dtype=np.int16)
    lib().oracle_float_to_16bit_pcm(_p(x, C.c_float), x.size, _p(out, C.c_int16))
    return out


def resample_len(n: int, rate: float) -> int:
    """App.tsx:22-24.  Raises ValueError where JS throws RangeError."""
    st = C.c_int(0)
    r = lib().oracle_resample_len(n, float(rate), C.byref(st))
    if st.value:
        raise ValueError("RangeError: Invalid typed array length")
    return int(r)


def index_table(n_in: int, rate: float, n_out: int | None = None) -> np.ndarray:
    """App.tsx:26-30 pick indices (uint32, 0xFFFFFFFF = outside the input)."""
    if n_out is None:
        n_out = resample_len(n_in, rate)
    idx = np.empty(n_out, dtype=np.uint32)
    lib().oracle_index_table(n_in, float(rate), n_out, _p(idx, C.c_uint32))
    return idx


def resample_to_16k(x: np.ndarray, rate: float) -> np.ndarray:
    """App.tsx:18-32."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    n_out = resample_len(x.size, rate)
    out = np.empty(n_out, dtype=np.float32)
    lib().oracle_resample_to_16k(_p(x, C.c_float), x.size, float(rate), _p(out, C.c_float))
    return out


def frames(x: np.ndarray, n_in: int, rate: float, threads: int = 1, pitch_out: int | None = None) -> np.ndarray:
    """App.tsx:281-282 per row of a [rows, pitch_in] float32 matrix -> [rows, n_out] int16."""
    assert x.ndim == 2 and x.dtype == np.float32 and x.strides[1] == 4
    rows, pitch_in = x.shape[0], x.strides[0] // 4
    n_out = resample_len(n_in, rate)
    po = n_out if pitch_out is None else pitch_out
    out = np.zeros((rows, po), dtype=np.int16)
    if rows and n_out:
        if threads <= 1:
            lib().oracle_frames(_p(x, C.c_float), pitch_in, rows, n_in, float(rate), _p(out, C.c_int16), po)
        else:
            lib().oracle_frames_mt(_p(x, C.c_float), pitch_in, rows, n_in, float(rate), _p(out, C.c_int16), po, threads)
    return out[:, :n_out] if pitch_out is None else out


def level_meter(x: np.ndarray, quantum: int = 128, prev: float = 0.0) -> np.ndarray:
    """App.tsx:267-271 over consecutive quanta of one session."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    nq = x.size // quantum
    lv = np.empty(nq, dtype=np.float64)
    lib().oracle_level_meter(_p(x, C.c_float), nq, quantum, float(prev), _p(lv, C.c_double))
    return lv


def aggregator_plan(arrival_ms: np.ndarray, desired_ms: float = 60.0, last_send: float = 0.0):
    """App.tsx:273-288 with an injected clock -> (flush_after[uint8], new last_send)."""
    a = np.ascontiguousarray(arrival_ms, dtype=np.float64)
    fl = np.zeros(a.size, dtype=np.uint8)
    ls = C.c_double(last_send)
    lib().oracle_aggregator_plan(_p(a, C.c_double), a.size, float(desired_ms), C.byref(ls), _p(fl, C.c_uint8))
    return fl, ls.value


def keepalive_frame() -> np.ndarray:
    out = np.empty(1600, dtype=np.int16)
    lib().oracle_keepalive_frame(_p(out, C.c_int16))
    return out


def synth_fill(rows: int, n: int, rate: float, seed: int = 0x5EED, row0: int = 0, threads: int = 1) -> np.ndarray:
    """The benchmark's synthetic audio generated on the host: same counter hash as
    the device generator (identical noise bits), host libm sine (last-bit differences possible)."""
    out = np.empty((rows, n), dtype=np.float32)
    if rows and n:
        lib().oracle_synth_fill(_p(out, C.c_float), n, rows, n, float(rate), seed, row0, threads)
    return out


def keepalive_timer(start_ms: float, polls: np.ndarray, socket_open: np.ndarray, period_ms: float = 2000.0) -> np.ndarray:
    """App.tsx:291-296 for one session: which of the polls send a keep-alive frame (uint8 per poll)."""
    polls = np.ascontiguousarray(polls, dtype=np.float64)
    op = np.ascontiguousarray(socket_open, dtype=np.uint8)
    sent = np.zeros(polls.size, dtype=np.uint8)
    lib().oracle_keepalive_timer(float(start_ms), float(period_ms), _p(polls, C.c_double), polls.size, _p(op, C.c_uint8), _p(sent, C.c_uint8))
    return sent
