"""GPU test of pcmb200_frames_encode_noalloc, the entry point for callers that must not allocate
device memory while running (a real-time tick thread that cannot afford cudaMalloc's device-wide
stall; a process whose GPU memory already belongs to another framework's allocator).

It is called through the C ABI on device buffers and must produce exactly the bytes of the existing
pcmb200_frames_encode and of the CPU oracle (tolerance 0: integer output), on two calls in a row --
the second call reuses
the library's built-in table that the first one filled, then a different frame shape overwrites it.
The file sorts last so that a run stopped at the first failure has already run every other test; it
also passes alone in a fresh process (it sets up everything it needs itself)."""
import ctypes as C

import numpy as np
import pytest

import oracle
import pcm_frontend
from pcm_frontend import MEM_DEVICE, _capi

pytestmark = pytest.mark.gpu

MAX_OUT = 16384


def _noalloc(in_ptr, pitch_in, out_ptr, pitch_out, rows, n_in, rate):
    got = C.c_size_t(0)
    st = pcm_frontend.load().pcmb200_frames_encode_noalloc(int(in_ptr), pitch_in, int(out_ptr), pitch_out, rows, n_in,
                                                         float(rate), None, 1, C.byref(got))
    _capi.check(st, None)
    return int(got.value)


def _both(ctx, rng, rows, n_in, rate, pitch_pad=0):
    n_out = pcm_frontend.resample_len(n_in, rate)
    pitch_in, pitch_out = n_in + pitch_pad, n_out + pitch_pad
    x = rng.uniform(-1.3, 1.3, size=(rows, pitch_in)).astype(np.float32)
    x[0, :4] = [np.nan, np.inf, -np.inf, -0.0]
    d_in = ctx.device_buffer(x.nbytes)
    d_ref = ctx.device_buffer(rows * pitch_out * 2)
    d_new = ctx.device_buffer(rows * pitch_out * 2)
    d_in.upload(x)
    d_ref.memset(0x11)
    d_new.memset(0x11)
    ctx.frames_encode_ptr(d_in.ptr, pitch_in, d_ref.ptr, pitch_out, rows, n_in, rate, MEM_DEVICE)
    ctx.sync()
    assert _noalloc(d_in.ptr, pitch_in, d_new.ptr, pitch_out, rows, n_in, rate) == n_out
    ref = d_ref.download(np.int16, rows * pitch_out)
    new = d_new.download(np.int16, rows * pitch_out)
    for d in (d_in, d_ref, d_new):
        d.free()
    # and against the CPU oracle, not only against the other CUDA entry point: the frames must be the
    # oracle's bytes and the padding between rows must still hold the 0x11 fill
    want = np.full((rows, pitch_out), 0x1111, dtype=np.int16)
    want[:, :n_out] = oracle.frames(x, n_in, rate)
    assert np.array_equal(new.reshape(rows, pitch_out), want), (rows, n_in, rate, pitch_pad)
    return ref, new


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    rng = np.random.RandomState(2026)
    with pcm_frontend.Context(0) as ctx:
        # two calls in a row with the same frame shape (the second reuses the stored table) ...
        for _ in range(2):
            ref, new = _both(ctx, rng, 777, 882, 44100)
            assert np.array_equal(ref, new)
        # ... then other shapes overwrite the shared table: integer ratio, padded rows, upsampling
        for rows, n_in, rate, pad in ((513, 960, 48000, 0), (64, 2944, 48000, 3), (300, 160, 8000, 1),
                                      (5, 441, 22050, 0), (2, 320, 16000, 0), (1, 44100, 44100, 0)):
            ref, new = _both(ctx, rng, rows, n_in, rate, pad)
            assert np.array_equal(ref, new), (rows, n_in, rate, pad)
        ref, new = _both(ctx, rng, 777, 882, 44100)      # back to the first shape
        assert np.array_equal(ref, new)


def test_noalloc_rejects_frames_beyond_its_table():
    with pcm_frontend.Context(0) as ctx:
        n_in = (MAX_OUT + 1) * 3
        d_in = ctx.device_buffer(n_in * 4)
        d_out = ctx.device_buffer((MAX_OUT + 1) * 2)
        with pytest.raises(pcm_frontend.PcmError) as ei:
            _noalloc(d_in.ptr, n_in, d_out.ptr, MAX_OUT + 1, 1, n_in, 48000)
        assert ei.value.status == _capi.ERR_CAPACITY and str(MAX_OUT) in str(ei.value)
        # exactly at the limit is accepted
        assert _noalloc(d_in.ptr, MAX_OUT * 3, d_out.ptr, MAX_OUT, 1, MAX_OUT * 3, 48000) == MAX_OUT
        with pytest.raises(pcm_frontend.PcmRangeError):
            _noalloc(d_in.ptr, 16, d_out.ptr, 16, 1, 16, 0.0)
        d_in.free(); d_out.free()
