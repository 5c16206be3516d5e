"""Size-independent properties at BASELINE's full sizes (the oracle cannot sweep 10^9 samples in
seconds): the composition law of the reference, monotonicity of the encoder, determinism."""
import numpy as np
import pytest

from pcm_frontend import MEM_DEVICE

pytestmark = pytest.mark.gpu


def test_full_size_composition_law(ctx):
    """App.tsx:281-282 says frame = floatTo16BitPCM(resampleTo16k(buf)).  At config 4's size
    (2^20 x 960) the FUSED kernel must equal the two stand-alone functions applied one after the
    other -- three different kernels (1b vs the f32 pick kernel + the encode-only kernel)."""
    rows, n_in, n_out = 1 << 20, 960, 320
    d_in = ctx.device_buffer(rows * n_in * 4)
    ctx.synth_fill(d_in.ptr, n_in, rows, n_in, 48000, 0x5EED, 0)
    d_fused = ctx.device_buffer(rows * n_out * 2)
    ctx.frames_encode_ptr(d_in.ptr, n_in, d_fused.ptr, n_out, rows, n_in, 48000, MEM_DEVICE)
    d_mid = ctx.device_buffer(rows * n_out * 4)
    # the batch is one flat stream whose picks are every third float, so it can be resampled as one buffer
    got = ctx.resample_ptr(d_in.ptr, rows * n_in, 48000, d_mid.ptr, rows * n_out, MEM_DEVICE)
    assert got == rows * n_out
    d_two = ctx.device_buffer(rows * n_out * 2)
    ctx.encode_ptr(d_mid.ptr, d_two.ptr, rows * n_out, MEM_DEVICE)
    ctx.sync()
    a = d_fused.download(np.int16, rows * n_out)
    b = d_two.download(np.int16, rows * n_out)
    assert np.array_equal(a, b)
    assert a.min() == -32768 and a.max() == 32767            # both clamps were exercised
    # determinism: a second run gives the same bytes
    d_fused.memset(0)
    ctx.frames_encode_ptr(d_in.ptr, n_in, d_fused.ptr, n_out, rows, n_in, 48000, MEM_DEVICE)
    ctx.sync()
    assert np.array_equal(d_fused.download(np.int16, rows * n_out), a)
    for d in (d_in, d_fused, d_mid, d_two):
        d.free()


def test_encoder_is_monotonic_over_all_finite_floats(ctx):
    """s -> int16 is non-decreasing over the ordered finite floats (the clamp, the two scale factors
    and the truncation never cross); checked on every 64th float of the whole finite range."""
    pos = np.arange(0, 0x7F800000, 64, dtype=np.uint32)                   # +0 ... just below +Inf, ascending
    neg = (pos[::-1] | np.uint32(0x80000000))                             # most negative ... -0, ascending
    x = np.concatenate([neg, pos]).view(np.float32)
    y = ctx.floatTo16BitPCM(x).astype(np.int32)
    assert (np.diff(y) >= 0).all()
    assert y[0] == -32768 and y[-1] == 32767 and y[neg.size - 1] == 0 and y[neg.size] == 0
