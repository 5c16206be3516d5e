"""GPU parity of the fused level meter (App.tsx:267-271) against the oracle:
binary64 results must be BIT-identical (sequential accumulation order is preserved on the device)."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _oracle_levels(x, n_in, state, quantum=128):
    nq = n_in // quantum
    lv = np.empty((x.shape[0], nq), dtype=np.float64)
    for r in range(x.shape[0]):
        lv[r] = oracle.level_meter(x[r, : nq * quantum], quantum, state[r])
    return lv


@pytest.mark.parametrize("rate,n_in,rows", [(48000, 2944, 700), (48000, 960, 1031), (44100, 882, 517), (16000, 320, 300),
                                             (48000, 128, 64), (44100, 2646, 100), (48000, 96000, 3), (32000, 641, 50)])
def test_fused_meter_bitwise(ctx, rate, n_in, rows):
    rng = np.random.RandomState(n_in + rows)
    n_out = oracle.resample_len(n_in, rate)
    nq = n_in // 128
    x = (rng.uniform(-1, 1, size=(rows, n_in)) * rng.uniform(0.01, 1.0, size=(rows, 1))).astype(np.float32)
    if nq:
        x[0, 5] = np.float32(np.nan)            # NaN poisons that session's level for good (Math.max semantics)
        x[1, 3] = np.float32(np.inf)
        if rows > 8:
            # a quantum of denormals, zeros and the smallest normals: the sums stay near 2^-250, where
            # a wrong widening of a denormal would change the rounded result
            x[3, :128] = (rng.randint(0, 3, 128) * rng.uniform(0, 1.2e-38, 128)).astype(np.float32)
            x[3, 7] = np.float32(1e-40)
            # around the 2^-60 threshold of the fast widening, with zeros before / between / after
            x[5, :128] = 0.0
            x[5, 17] = np.float32(2.0 ** -60)
            x[5, 90] = np.float32(-(2.0 ** -60) * 1.0000001)
            x[6, :128] = (rng.randint(0, 2, 128) * 2.0 ** rng.uniform(-63, -57, 128)).astype(np.float32)
            x[7, :128] = 0.0
            x[7, 127] = np.float32(3.0e-19)        # first nonzero sample is the last of the quantum
            x[4, :128] = 0.0                        # digital silence -> exactly 0
            if n_in > 130:
                x[4, 130] = np.float32(-1e-45)
    state = rng.uniform(0, 1, rows)
    state[2 % rows] = 0.0
    d_in = ctx.device_buffer(x.nbytes)
    d_out = ctx.device_buffer(max(rows * n_out * 2, 16))
    d_state = ctx.device_buffer(rows * 8)
    d_lv = ctx.device_buffer(max(rows * nq * 8, 16))
    d_in.upload(x)
    d_state.upload(state)
    got_n = ctx.frames_encode_meter_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate, d_state.ptr, d_lv.ptr)
    ctx.sync()
    assert got_n == n_out
    pcm = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
    assert np.array_equal(pcm, oracle.frames(x, n_in, rate))          # the PCM is unchanged by the fusion
    want = _oracle_levels(x, n_in, state)
    got = d_lv.download(np.float64, rows * nq).reshape(rows, nq)
    assert got.view(np.uint64).tolist() == want.view(np.uint64).tolist() if rows * nq < 5000 else \
        np.array_equal(got.view(np.uint64), want.view(np.uint64))
    new_state = d_state.download(np.float64, rows)
    expect_state = want[:, -1] if nq else state
    assert np.array_equal(new_state.view(np.uint64), expect_state.view(np.uint64))
    # second frame continues the decay chain from the stored state; no per-quantum trace this time
    x2 = (x * 0.1).astype(np.float32)
    d_in.upload(x2)
    ctx.frames_encode_meter_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate, d_state.ptr, None)
    ctx.sync()
    want2 = _oracle_levels(x2, n_in, expect_state)
    s2 = d_state.download(np.float64, rows)
    assert np.array_equal(s2.view(np.uint64), (want2[:, -1] if nq else expect_state).view(np.uint64))
    for d in (d_in, d_out, d_state, d_lv):
        d.free()


def test_meter_fallback_kernel_on_pitched_rows(ctx):
    rng = np.random.RandomState(9)
    rows, n_in, pitch = 77, 960, 1000
    x = np.zeros((rows, pitch), np.float32)
    x[:, :n_in] = rng.uniform(-1, 1, size=(rows, n_in))
    state = np.zeros(rows)
    d_in, d_out = ctx.device_buffer(x.nbytes), ctx.device_buffer(rows * 320 * 2)
    d_state, d_lv = ctx.device_buffer(rows * 8), ctx.device_buffer(rows * 7 * 8)
    d_in.upload(x); d_state.upload(state)
    ctx.frames_encode_meter_ptr(d_in.ptr, pitch, d_out.ptr, 320, rows, n_in, 48000, d_state.ptr, d_lv.ptr)
    ctx.sync()
    want = _oracle_levels(x, n_in, state)
    assert np.array_equal(d_lv.download(np.float64, rows * 7).reshape(rows, 7).view(np.uint64), want.view(np.uint64))
    assert np.array_equal(d_out.download(np.int16, rows * 320).reshape(rows, 320), oracle.frames(x, n_in, 48000))


def test_host_array_convenience(ctx):
    rng = np.random.RandomState(3)
    x = rng.uniform(-1, 1, size=(50, 2944)).astype(np.float32)
    state = rng.uniform(0, 1, 50)
    pcm, levels, new_state = ctx.frames_encode_meter(x, 48000, state)
    assert np.array_equal(pcm, oracle.frames(x, 2944, 48000))
    want = _oracle_levels(x, 2944, state)
    assert np.array_equal(levels.view(np.uint64), want.view(np.uint64))
    assert np.array_equal(new_state.view(np.uint64), want[:, -1].view(np.uint64))


@pytest.mark.parametrize("rate,n_in,rows,quantum", [
    (44100, 2816, 333, 128),     # pick table + kernel 2d (n_in % 4 == 0), 22 quanta: one frame per stage
    (48000, 4096, 97, 128),      # 32 quanta: a stage's lanes exactly full
    (48000, 4224, 41, 128),      # 33 quanta: more than a warp of quanta per frame -> kernel 2c / 2b take it
    (16000, 2048, 130, 128),     # ratio 1
    (96000, 3840, 77, 128),      # ratio 6, whole 8-sample chunks
    (8000, 1024, 50, 128),       # upsampling (table), n_out = 2 * n_in
    (48000, 960, 1, 128), (48000, 960, 3, 128), (48000, 2944, 1, 128), (48000, 2944, 2, 128),   # fewer rows than a stage
    (48000, 1920, 211, 64),      # other quanta: 64 ...
    (48000, 3072, 59, 256),      # ... and 256 samples
    (44100, 3528, 35, 84),       # a quantum that is not a power of two (sum / quantum is a real division)
])
def test_meter_shapes_of_kernel_2d(ctx, rate, n_in, rows, quantum):
    """Shapes that steer kernel 2d (warp pairs) through its corners -- tables, full and partial stages, a
    single row, other quantum sizes -- and past its limits into the fallbacks; PCM, per-quantum trace
    and final state bit-identical to the oracle, with and without the trace (the no-trace path runs the
    level chain in its parallel form)."""
    rng = np.random.RandomState(n_in * 7 + rows + quantum)
    n_out = oracle.resample_len(n_in, rate)
    nq = n_in // quantum
    x = (rng.uniform(-1, 1, size=(rows, n_in)) * rng.uniform(0.001, 1.0, size=(rows, 1))).astype(np.float32)
    x[0, :quantum] = 0.0                                   # a silent first quantum
    if rows > 2:
        x[2, quantum + 1] = np.float32(np.nan)             # poisons row 2 from its second quantum on
        x[1, 5] = np.float32(1e-41)                        # a denormal among ordinary samples
    state = rng.uniform(0, 2, rows)
    state[rows // 2] = np.nan if rows > 4 else state[rows // 2]
    want = np.empty((rows, nq))
    for r in range(rows):
        want[r] = oracle.level_meter(x[r, : nq * quantum], quantum, state[r])
    d_in, d_out = ctx.device_buffer(x.nbytes), ctx.device_buffer(max(rows * n_out * 2, 16))
    d_state, d_lv = ctx.device_buffer(rows * 8), ctx.device_buffer(rows * nq * 8)
    d_in.upload(x)
    for with_trace in (True, False):
        d_state.upload(state)
        d_out.memset(0x22)
        ctx.frames_encode_meter_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate, d_state.ptr,
                                    d_lv.ptr if with_trace else None, quantum)
        ctx.sync()
        assert np.array_equal(d_out.download(np.int16, rows * n_out).reshape(rows, n_out), oracle.frames(x, n_in, rate))
        if with_trace:
            got = d_lv.download(np.float64, rows * nq).reshape(rows, nq)
            assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
        assert np.array_equal(d_state.download(np.float64, rows).view(np.uint64), want[:, -1].view(np.uint64)), with_trace
    for d in (d_in, d_out, d_state, d_lv):
        d.free()
