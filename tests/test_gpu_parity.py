"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle and the
committed golden fixtures.  Integer / byte work: the bar is BIT-EXACT (0 LSB), for the resampled
path too (the decimator has no filter arithmetic)."""
import json
import os

import numpy as np
import pytest

import oracle
import pcm_frontend
from pcm_frontend import MEM_DEVICE, MEM_HOST

pytestmark = pytest.mark.gpu

RATES = [8000, 11025, 16000, 22050, 24000, 32000, 44100, 48000, 88200, 96000, 192000]


def _bits(hexes):
    return np.array([int(h, 16) for h in hexes], dtype=np.uint32).view(np.float32)


def _audio(rng, rows, n, specials=True):
    x = (0.7 * np.sin(np.arange(n)[None, :] * (0.01 + 0.003 * np.arange(rows)[:, None]))
         + rng.uniform(-0.6, 0.6, size=(rows, n))).astype(np.float32)
    if specials and n >= 8 and rows:
        sp = np.array([np.nan, np.inf, -np.inf, -0.0, 1e-40, -1e-40, 1.0, -1.0], dtype=np.float32)
        x[0, :8] = sp
        x[-1, -8:] = sp[::-1]
    return x


# ---- floatTo16BitPCM ---------------------------------------------------------------------------
def test_value_table_and_hazard_sets(ctx, golden_dir):
    kat = json.load(open(os.path.join(golden_dir, "encode_kat.json")))
    x = _bits([r["bits"] for r in kat["value_table"]])
    want = np.array([r["int16"] for r in kat["value_table"]], dtype=np.int16)
    assert np.array_equal(ctx.floatTo16BitPCM(x), want)
    for key in ("hazard", "random_bits"):
        x = _bits(kat[key]["bits"])
        assert np.array_equal(ctx.floatTo16BitPCM(x), np.array(kat[key]["int16"], dtype=np.int16)), key


def test_module_level_mirror_names():
    x = np.array([1.0, -1.0, 0.5, -0.5, np.nan], dtype=np.float32)
    assert pcm_frontend.floatTo16BitPCM(x).tolist() == [32767, -32768, 16383, -16384, 0]
    assert pcm_frontend.resampleTo16k(x, 16000) is x                      # App.tsx:22 identity
    assert pcm_frontend.resampleTo16k(np.arange(6, dtype=np.float32), 48000).tolist() == [0.0, 3.0]


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 31, 255, 256, 257, 4097, 48000, (1 << 22) + 13])
def test_encode_lengths_vs_oracle(ctx, n):
    rng = np.random.RandomState(n & 0xFFFF)
    x = rng.uniform(-1.3, 1.3, n).astype(np.float32)
    got = ctx.floatTo16BitPCM(x)
    assert got.dtype == np.int16 and got.size == n
    assert np.array_equal(got, oracle.float_to_16bit_pcm(x))


def test_encode_does_not_modify_input_and_rejects_other_dtypes(ctx):
    x = np.linspace(-2, 2, 1000).astype(np.float32)
    keep = x.copy()
    ctx.floatTo16BitPCM(x)
    assert x.tobytes() == keep.tobytes()
    with pytest.raises(TypeError):
        ctx.floatTo16BitPCM(np.zeros(4, dtype=np.float64))


def test_encode_device_resident_all_alignments(ctx):
    """Device pointers at every 4-byte input offset / 2-byte output offset: the aligned ones take
    the 256-bit kernel, the others the general kernel; all must be bit-exact."""
    rng = np.random.RandomState(7)
    n = 5000
    x = rng.uniform(-1.2, 1.2, n + 16).astype(np.float32)
    d_in = ctx.device_buffer(x.nbytes)
    d_out = ctx.device_buffer((n + 32) * 2)
    d_in.upload(x)
    for io in (0, 1, 3, 8):
        for oo in (0, 1, 2, 8):
            d_out.memset(0x55)
            ctx.encode_ptr(d_in.offset(4 * io), d_out.offset(2 * oo), n, MEM_DEVICE)
            ctx.sync()
            raw = d_out.download(np.int16, n + 32)
            assert np.array_equal(raw[oo:oo + n], oracle.float_to_16bit_pcm(x[io:io + n])), (io, oo)
            assert (raw[:oo] == 0x5555).all() and (raw[oo + n:] == 0x5555).all(), "wrote outside the output"


# ---- resampleTo16k -------------------------------------------------------------------------------
@pytest.mark.parametrize("rate", RATES + [12000, 15999, 16001, 37800, 47999.5, 176400, 1e6])
def test_resample_vs_oracle_bitwise(ctx, rate):
    rng = np.random.RandomState(int(rate) & 0xFFFF)
    for n in (0, 1, 2, 128, 441, 882, 960, 2944, 4800, 44100):
        x = rng.uniform(-1, 1, n).astype(np.float32)
        if n > 4:
            x[1] = np.float32(np.nan)
            x.view(np.uint32)[2] = 0x7FA12345          # NaN with a payload must survive the copy
        got = ctx.resampleTo16k(x, rate)
        want = oracle.resample_to_16k(x, rate)
        assert got.dtype == np.float32 and got.tobytes() == want.tobytes(), (rate, n)


def test_resample_range_errors(ctx):
    x = np.zeros(16, np.float32)
    for rate in (0.0, -1.0):
        with pytest.raises(pcm_frontend.PcmRangeError):
            ctx.resampleTo16k(x, rate)
    assert ctx.resampleTo16k(x, float("nan")).size == 0
    assert ctx.resampleTo16k(x, float("inf")).size == 0


# ---- fused frames (App.tsx:281-282) ----------------------------------------------------------------
def test_frames_golden(ctx, golden_dir):
    z = np.load(os.path.join(golden_dir, "frames_kat.npz"))
    for m in json.loads(bytes(z["meta"]).decode()):
        x = np.ascontiguousarray(z[m["key"] + "_in"]).reshape(m["rows"], m["n_in"])
        got = ctx.frames_encode(x, m["rate"])
        assert got.shape == (m["rows"], m["n_out"])
        assert np.array_equal(got, z[m["key"] + "_out"].reshape(m["rows"], m["n_out"])), m


@pytest.mark.parametrize("rate", RATES + [12000, 80000, 37800])
@pytest.mark.parametrize("ms", [10, 20, 60, 100])
def test_frames_host_vs_oracle(ctx, rate, ms):
    rng = np.random.RandomState(int(rate + ms) & 0xFFFF)
    n_in = int(rate * ms // 1000)
    for rows in (1, 3, 257):
        x = _audio(rng, rows, n_in)
        got = ctx.frames_encode(x, rate)
        assert np.array_equal(got, oracle.frames(x, n_in, rate)), (rate, ms, rows)


def test_frames_quantum_multiples_and_tail_drop(ctx):
    """F4: frames are multiples of 128 samples; 23 quanta at 48 kHz -> 981 outputs, tail dropped."""
    rng = np.random.RandomState(23)
    for rate in (48000, 44100, 96000, 32000):
        for k in (1, 7, 23, 47):
            x = _audio(rng, 5, 128 * k)
            got = ctx.frames_encode(x, rate)
            assert got.shape[1] == oracle.resample_len(128 * k, rate)
            assert np.array_equal(got, oracle.frames(x, 128 * k, rate)), (rate, k)
    assert ctx.frames_encode(_audio(rng, 2, 2944), 48000).shape == (2, 981)


def test_frames_pitched_host_buffers(ctx):
    rng = np.random.RandomState(5)
    for rate, n_in in ((48000, 960), (44100, 882), (16000, 320), (32000, 641)):
        big = _audio(rng, 9, n_in + 13)
        view = big[:, 3:3 + n_in]                      # row pitch != n_in, start misaligned
        n_out = oracle.resample_len(n_in, rate)
        outbig = np.full((9, n_out + 5), 0x5555, dtype=np.int16)
        got = ctx.frames_encode(view, rate, out=outbig)
        assert np.array_equal(got, oracle.frames(np.ascontiguousarray(view), n_in, rate))
        assert (outbig[:, n_out:] == 0x5555).all()


@pytest.mark.parametrize("rate,n_in", [(48000, 960), (48000, 2944), (44100, 882), (16000, 320), (96000, 1920),
                                        (32000, 640), (64000, 1280), (8000, 160), (80000, 1600), (192000, 3840),
                                        (128000, 2560), (192000, 128 * 70), (80000, 128 * 30), (128000, 128 * 47)])
def test_frames_device_resident_pitches(ctx, rate, n_in):
    """Device-resident batches: dense rows, padded aligned rows, odd pitches (general kernel)."""
    rng = np.random.RandomState(n_in)
    rows = 300
    n_out = oracle.resample_len(n_in, rate)
    for pitch_in, pitch_out, io, oo in ((n_in, n_out, 0, 0), ((n_in + 7) // 8 * 8 + 8, (n_out + 7) // 8 * 8 + 8, 0, 0),
                                        (n_in + 3, n_out + 1, 1, 1), (n_in + 2, n_out + 2, 0, 2)):
        x = np.zeros((rows, pitch_in), dtype=np.float32)
        x[:, :n_in] = _audio(rng, rows, n_in)
        d_in = ctx.device_buffer(x.nbytes + 64)
        d_out = ctx.device_buffer(rows * pitch_out * 2 + 64)
        d_in.upload(x, 4 * io)
        d_out.memset(0x55)
        got_n = ctx.frames_encode_ptr(d_in.offset(4 * io), pitch_in, d_out.offset(2 * oo), pitch_out, rows, n_in, rate)
        ctx.sync()
        assert got_n == n_out
        raw = d_out.download(np.int16, rows * pitch_out, 2 * oo).reshape(rows, pitch_out)
        assert np.array_equal(raw[:, :n_out], oracle.frames(x, n_in, rate)), (rate, pitch_in, pitch_out, io, oo)
        assert (raw[:, n_out:] == 0x5555).all(), "wrote into the row padding"
        d_in.free(); d_out.free()


def test_forced_general_kernel_matches_fast_kernel(ctx):
    rng = np.random.RandomState(11)
    x = _audio(rng, 513, 960)
    fast = ctx.frames_encode(x, 48000)
    ctx.tune(path=1)
    try:
        slow = ctx.frames_encode(x, 48000)
    finally:
        ctx.tune(path=0)
    assert np.array_equal(fast, slow) and np.array_equal(fast, oracle.frames(x, 960, 48000))


@pytest.mark.parametrize("unroll", [1, 2, 4])
@pytest.mark.parametrize("block", [128, 256, 512])
def test_tuning_variants_are_bit_identical(ctx, unroll, block):
    rng = np.random.RandomState(unroll * 1000 + block)
    ctx.tune(unroll=unroll, block=block)
    try:
        for rate, n_in in ((16000, 320), (32000, 640), (48000, 960), (64000, 1280), (96000, 1920), (44100, 882)):
            x = _audio(rng, 1031, n_in)
            assert np.array_equal(ctx.frames_encode(x, rate), oracle.frames(x, n_in, rate)), (rate, unroll, block)
    finally:
        ctx.tune(unroll=0, block=256)


def test_zero_rows_and_zero_length(ctx):
    assert ctx.frames_encode(np.zeros((0, 960), np.float32), 48000).shape == (0, 320)
    assert ctx.frames_encode(np.zeros((4, 0), np.float32), 48000).shape == (4, 0)
    assert ctx.frames_encode(np.zeros((4, 2), np.float32), 48000).shape == (4, 0)   # floor(2/3) = 0
    with pytest.raises(pcm_frontend.PcmRangeError):
        ctx.frames_encode(np.zeros((4, 8), np.float32), 0.0)


def test_host_pipeline_many_slices(ctx):
    """Small slices force many H2D | kernel | D2H rounds over the three streams."""
    rng = np.random.RandomState(99)
    x = _audio(rng, 4099, 960)
    ctx.tune(slice_bytes=64 * 1024)
    try:
        got = ctx.frames_encode(x, 48000)
        got2 = ctx.frames_encode(_audio(rng, 777, 882), 44100)
    finally:
        ctx.tune(slice_bytes=32 << 20)
    assert np.array_equal(got, oracle.frames(x, 960, 48000))
    assert got2.shape == (777, 320)


def test_pinned_buffers_roundtrip(ctx):
    rng = np.random.RandomState(42)
    pin_in = ctx.pinned((2048, 960), np.float32)
    pin_out = ctx.pinned((2048, 320), np.int16)
    pin_in.array[:] = _audio(rng, 2048, 960)
    got = ctx.frames_encode(pin_in.array, 48000, out=pin_out.array)
    assert np.array_equal(got, oracle.frames(pin_in.array, 960, 48000))
    pin_in.free(); pin_out.free()


# ---- BASELINE.json configs -----------------------------------------------------------------------
def _synth_rows(ctx, rows, n, rate, row0=0):
    d = ctx.device_buffer(rows * n * 4)
    ctx.synth_fill(d.ptr, n, rows, n, rate, 0x5EED, row0)
    ctx.sync()
    return d


def test_config2_1024_streams_16k_encode_only(ctx):
    rows, n = 1024, 320
    d_in = _synth_rows(ctx, rows, n, 16000)
    x = d_in.download(np.float32, rows * n).reshape(rows, n)
    assert np.isfinite(x).all() and x.max() > 1.0 and x.min() < -1.0      # both clamps exercised
    d_out = ctx.device_buffer(rows * n * 2)
    ctx.frames_encode_ptr(d_in.ptr, n, d_out.ptr, n, rows, n, 16000)
    ctx.sync()
    got = d_out.download(np.int16, rows * n).reshape(rows, n)
    assert np.array_equal(got, oracle.frames(x, n, 16000, threads=4))
    assert got.max() == 32767 and got.min() == -32768


@pytest.mark.parametrize("ms", [20, 60])
def test_config3_65536_streams_48k(ctx, ms):
    rows, n_in = 65536, 48 * ms
    n_out = n_in // 3
    d_in = _synth_rows(ctx, rows, n_in, 48000)
    d_out = ctx.device_buffer(rows * n_out * 2)
    ctx.frames_encode_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, 48000)
    ctx.sync()
    x = d_in.download(np.float32, rows * n_in).reshape(rows, n_in)
    got = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
    assert np.array_equal(got, oracle.frames(x, n_in, 48000, threads=8))


def test_config4_full_size_properties(ctx):
    """2^20 sessions x 960 samples (4.0 GB in, 0.67 GB out), BASELINE configs[3] at full single-GPU
    size: (1) EVERY row bit-exact vs the oracle (the multi-threaded oracle sweeps the 10^9 samples in
    well under a second per chunk), (2) every shard of a session-sharded run reproduces the
    corresponding rows of the whole run."""
    import os
    rows, n_in, n_out = 1 << 20, 960, 320
    d_in = _synth_rows(ctx, rows, n_in, 48000)
    d_out = ctx.device_buffer(rows * n_out * 2)
    ctx.frames_encode_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, 48000)
    ctx.sync()
    whole = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
    threads = os.cpu_count() or 1
    step = 1 << 17
    for r0 in range(0, rows, step):                        # all 1 048 576 rows, 128 Ki rows at a time
        x = d_in.download(np.float32, step * n_in, r0 * n_in * 4).reshape(step, n_in)
        want = oracle.frames(x, n_in, 48000, threads=threads)
        assert np.array_equal(whole[r0:r0 + step], want), r0
    # session-affine shards (8 ranks) on the same data give the same bytes
    d_out.memset(0)
    for rank in range(8):
        first, count = pcm_frontend.shard_sessions(rows, 8, rank)
        ctx.frames_encode_ptr(d_in.offset(first * n_in * 4), n_in, d_out.offset(first * n_out * 2), n_out, count, n_in, 48000)
    ctx.sync()
    sharded = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
    assert np.array_equal(sharded, whole)
    d_in.free(); d_out.free()


def test_launch_counter_counts_kernels(ctx):
    before = ctx.launches
    ctx.floatTo16BitPCM(np.zeros(1000, np.float32))
    assert ctx.launches == before + 1


@pytest.mark.parametrize("rate,n_in", [(16000, 320), (48000, 960), (44100, 882)])
def test_tick_graph_replay(ctx, rate, n_in):
    """A recorded tick replays on new data in the same buffers and stays bit-exact."""
    rng = np.random.RandomState(n_in + 1)
    rows = 1024
    n_out = oracle.resample_len(n_in, rate)
    d_in = ctx.device_buffer(rows * n_in * 4)
    d_out = ctx.device_buffer(rows * n_out * 2)
    g = ctx.tick_graph(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate)
    before = ctx.launches
    for _ in range(3):
        x = _audio(rng, rows, n_in)
        d_in.upload(x)
        g.launch()
        ctx.sync()
        got = d_out.download(np.int16, rows * n_out).reshape(rows, n_out)
        assert np.array_equal(got, oracle.frames(x, n_in, rate))
    assert ctx.launches == before + 3
    g.destroy()
    d_in.free(); d_out.free()


@pytest.mark.parametrize("rate,n_in", [(44100, 882), (22050, 441), (11025, 1103), (8000, 160), (44100, 2646),
                                        (88200, 1764), (48000, 960), (80000, 1600), (37800, 756),
                                        (48000, 2944), (44100, 1000), (96000, 1927), (48000, 128 * 47), (16000, 333)])
def test_tile_staged_kernel_matches_general_kernel(ctx, rate, n_in):
    """Kernel 2b (TMA bulk copy -> shared memory -> table picks) against kernel 2 and the oracle, on
    dense device batches whose row count is not a multiple of the tile height."""
    rng = np.random.RandomState(n_in * 7)
    n_out = oracle.resample_len(n_in, rate)
    for rows in (1, 37, 1031, 4099):
        x = _audio(rng, rows, n_in)
        want = oracle.frames(x, n_in, rate)
        d_in = ctx.device_buffer(x.nbytes)
        d_out = ctx.device_buffer(rows * n_out * 2 + 64)
        d_in.upload(x)
        for tune in (dict(path=2, tile=1, stages=2), dict(path=2, tile=1, stages=3, tile_kb=12),
                     dict(path=2, tile=1, stages=1, tile_kb=56), dict(path=1, tile=0)):
            ctx.tune(**tune)
            try:
                d_out.memset(0x55)
                ctx.frames_encode_ptr(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate)
                ctx.sync()
            finally:
                ctx.tune(path=0, tile=1, stages=2, tile_kb=28)
            raw = d_out.download(np.int16, rows * n_out + 32)
            assert np.array_equal(raw[: rows * n_out].reshape(rows, n_out), want), (rate, rows, tune)
            assert (raw[rows * n_out:] == 0x5555).all()
        d_in.free(); d_out.free()


def test_two_contexts_from_two_threads(ctx):
    """A context is not thread-safe, but two contexts on one device used from two threads are fine."""
    import threading

    import pcm_frontend
    rng = np.random.RandomState(5)
    x = _audio(rng, 2048, 960)
    want = oracle.frames(x, 960, 48000)
    errors = []

    def work():
        try:
            with pcm_frontend.Context(0) as c:
                for _ in range(20):
                    if not np.array_equal(c.frames_encode(x, 48000), want):
                        errors.append("mismatch")
        except Exception as e:   # noqa: BLE001
            errors.append(repr(e))
    ts = [threading.Thread(target=work) for _ in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors
