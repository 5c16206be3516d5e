"""GPU tests of capacity-up-front (pcmb200_ctx_reserve) and of the table lifetime of
recorded ticks: after a reserve the host-resident frame path must not allocate, free or synchronise
on its own; a tick graph must survive the eviction of its pick table from the context's cache."""
import numpy as np
import pytest

import oracle
import pcm_frontend

pytestmark = pytest.mark.gpu


def test_no_allocation_after_reserve():
    rng = np.random.RandomState(11)
    with pcm_frontend.Context(0) as ctx:
        ctx.reserve(max_rows=50000, max_n_in=2944, max_n_out=2944, max_rate_classes=6)
        base = ctx.alloc_stats
        assert base["device_mallocs"] >= 6 + 6            # 3 slots x (in, out) + 6 tables
        # shapes within the reserve: integer ratios, the 44.1 kHz family (tables), upsampling, a batch
        # far larger than one slice (cut into slices instead of growing), the stand-alone functions
        for rows, n_in, rate in ((40000, 960, 48000), (3000, 882, 44100), (500, 2646, 44100), (700, 441, 22050),
                                 (900, 160, 8000), (64, 2944, 48000), (20000, 320, 16000), (333, 1764, 88200),
                                 (45000, 2944, 48000)):
            x = rng.uniform(-1.2, 1.2, size=(rows, n_in)).astype(np.float32)
            got = ctx.frames_encode(x, rate)
            assert np.array_equal(got, oracle.frames(x, n_in, rate, threads=8)), (rows, n_in, rate)
        y = rng.uniform(-1, 1, 2944).astype(np.float32)
        assert np.array_equal(ctx.floatTo16BitPCM(y), oracle.float_to_16bit_pcm(y))
        assert np.array_equal(ctx.resampleTo16k(y, 44100).view(np.uint32), oracle.resample_to_16k(y, 44100).view(np.uint32))
        assert ctx.alloc_stats == base, (base, ctx.alloc_stats)


def test_without_reserve_the_pool_grows_on_demand():
    with pcm_frontend.Context(0) as ctx:
        assert ctx.alloc_stats["device_mallocs"] == 0
        x = np.zeros((100, 882), np.float32)
        ctx.frames_encode(x, 44100)
        first = ctx.alloc_stats
        assert first["device_mallocs"] >= 3                # a table + one slot's two buffers
        ctx.frames_encode(x, 44100)
        assert ctx.alloc_stats == first                    # same shape again: nothing new


def test_tick_graph_survives_table_eviction():
    """a recorded 44.1 kHz tick used to keep a pointer into the context's table cache, which
    evicts (and frees) after 32 other shapes.  The graph now owns its table."""
    rng = np.random.RandomState(5)
    rows, n_in, rate = 256, 882, 44100
    n_out = oracle.resample_len(n_in, rate)
    with pcm_frontend.Context(0) as ctx:
        d_in, d_out = ctx.device_buffer(rows * n_in * 4), ctx.device_buffer(rows * n_out * 2)
        g = ctx.tick_graph(d_in.ptr, n_in, d_out.ptr, n_out, rows, n_in, rate)
        x = rng.uniform(-1, 1, size=(rows, n_in)).astype(np.float32)
        d_in.upload(x)
        g.launch(); ctx.sync()
        want = oracle.frames(x, n_in, rate)
        assert np.array_equal(d_out.download(np.int16, rows * n_out).reshape(rows, n_out), want)
        # 40 other table shapes push everything older out of the 32-entry cache
        for k in range(40):
            n2 = 400 + 7 * k
            z = rng.uniform(-1, 1, size=(8, n2)).astype(np.float32)
            assert np.array_equal(ctx.frames_encode(z, 44100), oracle.frames(z, n2, 44100))
        d_out.memset(0)
        x2 = rng.uniform(-1, 1, size=(rows, n_in)).astype(np.float32)
        d_in.upload(x2)
        g.launch(); ctx.sync()
        assert np.array_equal(d_out.download(np.int16, rows * n_out).reshape(rows, n_out), oracle.frames(x2, n_in, rate))
        g.destroy()
