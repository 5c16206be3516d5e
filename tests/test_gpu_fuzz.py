"""Seeded fuzz of the kernel dispatch: random rates, frame lengths, row counts, pitches and pointer
alignments, device-resident and host-resident, every case bit-exact against the oracle and with
canaries around the output."""
import numpy as np
import pytest

import oracle
from pcm_frontend import MEM_DEVICE

pytestmark = pytest.mark.gpu

RATES = [8000, 11025, 12000, 16000, 22050, 24000, 32000, 37800, 44100, 48000, 64000, 88200, 96000, 176400, 192000,
         15999.5, 48000.25]


def _case(rng):
    rate = RATES[rng.randint(len(RATES))]
    kind = rng.randint(4)
    if kind == 0:
        n_in = int(rate * rng.choice([10, 20, 40, 60, 100]) // 1000)
    elif kind == 1:
        n_in = 128 * int(rng.randint(1, 48))
    elif kind == 2:
        n_in = int(rng.randint(1, 5000))
    else:
        n_in = int(rng.choice([1, 2, 3, 7, 8, 9, 15, 16, 17, 255, 256, 257]))
    rows = int(rng.choice([1, 2, 3, 5, 31, 64, 100, 257, 1000, 4097]))
    if rows * n_in > 6_000_000:
        rows = max(1, 6_000_000 // n_in)
    return rate, n_in, rows


@pytest.mark.parametrize("seed", range(12))
def test_fuzz_device_resident(ctx, seed):
    rng = np.random.RandomState(1000 + seed)
    for _ in range(20):
        rate, n_in, rows = _case(rng)
        n_out = oracle.resample_len(n_in, rate)
        pad_in, pad_out = int(rng.choice([0, 0, 0, 1, 3, 8, 24])), int(rng.choice([0, 0, 0, 1, 2, 8]))
        off_in, off_out = int(rng.choice([0, 0, 1, 2, 4, 8])), int(rng.choice([0, 0, 1, 2, 8]))
        pitch_in, pitch_out = n_in + pad_in, max(n_out + pad_out, 1)
        x = rng.uniform(-1.25, 1.25, size=(rows, pitch_in)).astype(np.float32)
        x[rng.randint(rows), rng.randint(pitch_in)] = np.float32(np.nan)
        d_in = ctx.device_buffer(x.nbytes + 64)
        d_out = ctx.device_buffer(rows * pitch_out * 2 + 64)
        d_in.upload(x, 4 * off_in)
        d_out.memset(0x55)
        got_n = ctx.frames_encode_ptr(d_in.offset(4 * off_in), pitch_in, d_out.offset(2 * off_out), pitch_out, rows, n_in, rate,
                                      MEM_DEVICE)
        ctx.sync()
        assert got_n == n_out
        raw = d_out.download(np.int16, rows * pitch_out + 16)
        body = raw[off_out: off_out + rows * pitch_out].reshape(rows, pitch_out)
        want = oracle.frames(x, n_in, rate)
        tag = (rate, n_in, rows, pad_in, pad_out, off_in, off_out)
        assert np.array_equal(body[:, :n_out], want), tag
        assert (body[:, n_out:] == 0x5555).all() or rows == 0, tag          # row padding untouched
        assert (raw[:off_out] == 0x5555).all() and (raw[off_out + rows * pitch_out:] == 0x5555).all(), tag
        d_in.free(); d_out.free()


@pytest.mark.parametrize("seed", range(4))
def test_fuzz_host_resident(ctx, seed):
    rng = np.random.RandomState(2000 + seed)
    for _ in range(15):
        rate, n_in, rows = _case(rng)
        x = rng.uniform(-1.25, 1.25, size=(rows, n_in + int(rng.choice([0, 0, 5])))).astype(np.float32)[:, :n_in]
        got = ctx.frames_encode(x, rate)
        assert np.array_equal(got, oracle.frames(np.ascontiguousarray(x), n_in, rate)), (rate, n_in, rows)


@pytest.mark.parametrize("seed", range(6))
def test_fuzz_fused_meter(ctx, seed):
    """Random shapes through the fused frames + meter call: any rate, 1-40 quanta of 32-256 samples per frame, a few
    to a few thousand rows, three ticks with the level state carried over; rows sprinkled with NaN, Inf, zeros,
    denormals and values around the 2^-60 threshold of the integer widening.  PCM, the level after every quantum and
    the state must be the oracle's, bit for bit (whichever meter kernel the shape lands on)."""
    rng = np.random.RandomState(3000 + seed)
    for _ in range(10):
        rate = RATES[rng.randint(len(RATES))]
        quantum = int(rng.choice([32, 64, 128, 128, 128, 256]))
        nq = int(rng.choice([1, 2, 3, 7, 8, 15, 16, 23, 24, 31, 32, 33, 40]))
        n_in = nq * quantum
        rows = int(rng.choice([1, 2, 5, 31, 64, 100, 257, 1000, 3000]))
        if rows * n_in > 4_000_000:
            rows = max(1, 4_000_000 // n_in)
        state = np.zeros(rows, dtype=np.float64)
        for tick in range(3):
            x = (rng.uniform(-1.2, 1.2, size=(rows, n_in)) * rng.uniform(1e-3, 1.0, size=(rows, 1))).astype(np.float32)
            for _k in range(min(rows, 6)):                                   # special values somewhere
                r, c = rng.randint(rows), rng.randint(n_in)
                x[r, c] = np.float32(rng.choice([np.nan, np.inf, -np.inf, 0.0, -0.0, 1e-40, -3e-39, 2.0 ** -60, 2.0 ** -61, 1e-30]))
            if rows > 2:
                q0 = rng.randint(nq) * quantum
                x[1, q0:q0 + quantum] = 0.0                                  # a silent quantum
                x[2, q0:q0 + quantum] = (rng.randint(0, 2, quantum) * 2.0 ** rng.uniform(-64, -56, quantum)).astype(np.float32)
            pcm, levels, new_state = ctx.frames_encode_meter(x, rate, state, quantum)
            tag = (rate, quantum, nq, rows, tick)
            assert np.array_equal(pcm, oracle.frames(x, n_in, rate)), tag
            for r in range(rows):
                want = oracle.level_meter(x[r], quantum, state[r])
                assert np.array_equal(levels[r].view(np.uint64), want.view(np.uint64)), tag + (r,)
                assert new_state[r:r + 1].view(np.uint64)[0] == want[-1:].view(np.uint64)[0], tag + (r,)
            state = new_state
