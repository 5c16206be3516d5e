"""The premise of the aggregator's staged-output layout (agg_kernels.cuh: AggAppendArgs), checked on
the CPU against the oracle alone: the reference's pick sequence floor(idx), idx += ratio (App.tsx:26-30) does not depend
on the length of the buffer it runs over, so

  * the pick table of a short frame is a prefix of the pick table of a longer one,
  * the outputs whose pick falls inside quantum j form one contiguous range [out_lo[j], out_lo[j+1]) that is the same
    for every frame length that contains quantum j, and
  * converting each quantum when it arrives (picking its range out of the 128 samples alone) and concatenating gives the
    bytes floatTo16BitPCM(resampleTo16k(whole buffer, rate)) gives, for every frame length -- up to the frame's own
    output count floor(n / ratio); outputs whose pick lies at or beyond the end of the frame read `undefined` in the
    reference and encode as 0.

No GPU, no library: if this ever failed, the CUDA design would be wrong however well it matched itself."""
import numpy as np
import pytest

import oracle

RATES = [8000, 11025, 12345.678, 16000, 22050, 24000, 32000, 44100, 47999, 48000, 48001, 88200, 96000, 176400, 192000]
Q = 128
OUTSIDE = 0xFFFFFFFF


def out_lo_table(rate, max_q):
    """out_lo[j] = first output whose pick is >= j * Q, from the pick table of the longest frame."""
    n_out = oracle.resample_len(max_q * Q, rate)
    picks = oracle.index_table(max_q * Q, rate, n_out).astype(np.int64)
    picks[picks == OUTSIDE] = max_q * Q                      # beyond the buffer: past every quantum
    assert (np.diff(picks) >= 0).all()                       # picks never decrease with the output index
    return np.searchsorted(picks, np.arange(max_q + 1) * Q, side="left"), picks, n_out


@pytest.mark.parametrize("rate", RATES)
def test_pick_table_of_a_short_frame_is_a_prefix(rate):
    max_q = 40
    _, full, _ = out_lo_table(rate, max_q)
    for nq in (1, 2, 7, 21, 23, 39):
        n_out = oracle.resample_len(nq * Q, rate)
        short = oracle.index_table(nq * Q, rate, n_out).astype(np.int64)
        inside = short != OUTSIDE
        assert np.array_equal(short[inside], full[:n_out][inside]), (rate, nq)
        # a pick the short frame calls outside is, in the long table, at or beyond the short frame's end
        assert (full[:n_out][~inside] >= nq * Q).all(), (rate, nq)


@pytest.mark.parametrize("rate", RATES)
def test_convert_on_arrival_equals_convert_at_flush(rate):
    rng = np.random.RandomState(int(rate) % 977)
    max_q = 40
    out_lo, picks, _ = out_lo_table(rate, max_q)
    x = (rng.uniform(-1.1, 1.1, size=(max_q, Q)) * rng.uniform(0.05, 1.0, size=(max_q, 1))).astype(np.float32)
    x[3, 5] = np.nan; x[4, 0] = np.inf; x[4, 1] = -np.inf     # the encoder's special cases travel through the staging too
    # every quantum converted alone, as the append kernel does: its range of outputs, picked inside its 128 samples
    staged = []
    for j in range(max_q):
        sel = picks[out_lo[j]:out_lo[j + 1]] - j * Q
        assert ((sel >= 0) & (sel < Q)).all(), (rate, j)
        staged.append(oracle.float_to_16bit_pcm(x[j][sel]))
    for nq in (1, 2, 6, 7, 21, 22, 23, 24, 39, 40):
        want = oracle.frames(np.ascontiguousarray(x[:nq].reshape(1, -1)), nq * Q, rate)[0]
        got = np.concatenate(staged[:nq]) if nq else np.zeros(0, np.int16)
        n_out = want.size
        # The frame has newLen = floor(n / ratio) outputs (App.tsx:24), which can be one fewer than the outputs picked
        # inside its quanta (11.025 kHz, one quantum: 185 outputs, 186 picks below 128) -- the flush then simply does not
        # use the last staged output -- or more (picks at or beyond the end of the frame, which the reference reads as
        # undefined -> NaN -> 0).
        m = min(got.size, n_out)
        assert np.array_equal(want[:m], got[:m]), (rate, nq)
        assert not want[m:].any(), (rate, nq)
        assert got.size == out_lo[nq] and got.size - n_out <= 1, (rate, nq, got.size, n_out)


def test_pairs_per_quantum_bound():
    """The append kernel deals out aligned output PAIRS; the host sizes the deal by the most pairs any quantum position
    can touch (pcm_aggregator.inl).  Restated here: that bound covers every position at every rate."""
    for rate in RATES:
        out_lo, _, _ = out_lo_table(rate, 64)
        lo, hi = out_lo[:-1], out_lo[1:]
        pairs = np.where(hi > lo, (hi + 1) // 2 - lo // 2, 0)
        bound = max(2, int(pairs.max()))
        per_q = hi - lo
        assert bound * 2 >= per_q.max() and bound <= per_q.max() // 2 + 2, (rate, bound, per_q.max())
