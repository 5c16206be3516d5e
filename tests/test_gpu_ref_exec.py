"""The CUDA path against outputs of the reference's own source text (tests/golden/ref_exec.npz, written by
tests/golden/make_ref_exec.py: App.tsx's functions executed from their text by tests/golden/mini_js.py).
No oracle in between: what the library returns through the C ABI is compared with what the reference's
code produced, bit for bit."""
import json
import os

import numpy as np
import pytest

import pcm_frontend

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def vec():
    return np.load(os.path.join(GOLDEN, "ref_exec.npz")), json.load(open(os.path.join(GOLDEN, "ref_exec.json")))


def test_encoder_vs_reference_text(ctx, vec):
    z, _ = vec
    assert np.array_equal(ctx.floatTo16BitPCM(z["enc_bits"].view(np.float32)), z["enc_out"])
    assert np.array_equal(ctx.floatTo16BitPCM(z["enc_audio_in"]), z["enc_audio_out"])


def test_frames_vs_reference_text(ctx, vec):
    """floatTo16BitPCM(resampleTo16k(buf, rate)) (App.tsx:281-282): fused call and the two separate calls."""
    z, rec = vec
    for c in rec["frame_cases"]:
        x, want = z[c["key"] + "_in"], z[c["key"] + "_out"]
        fused = ctx.frames_encode(x.reshape(1, -1), c["rate"])[0]
        assert fused.size == c["n_out"] and np.array_equal(fused, want), c
        two = ctx.floatTo16BitPCM(ctx.resampleTo16k(x, c["rate"]))
        assert np.array_equal(two, want), c
        # the same frame as every row of a batch large enough to reach the tiled kernels
        batch = np.repeat(x.reshape(1, -1), 300, axis=0)
        got = ctx.frames_encode(batch, c["rate"])
        assert (got == want[None, :]).all(), c


def test_sessions_vs_reference_text(ctx, vec):
    """The onmessage handler (App.tsx:264-289) against the device aggregator: the five recorded sessions are
    pushed interleaved in arrival order, in pushes of uneven size; frames, their order and the level after each
    session's last flushed quantum must be the reference's."""
    z, rec = vec
    for rate in sorted({s["rate"] for s in rec["sessions"]}):
        group = [s for s in rec["sessions"] if s["rate"] == rate]
        agg = pcm_frontend.Aggregator(ctx, len(group), rate, max_frame_quanta=96, max_push_quanta=512, max_out_samples=1 << 18)
        agg.set_meter(True)
        events = []
        for k, s in enumerate(group):
            agg.reset_session(k, s["t0"])
            events += [(t, k, q) for q, t in enumerate(z[s["key"] + "_arrival"])]
        events.sort()
        got = {k: [] for k in range(len(group))}
        i, sizes, n_push = 0, [1, 5, 23, 64, 2, 200], 0
        while i < len(events):
            chunk = events[i:i + sizes[n_push % len(sizes)]]
            i += len(chunk)
            n_push += 1
            frames = agg.push(np.array([e[1] for e in chunk]), np.array([e[0] for e in chunk]),
                              np.stack([z[group[e[1]]["key"] + "_in"][e[2]] for e in chunk]))
            for k, pcm in frames:
                got[k].append(pcm)
        levels = agg.levels(0, len(group))
        agg.close()
        for k, s in enumerate(group):
            lens, blob = z[s["key"] + "_frame_len"], z[s["key"] + "_frames"]
            assert len(got[k]) == lens.size == s["flushes"], (s["key"], len(got[k]))
            off = 0
            for f, pcm in enumerate(got[k]):
                assert pcm.size == lens[f] and np.array_equal(pcm, blob[off:off + lens[f]]), (s["key"], f)
                off += int(lens[f])
            last = np.nonzero(z[s["key"] + "_flush"])[0][-1]
            want = z[s["key"] + "_levels"][last:last + 1]
            assert levels[k:k + 1].view(np.uint64)[0] == want.view(np.uint64)[0], (s["key"], levels[k], want)


@pytest.mark.parametrize("quanta_per_row", [1, 4, 8, 23])
def test_level_meter_vs_reference_text(ctx, vec, quanta_per_row):
    """App.tsx:267-271 through the fused frames + meter call: every session's quanta as consecutive ticks of
    `quanta_per_row` quanta, the level state carried from tick to tick; the level after EVERY quantum must be
    the reference's (binary64, bit for bit, NaN and Infinity included)."""
    z, rec = vec
    sess = rec["sessions"]
    n_ticks = min(s["quanta"] for s in sess) // quanta_per_row
    state = np.zeros(len(sess), dtype=np.float64)
    for tick in range(n_ticks):
        lo = tick * quanta_per_row
        batch = np.stack([z[s["key"] + "_in"][lo:lo + quanta_per_row].reshape(-1) for s in sess])
        _, levels, state = ctx.frames_encode_meter(batch, 48000, state)
        for k, s in enumerate(sess):
            want = z[s["key"] + "_levels"][lo:lo + quanta_per_row]
            assert np.array_equal(levels[k].view(np.uint64), want.view(np.uint64)), (s["key"], tick)
            assert state[k:k + 1].view(np.uint64)[0] == want[-1:].view(np.uint64)[0]


def test_keepalive_vs_reference_text(vec):
    _, rec = vec
    ka = rec["keepalive"]
    frame = pcm_frontend.keepalive_frame()
    assert frame.dtype == np.int16 and frame.size == ka["samples"] == 1600 and not frame.any()
    sched = pcm_frontend.KeepAliveScheduler(2)
    assert sched.period == ka["period_ms"]                      # the default period is the reference's 2000 ms
