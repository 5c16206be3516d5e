"""CPU tests: the C oracle (oracle/pcm_oracle.c) against the committed golden fixtures
(tests/golden/, produced by the independent pure-Python restatement make_golden.py) and
against the hand-stated value table of SURVEY.md 8a-1.  No GPU needed."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle


def _bits(hexes):
    return np.array([int(h, 16) for h in hexes], dtype=np.uint32).view(np.float32)


@pytest.fixture(scope="module")
def enc_kat(golden_dir):
    with open(os.path.join(golden_dir, "encode_kat.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def idx_kat(golden_dir):
    with open(os.path.join(golden_dir, "index_kat.json")) as fh:
        return json.load(fh)["cases"]


def test_value_table(enc_kat):
    for row in enc_kat["value_table"]:
        got = oracle.float_to_16bit_pcm(_bits([row["bits"]]))[0]
        assert got == row["int16"], row


def test_hazard_and_random_bits(enc_kat):
    for key in ("hazard", "random_bits"):
        x = _bits(enc_kat[key]["bits"])
        want = np.array(enc_kat[key]["int16"], dtype=np.int16)
        assert np.array_equal(oracle.float_to_16bit_pcm(x), want), key


def test_truncation_not_rounding():
    x = np.array([0.5, -0.5, 0.99999994, 2.0 ** -15, -(2.0 ** -15)], dtype=np.float32)
    assert oracle.float_to_16bit_pcm(x).tolist() == [16383, -16384, 32766, 0, -1]


def test_to_int16_wraps_like_ecmascript():
    L = oracle.lib()
    assert L.oracle_to_int16(32768.0) == -32768
    assert L.oracle_to_int16(65536.0 + 5.9) == 5
    assert L.oracle_to_int16(-32769.0) == 32767
    assert L.oracle_to_int16(float("nan")) == 0
    assert L.oracle_to_int16(float("inf")) == 0
    assert L.oracle_to_int16(-0.0) == 0


def test_index_tables(idx_kat):
    for c in idx_kat:
        n_out = oracle.resample_len(c["n_in"], c["rate"])
        assert n_out == c["n_out"], c
        if c.get("identity"):
            continue
        idx = oracle.index_table(c["n_in"], c["rate"])
        assert hashlib.sha256(idx.astype("<u4").tobytes()).hexdigest() == c["sha256_u32le"], c
        assert idx[:8].tolist() == c["first"]
        exact = (np.arange(n_out, dtype=np.int64) * c["rate"]) // 16000
        diff = np.nonzero(idx.astype(np.int64) != exact)[0]
        assert diff.size == c["n_differs_from_rational"]
        if diff.size:
            assert diff[0] == c["first_differs_at"] and idx[diff[0]] == c["pick_at_first_differs"]


def test_44k1_running_sum_drift_pin():
    # at 44.1 kHz the running double sum picks sample 1322 (not 1323) at i=480
    idx = oracle.index_table(44100, 44100)
    assert idx[480] == 1322 and (480 * 44100) // 16000 == 1323


def test_identity_at_16k_and_upsample_and_empty():
    x = np.array([0.1, np.nan, -0.2, np.inf], dtype=np.float32)
    assert oracle.resample_to_16k(x, 16000).tobytes() == x.tobytes()
    assert oracle.index_table(4, 8000).tolist() == [0, 0, 1, 1, 2, 2, 3, 3]
    assert oracle.resample_to_16k(np.zeros(0, np.float32), 48000).size == 0
    assert oracle.float_to_16bit_pcm(np.zeros(0, np.float32)).size == 0
    assert oracle.resample_len(2944, 48000) == 981          # F4: 23 quanta -> 981, tail dropped
    assert oracle.resample_len(882, 44100) == 320


def test_resample_len_range_errors():
    with pytest.raises(ValueError):
        oracle.resample_len(100, 0.0)        # n/0 = Infinity -> RangeError
    with pytest.raises(ValueError):
        oracle.resample_len(100, -48000.0)   # negative length -> RangeError
    assert oracle.resample_len(100, float("nan")) == 0
    assert oracle.resample_len(100, float("inf")) == 0
    assert oracle.resample_len(0, 0.0) == 0  # 0/0 = NaN -> length 0


def test_frames_golden(golden_dir):
    z = np.load(os.path.join(golden_dir, "frames_kat.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    for m in meta:
        x, want = z[m["key"] + "_in"], z[m["key"] + "_out"]
        got = oracle.frames(np.ascontiguousarray(x).reshape(m["rows"], m["n_in"]), m["n_in"], m["rate"])
        assert got.shape == (m["rows"], m["n_out"])
        assert np.array_equal(got, want.reshape(m["rows"], m["n_out"])), m
        got_mt = oracle.frames(np.ascontiguousarray(x).reshape(m["rows"], m["n_in"]), m["n_in"], m["rate"], threads=2)
        assert np.array_equal(got_mt, got)


def test_level_meter_matches_numpy_double():
    rng = np.random.RandomState(1)
    x = rng.uniform(-1, 1, 128 * 9).astype(np.float32)
    lv = oracle.level_meter(x)
    prev = 0.0
    for q in range(9):
        s = 0.0
        for v in x[q * 128:(q + 1) * 128]:
            s += float(v) * float(v)
        prev = max(np.sqrt(s / 128), prev * 0.85)
        assert lv[q] == prev


def test_aggregator_plan_and_keepalive():
    # 48 kHz: a quantum every 128/48 ms; the 23rd quantum (61.33 ms) is the first with elapsed >= 60
    arr = (np.arange(1, 70) * 128 / 48.0)
    fl, last = oracle.aggregator_plan(arr, 60.0, 0.0)
    assert np.nonzero(fl)[0].tolist() == [22, 45, 68]
    assert last == arr[68]
    ka = oracle.keepalive_frame()
    assert ka.size == 1600 and not ka.any()
