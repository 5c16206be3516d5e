"""The third, integer-only derivation of the path (tests/golden/make_golden.py: int_float_to_16bit_pcm_bits,
soft_add_f64, int_pick_indices -- no hardware floating point in the arithmetic) against the committed
fixtures and against the C oracle.  Three restatements that share no arithmetic must agree: Python
floats (the fixtures), C doubles (the oracle), Python integers (this file).  CPU only."""
import hashlib
import importlib.util
import json
import os
import struct

import numpy as np

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mg)


def test_integer_encoder_reproduces_the_committed_vectors():
    kat = json.load(open(os.path.join(HERE, "golden", "encode_kat.json")))
    for row in kat["value_table"]:
        assert mg.int_float_to_16bit_pcm_bits(int(row["bits"], 16)) == row["int16"], row
    for key in ("hazard", "random_bits"):
        for b, want in zip(kat[key]["bits"], kat[key]["int16"]):
            assert mg.int_float_to_16bit_pcm_bits(int(b, 16)) == want, (key, b)


def test_integer_encoder_agrees_with_the_c_oracle_on_fresh_patterns():
    rng = np.random.RandomState(123)
    bits = np.concatenate([rng.randint(0, 2**32, size=150_000, dtype=np.uint64).astype(np.uint32),
                           # dense around the clamp, zero, the denormal range and both infinities
                           np.arange(0x3F7FFF00, 0x3F800100, dtype=np.uint32), np.arange(0xBF7FFF00, 0xBF800100, dtype=np.uint32),
                           np.arange(0, 0x200, dtype=np.uint32), np.arange(0x80000000, 0x80000200, dtype=np.uint32),
                           np.arange(0x007FFF00, 0x00800100, dtype=np.uint32), np.arange(0x7F7FFF00, 0x7F800100, dtype=np.uint32),
                           np.arange(0xFF7FFF00, 0xFF800100, dtype=np.uint32)])
    got = oracle.float_to_16bit_pcm(bits.view(np.float32))
    want = np.array([mg.int_float_to_16bit_pcm_bits(int(b)) for b in bits], dtype=np.int16)
    assert np.array_equal(got, want)


def test_soft_add_is_ieee_addition():
    rng = np.random.RandomState(5)
    xs = np.concatenate([rng.uniform(0, 1e6, 4000), 2.0 ** rng.uniform(-1070, 1000, 4000), [0.0, 5e-324, 2.2250738585072014e-308]])
    ys = np.concatenate([rng.uniform(0, 3, 4000), 2.0 ** rng.uniform(-1070, 1000, 4000), [0.0, 5e-324, 2.2250738585072014e-308]])
    for x, y in zip(xs, ys):
        s = x + y
        if s == float("inf"):
            continue
        xb, yb, sb = (struct.unpack("<Q", struct.pack("<d", v))[0] for v in (x, y, s))
        assert mg.soft_add_f64(xb, yb) == sb, (x, y)


def test_integer_pick_indices_reproduce_the_committed_tables():
    cases = json.load(open(os.path.join(HERE, "golden", "index_kat.json")))["cases"]
    checked = 0
    for c in cases:
        if c.get("identity") or c["n_out"] > 20000:
            continue
        idx = mg.int_pick_indices(c["n_in"], c["rate"])
        assert len(idx) == c["n_out"]
        assert hashlib.sha256(np.asarray(idx, dtype="<u4").tobytes()).hexdigest() == c["sha256_u32le"], c
        assert np.array_equal(np.asarray(idx, dtype=np.uint32), oracle.index_table(c["n_in"], c["rate"]))
        checked += 1
    assert checked >= 80
    # the SURVEY.md F3 pin, from integers alone: at 44.1 kHz output 480 picks input 1322, not 1323
    assert mg.int_pick_indices(2646, 44100)[480] == 1322 and (480 * 44100) // 16000 == 1323


def test_raw_frame_fixtures_match_the_npz():
    raw = json.load(open(os.path.join(HERE, "golden", "frames_kat_raw.json")))["cases"]
    fin = np.fromfile(os.path.join(HERE, "golden", "frames_kat_in.f32"), dtype="<f4")
    fout = np.fromfile(os.path.join(HERE, "golden", "frames_kat_out.s16"), dtype="<i2")
    z = np.load(os.path.join(HERE, "golden", "frames_kat.npz"))
    for c in raw:
        x = fin[c["in_offset"]: c["in_offset"] + c["rows"] * c["n_in"]].reshape(c["rows"], c["n_in"])
        y = fout[c["out_offset"]: c["out_offset"] + c["rows"] * c["n_out"]].reshape(c["rows"], c["n_out"])
        assert np.array_equal(x.view(np.uint32), z[c["key"] + "_in"].view(np.uint32)) and np.array_equal(y, z[c["key"] + "_out"])
