#!/usr/bin/env python3
"""Generates the committed golden fixtures in tests/golden/.

The reference (TypeScript, browser) cannot be executed in the build container: there is
no JavaScript engine, so these vectors are NOT outputs of the reference itself ("parity
unpinned").  They are produced by a second, independent restatement of the reference's
semantics in pure-Python floats -- Python's float is IEEE-754 binary64 exactly like a JS
Number, `math.floor`/`math.trunc` match Math.floor/ToInt16 truncation -- written from
    /root/reference/apps/web/src/App.tsx:7-16   (floatTo16BitPCM)
    /root/reference/apps/web/src/App.tsx:18-32  (resampleTo16k)
It shares no code with oracle/pcm_oracle.c (C) or the CUDA kernels; the tests require all
three to agree.  Run:  python tests/golden/make_golden.py   (rewrites the fixtures).
Since then tests/golden/make_ref_exec.py has executed the reference's own source text (with
tests/golden/mini_js.py) on every input held here and found every expectation below reproduced.
"""
import hashlib
import json
import math
import os
import struct

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def f32(bits: int) -> float:
    return struct.unpack("<f", struct.pack("<I", bits & 0xFFFFFFFF))[0]


def bits_of(x: float) -> int:
    return struct.unpack("<I", struct.pack("<f", x))[0]


def to_int16(v: float) -> int:
    """ECMA-262 ToInt16."""
    if v != v or v in (math.inf, -math.inf):
        return 0
    i = math.trunc(v)
    i %= 65536
    return i - 65536 if i >= 32768 else i


def js_float_to_16bit_pcm(samples):
    out = []
    for s in samples:            # s is a Python float holding an exact float32 value
        if s > 1:
            s = 1
        elif s < -1:
            s = -1
        out.append(to_int16(s * 0x8000 if s < 0 else s * 0x7FFF))
    return out


def js_resample_len(n, rate):
    if rate == 16000:
        return n
    ratio = rate / 16000
    return math.floor(n / ratio)


def js_pick_indices(n, rate):
    ratio = rate / 16000
    new_len = math.floor(n / ratio)
    idx, pos = [], 0.0
    for _ in range(new_len):
        idx.append(math.floor(pos))
        pos += ratio
    return idx


def js_resample_to_16k(samples, rate):
    if rate == 16000:
        return samples
    return [samples[i] if i < len(samples) else math.nan for i in js_pick_indices(len(samples), rate)]


# ---- 0. a THIRD derivation, integers only (no hardware floating point anywhere) ----------------
# js_float_to_16bit_pcm above and oracle/pcm_oracle.c both lean on the machine's IEEE-754 unit.  These
# two functions redo the same semantics with Python integers: the float32 is taken apart into sign,
# exponent and significand, the product with 0x7fff / 0x8000 is an exact integer times a power of two,
# and ToInt16's truncation is a shift.  The binary64 running sum of resampleTo16k is redone with a
# software round-to-nearest-even addition.  tests/test_golden_integer.py checks that all three agree
# on the committed fixtures.
def int_float_to_16bit_pcm_bits(bits: int) -> int:
    """floatTo16BitPCM (App.tsx:7-16) of one float32 given as its bit pattern, integers only."""
    sign, e, frac = bits >> 31, (bits >> 23) & 0xFF, bits & 0x7FFFFF
    if e == 0xFF:
        if frac:                       # NaN: both comparisons false, NaN * 0x7fff = NaN, ToInt16(NaN) = 0
            return 0
        return -32768 if sign else 32767           # +-Inf clamps to +-1
    if e == 0:
        m, ex = frac, -149                          # denormal: frac * 2^-149
    else:
        m, ex = frac | 0x800000, e - 150           # (1.frac) * 2^(e-127) = m * 2^(e-150)
    # |s| > 1  <=>  m * 2^ex > 1: clamp (App.tsx:11-12)
    if ex >= 0 or m > (1 << -ex):
        return -32768 if sign else 32767
    c = 0x8000 if (sign and m != 0) else 0x7FFF     # `s < 0` is false for -0 (App.tsx:13)
    p = m * c                                       # exact: |v| = p * 2^ex with ex < 0
    t = p >> -ex                                    # truncate toward zero (ToInt16 of a finite |v| <= 32768)
    return -t if sign else t


def soft_add_f64(a_bits: int, b_bits: int) -> int:
    """IEEE-754 binary64 addition of two non-negative finite doubles, round to nearest even, integers only."""
    def unpack(x):
        e, f = (x >> 52) & 0x7FF, x & ((1 << 52) - 1)
        return (f, -1074) if e == 0 else (f | (1 << 52), e - 1075)
    (ma, ea), (mb, eb) = unpack(a_bits), unpack(b_bits)
    e = min(ea, eb)
    m = (ma << (ea - e)) + (mb << (eb - e))         # exact sum = m * 2^e
    if m == 0:
        return 0
    shift = max(m.bit_length() - 53, -1074 - e if e < -1074 else 0)
    if shift > 0:
        rem, half = m & ((1 << shift) - 1), 1 << (shift - 1)
        m >>= shift
        e += shift
        if rem > half or (rem == half and (m & 1)):
            m += 1
            if m.bit_length() > 53:
                m >>= 1
                e += 1
    while m.bit_length() < 53 and e > -1074:        # normalise (exact)
        m <<= 1
        e -= 1
    if m.bit_length() < 53:                          # subnormal result
        return m
    return ((e + 1075) << 52) | (m & ((1 << 52) - 1))


def int_pick_indices(n: int, rate: int):
    """resampleTo16k's pick indices (App.tsx:23-30) with the running sum done by soft_add_f64.  The
    two divisions (rate / 16000 and n / ratio) are single correctly-rounded operations that Python's
    integer-to-float true division reproduces; what accumulates -- the sum -- is done in integers."""
    ratio = rate / 16000
    new_len = math.floor(n / ratio)
    rb = struct.unpack("<Q", struct.pack("<d", ratio))[0]
    pos, idx = 0, []
    for _ in range(new_len):
        e, f = (pos >> 52) & 0x7FF, pos & ((1 << 52) - 1)
        m, ex = (f, -1074) if e == 0 else (f | (1 << 52), e - 1075)
        idx.append(m << ex if ex >= 0 else m >> -ex)           # Math.floor of a non-negative double
        pos = soft_add_f64(pos, rb)
    return idx


# ---- 1. hand-stated value table + computed hazard cases ---------------
VALUE_TABLE = [  # (description, float32 bit pattern, expected int16) -- expectations written by hand
    ("1.0", 0x3F800000, 32767), ("-1.0", 0xBF800000, -32768),
    ("0.5", 0x3F000000, 16383), ("-0.5", 0xBF000000, -16384),
    ("1.5 clamps", 0x3FC00000, 32767), ("-1.5 clamps", 0xBFC00000, -32768),
    ("+Inf", 0x7F800000, 32767), ("-Inf", 0xFF800000, -32768),
    ("NaN", 0x7FC00000, 0), ("-NaN", 0xFFC00000, 0), ("sNaN payload", 0x7F800001, 0),
    ("+0", 0x00000000, 0), ("-0", 0x80000000, 0),
    ("+denormal 1e-40", bits_of(1e-40), 0), ("-denormal 1e-40", bits_of(-1e-40), 0),
    ("min denormal", 0x00000001, 0), ("-min denormal", 0x80000001, 0),
    ("nextbelow(1)", 0x3F7FFFFF, 32766), ("-nextbelow(1)", 0xBF7FFFFF, -32767),
    ("nextabove(1)", 0x3F800001, 32767), ("-nextabove(1)", 0xBF800001, -32768),
    ("2^-15", 0x38000000, 0), ("-2^-15", 0xB8000000, -1),
    ("FLT_MAX", 0x7F7FFFFF, 32767), ("-FLT_MAX", 0xFF7FFFFF, -32768),
    ("1/32767 rounded", bits_of(1 / 32767), None), ("0.25", 0x3E800000, 8191), ("-0.25", 0xBE800000, -8192),
]


def make_encode_kat():
    rows = []
    for name, b, want in VALUE_TABLE:
        got = js_float_to_16bit_pcm([f32(b)])[0]
        if want is not None:
            assert got == want, (name, got, want)
        rows.append({"name": name, "bits": f"{b:08x}", "int16": got})
    # hazard set: float32 values around k/32767 where an fp32 product rounds to an integer
    hazard = []
    for k in list(range(1, 257)) + list(range(32767 - 256, 32768)) + list(range(1000, 32767, 997)):
        centre = bits_of(k / 32767)
        for d in range(-4, 5):
            b = centre + d
            hazard.append({"bits": f"{b:08x}", "int16": js_float_to_16bit_pcm([f32(b)])[0]})
            nb = b | 0x80000000
            hazard.append({"bits": f"{nb:08x}", "int16": js_float_to_16bit_pcm([f32(nb)])[0]})
    rng = np.random.RandomState(0x5EED)
    rnd_bits = rng.randint(0, 2**32, size=4096, dtype=np.uint64).astype(np.uint32)
    rnd = [{"bits": f"{int(b):08x}", "int16": js_float_to_16bit_pcm([f32(int(b))])[0]} for b in rnd_bits]
    def cols(items):  # column form keeps the fixture small
        return {"bits": [i["bits"] for i in items], "int16": [i["int16"] for i in items]}

    with open(os.path.join(HERE, "encode_kat.json"), "w") as fh:
        json.dump({"source": "App.tsx:7-16 restated in Python floats (make_golden.py)",
                   "value_table": rows, "hazard": cols(hazard), "random_bits": cols(rnd)}, fh,
                  separators=(",", ":"))


# ---- 2. pick-index tables ---------------------------------------------------------------
RATES = [8000, 11025, 16000, 22050, 24000, 32000, 44100, 48000, 88200, 96000, 192000]


def make_index_kat():
    cases = []
    for rate in RATES:
        lens = sorted({rate * ms // 1000 for ms in (10, 20, 40, 60, 100, 1000)} | {128 * k for k in (1, 7, 23, 47)})
        for n in lens:
            if rate == 16000:
                cases.append({"rate": rate, "n_in": n, "n_out": n, "identity": True})
                continue
            idx = js_pick_indices(n, rate)
            exact = [(i * rate) // 16000 for i in range(len(idx))]
            mism = [i for i, (a, b) in enumerate(zip(idx, exact)) if a != b]
            h = hashlib.sha256(np.asarray(idx, dtype="<u4").tobytes()).hexdigest()
            cases.append({"rate": rate, "n_in": n, "n_out": len(idx), "sha256_u32le": h,
                          "first": idx[:8], "last": idx[-4:] if idx else [],
                          "n_differs_from_rational": len(mism),
                          "first_differs_at": mism[0] if mism else None,
                          "pick_at_first_differs": idx[mism[0]] if mism else None})
    # the pins quoted in SURVEY.md F3
    one_s = next(c for c in cases if c["rate"] == 44100 and c["n_in"] == 44100)
    assert one_s["first_differs_at"] == 480 and one_s["pick_at_first_differs"] == 1322, one_s
    assert one_s["n_differs_from_rational"] == 64, one_s
    with open(os.path.join(HERE, "index_kat.json"), "w") as fh:
        json.dump({"source": "App.tsx:23-30 restated in Python floats (make_golden.py)", "cases": cases}, fh, separators=(",", ":"))


# ---- 3. whole frames: resample + encode on seeded audio -------------------------------------
def make_frame_kat():
    rng = np.random.RandomState(20261106)
    blobs = {}
    meta = []
    for rate, n_in, rows in [(48000, 960, 3), (44100, 882, 3), (48000, 2944, 2), (8000, 160, 2),
                             (16000, 320, 2), (22050, 441, 2), (96000, 1920, 1), (32000, 640, 1),
                             (192000, 3840, 1), (11025, 1103, 1), (48000, 0, 1), (44100, 1, 1)]:
        t = np.arange(max(n_in, 1))[None, :] / rate
        x = (0.9 * np.sin(2 * np.pi * (110 + 7 * np.arange(rows))[:, None] * t)
             + 0.4 * rng.uniform(-1, 1, size=(rows, max(n_in, 1)))).astype(np.float32)[:, :n_in]
        if n_in >= 16:  # sprinkle special values
            x[0, 3] = np.float32(np.nan); x[0, 6] = np.float32(np.inf); x[0, 9] = np.float32(-np.inf)
            x[0, 12] = np.float32(-0.0); x[0, 15] = np.float32(1e-40)
        out = []
        for r in range(rows):
            res = js_resample_to_16k([float(v) for v in x[r]], rate)
            out.append(js_float_to_16bit_pcm(res))
        key = f"r{rate}_n{n_in}"
        blobs[key + "_in"] = x
        blobs[key + "_out"] = np.asarray(out, dtype=np.int16).reshape(rows, -1)
        meta.append({"key": key, "rate": rate, "n_in": n_in, "rows": rows, "n_out": js_resample_len(n_in, rate)})
    blobs["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "frames_kat.npz"), **blobs)
    write_frames_raw(meta, blobs)


def write_frames_raw(meta, blobs):
    """The same frame fixtures as flat little-endian files, for tools/pin_with_node.mjs (Node has no npz
    reader): frames_kat_in.f32 / frames_kat_out.s16 hold the rows back to back in `meta` order and
    frames_kat_raw.json says where each case starts."""
    index, fin, fout = [], bytearray(), bytearray()
    for m in meta:
        x, y = blobs[m["key"] + "_in"], blobs[m["key"] + "_out"]
        index.append(dict(m, in_offset=len(fin) // 4, out_offset=len(fout) // 2))
        fin += np.ascontiguousarray(x, dtype="<f4").tobytes()
        fout += np.ascontiguousarray(y, dtype="<i2").tobytes()
    open(os.path.join(HERE, "frames_kat_in.f32"), "wb").write(bytes(fin))
    open(os.path.join(HERE, "frames_kat_out.s16"), "wb").write(bytes(fout))
    json.dump({"cases": index}, open(os.path.join(HERE, "frames_kat_raw.json"), "w"), separators=(",", ":"))


if __name__ == "__main__":
    make_encode_kat()
    make_index_kat()
    make_frame_kat()
    print("golden fixtures written to", HERE)
