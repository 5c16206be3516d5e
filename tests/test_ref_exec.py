"""The C oracle against outputs of the reference's own source text.

tests/golden/ref_exec.npz was written by tests/golden/make_ref_exec.py, which slices floatTo16BitPCM,
resampleTo16k, the worklet onmessage handler and the keep-alive timer out of the reference's App.tsx at run
time and executes that text with tests/golden/mini_js.py (no JavaScript engine exists in the image).  Three
groups of tests:
  1. the interpreter's own semantics, rule by rule -- the trust placed in the vectors is the trust placed in these;
  2. the C oracle (oracle/pcm_oracle.c) must reproduce every vector;
  3. where the reference tree is present (the build container, not the GPU box) the vectors are regenerated
     from the live text and compared with the committed ones.
"""
import json
import math
import os
import sys

import numpy as np
import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
sys.path.insert(0, GOLDEN)
import mini_js as js  # noqa: E402


@pytest.fixture(scope="module")
def vec():
    return np.load(os.path.join(GOLDEN, "ref_exec.npz")), json.load(open(os.path.join(GOLDEN, "ref_exec.json")))


# ---- 1. interpreter semantics -------------------------------------------------------------------------
def run(src, **env):
    it = js.Interpreter(env)
    it.run(src)
    return it


def same(a, b):
    """Equality that tells -0 from +0 and treats NaN as equal to NaN."""
    if isinstance(a, float) and isinstance(b, float):
        if a != a or b != b:
            return a != a and b != b
        return a == b and math.copysign(1.0, a) == math.copysign(1.0, b)
    return a is b or a == b


@pytest.mark.parametrize("expr, want", [
    ("0x8000", 32768.0), ("0x7fff", 32767.0), ("1 / 0", math.inf), ("-1 / 0", -math.inf), ("0 / 0", math.nan),
    ("1 / (-0)", -math.inf),
    ("0.1 + 0.2", 0.30000000000000004), ("44100 / 16000", 2.75625), ("1e308 * 10", math.inf),
    ("5 % 3", 2.0), ("-5 % 3", -2.0), ("5 % 0", math.nan),
    ("NaN > 1", False), ("NaN < -1", False), ("NaN < 0", False), ("NaN === NaN", False), ("NaN !== NaN", True),
    ("undefined > 1", False), ("undefined * 2", math.nan), ("-0 < 0", False), ("0 === -0", True),
    ("1 < 2 ? 10 : 20", 10.0), ("NaN ? 1 : 2", 2.0), ("0 || 7", 7.0), ("3 && 0", 0.0), ("!0", True),
    ("2 + 3 * 4", 14.0), ("(2 + 3) * 4", 20.0), ("8 - 3 - 2", 3.0), ("2 * 3 / 4", 1.5), ("- 3 * 2", -6.0),
    ("1 < 2 === true", True), ("null == undefined", True), ("null === undefined", False),
    ("Math.floor(2.9)", 2.0), ("Math.floor(-2.1)", -3.0), ("Math.floor(-0)", -0.0), ("Math.floor(NaN)", math.nan),
    ("Math.floor(Infinity)", math.inf), ("Math.floor(-0.5)", -1.0),
    ("Math.sqrt(2)", math.sqrt(2)), ("Math.sqrt(-1)", math.nan), ("Math.sqrt(-0)", -0.0), ("Math.sqrt(Infinity)", math.inf),
    ("Math.max(1, 2)", 2.0), ("Math.max(NaN, 2)", math.nan), ("Math.max(2, NaN)", math.nan), ("Math.max(-0, 0)", 0.0),
    ("Math.max(0, -0)", 0.0), ("Math.max()", -math.inf), ("Math.max(1, undefined)", math.nan), ("Math.min(-0, 0)", -0.0),
])
def test_expression_semantics(expr, want):
    it = run(f"const r = {expr};")
    assert same(it.get("r"), want), (expr, it.get("r"), want)


@pytest.mark.parametrize("value, want", [
    (0.0, 0), (1.9, 1), (-1.9, -1), (32767.0, 32767), (32768.0, -32768), (-32768.0, -32768), (-32769.0, 32767),
    (65536.0, 0), (65535.0, -1), (98304.5, -32768), (math.nan, 0), (math.inf, 0), (-math.inf, 0), (-0.0, 0),
    (-0.99, 0), (1e20, int(math.trunc(1e20)) % 65536 - (65536 if int(math.trunc(1e20)) % 65536 >= 32768 else 0)),
])
def test_int16array_store_is_toint16(value, want):
    a = js.Int16Array(1.0)
    a.js_set(0.0, value)
    assert int(a.a[0]) == want


def test_float32array_store_rounds_to_nearest_even_and_reads_undefined_outside():
    a = js.Float32Array(3.0)
    a.js_set(0.0, 1.0 + 2.0 ** -24)            # exactly half way between 1 and nextafter(1): ties to even -> 1
    a.js_set(1.0, 1.0 + 2.0 ** -24 + 2.0 ** -40)
    a.js_set(2.0, 1e39)                        # overflows binary32
    assert a.a[0] == np.float32(1.0) and a.a[1] == np.nextafter(np.float32(1.0), np.float32(2.0)) and np.isinf(a.a[2])
    a.js_set(7.0, 5.0)                         # outside: dropped
    assert a.a.size == 3
    assert a.js_get(3.0) is js.undefined and a.js_get(-1.0) is js.undefined and a.js_get(0.5) is js.undefined
    a.js_set(0.0, js.undefined)                # ToNumber(undefined) = NaN
    assert np.isnan(a.a[0])
    assert a.js_get("length") == 3.0


def test_typed_array_lengths_follow_toindex():
    assert js.Float32Array(math.nan).a.size == 0
    assert js.Float32Array(2.9).a.size == 2
    assert js.Float32Array(-0.5).a.size == 0
    for bad in (-1.0, math.inf, -math.inf):
        with pytest.raises(js.JSRangeError):
            js.Float32Array(bad)
    big = js.Float32Array(4.0)
    big._set(js.Float32Array(np.array([1, 2], np.float32)), 2.0)
    assert big.a.tolist() == [0, 0, 1, 2]
    with pytest.raises(js.JSRangeError):
        big._set(js.Float32Array(2.0), 3.0)


def test_statements_closures_and_typescript_annotations():
    it = run("""
        function f(a: Float32Array, n: number): Float32Array {
          let s = 0;
          for (let i = 0; i < n; i++) { if (i === 2) s += 10; else if (i > 3) s -= 1; else s += a[i]; }
          return s;
        }
        const arr = new Float32Array(6); arr[0] = 0.5; arr[1] = 0.25; arr[3] = 4;
        const r = f(arr, 6);
        const box = { v: 1, w: 2 };
        box.v = box.w * 3;
        let k = 5; const post = k++; const pre = ++k;
        const g = (e: Wrapper<Float32Array>) => { const d = e.data; return d.length; };
        const h = (x) => x * 2;
        const r2 = g({ data: arr }) + h(4);
    """)
    assert it.get("r") == 0.5 + 0.25 + 10 + 4 - 2
    assert it.get("box").js_get("v") == 6.0 and it.get("post") == 5.0 and it.get("pre") == 7.0 and it.get("r2") == 14.0
    with pytest.raises(js.JSTypeError):
        run("const c = 1; c = 2;")
    with pytest.raises(js.JSTypeError):
        run("const o = null; o.x;")


# ---- 2. the C oracle reproduces the reference text's outputs ----------------------------------------------
def test_oracle_encoder_vs_reference_text(vec):
    z, _ = vec
    assert np.array_equal(oracle.float_to_16bit_pcm(z["enc_bits"].view(np.float32)), z["enc_out"])
    assert np.array_equal(oracle.float_to_16bit_pcm(z["enc_audio_in"]), z["enc_audio_out"])
    assert len(np.unique(z["enc_audio_out"])) > 10000           # not a degenerate vector


def test_oracle_frames_vs_reference_text(vec):
    z, rec = vec
    for c in rec["frame_cases"]:
        x = z[c["key"] + "_in"].reshape(1, -1)
        got = oracle.frames(x, c["n_in"], c["rate"])[0]
        assert got.size == c["n_out"] and np.array_equal(got, z[c["key"] + "_out"]), c


def test_oracle_sessions_vs_reference_text(vec):
    """The onmessage handler (App.tsx:264-289): flush decisions, frames and the level after every quantum."""
    z, rec = vec
    for s in rec["sessions"]:
        k = s["key"]
        q, t = z[k + "_in"], z[k + "_arrival"]
        flush, last = oracle.aggregator_plan(t, s["desired_ms"], s["t0"])
        assert np.array_equal(flush, z[k + "_flush"]), k
        assert int(flush.sum()) == s["flushes"] > 0
        assert last == t[np.nonzero(flush)[0][-1]]
        lens, blob, start, off = z[k + "_frame_len"], z[k + "_frames"], 0, 0
        for f, qi in enumerate(np.nonzero(flush)[0]):
            buf = q[start:qi + 1].reshape(1, -1)
            got = oracle.frames(buf, buf.shape[1], s["rate"])[0]
            assert got.size == lens[f] and np.array_equal(got, blob[off:off + lens[f]]), (k, f)
            start, off = qi + 1, off + int(lens[f])
        assert off == blob.size
        lv = oracle.level_meter(q.reshape(-1))
        assert np.array_equal(lv.view(np.uint64), z[k + "_levels"].view(np.uint64)), k
    clean = z[rec["sessions"][-1]["key"] + "_levels"]
    assert np.isfinite(clean).all() and clean.max() > 0.3
    assert np.isnan(z[rec["sessions"][0]["key"] + "_levels"][-1])          # Math.max keeps the NaN (App.tsx:271)


def test_oracle_keepalive_vs_reference_text(vec):
    _, rec = vec
    ka = rec["keepalive"]
    frame = oracle.keepalive_frame()
    assert ka["samples"] == frame.size == 1600 and ka["all_zero"] and not frame.any() and ka["bytes"] == 3200
    assert ka["period_ms"] == 2000.0 and not ka["sent_when_closed"] and not ka["sent_without_socket"]
    # the oracle's timer fires on the same grid: polls at 1999 / 2000 / 3999 / 4000 ms after start
    sent = oracle.keepalive_timer(0.0, np.array([1999.0, 2000.0, 3999.0, 4000.0]), np.ones(4, np.uint8), ka["period_ms"])
    assert sent.tolist() == [0, 1, 0, 1]


# ---- 3. regenerate from the live reference text (build container only) ---------------------------------------
REF_ROOT = os.environ.get("PCMB200_REFERENCE_ROOT", "/root/reference")
APP = os.path.join(REF_ROOT, "apps", "web", "src", "App.tsx")


@pytest.mark.skipif(not os.path.exists(APP), reason="the reference tree is not on this machine")
def test_committed_vectors_are_what_the_reference_text_produces(vec):
    import make_ref_exec as gen

    z, rec = vec
    slices = gen.slice_reference(APP)
    for name, s in slices.items():
        assert s["sha256"] == rec["slices"][name]["sha256"], f"{name}: App.tsx changed since the vectors were made"
        assert s["lines"] == rec["slices"][name]["lines"]
    ref = gen.Reference(slices)
    assert np.array_equal(ref.encode(z["enc_bits"][:2048].view(np.float32)), z["enc_out"][:2048])
    c = rec["frame_cases"][1]
    assert np.array_equal(ref.frame(z[c["key"] + "_in"], c["rate"]), z[c["key"] + "_out"])
    s = rec["sessions"][1]
    flush, frames, levels = ref.session(z[s["key"] + "_in"][:40], z[s["key"] + "_arrival"][:40], s["rate"], s["t0"])
    assert np.array_equal(flush, z[s["key"] + "_flush"][:40])
    assert np.array_equal(levels.view(np.uint64), z[s["key"] + "_levels"][:40].view(np.uint64))
    assert np.array_equal(np.concatenate(frames), z[s["key"] + "_frames"][:sum(f.size for f in frames)])


# ---- 1b. differential test of the expression grammar: random trees, printed with MINIMAL parentheses (so that the
# interpreter's precedence and associativity decide the meaning) against a direct evaluation of the tree with numpy
# binary64 scalars (whose division, remainder and comparisons follow IEEE-754 without Python's exceptions) -----------
_PREC = {"?:": 1, "||": 2, "&&": 3, "===": 4, "!==": 4, "<": 5, ">": 5, "<=": 5, ">=": 5, "+": 6, "-": 6, "*": 7, "/": 7, "%": 7, "neg": 8, "!": 8}


def _gen_num(rng, depth):
    if depth <= 0 or rng.rand() < 0.25:
        kind = rng.randint(6)
        if kind == 0:
            return ("lit", str(int(rng.randint(0, 1000))))
        if kind == 1:
            return ("lit", f"{rng.uniform(0, 50):.6f}")
        if kind == 2:
            return ("lit", hex(int(rng.randint(0, 65536))))
        if kind == 3:
            return ("lit", rng.choice(["NaN", "Infinity", "0", "1e308", "5e-324", "0.5"]))
        return ("lit", str(int(rng.randint(0, 10))))
    kind = rng.randint(10)
    if kind < 6:
        return ("bin", rng.choice(["+", "-", "*", "/", "%"]), _gen_num(rng, depth - 1), _gen_num(rng, depth - 1))
    if kind == 6:
        return ("neg", _gen_num(rng, depth - 1))
    if kind == 7:
        return ("cond", _gen_bool(rng, depth - 1), _gen_num(rng, depth - 1), _gen_num(rng, depth - 1))
    return ("call", rng.choice(["Math.floor", "Math.sqrt", "Math.max", "Math.min"]), _gen_num(rng, depth - 1), _gen_num(rng, depth - 1))


def _gen_bool(rng, depth):
    kind = rng.randint(6)
    if depth <= 0 or kind < 3:
        return ("cmp", rng.choice(["<", ">", "<=", ">=", "===", "!=="]), _gen_num(rng, depth - 1), _gen_num(rng, depth - 1))
    if kind == 3:
        return ("not", _gen_bool(rng, depth - 1))
    return ("logic", rng.choice(["&&", "||"]), _gen_bool(rng, depth - 1), _gen_bool(rng, depth - 1))


def _to_js(t, parent=0, right=False):
    k = t[0]
    if k == "lit":
        return t[1]
    if k == "call":
        args = [_to_js(t[2])] if t[1] in ("Math.floor", "Math.sqrt") else [_to_js(t[2]), _to_js(t[3])]
        return f"{t[1]}({', '.join(args)})"
    if k == "neg":
        inner = _to_js(t[1], _PREC["neg"])
        s, p = ("- " + inner), _PREC["neg"]
    elif k == "not":
        s, p = "!" + _to_js(t[1], _PREC["!"]), _PREC["!"]
    elif k == "cond":
        p = _PREC["?:"]
        s = f"{_to_js(t[1], p + 1)} ? {_to_js(t[2], p)} : {_to_js(t[3], p)}"
    else:
        op = t[1]
        p = _PREC[op]
        s = f"{_to_js(t[2], p)} {op} {_to_js(t[3], p, right=True)}"          # left-associative: the right child needs > p
    need = p < parent or (p == parent and right) or (k == "cond" and parent >= p and parent != 0)
    return f"({s})" if need else s


def _eval_tree(t):
    k = t[0]
    with np.errstate(all="ignore"):
        if k == "lit":
            return np.float64(int(t[1], 16)) if t[1].startswith("0x") else np.float64({"NaN": "nan", "Infinity": "inf"}.get(t[1], t[1]))
        if k == "neg":
            return -_eval_tree(t[1])
        if k == "not":
            return not _eval_tree(t[1])
        if k == "cond":
            return _eval_tree(t[2]) if _eval_tree(t[1]) else _eval_tree(t[3])
        if k == "logic":
            a = _eval_tree(t[2])
            return (a and _eval_tree(t[3])) if t[1] == "&&" else (a or _eval_tree(t[3]))
        if k == "call":
            a = _eval_tree(t[2])
            if t[1] == "Math.floor":
                return np.floor(a)
            if t[1] == "Math.sqrt":
                return np.sqrt(a)
            b = _eval_tree(t[3])
            if np.isnan(a) or np.isnan(b):
                return np.float64("nan")
            return np.maximum(a, b) if t[1] == "Math.max" else np.minimum(a, b)
        a, b = _eval_tree(t[2]), _eval_tree(t[3])
        if k == "cmp":
            return bool({"<": a < b, ">": a > b, "<=": a <= b, ">=": a >= b, "===": a == b, "!==": a != b}[t[1]])
        return {"+": a + b, "-": a - b, "*": a * b, "/": np.divide(a, b), "%": np.fmod(a, b)}[t[1]]


@pytest.mark.parametrize("seed", range(8))
def test_expression_grammar_against_direct_evaluation(seed):
    rng = np.random.RandomState(500 + seed)
    for _ in range(150):
        tree = _gen_num(rng, 5)
        src = _to_js(tree)
        got = run(f"const r = {src};").get("r")
        want = float(_eval_tree(tree))
        ok = (got != got and want != want) or got == want           # (Math.max / fmod sign-of-zero corners are covered above)
        assert ok, (src, got, want)
