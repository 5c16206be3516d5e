#!/usr/bin/env python3
"""Runs the REFERENCE'S OWN SOURCE TEXT on the committed input vectors and writes what it produced.

    python tests/golden/make_ref_exec.py [/path/to/reference]      (default /root/reference)

No JavaScript engine exists in the build image, so the reference (apps/web/src/App.tsx, TypeScript)
was never executed by Node here.  This script does the next best thing that is possible in the
image: it slices the four pieces of the audio path out of App.tsx AT RUN TIME (located by their
first line, extent found by bracket matching; nothing of the reference is committed) and executes
that text with tests/golden/mini_js.py, a small interpreter for the JavaScript subset those pieces use:

    floatTo16BitPCM        App.tsx:7-16
    resampleTo16k          App.tsx:18-32
    worklet.port.onmessage App.tsx:264-289   (level meter, aggregator, flush; host objects injected)
    keep-alive timer       App.tsx:292-296   (window.setInterval captured)

Then it (1) checks that the reference text reproduces every expectation already committed under
tests/golden/ (encode_kat.json, index_kat.json, frames_kat.npz -- until now outputs of restatements),
and (2) writes fresh vectors, outputs of the reference text itself, to ref_exec.npz / ref_exec.json.
tests/test_ref_exec.py holds the C oracle to them; tests/test_gpu_ref_exec.py holds the CUDA path to them.

What this pins and what it does not: the program that ran is the reference's; the language semantics
(binary64 arithmetic, typed-array conversions, Math.*) are mini_js.py's implementation of ECMA-262,
unit-tested in tests/test_ref_exec.py.  `node tools/pin_with_node.mjs` remains the stronger check.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import mini_js as js  # noqa: E402

ANCHORS = {  # name -> (first-line anchor, line range SURVEY.md cites, how the slice ends)
    "floatTo16BitPCM": ("function floatTo16BitPCM(", (7, 16), "}"),
    "resampleTo16k": ("function resampleTo16k(", (18, 32), "}"),
    "onmessage": ("worklet.port.onmessage = (", (264, 289), ";"),
    "keepalive": ("keepaliveTimer.current = window.setInterval(", (292, 296), ";"),
}


def slice_reference(app_tsx: str) -> dict:
    """Cuts the four statements out of App.tsx: from the anchor line to the bracket that closes it."""
    text = open(app_tsx, encoding="utf-8").read()
    lines = text.split("\n")
    out = {}
    for name, (anchor, cited, end) in ANCHORS.items():
        hits = [k for k, ln in enumerate(lines) if ln.strip().startswith(anchor)]
        if len(hits) != 1:
            raise SystemExit(f"{name}: anchor {anchor!r} found {len(hits)} times in {app_tsx}")
        start = sum(len(ln) + 1 for ln in lines[:hits[0]])
        depth, pos, seen_open = 0, start, False
        in_str = None
        while pos < len(text):
            c = text[pos]
            if in_str:
                if c == "\\":
                    pos += 1
                elif c == in_str:
                    in_str = None
            elif c in "\"'`":
                in_str = c
            elif c == "/" and text[pos + 1] == "/":
                pos = text.index("\n", pos)
                continue
            elif c in "([{":
                depth += 1
                seen_open = True
            elif c in ")]}":
                depth -= 1
                if depth == 0 and seen_open and end == "}" and c == "}":
                    pos += 1
                    break
            elif c == ";" and depth == 0 and end == ";":
                pos += 1
                break
            pos += 1
        src = text[start:pos]
        first = hits[0] + 1
        last = first + src.count("\n")
        out[name] = {"src": src, "lines": [first, last], "cited": list(cited),
                     "sha256": hashlib.sha256(src.encode()).hexdigest()}
    return out


class Reference:
    """The reference's audio path, executed from its text."""

    def __init__(self, slices: dict):
        self.slices = slices
        self.sent = []            # what wsRef.current.send(...) received: bytes objects
        self.clock = 0.0
        self.level = 0.0          # React state behind setLevel (App.tsx:141: useState(0))
        self.intervals = []
        self.ws = js.JSObject(readyState=1.0, send=self._send)
        self.env = {
            "worklet": js.JSObject(port=js.JSObject()),
            # App.tsx:149-157: { buf: new Float32Array(0), lastSend: performance.now(), desiredMs: 60 }
            "aggRef": js.JSObject(current=js.JSObject(buf=js.Float32Array(0.0), lastSend=0.0, desiredMs=60.0)),
            "ctx": js.JSObject(sampleRate=48000.0),
            "wsRef": js.JSObject(current=self.ws),
            "WebSocket": js.JSObject(CONNECTING=0.0, OPEN=1.0, CLOSING=2.0, CLOSED=3.0),
            "performance": js.JSObject(now=lambda: self.clock),
            "setLevel": self._set_level,
            "keepaliveTimer": js.JSObject(current=None),
            "window": js.JSObject(setInterval=self._set_interval),
        }
        self.interp = js.Interpreter(self.env)
        for name in ("floatTo16BitPCM", "resampleTo16k", "onmessage", "keepalive"):
            self.interp.run(slices[name]["src"])
        self.f_encode = self.interp.get("floatTo16BitPCM")
        self.f_resample = self.interp.get("resampleTo16k")
        self.on_message = self.env["worklet"].js_get("port").js_get("onmessage")

    # host objects ---------------------------------------------------------------------------
    def _send(self, buf):
        self.sent.append(buf.tobytes())
        return js.undefined

    def _set_level(self, updater):
        # React's functional update: setLevel(prev => next)
        self.level = updater(self.level) if callable(updater) else updater
        return js.undefined

    def _set_interval(self, fn, period):
        self.intervals.append((fn, period))
        return float(len(self.intervals))

    # the four entry points ---------------------------------------------------------------------
    def encode(self, x: np.ndarray) -> np.ndarray:
        return self.f_encode(js.Float32Array(np.asarray(x, dtype=np.float32))).a.copy()

    def resample(self, x: np.ndarray, rate: float) -> np.ndarray:
        return self.f_resample(js.Float32Array(np.asarray(x, dtype=np.float32)), float(rate)).a.copy()

    def frame(self, x: np.ndarray, rate: float) -> np.ndarray:
        """App.tsx:281-282."""
        return self.f_encode(self.f_resample(js.Float32Array(np.asarray(x, dtype=np.float32)), float(rate))).a.copy()

    def session(self, quanta: np.ndarray, arrival_ms: np.ndarray, rate: float, t0: float, desired_ms: float = 60.0):
        """Feeds one capture session through the onmessage handler, one quantum per message, with
        performance.now() returning arrival_ms[i] during message i.  Returns (flush flags, frames, levels)."""
        agg = self.env["aggRef"].js_get("current")
        agg.js_set("buf", js.Float32Array(0.0))
        agg.js_set("lastSend", float(t0))
        agg.js_set("desiredMs", float(desired_ms))
        self.env["ctx"].js_set("sampleRate", float(rate))
        self.level = 0.0
        self.sent = []
        flush, levels = [], []
        for q, t in zip(quanta, arrival_ms):
            self.clock = float(t)
            before = len(self.sent)
            self.on_message(js.JSObject(data=js.Float32Array(np.asarray(q, dtype=np.float32))))
            flush.append(len(self.sent) > before)
            levels.append(self.level)
        frames = [np.frombuffer(b, dtype="<i2").copy() for b in self.sent]
        return np.asarray(flush, dtype=np.uint8), frames, np.asarray(levels, dtype=np.float64)


def bits_to_f32(hex_list):
    return np.asarray([int(h, 16) for h in hex_list], dtype=np.uint32).view(np.float32)


def special_audio(rng, rows, n, rate):
    t = np.arange(max(n, 1))[None, :] / rate
    x = (0.8 * np.sin(2 * np.pi * (180 + 11 * np.arange(rows))[:, None] * t) + 0.5 * rng.uniform(-1, 1, size=(rows, max(n, 1))))
    return x.astype(np.float32)[:, :n]


def main():
    ref_root = sys.argv[1] if len(sys.argv) > 1 else os.environ.get("PCMB200_REFERENCE_ROOT", "/root/reference")
    app = os.path.join(ref_root, "apps", "web", "src", "App.tsx")
    slices = slice_reference(app)
    ref = Reference(slices)
    record = {"what": "outputs of the reference's source text (App.tsx) executed by tests/golden/mini_js.py",
              "app_tsx_sha256": hashlib.sha256(open(app, "rb").read()).hexdigest(),
              "slices": {k: {kk: v[kk] for kk in ("lines", "cited", "sha256")} for k, v in slices.items()},
              "checked_against_committed": {}}
    for name, s in slices.items():
        assert s["lines"] == s["cited"], (name, s["lines"], s["cited"])

    # ---- 1. the committed expectations (until now outputs of restatements) ----------------------------
    kat = json.load(open(os.path.join(HERE, "encode_kat.json")))
    n_enc = 0
    vt_bits = [r["bits"] for r in kat["value_table"]]
    got = ref.encode(bits_to_f32(vt_bits))
    assert got.tolist() == [r["int16"] for r in kat["value_table"]], "value table"
    n_enc += len(vt_bits)
    for key in ("hazard", "random_bits"):
        got = ref.encode(bits_to_f32(kat[key]["bits"]))
        assert got.tolist() == kat[key]["int16"], key
        n_enc += len(kat[key]["bits"])
    record["checked_against_committed"]["encode_kat.json"] = {"samples": n_enc, "equal": True}

    idx = json.load(open(os.path.join(HERE, "index_kat.json")))
    n_idx = 0
    for c in idx["cases"]:
        ramp = np.arange(c["n_in"], dtype=np.float32)          # exact below 2^24: the output IS the pick index
        assert c["n_in"] < (1 << 24)
        out = ref.resample(ramp, c["rate"])
        assert out.size == c["n_out"], c
        if c.get("identity"):
            assert np.array_equal(out, ramp)
            continue
        picks = out.astype(np.uint32)
        assert np.array_equal(picks.astype(np.float32), out)
        assert hashlib.sha256(picks.astype("<u4").tobytes()).hexdigest() == c["sha256_u32le"], c
        n_idx += picks.size
    record["checked_against_committed"]["index_kat.json"] = {"cases": len(idx["cases"]), "picks": n_idx, "equal": True}

    z = np.load(os.path.join(HERE, "frames_kat.npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    n_fr = 0
    for m in meta:
        x, want = z[m["key"] + "_in"], z[m["key"] + "_out"]
        for r in range(m["rows"]):
            got = ref.frame(x[r], m["rate"])
            assert np.array_equal(got, want[r]), m
            n_fr += 1
    record["checked_against_committed"]["frames_kat.npz"] = {"frames": n_fr, "equal": True}

    # ---- 2. fresh vectors: outputs of the reference text ---------------------------------------------------
    blobs = {}
    rng = np.random.RandomState(20261107)
    enc_bits = rng.randint(0, 2 ** 32, size=16384, dtype=np.uint64).astype(np.uint32)
    enc_bits[:8] = [0x7FC00000, 0xFFC00000, 0x7F800000, 0xFF800000, 0x80000000, 0x00000001, 0x3F800000, 0xBF800000]
    blobs["enc_bits"] = enc_bits
    blobs["enc_out"] = ref.encode(enc_bits.view(np.float32))
    # audio-like values (|x| <= 1.3) so the non-clamping branch is the common one
    enc_audio = (rng.uniform(-1.3, 1.3, size=16384)).astype(np.float32)
    blobs["enc_audio_in"] = enc_audio
    blobs["enc_audio_out"] = ref.encode(enc_audio)

    # one-shot frames (resample + encode) at rates the committed cases do not hold
    frame_cases = []
    for rate, n_in in [(48000, 2880), (44100, 2646), (44100, 2944), (24000, 480), (88200, 1764), (12000, 240), (48000, 127), (37800, 756)]:
        x = special_audio(rng, 1, n_in, rate)[0]
        x[5 % max(n_in, 1)] = np.float32(np.nan)
        key = f"frame_r{rate}_n{n_in}"
        blobs[key + "_in"] = x
        blobs[key + "_out"] = ref.frame(x, rate)
        frame_cases.append({"key": key, "rate": rate, "n_in": n_in, "n_out": int(blobs[key + "_out"].size)})
    record["frame_cases"] = frame_cases

    # capture sessions through the onmessage handler: meter + aggregator + flush
    sessions = []
    for rate, nq, jitter, clean in [(48000, 160, 0.0, False), (44100, 150, 0.4, False), (16000, 96, 0.0, False),
                                    (96000, 200, 0.2, False), (48000, 120, 3.0, True)]:
        q = special_audio(rng, nq, 128, rate)
        q[40, :] = 0.0                        # silence
        q[41, :] = np.float32(1e-30)          # tiny but normal squares
        if nq > 100:
            q[100, :] *= np.float32(3.0)      # clamps
        if not clean:                         # Math.max keeps a NaN level for good (App.tsx:271): the last
            q[60, 7] = np.float32(np.nan)     # session stays clean so that a full level record exists too
            q[42, 0] = np.float32(np.inf)
        period = 128.0 / rate * 1000.0
        t0 = 1000.0 + 7.0 * len(sessions)
        arrival = np.sort(t0 + period * np.arange(1, nq + 1) + jitter * rng.uniform(0, 1, size=nq))
        flush, frames, levels = ref.session(q, arrival, rate, t0)
        key = f"sess{len(sessions)}"
        blobs[key + "_in"] = q
        blobs[key + "_arrival"] = arrival
        blobs[key + "_flush"] = flush
        blobs[key + "_levels"] = levels
        blobs[key + "_frame_len"] = np.asarray([f.size for f in frames], dtype=np.uint32)
        blobs[key + "_frames"] = np.concatenate(frames) if frames else np.zeros(0, dtype=np.int16)
        sessions.append({"key": key, "rate": rate, "quanta": nq, "t0": t0, "desired_ms": 60.0,
                         "flushes": int(flush.sum()), "leftover_quanta": int(nq - (np.nonzero(flush)[0].max() + 1) if flush.any() else nq)})
    record["sessions"] = sessions

    # keep-alive: the callback and period handed to window.setInterval (App.tsx:292-296)
    assert len(ref.intervals) == 1
    fn, period = ref.intervals[0]
    ref.sent = []
    fn()                                              # socket OPEN
    assert len(ref.sent) == 1
    ka = np.frombuffer(ref.sent[0], dtype="<i2")
    ref.ws.js_set("readyState", 3.0)                 # CLOSED: nothing is sent
    fn()
    assert len(ref.sent) == 1
    ref.env["wsRef"].js_set("current", None)         # no socket: nothing is sent
    fn()
    assert len(ref.sent) == 1
    ref.env["wsRef"].js_set("current", ref.ws)
    ref.ws.js_set("readyState", 1.0)
    record["keepalive"] = {"period_ms": period, "samples": int(ka.size), "all_zero": bool((ka == 0).all()),
                           "bytes": len(ref.sent[0]), "sent_when_closed": False, "sent_without_socket": False}

    record["interpreter_steps"] = ref.interp.steps
    np.savez_compressed(os.path.join(HERE, "ref_exec.npz"), **blobs)
    json.dump(record, open(os.path.join(HERE, "ref_exec.json"), "w"), indent=1)
    print(json.dumps(record["checked_against_committed"]), "\nwrote ref_exec.npz / ref_exec.json;", ref.interp.steps, "interpreter steps")


if __name__ == "__main__":
    main()
