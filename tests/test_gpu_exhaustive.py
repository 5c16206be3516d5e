"""Exhaustive known-answer test of the encoder: ALL 2^32 float32 bit patterns through the CUDA
path (C ABI, host buffers) against the oracle's binary64 arithmetic.  Closes hard part H1 of
SURVEY.md: the device code uses a round-toward-zero fp32 product instead of binary64."""
import ctypes as C
import hashlib
import json
import os
import threading

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _oracle_mt(x, out, threads=8):
    L = oracle.lib()
    n = x.size
    step = (n + threads - 1) // threads

    def work(t):
        a, b = t * step, min(n, (t + 1) * step)
        if a < b:
            L.oracle_float_to_16bit_pcm(x[a:b].ctypes.data_as(C.POINTER(C.c_float)), b - a,
                                        out[a:b].ctypes.data_as(C.POINTER(C.c_int16)))
    ts = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]


def test_all_float32_bit_patterns(ctx):
    chunk = 1 << 26
    pin_in = ctx.pinned(chunk, np.uint32)
    pin_out = ctx.pinned(chunk, np.int16)
    want = np.empty(chunk, dtype=np.int16)
    base = np.arange(chunk, dtype=np.uint32)
    digest = hashlib.sha256()
    for c in range((1 << 32) // chunk):
        np.add(base, np.uint32(c * chunk), out=pin_in.array, casting="unsafe")
        x = pin_in.array.view(np.float32)
        ctx.encode_ptr(pin_in.ptr, pin_out.ptr, chunk, 0)          # MEM_HOST: synchronous
        _oracle_mt(x, want)
        if not np.array_equal(pin_out.array, want):
            bad = np.nonzero(pin_out.array != want)[0][:5]
            raise AssertionError(f"chunk {c}: first mismatches at bits "
                                 f"{[hex(int(pin_in.array[i])) for i in bad]} got {pin_out.array[bad]} want {want[bad]}")
        digest.update(pin_out.array.astype("<i2", copy=False).tobytes())
    pin_in.free(); pin_out.free()
    # the committed digest of the whole sweep (tools/make_sweep_digest.py): the same number
    # tools/pin_with_node.mjs expects from the reference's own JavaScript
    here = os.path.dirname(os.path.abspath(__file__))
    committed = json.load(open(os.path.join(here, "golden", "encode_sweep_sha256.json")))["sha256"]
    assert digest.hexdigest() == committed
