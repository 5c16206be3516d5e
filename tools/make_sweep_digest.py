#!/usr/bin/env python3
"""SHA-256 of floatTo16BitPCM (App.tsx:7-16) over ALL 2^32 float32 bit patterns, in ascending bit-pattern
order, results as little-endian int16 (8 GiB hashed).  Computed with the C oracle; written to
tests/golden/encode_sweep_sha256.json.  The GPU exhaustive test and tools/pin_with_node.mjs (the
reference's own JS, when a Node exists) must produce the same digest."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402

h = hashlib.sha256()
CH = 1 << 24
for c in range(1 << 8):
    bits = np.arange(c * CH, (c + 1) * CH, dtype=np.uint64).astype(np.uint32)
    h.update(oracle.float_to_16bit_pcm(bits.view(np.float32)).astype("<i2").tobytes())
out = {"what": "sha256 over int16le results of floatTo16BitPCM for float32 bit patterns 0x00000000..0xFFFFFFFF ascending",
       "source": "oracle/pcm_oracle.c (tools/make_sweep_digest.py)", "sha256": h.hexdigest()}
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "encode_sweep_sha256.json"), "w"), indent=1)
print(out["sha256"])
