"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/pcm_b200.h
declares, and its host-side frame arithmetic agrees with the oracle and the golden fixtures.
No compute call is made here (no GPU)."""
import ctypes as C
import hashlib
import json
import os
import re

import numpy as np
import pytest

import oracle
import pcm_frontend
from pcm_frontend import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "pcm_b200.h")).read()
    return re.findall(r"PCMB200_API\s+[\w\s\*]+?\b(pcmb200_\w+)\s*\(", text)


def test_library_exports_every_declared_symbol():
    L = pcm_frontend.load()
    names = _header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/pcm_b200.h but not exported"
    # and the Python declaration table covers exactly the header
    assert sorted(set(names)) == sorted(_capi.SYMBOLS)


def test_version_and_status_strings():
    L = pcm_frontend.load()
    assert L.pcmb200_version() >= 1000
    assert L.pcmb200_status_string(0) == b"ok"
    assert b"RangeError" in L.pcmb200_status_string(_capi.ERR_RANGE)
    assert b"no CPU fallback" in L.pcmb200_status_string(_capi.ERR_NO_DEVICE)


def test_no_device_fails_loudly():
    n = C.c_int(-1)
    st = pcm_frontend.load().pcmb200_device_count(C.byref(n))
    if st == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pcm_frontend.PcmError) as ei:
        pcm_frontend.Context(0)
    assert ei.value.status == _capi.ERR_NO_DEVICE


def test_resample_len_matches_oracle_and_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "index_kat.json")))["cases"]
    for c in cases:
        assert pcm_frontend.resample_len(c["n_in"], c["rate"]) == c["n_out"] == oracle.resample_len(c["n_in"], c["rate"])
    for rate in (0.0, -48000.0):
        with pytest.raises(pcm_frontend.PcmRangeError):
            pcm_frontend.resample_len(100, rate)
    assert pcm_frontend.resample_len(100, float("nan")) == 0
    assert pcm_frontend.resample_len(100, float("inf")) == 0
    assert pcm_frontend.resample_len(0, 48000) == 0
    rng = np.random.RandomState(3)
    for _ in range(300):
        n, rate = int(rng.randint(0, 200000)), float(rng.choice([8000, 11025, 12000, 15999, 16001, 22050, 37800, 44100, 47999.5, 48000, 96000, 176400]))
        assert pcm_frontend.resample_len(n, rate) == oracle.resample_len(n, rate)


def test_index_table_matches_oracle_and_golden(golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "index_kat.json")))["cases"]
    for c in cases:
        if c.get("identity"):
            continue
        idx = pcm_frontend.index_table(c["n_in"], c["rate"])
        assert hashlib.sha256(idx.astype("<u4").tobytes()).hexdigest() == c["sha256_u32le"], c
        assert np.array_equal(idx, oracle.index_table(c["n_in"], c["rate"]))
    assert pcm_frontend.index_table(44100, 44100)[480] == 1322
    # a pick outside the input is flagged, not clamped
    assert pcm_frontend.index_table(2, 48000, 3).tolist() == [0, 0xFFFFFFFF, 0xFFFFFFFF]


def test_shard_sessions_is_a_partition():
    for n in (0, 1, 7, 1024, 65536, 2 ** 20, 2 ** 20 + 3):
        for world in (1, 2, 3, 4, 8):
            spans = [pcm_frontend.shard_sessions(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == n
            for (a, ca), (b, _) in zip(spans, spans[1:]):
                assert a + ca == b
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        pcm_frontend.shard_sessions(10, 2, 2)


def test_keepalive_frame_and_wire_endianness():
    ka = pcm_frontend.keepalive_frame()
    assert ka.dtype == np.int16 and ka.size == 1600 and not ka.any()
    assert ka.tobytes() == oracle.keepalive_frame().tobytes()          # App.tsx:291-296
    assert pcm_frontend.load().pcmb200_wire_is_little_endian() == 1
    n = C.c_size_t(0)
    small = np.empty(10, np.int16)
    assert pcm_frontend.load().pcmb200_keepalive_frame(small.ctypes.data, small.size, C.byref(n)) == _capi.ERR_CAPACITY


def test_header_is_plain_c_and_links(tmp_path):
    """include/pcm_b200.h compiles as C11 and the example program links against the library and
    runs (without a GPU it stops at context creation, loudly, with exit code 0 by design)."""
    import subprocess
    lib = os.path.join(ROOT, "voice-enabled-browser-automation_b200", "lib")
    for name, ok_line in (("frames_c_abi", "frames of 320 samples"), ("aggregator_c_abi", "3 frames from 3 connections")):
        src = os.path.join(ROOT, "voice-enabled-browser-automation_b200", "examples", name + ".c")
        exe = str(tmp_path / name)
        subprocess.run(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), src,
                        "-L", lib, "-lpcm_b200", f"-Wl,-rpath,{lib}", "-lm", "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        assert out.returncode == 0, name + ": " + out.stdout + out.stderr
        assert ok_line in out.stdout or "no device" in out.stdout, out.stdout


@pytest.mark.gpu
def test_c_examples_run_on_the_gpu(tmp_path):
    """The same two plain-C programs on a box with a device: they must get all the way through."""
    import subprocess
    lib = os.path.join(ROOT, "voice-enabled-browser-automation_b200", "lib")
    for name, ok_line in (("frames_c_abi", "frames of 320 samples"), ("aggregator_c_abi", "3 frames from 3 connections")):
        src = os.path.join(ROOT, "voice-enabled-browser-automation_b200", "examples", name + ".c")
        exe = str(tmp_path / name)
        subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), src, "-L", lib, "-lpcm_b200", f"-Wl,-rpath,{lib}", "-lm",
                        "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
        assert out.returncode == 0 and ok_line in out.stdout, name + ": " + out.stdout + out.stderr
