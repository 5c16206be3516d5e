"""The N-API shim (voice-enabled-browser-automation_b200/node/addon.cc) under a mock Node-API host
(tests/napi_mock/mock_host.cc): Node itself is not installed on this image, so the mock implements
the napi_* functions the shim uses, dlopens the built .node file as Node would and drives the
exported JavaScript-facing functions.  CPU: module load must throw (no CPU fallback).  GPU: the
reference's signatures, error mapping, pinned allocators and the async entry point."""
import ctypes as C
import os
import subprocess

import pytest

import pcm_frontend

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NODE = os.path.join(ROOT, "voice-enabled-browser-automation_b200", "lib", "pcm_b200_napi.node")
SRC = os.path.join(ROOT, "tests", "napi_mock", "mock_host.cc")


def _host(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("napi") / "mock_host")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-rdynamic", SRC, "-ldl", "-lpthread", "-o", exe], check=True)
    if not os.path.exists(NODE):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "voice-enabled-browser-automation_b200"), "napi"], check=True)
    return exe


def _has_gpu():
    n = C.c_int(0)
    return pcm_frontend.load().pcmb200_device_count(C.byref(n)) == 0 and n.value > 0


def test_addon_load_fails_loudly_without_a_device(tmp_path_factory):
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    out = subprocess.run([_host(tmp_path_factory), NODE, "--expect-no-device"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok   module load throws without a CUDA device" in out.stdout


@pytest.mark.gpu
def test_addon_under_mock_napi_host(tmp_path_factory):
    out = subprocess.run([_host(tmp_path_factory), NODE], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ALL OK" in out.stdout, out.stdout + out.stderr
    assert "FAIL" not in out.stdout
