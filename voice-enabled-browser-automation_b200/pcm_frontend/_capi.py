"""ctypes declarations of every entry point in include/pcm_b200.h."""
from __future__ import annotations

import ctypes as C
import os

MEM_HOST = 0
MEM_DEVICE = 1

OK, ERR_INVALID, ERR_RANGE, ERR_CUDA, ERR_NO_DEVICE, ERR_NOMEM, ERR_CAPACITY = range(7)


class PcmError(RuntimeError):
    """A libpcm_b200 call failed (status, message from pcmb200_last_error)."""

    def __init__(self, status: int, message: str):
        super().__init__(f"pcm_b200 status {status}: {message}")
        self.status = status


class PcmRangeError(PcmError, ValueError):
    """Where the reference throws RangeError (invalid typed array length)."""


_vp, _sz, _u32, _u64, _dbl, _int = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_double, C.c_int
_P = C.POINTER

# name -> (restype, argtypes): the single source of truth for the Python side of the ABI
SYMBOLS = {
    "pcmb200_version": (_int, []),
    "pcmb200_status_string": (C.c_char_p, [_int]),
    "pcmb200_device_count": (_int, [_P(_int)]),
    "pcmb200_ctx_create": (_int, [_int, _P(_vp)]),
    "pcmb200_ctx_destroy": (_int, [_vp]),
    "pcmb200_last_error": (C.c_char_p, [_vp]),
    "pcmb200_sync": (_int, [_vp]),
    "pcmb200_launch_count": (_int, [_vp, _P(_u64)]),
    "pcmb200_device_info": (_int, [_vp, C.c_char_p, _sz, _P(_int), _P(_sz)]),
    "pcmb200_ctx_reserve": (_int, [_vp, _u32, _u32, _u32, _u32]),
    "pcmb200_alloc_stats": (_int, [_vp, _P(_u64), _P(_u64), _P(_u64)]),
    "pcmb200_resample_len": (_int, [_sz, _dbl, _P(_sz)]),
    "pcmb200_index_table": (_int, [_sz, _dbl, _sz, _vp]),
    "pcmb200_device_alloc": (_int, [_vp, _sz, _P(_vp)]),
    "pcmb200_device_free": (_int, [_vp, _vp]),
    "pcmb200_pinned_alloc": (_int, [_vp, _sz, _P(_vp)]),
    "pcmb200_pinned_alloc_write_combined": (_int, [_vp, _sz, _P(_vp)]),
    "pcmb200_pinned_free": (_int, [_vp, _vp]),
    "pcmb200_memcpy_h2d": (_int, [_vp, _vp, _vp, _sz]),
    "pcmb200_memcpy_d2h": (_int, [_vp, _vp, _vp, _sz]),
    "pcmb200_memset_device": (_int, [_vp, _vp, _int, _sz]),
    "pcmb200_encode_f32_s16": (_int, [_vp, _vp, _vp, _sz, _int]),
    "pcmb200_resample_to_16k": (_int, [_vp, _vp, _sz, _dbl, _vp, _sz, _P(_sz), _int]),
    "pcmb200_frames_encode": (_int, [_vp, _vp, _sz, _vp, _sz, _u32, _u32, _dbl, _int, _P(_sz)]),
    "pcmb200_synth_fill": (_int, [_vp, _vp, _sz, _u32, _u32, _dbl, _u64, _u64]),
    "pcmb200_pcie_probe": (_int, [_vp, _sz, _sz, _int, _P(C.c_float)]),
    "pcmb200_timer_start": (_int, [_vp]),
    "pcmb200_timer_stop": (_int, [_vp, _P(C.c_float)]),
    "pcmb200_tune": (_int, [_vp, C.c_char_p, C.c_longlong]),
    "pcmb200_frames_encode_meter": (_int, [_vp, _vp, _sz, _vp, _sz, _u32, _u32, _dbl, _u32, _vp, _vp, _P(_sz)]),
    "pcmb200_aggregator_create": (_int, [_vp, _u32, _dbl, _u32, _dbl, _u32, _u32, _sz, _P(_vp)]),
    "pcmb200_aggregator_destroy": (_int, [_vp, _vp]),
    "pcmb200_aggregator_reset_session": (_int, [_vp, _vp, _u32, _dbl]),
    "pcmb200_aggregator_reset_range": (_int, [_vp, _vp, _u32, _u32, _dbl]),
    "pcmb200_aggregator_push_device": (_int, [_vp, _vp, _u32, _vp, _vp, _vp, _vp, _sz, _vp, _vp, _vp, _u32]),
    "pcmb200_aggregator_wait": (_int, [_vp, _vp, _P(_u32), _P(_sz)]),
    "pcmb200_aggregator_set_meter": (_int, [_vp, _vp, _int]),
    "pcmb200_aggregator_get_levels": (_int, [_vp, _vp, _u32, _u32, _vp]),
    "pcmb200_aggregator_push": (_int, [_vp, _vp, _u32, _vp, _vp, _vp, _int, _vp, _sz, _int, _vp, _vp, _vp, _u32, _P(_u32)]),
    "pcmb200_keepalive_frame": (_int, [_vp, _sz, _P(_sz)]),
    "pcmb200_wire_is_little_endian": (_int, []),
    "pcmb200_egress_table": (_int, [_u32, _sz, _sz, _vp, _vp]),
    "pcmb200_egress_table_frames": (_int, [_u32, _vp, _vp, _vp, _vp]),
    "pcmb200_keepalive_due": (_int, [_dbl, _u32, _vp, _vp, _vp, _dbl, _vp, _u32, _P(_u32)]),
    "pcmb200_tick_graph_create": (_int, [_vp, _vp, _sz, _vp, _sz, _u32, _u32, _dbl, _P(_vp)]),
    "pcmb200_tick_graph_launch": (_int, [_vp, _vp]),
    "pcmb200_tick_graph_destroy": (_int, [_vp, _vp]),
    "pcmb200_frames_encode_noalloc": (_int, [_vp, _sz, _vp, _sz, _u32, _u32, _dbl, _vp, _int, _P(_sz)]),
}

_lib = None


def lib_path() -> str:
    here = os.path.dirname(os.path.abspath(__file__))
    return os.environ.get("PCMB200_LIB", os.path.join(os.path.dirname(here), "lib", "libpcm_b200.so"))


def load() -> C.CDLL:
    """dlopen libpcm_b200.so and declare every symbol.  Raises if the library is not built:
    there is no other implementation to fall back to."""
    global _lib
    if _lib is None:
        path = lib_path()
        if not os.path.exists(path):
            raise ImportError(
                f"{path} is missing: build it with `make -C voice-enabled-browser-automation_b200` "
                "(or __graft_entry__.build()).  pcm_frontend has no CPU fallback.")
        L = C.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)          # AttributeError if the .so does not export it
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(status: int, ctx_handle=None):
    if status == OK:
        return
    msg = load().pcmb200_last_error(ctx_handle)
    text = msg.decode(errors="replace") if msg else ""
    if not text:
        text = load().pcmb200_status_string(status).decode()
    raise (PcmRangeError if status == ERR_RANGE else PcmError)(status, text)
