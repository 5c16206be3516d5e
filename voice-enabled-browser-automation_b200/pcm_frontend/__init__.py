"""pcm_frontend -- Python host-side mirror of the reference's PCM front-end over libpcm_b200.so.

The reference's functions (apps/web/src/App.tsx) keep their names and argument meaning:

    floatTo16BitPCM(float32)             App.tsx:7-16   -> np.int16 array
    resampleTo16k(input, inputSampleRate) App.tsx:18-32  -> np.float32 array
    frames_encode(batch, rate)           App.tsx:281-282 over a [rows, n_in] batch

Everything is computed by the CUDA library through its C ABI (include/pcm_b200.h) with ctypes.
There is NO CPU fallback: importing works anywhere (so the ABI can be inspected on a CPU box),
but creating a Context without libpcm_b200.so or without a CUDA device raises.
"""
from __future__ import annotations

from ._capi import (  # noqa: F401
    MEM_DEVICE,
    MEM_HOST,
    PcmError,
    PcmRangeError,
    SYMBOLS,
    lib_path,
    load,
)
from .api import (  # noqa: F401
    Aggregator,
    Context,
    DeviceBuffer,
    PinnedBuffer,
    TickGraph,
    default_context,
    floatTo16BitPCM,
    frames_encode,
    index_table,
    keepalive_frame,
    resample_len,
    resampleTo16k,
    shard_sessions,
)

from .ingest import (  # noqa: F401,E402
    Connection,
    GpuGroupConverter,
    GpuQuantumAggregator,
    QuantumBatcher,
    KeepAliveScheduler,
    TickBatcher,
    classify_message,
)

__all__ = [
    "Connection", "GpuGroupConverter", "GpuQuantumAggregator", "QuantumBatcher", "KeepAliveScheduler", "TickBatcher", "classify_message",
    "Aggregator", "Context", "DeviceBuffer", "PinnedBuffer", "TickGraph", "PcmError", "PcmRangeError", "MEM_HOST", "MEM_DEVICE",
    "floatTo16BitPCM", "resampleTo16k", "frames_encode", "resample_len", "index_table", "keepalive_frame",
    "shard_sessions", "default_context", "load", "lib_path", "SYMBOLS",
]
