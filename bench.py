#!/usr/bin/env python3
"""bench.py -- PCM front-end throughput on B200 (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): BASELINE.json configs[3] -- 2^20 concurrent 48 kHz mono sessions,
20 ms frames (960 in -> 320 out per frame), nearest-sample decimation to 16 kHz + Int16 encode +
framing, i.e. floatTo16BitPCM(resampleTo16k(frame, 48000)) of App.tsx:281-282 for every session.
One STEP = one tick = one frame of every session of the rank's shard = one kernel launch.
Sessions are independent: multi-GPU is session-affine sharding, no collective on the data path
(torch.distributed, gloo backend, is used only for the start barrier and the max-over-ranks of the
timings; no NCCL communicator is ever created).

Prints ONE JSON line (rank 0).  `value` = input PCM samples/s, device-resident;
`e2e` = the same through pcmb200_frames_encode with pinned HOST buffers (H2D + kernel + D2H).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "voice-enabled-browser-automation_b200"))

RATE = 48000.0
METRIC = "pcm_input_samples_per_s"
UNIT = "samples/s"
FALLBACK_HBM_GBS = 6650.0     # B200_PROFILING.md fallback, used only if MEASURED_PEAKS.json is absent


# --------------------------------------------------------------------------------------------
def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--sessions", type=int, default=1 << 20, help="concurrent sessions (per GPU when --scaling weak)")
    ap.add_argument("--frame-ms", type=int, default=20)
    ap.add_argument("--rate", type=float, default=48000.0, help="capture rate of every session (Hz); BASELINE uses 48000")
    ap.add_argument("--scaling", choices=["weak", "strong"], default="weak",
                    help="weak: every GPU holds --sessions sessions; strong: --sessions are sharded over the GPUs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--wc", type=int, default=0, help="1: e2e input ring in write-combined pinned memory")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work budget of the cpu_baseline leg")
    ap.add_argument("--ref-sessions", type=int, default=1 << 16, help="sessions per step of --impl reference")
    ap.add_argument("--tune", default="", help="comma list key=value passed to pcmb200_tune (sweeps)")
    args = ap.parse_args(argv)
    global RATE
    RATE = float(args.rate)
    return args


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class Dist:
    """Barrier + max/sum over ranks.  World size 1 needs no torch at all."""

    def __init__(self, world, use_cuda, local_rank):
        self.world = world
        self.dist = None
        if world > 1:
            import torch
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            self.torch = torch
            # Control plane only (start barrier, MAX / SUM of a few doubles): gloo on CPU tensors.  The
            # data path has no exchange step, so no NCCL communicator is created -- which also keeps
            # stdout to the one JSON line (NCCL announces its version there).
            self.device = torch.device("cpu")
            dist.init_process_group("gloo")
            self.dist = dist

    def barrier(self):
        if self.dist:
            self.dist.barrier()

    def reduce(self, x: float, op: str) -> float:
        if not self.dist:
            return float(x)
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.dist:
            self.dist.barrier()
            self.dist.destroy_process_group()


def aggregate(dist: Dist, units_this_rank: float, ms_this_rank: float):
    """Whole-job throughput: units of ALL ranks / slowest rank's device time."""
    units = dist.reduce(units_this_rank, "sum")
    ms = dist.reduce(ms_this_rank, "max")
    return units, ms, (units / (ms * 1e-3) if ms > 0 else 0.0)


def sessions_for_rank(args, rank, world):
    """(first_session, count) of this rank (contiguous session blocks)."""
    import pcm_frontend
    if args.scaling == "weak":
        return rank * args.sessions, args.sessions
    return pcm_frontend.shard_sessions(args.sessions, world, rank)


def bind_to_gpu_numa(index):
    """Pin this rank to the CPUs NVML reports as local to GPU `index`, so that its pinned host ring is
    allocated on (and fed from) the GPU's own NUMA node.  Returns (cpus bound, previous affinity)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = ((os.cpu_count() or 64) + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        prev = os.sched_getaffinity(0)
        cpus = sorted(set(cpus) & prev)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus), prev
    except Exception:
        pass
    return 0, None


# ---- clocks ----------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with NVML while the timed regions run."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting",
               0x100: "display_clock_setting", 0x10: "sync_boost"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.in_region = index, [], False, False
        self.ok, self.max_mhz = False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = str(e)

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag:
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                try:
                    rs = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    rs = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((self.in_region, mhz, rs))
            except Exception:
                pass
            time.sleep(0.004)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        timed = [s for s in self.samples if s[0]] or self.samples
        mask = 0
        for _, _, rs in timed:
            mask |= rs
        reasons = sorted(name for bit, name in self.REASONS.items() if mask & bit)
        return {"sm_mhz": float(np.median([m for _, m, _ in timed])), "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples_in_timed_regions": len([s for s in self.samples if s[0]])}


# ---- synthetic audio on the host (reference arm; same recipe as pcmb200_synth_fill) --------------
def host_synth(rows, n, rate, seed=0x5EED, row0=0):
    rng = np.random.RandomState(seed & 0xFFFFFFFF)
    out = np.empty((rows, n), dtype=np.float32)
    k = np.arange(n, dtype=np.float64)[None, :]
    for a in range(0, rows, 8192):
        b = min(rows, a + 8192)
        s = (row0 + np.arange(a, b))[:, None]
        f = 110.0 + 7.0 * (s % 512)
        amp = np.where((s % 8) == 7, 1.2, 0.5)
        out[a:b] = (amp * np.sin(2 * np.pi * f * k / rate) + 0.05 * rng.uniform(-1, 1, size=(b - a, n))).astype(np.float32)
    return out


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def kernel_name(args):
    """The kernel pcmb200_frames_encode selects for a dense device-resident batch at RATE."""
    r = int(RATE // 16000)
    if "path=1" in args.tune or "tile=0" in args.tune and RATE % 16000 != 0:
        return "k_pick_encode"
    if RATE % 16000 != 0 or r not in (1, 2, 3, 4, 6) or "path=2" in args.tune:
        return "k_tile_pick (cp.async.bulk tile staging)"
    if r == 1 or "coalesced=0" in args.tune:
        return f"k_decim_int<R={r}>"
    return f"k_decim_coal<R={r}>"


def workload_name(args, n_in, n_out):
    return (f"BASELINE configs[3]: {args.sessions} concurrent {RATE / 1000:g} kHz mono sessions, {args.frame_ms} ms frames "
            f"({n_in} in -> {n_out} out), decimate to 16 kHz + int16 encode + framing")


# ---- CPU legs (oracle = the reference's algorithm restated; kind 'port') ---------------------------
def cpu_time_frames(x, n_in, rate, threads, reps):
    import ctypes as C
    import oracle
    L = oracle.lib()
    rows = x.shape[0]
    n_out = oracle.resample_len(n_in, rate)
    out = np.empty((rows, n_out), dtype=np.int16)
    pi, po = x.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_int16))
    t0 = time.perf_counter()
    for _ in range(reps):
        L.oracle_frames_mt(pi, n_in, rows, n_in, float(rate), po, n_out, threads)
    dt = time.perf_counter() - t0
    return dt, out


def cpu_baseline_leg(x, n_in, rate, cpu_seconds):
    """Oracle on this box's host cores over a bounded sample of the workload."""
    cores = os.cpu_count() or 1
    rows = x.shape[0]
    dt1, _ = cpu_time_frames(x[: max(1, rows // 16)], n_in, rate, 1, 1)          # calibrate, 1 thread
    per_sample_1t = dt1 / (max(1, rows // 16) * n_in)
    work_one_pass = per_sample_1t * rows * n_in                                   # CPU-seconds per pass
    reps = int(max(1, min(200, round(cpu_seconds / max(work_one_pass, 1e-9)))))
    dt, out = cpu_time_frames(x, n_in, rate, cores, reps)
    return {"value": rows * n_in * reps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "value_1thread": 1.0 / per_sample_1t,
            "sample": f"{rows} sessions x {n_in} samples x {reps} passes, oracle/pcm_oracle.c (C restatement of "
                      f"App.tsx:7-32,281-282; the reference is TypeScript and no JS engine exists on this image), "
                      f"{cores} threads, same input bits as the GPU run"}, out


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    import oracle
    oracle.build()
    n_in = int(RATE * args.frame_ms // 1000)
    n_out = oracle.resample_len(n_in, RATE)
    rows = args.ref_sessions
    cores = os.cpu_count() or 1
    x = host_synth(rows, n_in, RATE)
    for _ in range(args.warmup):
        cpu_time_frames(x, n_in, RATE, cores, 1)
    dt, _ = cpu_time_frames(x, n_in, RATE, cores, args.steps)
    value = rows * n_in * args.steps / dt
    sample = (f"each step = {rows} sessions x {n_in} samples (a bounded sample of the 2^20-session tick), "
              f"oracle/pcm_oracle.c port of App.tsx:7-32,281-282, {cores} host threads")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "synthetic sine+noise (host numpy, same recipe as the GPU arm)",
            "config": {"workload": workload_name(args, n_in, n_out), "reference_runtime":
                       "Node unavailable on this image: C port of the reference's JS (kind=port)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---- the CUDA arm ------------------------------------------------------------------------------------
def run_b200(args):
    rank, local_rank, world = dist_env()
    if world != args.gpus and world > 1:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    import pcm_frontend
    from pcm_frontend import MEM_DEVICE, MEM_HOST

    bound_cpus, prev_affinity = bind_to_gpu_numa(local_rank) if world > 1 else (0, None)
    ctx = pcm_frontend.Context(local_rank)          # raises if the CUDA library / device is missing
    dist = Dist(world, True, local_rank)
    if args.tune:
        ctx.tune(**{k: int(v) for k, v in (kv.split("=") for kv in args.tune.split(","))})

    n_in = int(RATE * args.frame_ms // 1000)
    n_out = pcm_frontend.resample_len(n_in, RATE)
    first, S = sessions_for_rank(args, rank, world)
    in_bytes, out_bytes = S * n_in * 4, S * n_out * 2
    algo_bytes = in_bytes + out_bytes                  # 4*N_in + 2*N_out per frame, FIR reads = 0

    # two distinct tick buffers, alternated; each tick (4.7 GB at 2^20 sessions) is >> the 126 MB L2
    n_buf = 2
    d_in = [ctx.device_buffer(in_bytes) for _ in range(n_buf)]
    d_out = [ctx.device_buffer(out_bytes) for _ in range(n_buf)]
    for b in range(n_buf):
        ctx.synth_fill(d_in[b].ptr, n_in, S, n_in, RATE, 0x5EED + b, first)
    ctx.sync()

    sampler = ClockSampler(local_rank)
    sampler.start()

    def step(i):
        b = i % n_buf
        ctx.frames_encode_ptr(d_in[b].ptr, n_in, d_out[b].ptr, n_out, S, n_in, RATE, MEM_DEVICE)

    # ---- device-resident: `value` and the roofline of the dominant (only) kernel -------------
    for i in range(max(args.warmup, 3)):
        step(i)
    ctx.sync()
    dist.barrier()
    l0 = ctx.launches
    sampler.in_region = True
    ctx.timer_start()
    for i in range(args.steps):
        step(i)
    ms = ctx.timer_stop()
    sampler.in_region = False
    launches = ctx.launches - l0
    dist.barrier()
    launches_all = int(dist.reduce(float(launches), "sum"))
    units, ms_max, value = aggregate(dist, float(S) * n_in * args.steps, ms)
    kernel_ms = ms / max(launches, 1)                  # the timed region holds nothing but these launches
    peak, peak_src = hbm_peak()
    achieved = algo_bytes * (args.steps / max(launches, 1)) / (kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src,
                "kernel": kernel_name(args),
                "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": kernel_ms}
    tr = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu capture
    if os.path.exists(tr):
        try:
            t = json.load(open(tr))
            if t.get("sessions") == S and t.get("n_in") == n_in:
                roofline["traffic"] = t["dram_bytes_per_launch"]
        except Exception:
            pass

    # ---- end to end: pinned host buffers through the C-ABI call, H2D + kernel + D2H per step ----
    e2e = None
    if not args.no_e2e:
        h_in = ctx.pinned((S, n_in), np.float32, write_combined=bool(args.wc))
        h_out = ctx.pinned((S, n_out), np.int16)
        ctx.memcpy_d2h(h_in.ptr, d_in[0].ptr, in_bytes)
        for _ in range(2):
            ctx.frames_encode_ptr(h_in.ptr, n_in, h_out.ptr, n_out, S, n_in, RATE, MEM_HOST)
        dist.barrier()
        e2e_steps = args.steps
        sampler.in_region = True
        ctx.timer_start()
        for _ in range(e2e_steps):
            ctx.frames_encode_ptr(h_in.ptr, n_in, h_out.ptr, n_out, S, n_in, RATE, MEM_HOST)   # returns when out is on the host
        ms_e = ctx.timer_stop()
        sampler.in_region = False
        dist.barrier()
        _, ms_e_max, e_val = aggregate(dist, float(S) * n_in * e2e_steps, ms_e)
        e2e = {"value": e_val, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": out_bytes,
               "ms_per_step": ms_e_max / e2e_steps, "steps": e2e_steps,
               "pcie_gbs": (in_bytes + out_bytes) * e2e_steps / (ms_e * 1e-3) / 1e9,
               "h2d_gbs": in_bytes * e2e_steps / (ms_e * 1e-3) / 1e9}
        # the device result of tick buffer 0 and the host-path result must be the same bytes
        chk = d_out[0].download(np.int16, min(S, 4096) * n_out)
        if not np.array_equal(chk, h_out.array[: min(S, 4096)].reshape(-1)):
            raise SystemExit("e2e output differs from the device-resident output")

    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only) ----------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle
        rows = min(S, 1 << 16)
        x = d_in[0].download(np.float32, rows * n_in).reshape(rows, n_in)
        cpu, want = cpu_baseline_leg(x, n_in, RATE, args.cpu_seconds)
        got = d_out[0].download(np.int16, rows * n_out).reshape(rows, n_out)
        if not np.array_equal(got, want):
            raise SystemExit("GPU output differs from the oracle on the cpu_baseline sample")
        cpu["parity_checked_rows"] = rows

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
                "data": "synthetic sine+noise generated on device (pcmb200_synth_fill, seed 0x5EED)",
                "config": {"workload": workload_name(args, n_in, n_out), "sessions_per_gpu": S,
                           "sessions_total": int(units / (n_in * args.steps)), "rate_hz": RATE,
                           "frame_ms": args.frame_ms, "n_in": n_in, "n_out": n_out,
                           "parallelism": f"session-affine shards x{world}, no collectives",
                           "numa": f"each rank bound to its GPU's {bound_cpus} local CPUs" if bound_cpus else "no CPU binding",
                           "l2": f"inputs larger than L2: {n_buf} alternating tick buffers of {algo_bytes / 1e9:.2f} GB each",
                           "tune": args.tune or "defaults",
                           "arithmetic": "f32 in, round-toward-zero fp32 product (exact), saturating s16 out; bit-exact vs the f64 reference"},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_all,
                "clocks": sampler.summary(),
                "output_samples_per_s": value * n_out / n_in, "device": ctx.device_info()["name"]}
        print(json.dumps(line), flush=True)
    dist.close()
    ctx.close()
    return 0


def main(argv=None):
    args = parse_args(argv)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
