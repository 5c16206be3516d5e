#!/usr/bin/env python3
"""bench.py -- PCM front-end throughput on B200 (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle port)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload): BASELINE.json configs[3] -- 2^20 concurrent 48 kHz mono sessions,
20 ms frames (960 in -> 320 out per frame), nearest-sample decimation to 16 kHz + Int16 encode +
framing, i.e. floatTo16BitPCM(resampleTo16k(frame, 48000)) of App.tsx:281-282 for every session.
One STEP = one tick = one frame of every session of the rank's shard = one kernel launch.
Default `--scaling strong`: the 2^20 sessions are sharded over the N ranks (2^20 / N per GPU), which is
what BASELINE configs[3] states; `--scaling weak` gives every GPU 2^20 sessions instead.
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
    ap.add_argument("--scaling", choices=["weak", "strong"], default="strong",
                    help="strong (default, BASELINE configs[3]): --sessions are sharded over the GPUs; "
                         "weak: every GPU holds --sessions sessions")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="do not let consecutive ticks overlap on the GPU (overlap_ticks=0)")
    ap.add_argument("--wc", type=int, default=0, help="1: e2e input ring in write-combined pinned memory")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work budget of the cpu_baseline leg")
    ap.add_argument("--ref-sessions", type=int, default=0,
                    help="sessions per step of --impl reference (0 = all --sessions: the full tick, like the CUDA arm)")
    ap.add_argument("--sustained-seconds", type=float, default=2.5,
                    help="length of the back-to-back `sustained` leg (0 = skip)")
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

    def gather(self, x: float):
        """Every rank's value, in rank order (used to separate per-rank time from the max over ranks)."""
        if not self.dist:
            return [float(x)]
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.device)
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(o.item()) for o in out]

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
    """Bind this rank to the CPUs NVML reports as local to GPU `index` -- but only when that is a real
    choice.  On a box whose GPUs all report the same affinity (one NUMA node; every 8-GPU box of this
    pool so far) the "local" set is every CPU, binding would do nothing, and the note says so.
    Returns (number of CPUs bound to or 0, note for the JSON line)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        words = ((os.cpu_count() or 64) + 63) // 64
        prev = os.sched_getaffinity(0)

        def local(i):
            mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(i), words)
            return {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1} & prev
        mine = local(index)
        if not mine:
            return 0, "NVML reports no local CPUs: no CPU binding"
        if len(mine) == len(prev):
            return 0, f"GPU {index} is local to all {len(prev)} CPUs of this process (single NUMA node): no CPU binding"
        os.sched_setaffinity(0, sorted(mine))
        return len(mine), f"rank bound to its GPU's {len(mine)} local CPUs (of {len(prev)})"
    except Exception as e:
        return 0, f"no CPU binding (NVML affinity unavailable: {type(e).__name__})"


# ---- clocks ----------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with NVML while the timed regions run."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting",
               0x100: "display_clock_setting", 0x10: "sync_boost"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.in_region = index, [], False, None
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
                try:
                    pw = self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((self.in_region, mhz, rs, pw))
            except Exception:
                pass
            time.sleep(0.004)

    def mark(self, region):
        """Name of the region the following samples belong to (None = outside every timed region)."""
        self.in_region = region

    def summary(self, region="timed"):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        timed = [s for s in self.samples if s[0] == region] or self.samples
        mask = 0
        for _, _, rs, _ in timed:
            mask |= rs
        reasons = sorted(name for bit, name in self.REASONS.items() if mask & bit)
        return {"sm_mhz": float(np.median([m for _, m, _, _ in timed])), "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "power_w_max": max((p for _, _, _, p in timed if p is not None), default=None),
                "samples_in_region": len([s for s in self.samples if s[0] == region])}


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
    if RATE % 16000 != 0 or r not in (1, 2, 3, 4, 5, 6, 8, 12) or "path=2" in args.tune:
        return "k_tile_pick (cp.async.bulk tile staging)"
    if r == 1 or "coalesced=0" in args.tune:
        return f"k_decim_int<R={r}>"
    return f"k_decim_coal<R={r}>"


def workload_name(args, n_in, n_out):
    return (f"BASELINE configs[3]: {args.sessions} concurrent {RATE / 1000:g} kHz mono sessions, {args.frame_ms} ms frames "
            f"({n_in} in -> {n_out} out), decimate to 16 kHz + int16 encode + framing")


def config_dict(args, world, n_in, n_out, sessions_per_gpu, sessions_total):
    """The `config` object: IDENTICAL for the CUDA arm and the reference arm at the same flags (what is
    run).  How an arm runs it (buffers, tuning, host threads ...) goes in the line's `arm` object."""
    return {"workload": workload_name(args, n_in, n_out), "sessions_total": int(sessions_total),
            "sessions_per_gpu": int(sessions_per_gpu), "rate_hz": RATE, "frame_ms": args.frame_ms,
            "n_in": n_in, "n_out": n_out, "scaling": args.scaling,
            "parallelism": f"session-affine shards x{world}, no collectives",
            "l2": "inputs larger than L2, no flush: the CUDA arm rotates >= 3 distinct tick buffers with a working set "
                  ">= 2 GB per GPU (L2 is 126 MB); the CPU arm's tick is far larger than the host's last-level cache"}


# ---- CPU legs (oracle = the reference's algorithm restated; kind 'port') ---------------------------
def cpu_time_frames(x, n_in, rate, threads, reps, out=None):
    import ctypes as C
    import oracle
    L = oracle.lib()
    rows = x.shape[0]
    n_out = oracle.resample_len(n_in, rate)
    if out is None:
        out = np.empty((rows, n_out), dtype=np.int16)
    pi, po = x.ctypes.data_as(C.POINTER(C.c_float)), out.ctypes.data_as(C.POINTER(C.c_int16))
    t0 = time.perf_counter()
    for _ in range(reps):
        L.oracle_frames_mt(pi, n_in, rows, n_in, float(rate), po, n_out, threads)
    dt = time.perf_counter() - t0
    return dt, out


def cpu_baseline_leg(x, n_in, rate, cpu_seconds):
    """Oracle on this box's host cores over the WHOLE tick (all rows of x), repeated until about
    `cpu_seconds` of CPU work are spent.  Returns (record, oracle output of every row)."""
    cores = os.cpu_count() or 1
    rows = x.shape[0]
    cal = max(1, rows // 64)
    dt1, _ = cpu_time_frames(x[:cal], n_in, rate, 1, 1)                           # calibrate, 1 thread
    per_sample_1t = dt1 / (cal * n_in)
    work_one_pass = per_sample_1t * rows * n_in                                   # CPU-seconds per pass
    reps = int(max(1, min(200, round(cpu_seconds / max(work_one_pass, 1e-9)))))
    cpu_time_frames(x, n_in, rate, cores, 1)                                      # warm (first touch of `out`)
    dt, out = cpu_time_frames(x, n_in, rate, cores, reps)
    return {"value": rows * n_in * reps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "value_1thread": 1.0 / per_sample_1t,
            "sample": f"{rows} sessions x {n_in} samples (the whole tick) x {reps} passes, oracle/pcm_oracle.c (C "
                      f"restatement of App.tsx:7-32,281-282; the reference is TypeScript and no JS engine exists on "
                      f"this image), {cores} threads, same input bits as the GPU run"}, out


def node_pin():
    """If a Node binary AND the reference's source tree are both present, run the reference's own
    JavaScript (sliced out of App.tsx at run time, tools/pin_with_node.mjs) on the committed vectors and
    time config 1 with it.  Neither exists on this image or on the GPU boxes so far."""
    import shutil
    import subprocess
    exe = shutil.which("node")
    ref = os.environ.get("PCMB200_REFERENCE_ROOT", "/root/reference")
    if not exe:
        return {"available": False, "why": "no `node` on PATH"}
    if not os.path.exists(os.path.join(ref, "apps", "web", "src", "App.tsx")):
        return {"available": False, "why": f"node found ({exe}) but no reference tree at {ref} (set PCMB200_REFERENCE_ROOT)"}
    try:
        out = subprocess.run([exe, os.path.join(ROOT, "tools", "pin_with_node.mjs"), ref, "--bench"], capture_output=True,
                             text=True, timeout=600)
        bench = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
        return {"available": True, "parity_pinned": out.returncode == 0 and "PARITY PINNED" in out.stdout,
                "config1": bench[-1] if bench else None, "stderr_tail": out.stderr[-300:]}
    except Exception as e:  # noqa: BLE001
        return {"available": True, "error": repr(e)}


def oracle_pin():
    """How the oracle that this arm times is tied to the reference: the committed record of the run in which the
    reference's own source text was executed on the golden vectors (tests/golden/make_ref_exec.py)."""
    try:
        rec = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_exec.json")))
        return {"by": "the reference's source text (App.tsx slices) executed by tests/golden/mini_js.py; oracle and CUDA "
                      "path held to its outputs in tests/test_ref_exec.py and tests/test_gpu_ref_exec.py",
                "app_tsx_sha256": rec.get("app_tsx_sha256"), "checked_against_committed": rec.get("checked_against_committed"),
                "production_js_engine": False}
    except Exception as e:  # noqa: BLE001
        return {"by": None, "error": repr(e)}


def run_reference(args):
    """The reference's own CPU path for the same workload: every step converts the full tick (all
    --sessions sessions) with the C port of App.tsx:7-32,281-282 on all host threads.  Rank 0 only."""
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    import oracle
    oracle.build()
    world = max(world, args.gpus, 1)
    n_in = int(RATE * args.frame_ms // 1000)
    n_out = oracle.resample_len(n_in, RATE)
    rows = args.ref_sessions or args.sessions
    cores = os.cpu_count() or 1
    x = oracle.synth_fill(rows, n_in, RATE, 0x5EED, 0, cores)
    out = np.empty((rows, n_out), dtype=np.int16)
    for _ in range(max(args.warmup, 1)):
        cpu_time_frames(x, n_in, RATE, cores, 1, out)
    dt, _ = cpu_time_frames(x, n_in, RATE, cores, args.steps, out)
    value = rows * n_in * args.steps / dt
    whole = "the whole tick" if rows == args.sessions else f"a bounded sample of the {args.sessions}-session tick"
    sample = (f"each step = {rows} sessions x {n_in} samples ({whole}), oracle/pcm_oracle.c port of "
              f"App.tsx:7-32,281-282, {cores} host threads")
    per_gpu = args.sessions if args.scaling == "weak" else -(-args.sessions // world)
    node = node_pin()
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "synthetic sine+noise (oracle_synth_fill: the device generator's recipe and noise bits, host libm sine)",
            "config": config_dict(args, world, n_in, n_out, per_gpu, args.sessions),
            "arm": {"runtime": "C port of the reference's JS (kind=port), all host threads; the reference itself is "
                               "single-threaded TypeScript for a browser",
                    "sessions_per_step": rows, "host_threads": cores, "node": node, "oracle_pin": oracle_pin()},
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

    bound_cpus, numa_note = bind_to_gpu_numa(local_rank) if world > 1 else (0, "single rank: no CPU binding")
    ctx = pcm_frontend.Context(local_rank)          # raises if the CUDA library / device is missing
    dist = Dist(world, True, local_rank)
    # Ticks are independent frames in rotating buffers, so consecutive launches may overlap on the GPU
    # (programmatic dependent launch: the next tick's blocks fill the previous tick's draining last wave).
    # That matters for small shards (a 131 072-session tick is 85 us: +6 %), not for the 2^20-session one.
    if not args.no_overlap:
        ctx.tune(overlap_ticks=1)
    if args.tune:
        ctx.tune(**{k: int(v) for k, v in (kv.split("=") for kv in args.tune.split(","))})

    n_in = int(RATE * args.frame_ms // 1000)
    n_out = pcm_frontend.resample_len(n_in, RATE)
    first, S = sessions_for_rank(args, rank, world)
    in_bytes, out_bytes = S * n_in * 4, S * n_out * 2
    algo_bytes = in_bytes + out_bytes                  # 4*N_in + 2*N_out per frame, FIR reads = 0

    # distinct tick buffers, rotated; the working set is kept >> the 126 MB L2 (a 2^20-session tick is
    # 4.7 GB; a 131 072-session shard is 0.59 GB, so small shards rotate more buffers)
    n_buf = int(min(8, max(3, -(-2_000_000_000 // max(algo_bytes, 1)))))
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
    sampler.mark("timed")
    ctx.timer_start()
    for i in range(args.steps):
        step(i)
    ms = ctx.timer_stop()
    sampler.mark(None)
    launches = ctx.launches - l0
    dist.barrier()
    launches_all = int(dist.reduce(float(launches), "sum"))
    per_rank_ms = dist.gather(ms)
    units, ms_max, value = aggregate(dist, float(S) * n_in * args.steps, ms)
    kernel_ms = ms / max(launches, 1)                  # the timed region holds nothing but these launches
    peak, peak_src = hbm_peak()
    achieved = algo_bytes * (args.steps / max(launches, 1)) / (kernel_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "traffic_source": None, "peak_source": peak_src,
                "kernel": kernel_name(args),
                "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": kernel_ms}
    tr = os.path.join(ROOT, "profiles", "traffic.json")   # dram bytes per launch from the committed ncu capture
    if os.path.exists(tr):
        try:
            t = json.load(open(tr))
            if t.get("sessions") == S and t.get("n_in") == n_in:
                roofline["traffic"] = t["dram_bytes_per_launch"]
                roofline["traffic_source"] = ("NOT measured in this run: dram__bytes_read.sum + dram__bytes_write.sum of "
                                              "the same kernel and shape from the committed ncu capture "
                                              "(profiles/traffic.json: " + str(t.get("source", "ncu --set full")) + ")")
        except Exception:
            pass

    # ---- sustained: >= 2 s of back-to-back ticks with their own clock record ---------------------
    sustained = None
    if args.sustained_seconds > 0:
        n_s = int(max(args.steps, args.sustained_seconds * 1e3 / max(ms / args.steps, 1e-6)))
        dist.barrier()
        sampler.mark("sustained")
        ctx.timer_start()
        for i in range(n_s):
            step(i)
        ms_s = ctx.timer_stop()
        sampler.mark(None)
        _, ms_s_max, v_s = aggregate(dist, float(S) * n_in * n_s, ms_s)
        gbs = algo_bytes * n_s / (ms_s * 1e-3) / 1e9
        sustained = {"seconds": ms_s_max * 1e-3, "steps": n_s, "ms_per_step": ms_s_max / n_s, "value": v_s,
                     "unit": UNIT, "achieved_gbs_rank0": gbs, "frac_rank0": gbs / peak,
                     "clocks": sampler.summary("sustained")}

    # ---- end to end: pinned host buffers through the C-ABI call, H2D + kernel + D2H per step ----
    e2e = None
    h_in = h_out = None
    if not args.no_e2e:
        h_in = ctx.pinned((S, n_in), np.float32, write_combined=bool(args.wc))
        h_out = ctx.pinned((S, n_out), np.int16)
        ctx.memcpy_d2h(h_in.ptr, d_in[0].ptr, in_bytes)
        for _ in range(2):
            ctx.frames_encode_ptr(h_in.ptr, n_in, h_out.ptr, n_out, S, n_in, RATE, MEM_HOST)
        dist.barrier()
        e2e_steps = args.steps
        sampler.mark("timed")
        ctx.timer_start()
        for _ in range(e2e_steps):
            ctx.frames_encode_ptr(h_in.ptr, n_in, h_out.ptr, n_out, S, n_in, RATE, MEM_HOST)   # returns when out is on the host
        ms_e = ctx.timer_stop()
        sampler.mark(None)
        dist.barrier()
        _, ms_e_max, e_val = aggregate(dist, float(S) * n_in * e2e_steps, ms_e)
        e2e = {"value": e_val, "unit": UNIT, "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": out_bytes,
               "ms_per_step": ms_e_max / e2e_steps, "steps": e2e_steps,
               "pcie_gbs": (in_bytes + out_bytes) * e2e_steps / (ms_e * 1e-3) / 1e9,
               "h2d_gbs": in_bytes * e2e_steps / (ms_e * 1e-3) / 1e9,
               "per_rank_ms": dist.gather(ms_e)}
        # the device result of tick buffer 0 and the host-path result must be the same bytes (all rows)
        step(0)
        ctx.sync()
        for r0 in range(0, S, 1 << 17):
            nr = min(1 << 17, S - r0)
            chk = d_out[0].download(np.int16, nr * n_out, r0 * n_out * 2)
            if not np.array_equal(chk, h_out.array[r0:r0 + nr].reshape(-1)):
                raise SystemExit("e2e output differs from the device-resident output")

    sampler.stop_flag = True
    sampler.join(timeout=2)

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only): the whole tick, every row checked
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import oracle
        if h_in is not None:
            x = h_in.array                                          # the tick's input bits, already on the host
        else:
            x = d_in[0].download(np.float32, S * n_in).reshape(S, n_in)
        cpu, want = cpu_baseline_leg(x, n_in, RATE, args.cpu_seconds)
        step(0)
        ctx.sync()
        for r0 in range(0, S, 1 << 17):
            nr = min(1 << 17, S - r0)
            got = d_out[0].download(np.int16, nr * n_out, r0 * n_out * 2).reshape(nr, n_out)
            if not np.array_equal(got, want[r0:r0 + nr]):
                raise SystemExit(f"GPU output differs from the oracle (rows {r0}..{r0 + nr})")
        if h_out is not None and not np.array_equal(h_out.array, want):
            raise SystemExit("e2e (host-resident) output differs from the oracle")
        cpu["parity_checked_rows"] = S

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_max / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
                "data": "synthetic sine+noise generated on device (pcmb200_synth_fill, seed 0x5EED)",
                "config": config_dict(args, world, n_in, n_out, S if args.scaling == "weak" else -(-args.sessions // world),
                                      int(units / (n_in * args.steps))),
                "arm": {"numa": numa_note,
                        "l2": f"inputs larger than L2: {n_buf} rotating tick buffers of {algo_bytes / 1e9:.2f} GB each",
                        "tune": args.tune or "defaults", "sessions_rank0": S,
                        "tick_overlap": "off" if args.no_overlap else "on: consecutive ticks may overlap on the GPU (programmatic "
                                        "dependent launch); kernel_ms is the timed region / launches",
                        "arithmetic": "f32 in, round-toward-zero fp32 product (exact), saturating s16 out; bit-exact vs the f64 reference"},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_all,
                "per_rank_ms": per_rank_ms, "sustained": sustained,
                "clocks": sampler.summary("timed"),
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
