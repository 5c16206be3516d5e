"""CPU test of the N>1 host logic with torch.distributed (gloo, world size 2): session sharding,
the whole-job aggregation bench.py uses (units summed over ranks / slowest rank's time), and that
the shards cover the sessions exactly once.  No GPU, no data-path collective (there is none)."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys, json
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "voice-enabled-browser-automation_b200"))
    import numpy as np
    import bench, oracle, pcm_frontend
    rank, local, world = bench.dist_env()
    d = bench.Dist(world, False, local)            # gloo on CPU
    args = bench.parse_args(["--gpus", str(world), "--sessions", "1001", "--scaling", sys.argv[1]])
    first, count = bench.sessions_for_rank(args, rank, world)
    # every rank converts ITS sessions only (oracle stands in for the device here) ...
    x = oracle.synth_fill(count, 96, 48000.0, row0=first)
    y = oracle.frames(x, 96, 48000.0)
    # ... and the job-level numbers come from a SUM of units and a MAX of times
    units, ms, value = bench.aggregate(d, float(count) * 96, 10.0 * (rank + 1))
    assert d.gather(10.0 * (rank + 1)) == [10.0, 20.0]           # per-rank times, in rank order
    cover = d.reduce(float(count), "sum")
    chk = d.reduce(float(np.int64(y.astype(np.int64).sum())), "sum")
    if rank == 0:
        print(json.dumps(dict(units=units, ms=ms, value=value, cover=cover, chk=chk, first=first, count=count)))
    d.close()
""")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(scaling):
    import json
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "-c" if False else "--no-python",
           sys.executable, "-c", WORKER.format(root=ROOT), scaling]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=240, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def test_strong_scaling_shards_cover_all_sessions_once():
    import numpy as np
    import bench
    import oracle
    r = _run("strong")
    assert r["cover"] == 1001 and r["first"] == 0 and r["count"] == 501
    assert r["units"] == 1001 * 96 and r["ms"] == 20.0           # MAX over ranks, not the mean
    assert abs(r["value"] - 1001 * 96 / 0.020) < 1e-6
    # the generator is keyed on the global session id, so the two shards are exactly the rows of the
    # unsharded batch: sharded result == whole-batch oracle
    whole = oracle.frames(oracle.synth_fill(1001, 96, 48000.0, row0=0), 96, 48000.0).astype(np.int64).sum()
    assert r["chk"] == float(whole)


def test_weak_scaling_gives_every_rank_the_full_session_count():
    r = _run("weak")
    assert r["cover"] == 2002 and r["count"] == 1001 and r["units"] == 2002 * 96
