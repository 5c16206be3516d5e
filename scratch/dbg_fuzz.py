import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/voice-enabled-browser-automation_b200")
import oracle, pcm_frontend
ctx = pcm_frontend.Context(0)
seed = 1
rng = np.random.RandomState(1000 + seed)
S, rate, q_ms = 97, 48000, 128e3 / 48000
nq = 150
audio = (rng.uniform(-1.05, 1.05, size=(S, nq, 128)) * rng.uniform(0.05, 1, size=(S, 1, 1))).astype(np.float32)
times = 4.0 * rng.rand(S, 1) + np.cumsum(q_ms + rng.uniform(-0.3, 0.3, size=(S, nq)), axis=1)
events = sorted((times[s, q], s, q) for s in range(S) for q in range(nq))
agg = pcm_frontend.Aggregator(ctx, S, rate, max_frame_quanta=32, max_push_quanta=4096, max_out_samples=30 * 984)
agg.set_meter(True)
agg.reset_range(0, S, 0.0)
TR = 7
fl, _ = oracle.aggregator_plan(times[TR], 60.0, 0.0)
print("oracle flush quanta of session", TR, np.nonzero(fl)[0][:6])
depth = [0]
def push(chunk):
    mine = [e[2] for e in chunk if e[1] == TR]
    try:
        frames = agg.push(np.array([e[1] for e in chunk], dtype=np.uint32), np.array([e[0] for e in chunk]),
                          np.stack([audio[e[1] % S, e[2]] for e in chunk]))
    except pcm_frontend.PcmError as err:
        print("  " * depth[0], "REJECT", len(chunk), "quanta of 7:", mine[:1], mine[-1:], str(err)[:70])
        if err.status == pcm_frontend._capi.ERR_INVALID:
            return push([e for e in chunk if e[1] < S])
        if len(chunk) == 1:
            raise
        depth[0] += 1
        push(chunk[:len(chunk) // 2]); push(chunk[len(chunk) // 2:])
        depth[0] -= 1
        return
    got7 = [pcm.size for s, pcm in frames if s == TR]
    print("  " * depth[0], "ok", len(chunk), "quanta of 7:", mine[:1], mine[-1:], "frames of 7:", got7, "nframes", len(frames))
i = 0
while i < len(events):
    n = int(rng.choice([1, 17, 97, 400, 1500, 4000]))
    chunk = list(events[i:i + n]); i += len(chunk)
    if rng.rand() < 0.3:
        k = int(rng.randint(0, len(chunk)))
        chunk.insert(k, (chunk[k][0], S + int(rng.randint(0, 5)), 0))
    push(chunk)
