// voice-glue.ts -- what apps/voice/src/server.ts gains to convert on the GPU (INTEGRATION.md section 4).
// Frames that arrive within one tick are grouped by (rate, length), converted with ONE framesEncode
// call per group, and each session's slice is passed to the unchanged DeepgramClient.sendAudio
// (apps/voice/src/deepgram.ts:55-58).  New glue, not a behaviour of the reference.
import { allocPinnedFloat32, allocPinnedInt16, framesEncodeAsync, resampleLen } from "./index";

type Sink = (pcm16: Buffer) => void;
type Group = { rate: number; nIn: number; frames: Float32Array[]; sinks: Sink[] };

export class FrameBatcher {
  private groups = new Map<string, Group>();
  private pinIn = allocPinnedFloat32(1 << 20);
  private pinOut = allocPinnedInt16(1 << 19);
  private timer: NodeJS.Timeout;
  private busy = false;

  constructor(opts: { tickMs?: number } = {}) {
    this.timer = setInterval(() => void this.flush(), opts.tickMs ?? 20);
  }

  /** One Float32 frame of one session (the bytes of a WS binary message). */
  push(_sessionId: string, rate: number, data: Buffer, sink: Sink): void {
    const f32 = new Float32Array(data.buffer, data.byteOffset, data.byteLength >> 2);
    const key = `${rate}:${f32.length}`;
    let g = this.groups.get(key);
    if (!g) this.groups.set(key, (g = { rate, nIn: f32.length, frames: [], sinks: [] }));
    g.frames.push(f32);
    g.sinks.push(sink);
  }

  /** Converts everything that arrived during the tick; the CUDA work runs off the event loop. */
  async flush(): Promise<void> {
    if (this.busy) return;              // the pinned blocks are still in use: keep collecting, flush next tick
    this.busy = true;
    const groups = [...this.groups.values()];
    this.groups.clear();
    for (const g of groups) {
      const rows = g.frames.length;
      const nOut = resampleLen(g.nIn, g.rate);
      if (rows * g.nIn > this.pinIn.length) this.pinIn = allocPinnedFloat32(2 * rows * g.nIn);
      if (rows * nOut > this.pinOut.length) this.pinOut = allocPinnedInt16(2 * rows * nOut);
      g.frames.forEach((f, r) => this.pinIn.set(f, r * g.nIn));
      const out = await framesEncodeAsync(this.pinIn, rows, g.nIn, g.rate, this.pinOut);
      g.sinks.forEach((sink, r) =>
        // copy out of the shared pinned block: sendAudio may keep the Buffer beyond this tick
        sink(Buffer.from(Buffer.from(out.buffer, out.byteOffset + r * nOut * 2, nOut * 2))));
    }
    this.busy = false;
  }

  /** The reference's keep-alive (App.tsx:291-296): 100 ms of silence every 2 s per open session. */
  startKeepAlive(sink: Sink): NodeJS.Timeout {
    return setInterval(() => sink(Buffer.alloc(1600 * 2)), 2000);
  }

  close(): void {
    clearInterval(this.timer);
    void this.flush();
  }
}
