// ingest.ts -- voice-server ingestion glue; the line-for-line TypeScript twin of
// pcm_frontend/ingest.py (that file is the tested one: no Node exists on the build or GPU boxes, so
// this file has never run -- keep the two in step).
//
// What apps/voice/src/server.ts:97-109, :260-284 gains when the PCM front-end runs on the server's GPU.
// The reference's handler relays every binary message, untouched, to DeepgramClient.sendAudio
// (apps/voice/src/deepgram.ts:55-58) and treats the rest as optional JSON control frames ("close",
// "context_update"; non-JSON text is ignored).  That stays for clients that send PCM16; a Float32
// sub-protocol is added:
//     text frame   {"type": "audio_format", "encoding": "f32le", "sample_rate": 48000}
//     binary frame raw little-endian Float32 samples of ONE flushed capture buffer (App.tsx:281)
// With `ws` ^8 a TEXT frame is also delivered as a Buffer, so `Buffer.isBuffer(data)` (server.ts:261) is
// true for both kinds and the reference's control branch is unreachable: use the `isBinary` flag that
// `ws` passes as the second argument of the "message" event.
import { allocPinnedFloat32, allocPinnedInt16, framesEncodeAsync, resampleLen } from "./index";

export type Kind = "audio" | "control" | "ignored";
export const OVERFLOW_POLICIES = ["drop_oldest", "drop_newest", "block"] as const;
export type Overflow = (typeof OVERFLOW_POLICIES)[number];

/** ["audio", bytes] for a binary frame; ["control", object] for a JSON text frame; ["ignored", null] otherwise (server.ts:281-283). */
export function classifyMessage(data: Buffer, isBinary: boolean): [Kind, any] {
  if (isBinary) return ["audio", data];
  try {
    const msg = JSON.parse(data.toString("utf8"));
    return msg !== null && typeof msg === "object" && !Array.isArray(msg) ? ["control", msg] : ["ignored", null];
  } catch {
    return ["ignored", null];
  }
}

/** Per-connection state (the ClientState of server.ts:88-95 plus the audio format). */
export class Connection {
  encoding: "pcm16" | "f32le" = "pcm16"; // the reference's clients send ready-made PCM16
  sampleRate = 16000;
  context: Record<string, any> = {};
  closed = false;
  protocolErrors = 0;
  constructor(
    readonly sessionId: string,
    public sendAudio: (chunk: Buffer | ArrayBuffer | Uint8Array) => void, // DeepgramClient.sendAudio, unchanged
    readonly closeUpstream?: () => void,
  ) {}

  /** server.ts:267-280 ("close", "context_update") + the audio_format announcement. */
  onControl(msg: any): string | null {
    const t = msg?.type;
    if (t === "close") {
      this.closed = true;
      this.closeUpstream?.();
      return "close";
    }
    if (t === "context_update" && msg.payload && typeof msg.payload === "object") {
      this.context = { ...this.context, ...msg.payload };
      return "context_update";
    }
    if (t === "audio_format") {
      const enc = msg.encoding, rate = msg.sample_rate;
      if ((enc !== "f32le" && enc !== "pcm16") || typeof rate !== "number" || !(rate > 0)) {
        this.protocolErrors += 1;
        return "bad_audio_format";
      }
      this.encoding = enc;
      this.sampleRate = rate;
      return "audio_format";
    }
    return null;
  }
}

type Pending = { conn: Connection; samples: Float32Array; tMs: number };

/** convertGroup() on the CUDA addon: gather rows into a pinned block, ONE framesEncodeAsync call, views of the pinned output. */
export class GpuGroupConverter {
  private pinIn = allocPinnedFloat32(1 << 24);
  private pinOut = allocPinnedInt16(1 << 24);

  async convertGroup(rate: number, nIn: number, frames: Float32Array[]): Promise<Buffer[]> {
    const rows = frames.length;
    const nOut = resampleLen(nIn, rate);
    // grown only here, between conversions: never while an async job holds the blocks
    if (rows * nIn > this.pinIn.length) this.pinIn = allocPinnedFloat32(2 * rows * nIn);
    if (rows * nOut > this.pinOut.length) this.pinOut = allocPinnedInt16(2 * rows * nOut);
    frames.forEach((f, r) => this.pinIn.set(f, r * nIn));
    const out = nOut ? await framesEncodeAsync(this.pinIn, rows, nIn, rate, this.pinOut) : this.pinOut;
    // egress table: row r is bytes [r * 2 * nOut, (r + 1) * 2 * nOut) of the block -- views, no copies
    // (valid until the next conversion; sendAudio hands the bytes to the socket layer before returning)
    return frames.map((_, r) => Buffer.from(out.buffer, out.byteOffset + r * nOut * 2, nOut * 2));
  }
}

/** Bounded cross-connection queue + per-tick conversion (see pcm_frontend/ingest.py TickBatcher). */
export class TickBatcher {
  private q: Pending[] = [];
  private next: Pending[] = []; // arrivals during a flush
  private bytes = 0;
  private busy = false;
  readonly stats = { frames: 0, dropped_oldest: 0, dropped_newest: 0, blocked: 0, protocol_errors: 0, relayed_pcm16: 0, flushes: 0, conversions: 0 };

  constructor(
    private converter: { convertGroup(rate: number, nIn: number, frames: Float32Array[]): Promise<Buffer[]> },
    private tickMs = 20,
    private maxFrames = 1 << 16,
    private maxBytes = 1 << 30,
    private overflow: Overflow = "drop_oldest",
    private maxFrameSamples = 1 << 20,
  ) {}

  private full(extraBytes: number): boolean {
    return this.q.length + this.next.length + 1 > this.maxFrames || this.bytes + extraBytes > this.maxBytes;
  }

  /** Returns false only under the "block" policy when the frame was refused: ws.pause() until the next flush. */
  offer(conn: Connection, kind: Kind, payload: any, nowMs: number): boolean {
    if (conn.closed || kind === "ignored") return true;
    if (kind === "control") {
      conn.onControl(payload);
      return true;
    }
    if (conn.encoding === "pcm16") {
      conn.sendAudio(payload); // the reference's path: relay the bytes untouched (server.ts:263)
      this.stats.relayed_pcm16 += 1;
      return true;
    }
    const buf = payload as Buffer;
    const nBytes = buf.byteLength;
    if (nBytes % 4 !== 0 || nBytes / 4 > this.maxFrameSamples) {
      conn.protocolErrors += 1;
      this.stats.protocol_errors += 1;
      return true;
    }
    // ws may hand out Buffers at any byte offset of its pool: a Float32Array view needs 4-byte alignment
    const aligned = buf.byteOffset % 4 === 0 ? buf : Buffer.from(buf);
    const item: Pending = { conn, samples: new Float32Array(aligned.buffer, aligned.byteOffset, nBytes >> 2), tMs: nowMs };
    if (this.full(nBytes)) {
      if (this.overflow === "drop_newest") {
        this.stats.dropped_newest += 1;
        return true;
      }
      if (this.overflow === "block") {
        this.stats.blocked += 1;
        return false;
      }
      while (this.q.length && this.full(nBytes)) {
        const old = this.q.shift()!;
        this.bytes -= old.samples.byteLength;
        this.stats.dropped_oldest += 1;
      }
      if (this.full(nBytes)) {
        this.stats.dropped_newest += 1; // nothing left to drop (everything is in the running flush)
        return true;
      }
    }
    (this.busy ? this.next : this.q).push(item);
    this.bytes += nBytes;
    this.stats.frames += 1;
    return true;
  }

  get pending(): number {
    return this.q.length + this.next.length;
  }

  due(nowMs: number): boolean {
    if (!this.q.length) return false;
    return nowMs - this.q[0].tMs >= this.tickMs || this.q.length >= this.maxFrames || this.bytes >= this.maxBytes;
  }

  /** Convert and deliver everything queued; the CUDA work runs off the event loop. */
  async flush(_nowMs = 0): Promise<number> {
    if (this.busy || !this.q.length) return 0;
    this.busy = true;
    const batch = this.q;
    this.q = [];
    let delivered = 0;
    try {
      const groups = new Map<string, Pending[]>();
      for (const it of batch) {
        const key = `${it.conn.sampleRate}:${it.samples.length}`; // arrival order is kept inside a group
        if (!groups.has(key)) groups.set(key, []);
        groups.get(key)!.push(it);
      }
      for (const items of groups.values()) {
        const live = items.filter((it) => !it.conn.closed);
        if (!live.length) continue;
        const views = await this.converter.convertGroup(live[0].conn.sampleRate, live[0].samples.length, live.map((it) => it.samples));
        this.stats.conversions += 1;
        live.forEach((it, r) => {
          it.conn.sendAudio(views[r]); // raw little-endian int16, 2 * N_out bytes, no header
          delivered += 1;
        });
      }
    } finally {
      this.bytes -= batch.reduce((a, it) => a + it.samples.byteLength, 0);
      this.q = this.next;
      this.next = [];
      this.busy = false;
      this.stats.flushes += 1;
    }
    return delivered;
  }
}

/** The reference's keep-alive timers (App.tsx:291-296, cleared at :309-312) for a fleet on an injected clock. */
export class KeepAliveScheduler {
  private start: Float64Array;
  private done: Float64Array;
  private open: Uint8Array;
  constructor(private n: number, private periodMs = 2000) {
    this.start = new Float64Array(n).fill(NaN);
    this.done = new Float64Array(n);
    this.open = new Uint8Array(n);
  }
  startSession(s: number, nowMs: number, socketOpen = true): void {
    this.start[s] = nowMs; this.done[s] = 0; this.open[s] = socketOpen ? 1 : 0;
  }
  stopSession(s: number): void {
    this.start[s] = NaN; this.open[s] = 0;
  }
  setOpen(s: number, socketOpen: boolean): void {
    this.open[s] = socketOpen ? 1 : 0;
  }
  /** Sessions that must be sent `new Int16Array(1600)` now (same arithmetic as pcmb200_keepalive_due). */
  poll(nowMs: number): number[] {
    const due: number[] = [];
    for (let s = 0; s < this.n; s++) {
      const el = nowMs - this.start[s];
      if (!(el >= 0)) continue;
      const ticks = Math.floor(el / this.periodMs);
      if (ticks > this.done[s]) {
        this.done[s] = ticks;
        if (this.open[s]) due.push(s);
      }
    }
    return due;
  }
}

/* Wiring inside apps/voice/src/server.ts (INTEGRATION.md section 4):
 *   const batcher = new TickBatcher(new GpuGroupConverter());
 *   setInterval(() => { if (batcher.due(performance.now())) void batcher.flush(); }, 5);
 *   wss.on("connection", (ws) => {
 *     const conn = new Connection(id, (chunk) => state.dg?.sendAudio(chunk), () => state.dg?.close());
 *     ws.on("message", (data, isBinary) => {
 *       const [kind, payload] = classifyMessage(data as Buffer, isBinary);
 *       if (!batcher.offer(conn, kind, payload, performance.now())) ws.pause();   // "block" policy
 *     });
 *   });
 */
