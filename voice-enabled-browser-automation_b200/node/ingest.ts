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
// and, with "framing": "quanta" in the announcement, raw 128-sample AudioWorklet quanta (one or more per message):
// the SERVER then runs the capture aggregator of App.tsx:148-157, :273-288 for every connection on the GPU
// (QuantumBatcher below; the arrival time of a message plays performance.now()).
// With `ws` ^8 a TEXT frame is also delivered as a Buffer, so `Buffer.isBuffer(data)` (server.ts:261) is
// true for both kinds and the reference's control branch is unreachable: use the `isBinary` flag that
// `ws` passes as the second argument of the "message" event.
import {
  aggregatorCreate, aggregatorDestroy, aggregatorPush, aggregatorReset, allocPinnedFloat32, allocPinnedInt16, framesEncodeAsync,
  resampleLen,
} from "./index";

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
  framing: "frames" | "quanta" = "frames"; // "quanta": raw AudioWorklet quanta, aggregated on the GPU
  slot = -1; // QuantumBatcher slot
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
      const framing = msg.framing ?? "frames";
      if ((framing !== "frames" && framing !== "quanta") || (framing === "quanta" && enc !== "f32le")) {
        this.protocolErrors += 1;
        return "bad_audio_format";
      }
      this.encoding = enc;
      this.sampleRate = rate;
      this.framing = framing;
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

/** One device-side capture aggregator for the connections of ONE capture rate (see pcm_frontend/ingest.py GpuQuantumAggregator). */
export class GpuQuantumAggregator {
  private handle: number;
  private sess: Uint32Array;
  private time: Float64Array;
  private x: Float32Array;
  private out: Int16Array;
  private fs: Uint32Array;
  private fl: Uint32Array;
  private fo: Float64Array;
  private free: number[] = [];
  private n = 0;
  private maxOut: number;

  constructor(readonly rate: number, maxSessions: number, readonly maxPush: number, readonly quantum = 128, maxFrameQuanta = 64) {
    const nOutMax = resampleLen(maxFrameQuanta * quantum, rate);
    this.maxOut = Math.min(maxPush * (Math.ceil(nOutMax / 8) * 8), Math.max(1 << 16, 2 * maxPush * quantum));
    this.handle = aggregatorCreate(maxSessions, rate, maxFrameQuanta, maxPush, this.maxOut);
    this.sess = new Uint32Array(maxPush);
    this.time = new Float64Array(maxPush);
    this.x = allocPinnedFloat32(maxPush * quantum);
    this.out = allocPinnedInt16(this.maxOut);
    this.fs = new Uint32Array(maxPush);
    this.fl = new Uint32Array(maxPush);
    this.fo = new Float64Array(maxPush);
    for (let s = maxSessions - 1; s >= 0; s--) this.free.push(s);
  }

  /** A slot for a new connection: empty buffer, lastSend = now (App.tsx:148-157).  -1 if all are taken. */
  open(nowMs: number): number {
    const slot = this.free.pop();
    if (slot === undefined) return -1;
    aggregatorReset(this.handle, slot, nowMs);
    return slot;
  }

  release(slot: number): void {
    if (slot >= 0) this.free.push(slot);
  }

  /** Stage the quanta of one message.  False if the tick is full. */
  add(slot: number, samples: Float32Array, nowMs: number): boolean {
    const k = samples.length / this.quantum;
    if (this.n + k > this.maxPush) return false;
    this.sess.fill(slot, this.n, this.n + k);
    this.time.fill(nowMs, this.n, this.n + k);
    this.x.set(samples, this.n * this.quantum);
    this.n += k;
    return true;
  }

  /** Run the staged quanta through the aggregator: [slot, frame bytes] in flush order (views, valid until the next push). */
  push(): [number, Buffer][] {
    const n = this.n;
    this.n = 0;
    return n ? this.pushRange(0, n, false) : [];
  }

  private pushRange(lo: number, hi: number, copy: boolean): [number, Buffer][] {
    const n = hi - lo;
    let nf: number;
    try {
      nf = aggregatorPush(this.handle, n, this.sess.subarray(lo, hi), this.time.subarray(lo, hi),
        this.x.subarray(lo * this.quantum, hi * this.quantum), this.out, this.fs, this.fl, this.fo);
    } catch (e: any) {
      if (e?.code !== "PCMB200_6" || n < 2) throw e; // 6 = capacity: a tick larger than the output block goes in halves
      const mid = lo + (n >> 1);
      return [...this.pushRange(lo, mid, true), ...this.pushRange(mid, hi, true)];
    }
    const frames: [number, Buffer][] = [];
    for (let f = 0; f < nf; f++) {
      const view = Buffer.from(this.out.buffer, this.out.byteOffset + this.fo[f] * 2, this.fl[f] * 2);
      frames.push([this.fs[f], copy ? Buffer.from(view) : view]);
    }
    return frames;
  }

  close(): void {
    aggregatorDestroy(this.handle);
  }
}

/** The "framing": "quanta" path (see pcm_frontend/ingest.py QuantumBatcher). */
export class QuantumBatcher {
  private aggs = new Map<number, GpuQuantumAggregator>();
  private conns = new Map<string, Connection>();
  private oldest: number | null = null;
  readonly stats = { quanta: 0, blocked: 0, protocol_errors: 0, no_slot: 0, frames: 0, flushes: 0 };

  constructor(private make: (rate: number) => GpuQuantumAggregator, readonly tickMs = 20, readonly quantum = 128) {}

  /** Returns false when the message did not fit the tick (back-pressure: ws.pause() until the next flush). */
  offer(conn: Connection, payload: Buffer, nowMs: number): boolean {
    if (conn.closed) return true;
    const nBytes = payload.length;
    if (nBytes === 0 || nBytes % (4 * this.quantum) !== 0) {
      conn.protocolErrors += 1;
      this.stats.protocol_errors += 1;
      return true;
    }
    let agg = this.aggs.get(conn.sampleRate);
    if (!agg) this.aggs.set(conn.sampleRate, (agg = this.make(conn.sampleRate)));
    if (conn.slot < 0) {
      conn.slot = agg.open(nowMs);
      if (conn.slot < 0) {
        this.stats.no_slot += 1;
        return true;
      }
      this.conns.set(`${conn.sampleRate}:${conn.slot}`, conn);
    }
    // a Float32Array needs 4-byte alignment; a Buffer from `ws` may sit at any offset of its pool: copy if it does
    const aligned = payload.byteOffset % 4 === 0 ? payload : Buffer.from(payload);
    if (!agg.add(conn.slot, new Float32Array(aligned.buffer, aligned.byteOffset, nBytes / 4), nowMs)) {
      this.stats.blocked += 1;
      return false;
    }
    this.stats.quanta += nBytes / (4 * this.quantum);
    if (this.oldest === null) this.oldest = nowMs;
    return true;
  }

  due(nowMs: number): boolean {
    return this.oldest !== null && nowMs - this.oldest >= this.tickMs;
  }

  flush(): number {
    let delivered = 0;
    for (const [rate, agg] of this.aggs) {
      for (const [slot, view] of agg.push()) {
        const conn = this.conns.get(`${rate}:${slot}`);
        if (conn && !conn.closed) {
          conn.sendAudio(view);
          delivered += 1;
        }
      }
    }
    this.oldest = null;
    this.stats.frames += delivered;
    this.stats.flushes += 1;
    return delivered;
  }

  /** The socket closed: its slot can serve another connection. */
  disconnect(conn: Connection): void {
    if (conn.slot >= 0) {
      this.conns.delete(`${conn.sampleRate}:${conn.slot}`);
      this.aggs.get(conn.sampleRate)?.release(conn.slot);
      conn.slot = -1;
    }
  }

  close(): void {
    for (const agg of this.aggs.values()) agg.close();
    this.aggs.clear();
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
