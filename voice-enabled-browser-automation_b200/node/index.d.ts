/** Same signature and result as the reference's helper (apps/web/src/App.tsx:7-16). */
export function floatTo16BitPCM(float32: Float32Array): Int16Array;
/** Same signature and result as the reference's helper (apps/web/src/App.tsx:18-32); returns `input` itself at 16 kHz. */
export function resampleTo16k(input: Float32Array, inputSampleRate: number): Float32Array;
/** floatTo16BitPCM(resampleTo16k(row, rate)) for `rows` dense rows of `nIn` samples (App.tsx:281-282), fused on the GPU. */
export function framesEncode(batch: Float32Array, rows: number, nIn: number, rate: number, out?: Int16Array): Int16Array;
/** Same on a worker thread: the event loop is not blocked; `batch` and `out` must not be touched until the promise settles. */
export function framesEncodeAsync(batch: Float32Array, rows: number, nIn: number, rate: number, out: Int16Array): Promise<Int16Array>;
/** Math.floor(n / (rate / 16000)) as the reference computes it (App.tsx:23-24); throws RangeError where `new Float32Array` would. */
export function resampleLen(n: number, rate: number): number;
/** Typed arrays over page-locked host memory (no staging copy on H2D / D2H). */
export function allocPinnedFloat32(length: number): Float32Array;
export function allocPinnedInt16(length: number): Int16Array;
/** The capture aggregator of App.tsx:148-157, :273-288 for many connections on the GPU (128-sample quanta, 60 ms). Returns a handle. */
export function aggregatorCreate(maxSessions: number, rate: number, maxFrameQuanta: number, maxPushQuanta: number, maxOutSamples: number): number;
/** Empty buffer, lastSend = nowMs for one session slot (a new connection). */
export function aggregatorReset(handle: number, slot: number, nowMs: number): void;
/** One push of n quanta (arrival order). Frame f of the result is out.subarray(frameOffset[f], frameOffset[f] + frameLen[f]) for session frameSession[f]; returns the number of frames. A rejected push (Error code "PCMB200_6": capacity) changes nothing. */
export function aggregatorPush(handle: number, n: number, sessions: Uint32Array, timesMs: Float64Array, samples: Float32Array, out: Int16Array,
  frameSession: Uint32Array, frameLen: Uint32Array, frameOffset: Float64Array): number;
export function aggregatorDestroy(handle: number): void;
