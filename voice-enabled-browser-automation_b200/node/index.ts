// Loads the native addon; there is no JavaScript fallback: if the addon or a CUDA device is
// missing, requiring this module throws.
import * as path from "path";

// eslint-disable-next-line @typescript-eslint/no-var-requires
const native = require(path.join(__dirname, "..", "lib", "pcm_b200_napi.node"));

export const floatTo16BitPCM: (float32: Float32Array) => Int16Array = native.floatTo16BitPCM;
export const resampleTo16k: (input: Float32Array, inputSampleRate: number) => Float32Array = native.resampleTo16k;
export const framesEncode: (batch: Float32Array, rows: number, nIn: number, rate: number, out?: Int16Array) => Int16Array =
  native.framesEncode;
export const framesEncodeAsync: (batch: Float32Array, rows: number, nIn: number, rate: number, out: Int16Array) => Promise<Int16Array> =
  native.framesEncodeAsync;
export const resampleLen: (n: number, rate: number) => number = native.resampleLen;
export const allocPinnedFloat32: (length: number) => Float32Array = native.allocPinnedFloat32;
export const allocPinnedInt16: (length: number) => Int16Array = native.allocPinnedInt16;
export const aggregatorCreate: (maxSessions: number, rate: number, maxFrameQuanta: number, maxPushQuanta: number, maxOutSamples: number) => number =
  native.aggregatorCreate;
export const aggregatorReset: (handle: number, slot: number, nowMs: number) => void = native.aggregatorReset;
export const aggregatorPush: (handle: number, n: number, sessions: Uint32Array, timesMs: Float64Array, samples: Float32Array, out: Int16Array,
  frameSession: Uint32Array, frameLen: Uint32Array, frameOffset: Float64Array) => number = native.aggregatorPush;
export const aggregatorDestroy: (handle: number) => void = native.aggregatorDestroy;
