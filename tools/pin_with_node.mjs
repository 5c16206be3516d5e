#!/usr/bin/env node
// pin_with_node.mjs -- pins the oracle to the REFERENCE'S OWN JavaScript.
//
//   node tools/pin_with_node.mjs [/path/to/reference] [--sweep] [--bench]
//
// Nothing of the reference is committed in this repository: at run time this script reads
// <reference>/apps/web/src/App.tsx, slices out the two module-private functions
//   floatTo16BitPCM  (App.tsx:7-16)      resampleTo16k  (App.tsx:18-32)
// by their `function NAME(` headers and matching braces, strips the four TypeScript type annotations,
// evaluates the text with `new Function`, and runs the reference's own code on the committed vectors:
//   tests/golden/encode_kat.json        value table, hazard set, 4096 random bit patterns  -> int16
//   tests/golden/index_kat.json         (rate, n_in) -> n_out and the SHA-256 of the pick indices
//   tests/golden/frames_kat_*.{f32,s16} whole frames: floatTo16BitPCM(resampleTo16k(row, rate))
//   tests/golden/encode_sweep_sha256.json   (--sweep) all 2^32 float32 bit patterns, SHA-256 of the int16 results
// Exit code 0 and "PARITY PINNED" only if every comparison is exact.  --bench additionally times
// config 1 (one 1 s 48 kHz buffer, BASELINE.json configs[0]) and prints ns per input sample.
//
// STATUS: there is no JavaScript engine in the build container or on the GPU boxes, so this script has
// never been executed by its authors; until someone runs it, parity stays "unpinned" (DESIGN.md 2).
import { createHash } from "node:crypto";
import { readFileSync } from "node:fs";
import { dirname, join } from "node:path";
import { fileURLToPath } from "node:url";

const ROOT = join(dirname(fileURLToPath(import.meta.url)), "..");
const args = process.argv.slice(2);
const refRoot = args.find((a) => !a.startsWith("--")) ?? "/root/reference";
const GOLDEN = join(ROOT, "tests", "golden");

// ---- slice the two functions out of the reference at run time --------------------------------------
function sliceFunction(src, name) {
  const start = src.indexOf(`function ${name}(`);
  if (start < 0) throw new Error(`function ${name} not found in App.tsx`);
  let i = src.indexOf("{", src.indexOf(")", start)); // the body's opening brace (after the parameter list)
  // the return-type annotation `): Int16Array {` sits between ')' and '{': indexOf("{") skips it
  let depth = 0;
  for (; i < src.length; i++) {
    if (src[i] === "{") depth++;
    else if (src[i] === "}" && --depth === 0) return src.slice(start, i + 1);
  }
  throw new Error(`unbalanced braces in ${name}`);
}
function stripTypes(fn) {
  // parameter and return annotations of the two signatures: `(float32: Float32Array): Int16Array`,
  // `(input: Float32Array, inputSampleRate: number): Float32Array`
  const open = fn.indexOf("("), close = fn.indexOf(")", open), brace = fn.indexOf("{", close);
  const params = fn.slice(open + 1, close).split(",").map((p) => p.split(":")[0].trim()).filter(Boolean).join(", ");
  return fn.slice(0, open + 1) + params + ") " + fn.slice(brace);
}
const appTsx = readFileSync(join(refRoot, "apps", "web", "src", "App.tsx"), "utf8");
const code = ["floatTo16BitPCM", "resampleTo16k"].map((n) => stripTypes(sliceFunction(appTsx, n))).join("\n");
const { floatTo16BitPCM, resampleTo16k } = new Function(`${code}\nreturn { floatTo16BitPCM, resampleTo16k };`)();

let failures = 0;
const fail = (msg) => { failures++; if (failures <= 20) console.error("MISMATCH:", msg); };

// ---- 1. encoder vectors ---------------------------------------------------------------------------
const enc = JSON.parse(readFileSync(join(GOLDEN, "encode_kat.json"), "utf8"));
const fromBits = (hexes) => {
  const u = new Uint32Array(hexes.map((h) => parseInt(h, 16)));
  return new Float32Array(u.buffer);
};
{
  const bits = enc.value_table.map((r) => r.bits);
  const got = floatTo16BitPCM(fromBits(bits));
  enc.value_table.forEach((r, i) => { if (got[i] !== r.int16) fail(`value table ${r.name}: ${got[i]} != ${r.int16}`); });
  for (const key of ["hazard", "random_bits"]) {
    const g = floatTo16BitPCM(fromBits(enc[key].bits));
    enc[key].int16.forEach((w, i) => { if (g[i] !== w) fail(`${key}[${i}] bits ${enc[key].bits[i]}: ${g[i]} != ${w}`); });
  }
  console.log(`encoder vectors: ${enc.value_table.length + enc.hazard.bits.length + enc.random_bits.bits.length} checked`);
}

// ---- 2. pick indices: resampleTo16k on a ramp returns the indices themselves ---------------------------
{
  const idx = JSON.parse(readFileSync(join(GOLDEN, "index_kat.json"), "utf8")).cases;
  for (const c of idx) {
    const ramp = new Float32Array(c.n_in);
    for (let i = 0; i < c.n_in; i++) ramp[i] = i;            // exact in float32 up to 2^24 > every n_in here
    const out = resampleTo16k(ramp, c.rate);
    if (c.identity) { if (out !== ramp) fail(`rate ${c.rate}: identity must return the same object`); continue; }
    if (out.length !== c.n_out) { fail(`rate ${c.rate} n_in ${c.n_in}: length ${out.length} != ${c.n_out}`); continue; }
    const u = new Uint32Array(out.length);
    for (let i = 0; i < out.length; i++) u[i] = out[i];
    const h = createHash("sha256").update(Buffer.from(u.buffer)).digest("hex");
    if (h !== c.sha256_u32le) fail(`rate ${c.rate} n_in ${c.n_in}: pick-index digest differs`);
  }
  console.log(`pick-index tables: ${idx.length} checked`);
}

// ---- 3. whole frames ---------------------------------------------------------------------------------
{
  const raw = JSON.parse(readFileSync(join(GOLDEN, "frames_kat_raw.json"), "utf8")).cases;
  const bi = readFileSync(join(GOLDEN, "frames_kat_in.f32")), bo = readFileSync(join(GOLDEN, "frames_kat_out.s16"));
  const fin = new Float32Array(bi.buffer.slice(bi.byteOffset, bi.byteOffset + bi.byteLength));
  const fout = new Int16Array(bo.buffer.slice(bo.byteOffset, bo.byteOffset + bo.byteLength));
  let rows = 0;
  for (const c of raw)
    for (let r = 0; r < c.rows; r++, rows++) {
      const x = fin.slice(c.in_offset + r * c.n_in, c.in_offset + (r + 1) * c.n_in);
      const got = floatTo16BitPCM(resampleTo16k(x, c.rate));           // App.tsx:281-282
      if (got.length !== c.n_out) { fail(`${c.key} row ${r}: length ${got.length} != ${c.n_out}`); continue; }
      for (let i = 0; i < c.n_out; i++)
        if (got[i] !== fout[c.out_offset + r * c.n_out + i]) { fail(`${c.key} row ${r} sample ${i}`); break; }
    }
  console.log(`whole frames: ${rows} rows checked`);
}

// ---- 4. all 2^32 bit patterns (optional: ~1 minute) ------------------------------------------------------
if (args.includes("--sweep")) {
  const want = JSON.parse(readFileSync(join(GOLDEN, "encode_sweep_sha256.json"), "utf8")).sha256;
  const h = createHash("sha256");
  const CH = 1 << 24, u = new Uint32Array(CH), f = new Float32Array(u.buffer);
  for (let c = 0; c < 256; c++) {
    const base = c * CH;
    for (let i = 0; i < CH; i++) u[i] = base + i;
    h.update(Buffer.from(floatTo16BitPCM(f).buffer));
  }
  const got = h.digest("hex");
  if (got !== want) fail(`2^32 sweep digest ${got} != ${want}`);
  console.log("2^32-pattern sweep:", got === want ? "digest matches" : "DIGEST DIFFERS");
}

// ---- config 1 timing (BASELINE.json configs[0]) -------------------------------------------------------------
if (args.includes("--bench")) {
  const n = 48000, x = new Float32Array(n);
  for (let i = 0; i < n; i++) x[i] = 0.5 * Math.sin((2 * Math.PI * 440 * i) / 48000);
  for (let k = 0; k < 200; k++) floatTo16BitPCM(resampleTo16k(x, 48000));
  const t = [];
  for (let k = 0; k < 1000; k++) {
    const t0 = process.hrtime.bigint();
    floatTo16BitPCM(resampleTo16k(x, 48000));
    t.push(Number(process.hrtime.bigint() - t0));
  }
  t.sort((a, b) => a - b);
  console.log(JSON.stringify({ config: 1, runtime: `node ${process.version}`, n_in: n, median_ns_per_call: t[500], ns_per_input_sample: t[500] / n }));
}

if (failures) {
  console.error(`${failures} mismatches: the oracle does NOT reproduce the reference`);
  process.exit(1);
}
console.log("PARITY PINNED: the committed vectors are exactly what the reference's own JavaScript produces");
