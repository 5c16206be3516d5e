// pcm_kernels.cuh -- device code of the B200 PCM front-end (sm_100a).
//
// Everything here is HBM-bound streaming work: one pass over Float32 frames in, Int16 frames
// out, 4*N_in + 2*N_out algorithmic bytes per frame and a handful of ALU ops per sample.  No
// tensor cores, no shared-memory reuse (the reference's decimator has zero taps: it PICKS
// samples, App.tsx:27-30), so the design rules are: 256-bit / 128-bit fully coalesced global
// accesses, enough independent loads in flight per SM to cover HBM latency, streaming cache
// hints so 4.7 GB ticks do not thrash L2, and (measured, see grid_for in pcm_b200.cu) one tile per
// block handed out by the hardware scheduler rather than a persistent grid-stride loop -- the
// loops below remain so that a capped grid still covers all tiles.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pcmb200 {

// ---- 256-bit global load (sm_100+: LDG.E.256), read-only path, no L1 allocation, L2 evict-first
struct __align__(32) f32x8 { float v[8]; };

__device__ __forceinline__ f32x8 ld_stream_256(const float *p)
{
    f32x8 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]),
                   "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}

// 128-bit store; `hint` is a launch-uniform tuning knob (0 default write-back, 1 streaming .cs,
// 2 L1::no_allocate).
__device__ __forceinline__ void st_128(int16_t *p, uint4 v, int hint)
{
    if (hint == 1)
        asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    else if (hint == 2)
        asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    else
        *reinterpret_cast<uint4 *>(p) = v;
}

// ---- the encoder (App.tsx:10-14), bit-exact without any binary64 arithmetic ------------------
// JS computes v = s * 0x7fff (or 0x8000) exactly-rounded in binary64 -- for a float32 s the
// product has <= 39 significant bits, so it is EXACT -- and the Int16Array store truncates toward
// zero.  A round-to-nearest fp32 product is wrong (it can round up to an integer the exact product
// never reaches).  A round-TOWARD-ZERO fp32 product is right: every integer below 2^24 is an fp32
// value, so rounding the exact product toward zero can never cross an integer, hence
// trunc(RZ32(s*c)) == trunc(s*c).  The clamp to [-1,1] commutes with the (monotonic) scaling and
// becomes a saturating s32->s16 pack; cvt.rzi maps NaN to 0 and +-Inf to INT_MAX/INT_MIN, which
// the saturating pack turns into 32767 / -32768, exactly what clamp-then-scale gives.
// Checked exhaustively over all 2^32 float32 bit patterns (tests/test_gpu_exhaustive.py).
// Must NOT be compiled with --use_fast_math / -ftz=true (denormal inputs must not flush).
__device__ __forceinline__ int enc_s32(float s)
{
    const float c = (s < 0.0f) ? 32768.0f : 32767.0f;   // NaN and -0 take the 0x7fff side, like JS `s < 0`
    return __float2int_rz(__fmul_rz(s, c));
}
__device__ __forceinline__ uint32_t pack_s16x2_sat(int lo, int hi)
{
    uint32_t d;
    asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(d) : "r"(hi), "r"(lo));   // d = {sat16(hi), sat16(lo)}
    return d;
}
__device__ __forceinline__ int16_t enc_s16(float s)
{
    return (int16_t)(pack_s16x2_sat(enc_s32(s), 0) & 0xFFFFu);
}

// ---- kernel 1: integer-ratio decimate + encode, 256-bit loads / 128-bit stores ----------------
// One work item ("chunk") = 8 consecutive output samples of one row = 8*R consecutive input floats
// (R = in_rate / 16000, an integer; R = 1 is encode-only).  A thread reads the chunk with R
// 256-bit loads (each warp-level load touches 32 whole 32-byte sectors), keeps every R-th float
// (the running sum idx += R is exact, so floor(idx) = i*R, App.tsx:27-29), encodes and writes one
// 128-bit store; a warp's stores cover 512 contiguous bytes.  U chunks per thread are issued
// before the first use, so R*U*32 bytes per thread are in flight.
//   DENSE: rows are back to back with n_in == 8*R*cpr, so chunk w starts at in + w*8*R and the
//   whole batch is one flat stream (no division).  Otherwise row = w / cpr.
// Requirements (checked by the launcher): in 32-byte aligned, pitch_in % 8 == 0, out 16-byte
// aligned, pitch_out % 8 == 0.  The n_out % 8 tail of each row is done by scalar code at the end.
template <int R, int U, bool DENSE>
__global__ void __launch_bounds__(512)
k_decim_int(const float *__restrict__ in, size_t pitch_in, int16_t *__restrict__ out, size_t pitch_out,
            uint32_t rows, uint32_t cpr, uint32_t tail, unsigned long long n_chunks, int hint)
{
    const unsigned long long tile_chunks = (unsigned long long)blockDim.x * U;
    const unsigned long long n_tiles = (n_chunks + tile_chunks - 1) / tile_chunks;

    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned long long w0 = tile * tile_chunks + threadIdx.x;
        const bool full = (tile + 1) * tile_chunks <= n_chunks;
        f32x8 v[U][R];
        int16_t *dst[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const unsigned long long w = w0 + (unsigned long long)u * blockDim.x;
            dst[u] = nullptr;
            if (full || w < n_chunks) {
                const float *src;
                if (DENSE) {
                    src = in + w * (8ull * R);
                    dst[u] = out + w * 8ull;
                } else {
                    const uint32_t w32 = (uint32_t)w;
                    const uint32_t r = w32 / cpr, c = w32 - r * cpr;
                    src = in + (size_t)r * pitch_in + (size_t)c * (8u * R);
                    dst[u] = out + (size_t)r * pitch_out + (size_t)c * 8u;
                }
#pragma unroll
                for (int k = 0; k < R; k++) v[u][k] = ld_stream_256(src + 8 * k);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (full || dst[u] != nullptr) {
                int e[8];
#pragma unroll
                for (int j = 0; j < 8; j++) e[j] = enc_s32(v[u][(j * R) / 8].v[(j * R) % 8]);
                uint4 o;
                o.x = pack_s16x2_sat(e[0], e[1]);
                o.y = pack_s16x2_sat(e[2], e[3]);
                o.z = pack_s16x2_sat(e[4], e[5]);
                o.w = pack_s16x2_sat(e[6], e[7]);
                st_128(dst[u], o, hint);
            }
        }
    }

    if (tail) {   // the last n_out % 8 samples of every row
        const unsigned long long total = (unsigned long long)rows * tail;
        for (unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
             t += (unsigned long long)gridDim.x * blockDim.x) {
            const size_t r = (size_t)(t / tail);
            const uint32_t o = cpr * 8u + (uint32_t)(t - (unsigned long long)r * tail);
            out[r * pitch_out + o] = enc_s16(__ldg(in + r * pitch_in + (size_t)o * R));
        }
    }
}

// ---- kernel 1b: integer-ratio decimate + encode for DENSE batches, fully coalesced -------------
// Same arithmetic as kernel 1, different data movement.  In kernel 1 a lane reads its own 8*R
// consecutive floats, so for R > 1 one warp-level load touches 32-byte pieces 32*R bytes apart
// and every 128-byte line is requested R times (with a quarter of its sectors each time).  Here a
// WARP owns 256 consecutive outputs = 256*R consecutive floats and reads them with R warp-level
// 256-bit loads that are each 1024 contiguous bytes (one request per line, all four sectors).  A
// lane then holds 8*R floats of which exactly 8 are picks, but not 8 consecutive ones, so the
// encoded int16 go through a 512-byte per-warp shared-memory tile and come back as one 128-bit,
// fully coalesced store per lane.  Needs the batch to be one flat stream: rows back to back,
// n_in == R * n_out (then the picks are the floats whose flat index is a multiple of R).
template <int R, int U>
__global__ void __launch_bounds__(512)
k_decim_coal(const float *__restrict__ in, int16_t *__restrict__ out, unsigned long long n_tiles, int hint)
{
    extern __shared__ __align__(16) int16_t smem_tiles[];
    // Programmatic dependent launch: when the NEXT launch in the stream carries the programmatic
    // serialisation attribute (tuning knob "overlap_ticks"), its blocks may start as soon as every block
    // of this grid has got this far, i.e. while this grid's last wave is still draining -- the tail of one
    // tick is filled with the head of the next.  Ticks are independent frames, nothing is waited for.
    // Without the attribute this is a no-op.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    int16_t *my = smem_tiles + (size_t)warp * (U * 256);
    constexpr int M = (8 + R - 1) / R;   // max picks in one 8-float group

    const unsigned long long stride = (unsigned long long)gridDim.x * warps_per_block * U;
    for (unsigned long long t0 = ((unsigned long long)blockIdx.x * warps_per_block + warp) * U; t0 < n_tiles; t0 += stride) {
        f32x8 v[U][R];
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (t0 + u < n_tiles) {
                const float *src = in + (t0 + u) * (256ull * R) + lane * 8;
#pragma unroll
                for (int k = 0; k < R; k++) v[u][k] = ld_stream_256(src + 256 * k);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (t0 + u < n_tiles) {
#pragma unroll
                for (int k = 0; k < R; k++) {
                    const int b = 256 * k + lane * 8;          // first float of this group within the tile
                    const int e0 = (R - b % R) % R;            // offset of the first pick in the group
                    const int j0 = (b + e0) / R;               // its output index within the tile
#pragma unroll
                    for (int m = 0; m < M; m++) {
                        float x = 0.0f;
                        bool have = false;
#pragma unroll
                        for (int e = 0; e < R; e++)
                            if (e + m * R < 8 && e0 == e) { x = v[u][k].v[(e + m * R) & 7]; have = true; }
                        if (have) my[u * 256 + j0 + m] = enc_s16(x);
                    }
                }
            }
        }
        __syncwarp();
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (t0 + u < n_tiles) {
                const uint4 o = *reinterpret_cast<const uint4 *>(my + u * 256 + lane * 8);
                st_128(out + (t0 + u) * 256ull + lane * 8, o, hint);
            }
        }
        __syncwarp();
    }
}

// ---- kernel 2: general pick + encode (any ratio, any alignment) -----------------------------------
// For rates whose ratio is not an integer (44.1 kHz family: the reference's running binary64 sum
// drifts from the rational index) the pick indices come from a table built on the host
// with the reference's own sequential loop and shared by every row of the batch (`lut`, n_out
// entries, L1/L2 resident).  lut == nullptr means idx = o * stride (integer ratio, used when the
// buffers do not meet kernel 1's alignment rules).
// A work item is a PAIR of adjacent outputs (o, o+1) of one row; lane i of a warp takes pair
// w0 + i, so one warp-level gather touches a ~64*ratio*4-byte contiguous span (every sector of it
// is used) and one warp-level store writes 128 contiguous bytes.  U pairs per thread in flight.
//   PAIR32: out rows are 4-byte aligned (base % 4 == 0, pitch_out even) -> 32-bit stores;
//   otherwise 16-bit stores.
template <int U, bool PAIR32>
__global__ void __launch_bounds__(512)
k_pick_encode(const float *__restrict__ in, size_t pitch_in, int16_t *__restrict__ out, size_t pitch_out,
              uint32_t n_in, uint32_t n_out, const uint32_t *__restrict__ lut, uint32_t stride,
              unsigned long long n_pairs)
{
    const uint32_t ppr = (n_out + 1u) >> 1;   // pairs per row
    const unsigned long long tile_pairs = (unsigned long long)blockDim.x * U;
    const unsigned long long n_tiles = (n_pairs + tile_pairs - 1) / tile_pairs;

    for (unsigned long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const unsigned long long w0 = tile * tile_pairs + threadIdx.x;
        float a[U], b[U];
        int16_t *dst[U];
        bool two[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const unsigned long long w = w0 + (unsigned long long)u * blockDim.x;
            dst[u] = nullptr;
            two[u] = false;
            a[u] = b[u] = 0.0f;
            if (w < n_pairs) {
                const unsigned long long r = w / ppr;
                const uint32_t o = (uint32_t)(w - r * ppr) * 2u;
                const float *row = in + (size_t)r * pitch_in;
                dst[u] = out + (size_t)r * pitch_out + o;
                two[u] = (o + 1u < n_out);
                uint32_t i0, i1 = 0xFFFFFFFFu;
                if (lut) {
                    i0 = __ldg(lut + o);
                    if (two[u]) i1 = __ldg(lut + o + 1);
                } else {
                    i0 = o * stride;
                    i1 = i0 + stride;
                }
                // a pick outside the input reads `undefined` in JS -> NaN -> 0 after the encoder
                a[u] = (i0 < n_in) ? __ldg(row + i0) : __int_as_float(0x7FC00000);
                b[u] = (two[u] && i1 < n_in) ? __ldg(row + i1) : __int_as_float(0x7FC00000);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (dst[u] != nullptr) {
                const uint32_t p = pack_s16x2_sat(enc_s32(a[u]), enc_s32(b[u]));
                if (PAIR32 && two[u]) {
                    *reinterpret_cast<uint32_t *>(dst[u]) = p;
                } else {
                    dst[u][0] = (int16_t)(p & 0xFFFFu);
                    if (two[u]) dst[u][1] = (int16_t)(p >> 16);
                }
            }
        }
    }
}

// ---- kernel 2b: tile-staged pick + encode for DENSE batches of short frames (any ratio) -----------
// The 44.1 kHz family needs table-driven picks at irregular offsets; doing that straight from global
// memory (kernel 2) costs sector-granular gathers.  Here a CTA walks over tiles of K whole frames
// (K * n_in * 4 bytes, contiguous because rows are dense).  One elected thread moves each tile
// global -> shared with a single bulk async copy (cp.async.bulk, the TMA engine's 1-D mode,
// SASS UBLKCP) that completes on an mbarrier; STAGES tiles are in flight per CTA, so the copy of
// tile i+1 runs while tile i is gathered.  All 256 threads then read their picks from shared
// memory through a copy of the pick table (also in shared memory), encode, and write one 128-bit
// streaming store per 8 outputs; consecutive lanes write consecutive 16-byte chunks.
// Requirements (launcher): in and out 16-byte aligned, rows dense (pitch_in == n_in, pitch_out ==
// n_out), K * n_in % 4 == 0, tile + table fit in shared memory.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Waits for the phase with the given parity.  try_wait suspends the thread in hardware for a while per
// attempt, so the loop is not a busy spin.  A pipeline bug here would otherwise hang the GPU until a
// watchdog somewhere else kills the process: after ~4 million failed attempts (seconds at most, against waits that
// are microseconds) the kernel traps, which surfaces as a CUDA error on the host instead of a hang.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    const uint32_t addr = smem_u32(bar);
    for (uint32_t tries = 0;; tries++) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x2000;\n"   // suspend up to ~8 us per attempt instead of spinning
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if (tries > (1u << 22)) asm volatile("trap;");
    }
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// Level meter of the reference's capture handler (App.tsx:267-271), fused into kernel 2b: while a
// tile of frames sits in shared memory, every whole `quantum` (128 samples, the AudioWorklet
// render quantum) of every frame gets rms = sqrt(sum(x*x) / quantum) in binary64, accumulated in
// sample order (x*x is exact in binary64, so one DFMA per sample reproduces JS's multiply-then-add
// rounding), and each session's level follows level = max(rms, level * 0.85) over its quanta
// (Math.max: NaN if either side is NaN).  0 extra DRAM read bytes; 8 bytes written per session
// (+ 8 per quantum if the per-quantum trace is requested).
struct MeterArgs {
    uint32_t quantum;     // samples per quantum (128)
    uint32_t nq;          // whole quanta per frame = n_in / quantum
    double *state;        // [rows] in/out: level of each session
    double *levels;       // [rows][nq] level after every quantum, or nullptr
};

// |x| widened to binary64 WITHOUT the F2F.F64.F32 conversion: on B200 that instruction runs on the
// XU pipe at 16 lanes/clk/SM, and at one conversion per input sample it (not HBM) bounded the
// fused meter (ncu: XU 112 % busy, FP64 13 %).  For a normal float the binary64 image is a pure
// re-packing of the same bits: exponent re-biased by 1023-127 = 896, 23 mantissa bits moved to the
// top of the 52.  Zero stays zero; denormals, Inf and NaN (rare in audio) take the real conversion.
// The sign is dropped because only x*x is needed.
__device__ __forceinline__ double widen_abs(float x)
{
    // branch-free (a branch per sample costs more than the arithmetic: BSSY/BSYNC pairs serialise
    // the four-sample step).  Denormals are first scaled by 2^24 (exact) into the normal range.
    const uint32_t a0 = __float_as_uint(x) & 0x7FFFFFFFu;
    const bool small = a0 < 0x00800000u;
    const uint32_t a = small ? __float_as_uint(__uint_as_float(a0) * 16777216.0f) : a0;   // |x| (* 2^24)
    uint32_t hi = (a >> 3) + (small ? 0x38000000u - (24u << 20) : 0x38000000u);
    if (a >= 0x7F800000u) hi = 0x7FF00000u | ((a & 0x007FFFFFu) >> 3);                     // Inf / NaN
    if (a == 0u) hi = 0u;                                                                   // +-0
    return __hiloint2double((int)hi, (int)(a << 29));
}

// sqrt(sum / quantum): for a power-of-two quantum the division is an exact scaling, so the (software)
// IEEE division is replaced by one multiplication with the same result.
__device__ __forceinline__ double rms_of(double sum, uint32_t quantum)
{
    const bool pow2 = (quantum & (quantum - 1u)) == 0u;
    const double inv = __hiloint2double((int)((1023u - (31u - (uint32_t)__clz((int)quantum))) << 20), 0);   // 2^-log2(quantum)
    const double mean = pow2 ? sum * inv : __ddiv_rn(sum, (double)quantum);
    return __dsqrt_rn(mean);
}

// The cheap form used in the meter's inner loop (the loop is issue-bound on the ALU pipe, so every
// instruction counts, and the shifts are written so that they land on the FMA pipe as IMADs):
//   * normal |x| >= 2^-60: exact re-packing;
//   * zero: comes out as 2^-127, i.e. adds 2^-254 to the sum instead of 0.  That cannot change a
//     correctly rounded sum that already holds a term >= 2^-120 (2^-254 is far below half an ulp),
//     and before the first such term the polluted partial sum (< 2^-246) is absorbed the same way
//     when the term arrives; a quantum of nothing but zeros ends below 2^-200 and is set to 0;
//   * 0 < |x| < 2^-60 (denormals included) would break that argument: flagged in `odd`;
//   * Inf / NaN come out finite but >= 2^128, pushing the sum to >= 2^250: tested once per quantum.
// A flagged quantum is redone with widen_abs.
__device__ __forceinline__ double widen_abs_fast(float x, bool &odd)
{
    const uint32_t b = __float_as_uint(x);
    const uint32_t t = b << 1;                                   // 2 * |x| bits: the sign is gone
    odd |= (t - 2u) < (2u * 0x21800000u - 2u);                   // 0 < |x| < 2^-60
    const uint32_t hi = __umulhi(t, 0x10000000u) + 0x38000000u;  // (|x| bits >> 3) + re-bias
    return __hiloint2double((int)hi, (int)(b << 29));
}

__device__ __forceinline__ double js_max(double a, double b)
{
    return (a != a || b != b) ? __longlong_as_double(0x7FF8000000000000ll) : (a > b ? a : b);
}


// rms = sqrt(sum(x*x) / Q) of one quantum that sits in shared memory (16-byte aligned), accumulated in
// sample order by ONE lane: LDS.128 + 4 x (2 IMAD, LEA.HI, VIADDMNMX, DFMA), fully unrolled for the
// 128-sample quantum.  widen_abs_fast's two escape hatches are tested once per quantum: a nonzero
// sample below 2^-60 (the running minimum of `2|x| bits - 2`, where zeros wrap to the top) or an
// Inf / NaN (an impossible sum) send the quantum through the exact widening instead.
__device__ __forceinline__ double quantum_rms(const float4 *src4, uint32_t Q)
{
    double sum = 0.0;
    uint32_t lowest = 0xFFFFFFFFu;
    auto four = [&](const float4 v) {
        const float xs[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t b = __float_as_uint(xs[j]);
            const uint32_t t2 = b << 1;                                       // sign gone
            lowest = min(lowest, t2 - 2u);
            const uint32_t hi = __umulhi(t2, 0x10000000u) + 0x38000000u;      // (|x| bits >> 3) + re-bias
            const double x = __hiloint2double((int)hi, (int)(b << 29));
            sum = __fma_rn(x, x, sum);
        }
    };
    if (Q == 128u) {
#pragma unroll 8
        for (uint32_t g = 0; g < 32u; g++) four(src4[g]);
    } else {
        for (uint32_t g = 0; g < (Q >> 2); g++) four(src4[g]);
    }
    bool odd = lowest < (2u * 0x21800000u - 2u);                              // 0 < |x| < 2^-60 somewhere
    odd |= !(sum < __hiloint2double(0x4F900000, 0));                          // >= 2^250 or NaN: an Inf / NaN sample
    if (sum < __hiloint2double(0x33700000, 0)) sum = 0.0;                     // < 2^-200: every sample was zero
    if (odd) {
        sum = 0.0;
        const float *src = reinterpret_cast<const float *>(src4);
        for (uint32_t k = 0; k < Q; k++) {
            const double x = widen_abs(src[k]);
            sum = __fma_rn(x, x, sum);
        }
    }
    return rms_of(sum, Q);
}

// The same quantity with the hardware conversion (F2F.F64.F32, exact for every float32: no escape hatches, 2 instead
// of 6 instructions per sample).  The conversion runs on the XU pipe (16 lanes/clk/SM), which kernels 2b-2d cannot
// afford -- there every sample of a batch that streams at HBM speed passes through it -- but the aggregator's kernels
// meter with ONE warp per frame while the copies are what bounds them: ncu showed their issue slots, not the XU pipe,
// as the cost of the integer widening (6 of every 6 instructions in that warp, the other warps waiting for it).
// Measured in kernel 2d as well: all samples this way 0.97 / 1.02 of peak (2944- / 960-sample frames) against 0.95 /
// 1.03 with the integer form, and a mix of the two gains nothing once the converted samples are watched for the
// tiny-value hazard too (which the zero pollution of the integer form makes necessary) -- so 2d keeps the integer form.
__device__ __forceinline__ double quantum_rms_cvt(const float4 *src4, uint32_t Q)
{
    double sum = 0.0;
    auto four = [&](const float4 v) {
        const double a = (double)v.x, b = (double)v.y, c = (double)v.z, d = (double)v.w;
        sum = __fma_rn(a, a, sum);
        sum = __fma_rn(b, b, sum);
        sum = __fma_rn(c, c, sum);
        sum = __fma_rn(d, d, sum);
    };
    if (Q == 128u) {
#pragma unroll 8
        for (uint32_t g = 0; g < 32u; g++) four(src4[g]);
    } else {
        for (uint32_t g = 0; g < (Q >> 2); g++) four(src4[g]);
    }
    return rms_of(sum, Q);
}

// One step of App.tsx:271, level = Math.max(rms, level * 0.85): NaN if either side is NaN, and a NaN
// level stays NaN.  NaN-ness is tracked with integer tests so a step costs one DMUL and one DSETP on
// the long-latency FP64 pipe.
__device__ __forceinline__ void level_step(double &lv, bool &poisoned, double rms)
{
    const uint32_t hi = (uint32_t)__double2hiint(rms) & 0x7FFFFFFFu;
    poisoned |= hi > 0x7FF00000u || (hi == 0x7FF00000u && __double2loint(rms) != 0);
    const double dec = __dmul_rn(lv, 0.85);
    lv = rms > dec ? rms : dec;
    if (poisoned) lv = __longlong_as_double(0x7FF8000000000000ll);
}

template <int STAGES, bool METER>
__global__ void __launch_bounds__(256)
k_tile_pick(const float *__restrict__ in, int16_t *__restrict__ out, uint32_t n_in, uint32_t n_out,
            uint32_t rows_per_tile, unsigned long long n_tiles, const uint32_t *__restrict__ lut, uint32_t stride,
            uint32_t tile_stride_bytes, int hint, MeterArgs meter)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [STAGES][tile_stride_bytes] | table u32[n_out] (only if lut) | mbarriers
    float *tiles = reinterpret_cast<float *>(smem_raw);
    uint32_t *tab = reinterpret_cast<uint32_t *>(smem_raw + (size_t)STAGES * tile_stride_bytes);
    const uint32_t tab_bytes = lut ? ((n_out * 4u + 15u) & ~15u) : 0u;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + (size_t)STAGES * tile_stride_bytes + tab_bytes);
    double *rms_s = reinterpret_cast<double *>(bars + ((STAGES + 1) & ~1));   // METER: [rows_per_tile * nq]

    const uint32_t tile_bytes = rows_per_tile * n_in * 4u;
    const uint32_t cpr = n_out >> 3;                             // 8-output chunks per row (rows_aligned)
    const bool rows_aligned = (n_out & 7u) == 0u;
    const uint32_t outs_per_tile = rows_per_tile * n_out;

    __shared__ float4 zero_granule;   // METER: what a lane reads before it starts / after it finishes
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        zero_granule = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (lut)
        for (uint32_t i = threadIdx.x; i < n_out; i += blockDim.x) tab[i] = __ldg(lut + i);
    __syncthreads();

    // prologue: fill the pipeline
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) {
            const unsigned long long t = blockIdx.x + (unsigned long long)s * gridDim.x;
            if (t < n_tiles) {
                mbar_expect_tx(&bars[s], tile_bytes);
                bulk_g2s(reinterpret_cast<unsigned char *>(tiles) + (size_t)s * tile_stride_bytes,
                         reinterpret_cast<const unsigned char *>(in) + t * (unsigned long long)tile_bytes, tile_bytes, &bars[s]);
            }
        }
    }

    uint32_t it = 0;
    for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x, it++) {
        const int s = it % STAGES;
        const uint32_t parity = (it / STAGES) & 1u;
        mbar_wait(&bars[s], parity);
        const float *tile = reinterpret_cast<const float *>(reinterpret_cast<const unsigned char *>(tiles) + (size_t)s * tile_stride_bytes);
        // Output of this tile: outs_per_tile consecutive int16 starting at `first`.  When frames are
        // a whole number of 8-sample chunks the tile starts 16-byte aligned; otherwise (the
        // reference's real frames: 23 quanta at 48 kHz -> 981 outputs) the aligned body is written
        // with 128-bit stores and the few samples before / after it one by one.
        const unsigned long long first = t * (unsigned long long)outs_per_tile;
        const unsigned long long body0 = (first + 7ull) & ~7ull;
        const unsigned long long body1 = (first + outs_per_tile) & ~7ull;
        const uint32_t n_body = body1 > body0 ? (uint32_t)((body1 - body0) >> 3) : 0u;
        const uint32_t head = (uint32_t)(body0 - first);
        for (uint32_t g = threadIdx.x; g < n_body; g += blockDim.x) {
            uint32_t idx[8];     // offset of each pick inside the tile (0xFFFFFFFF: outside its frame)
            if (rows_aligned) {
                const uint32_t rr = g / cpr, c = g - rr * cpr;
                const uint32_t base = rr * n_in;
                if (lut) {
                    const uint4 a = *reinterpret_cast<const uint4 *>(tab + 8u * c);
                    const uint4 b = *reinterpret_cast<const uint4 *>(tab + 8u * c + 4u);
                    idx[0] = a.x; idx[1] = a.y; idx[2] = a.z; idx[3] = a.w;
                    idx[4] = b.x; idx[5] = b.y; idx[6] = b.z; idx[7] = b.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; j++) idx[j] = (8u * c + j) * stride;
                }
#pragma unroll
                for (int j = 0; j < 8; j++) idx[j] = idx[j] < n_in ? base + idx[j] : 0xFFFFFFFFu;
            } else {
                // a chunk is 8 consecutive outputs of the tile's flat output; it may run over a frame boundary
                const uint32_t f = head + 8u * g;
                uint32_t rr = f / n_out, o = f - rr * n_out;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const uint32_t i = lut ? tab[o] : o * stride;
                    idx[j] = i < n_in ? rr * n_in + i : 0xFFFFFFFFu;
                    if (++o == n_out) { o = 0; rr++; }
                }
            }
            int e[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                // a pick outside the frame reads `undefined` in JS -> NaN -> 0
                const float x = (idx[j] != 0xFFFFFFFFu) ? tile[idx[j]] : __int_as_float(0x7FC00000);
                e[j] = enc_s32(x);
            }
            uint4 o4;
            o4.x = pack_s16x2_sat(e[0], e[1]);
            o4.y = pack_s16x2_sat(e[2], e[3]);
            o4.z = pack_s16x2_sat(e[4], e[5]);
            o4.w = pack_s16x2_sat(e[6], e[7]);
            st_128(out + body0 + (size_t)g * 8u, o4, hint);
        }
        if (!rows_aligned) {   // the < 8 + < 8 samples around the aligned body
            const uint32_t tail0 = (uint32_t)(body1 > body0 ? body1 - first : head), n_edge = head + (outs_per_tile - tail0);
            for (uint32_t k = threadIdx.x; k < n_edge && body1 > body0; k += blockDim.x) {
                const uint32_t f = k < head ? k : tail0 + (k - head);
                const uint32_t rr = f / n_out, o = f - rr * n_out;
                const uint32_t i = lut ? tab[o] : o * stride;
                out[first + f] = enc_s16(i < n_in ? tile[rr * n_in + i] : __int_as_float(0x7FC00000));
            }
            if (body1 <= body0)   // degenerate: the whole tile is shorter than one aligned chunk
                for (uint32_t f = threadIdx.x; f < outs_per_tile; f += blockDim.x) {
                    const uint32_t rr = f / n_out, o = f - rr * n_out;
                    const uint32_t i = lut ? tab[o] : o * stride;
                    out[first + f] = enc_s16(i < n_in ? tile[rr * n_in + i] : __int_as_float(0x7FC00000));
                }
        }
        if (METER) {
            // the session levels are needed only after the meter: fetch them now so the global-memory
            // latency hides behind it (threads 0..rows_per_tile-1 own one session of the tile each)
            double lv_in = 0.0;
            if (threadIdx.x < rows_per_tile) lv_in = meter.state[t * rows_per_tile + threadIdx.x];
            // One lane per quantum, strictly in sample order (binary64 rounding depends on it).  All
            // quanta start 128 floats apart, i.e. in the same bank, so lanes are SKEWED: at any step
            // the lanes that share a shared-memory access phase read different banks, while each
            // lane still walks its own quantum front to back.
            const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
            const uint32_t quanta = rows_per_tile * meter.nq;
            if ((n_in & 3u) == 0u && (meter.quantum & 3u) == 0u) {
                // 128-bit path: a quantum is G granules of 4 floats; an LDS.128 is served 8 lanes at a
                // time, so a skew of (lane & 7) granules is enough.  The next granule is fetched while
                // the current one goes through its four dependent DFMAs.
                const uint32_t sub = lane & 7u;
                const uint32_t octet = warp * 4u + (lane >> 3), octets = nwarps * 4u;   // packed: the loop is issue-bound
                const uint32_t G = meter.quantum >> 2;
                for (uint32_t base = octet * 8u; base < quanta; base += octets * 8u) {
                    const uint32_t qg = base + sub;
                    const bool valid = qg < quanta;
                    const uint32_t rr = valid ? qg / meter.nq : 0u, q = valid ? qg - rr * meter.nq : 0u;
                    const float4 *src4 = reinterpret_cast<const float4 *>(tile + rr * n_in + q * meter.quantum);
                    float4 cur = *((valid && sub == 0u) ? src4 : &zero_granule);
                    double sum = 0.0;
                    bool odd = false;
                    // No branches inside: a lane that has not started or has finished reads a granule
                    // of zeros instead (one address select per step, not four value selects).
                    for (uint32_t i = 0; i < G + 7u; i++) {
                        const uint32_t gn = i - sub + 1u;     // next granule; wraps to 0 one step before the start
                        const float4 nxt = *((valid && gn < G) ? src4 + gn : &zero_granule);
                        const double x0 = widen_abs_fast(cur.x, odd), x1 = widen_abs_fast(cur.y, odd),
                                     x2 = widen_abs_fast(cur.z, odd), x3 = widen_abs_fast(cur.w, odd);
                        sum = __fma_rn(x0, x0, sum);
                        sum = __fma_rn(x1, x1, sum);
                        sum = __fma_rn(x2, x2, sum);
                        sum = __fma_rn(x3, x3, sum);
                        cur = nxt;
                    }
                    // see widen_abs_fast: Inf / NaN inputs show up as an impossible sum; silence as ~0
                    odd |= !(sum < __hiloint2double(0x4F900000, 0));   // 2^250; also true for NaN
                    if (sum < __hiloint2double(0x33700000, 0)) sum = 0.0;   // < 2^-200: every sample was zero
                    if (odd) {   // rare: the quantum holds a denormal, Inf or NaN -> redo it exactly
                        sum = 0.0;
                        const float *src = reinterpret_cast<const float *>(src4);
                        for (uint32_t k = 0; k < meter.quantum; k++) {
                            const double x = widen_abs(src[k]);
                            sum = __fma_rn(x, x, sum);
                        }
                    }
                    if (valid) rms_s[qg] = rms_of(sum, meter.quantum);
                }
            } else {
                // scalar path (frames whose rows are not 16-byte aligned in the tile): skew by lane
                for (uint32_t base = warp * 32u; base < quanta; base += nwarps * 32u) {
                    const uint32_t qg = base + lane;
                    const bool valid = qg < quanta;
                    const uint32_t rr = valid ? qg / meter.nq : 0u, q = valid ? qg - rr * meter.nq : 0u;
                    const float *src = tile + rr * n_in + q * meter.quantum;
                    double sum = 0.0;
                    for (uint32_t i = 0; i < meter.quantum + 31u; i++) {
                        const uint32_t k = i - lane;
                        if (valid && k < meter.quantum) {   // unsigned: also false while i < lane
                            const double x = widen_abs(src[k]);
                            sum = __fma_rn(x, x, sum);
                        }
                    }
                    if (valid) rms_s[qg] = rms_of(sum, meter.quantum);
                }
            }
            __syncthreads();
            for (uint32_t rr = threadIdx.x; rr < rows_per_tile; rr += blockDim.x) {
                const unsigned long long row = t * rows_per_tile + rr;
                double lv = rr == threadIdx.x ? lv_in : meter.state[row];
                // Math.max(rms, level * 0.85): NaN if either is NaN, and a NaN level stays NaN
                // (NaN * 0.85).  NaN-ness is tracked with integer tests so each step costs one DMUL and
                // one DSETP on the (long-latency) FP64 pipe.
                bool poisoned = lv != lv;
                for (uint32_t q = 0; q < meter.nq; q++) {
                    const double rms = rms_s[rr * meter.nq + q];
                    const uint32_t hi = (uint32_t)__double2hiint(rms) & 0x7FFFFFFFu;
                    poisoned |= hi > 0x7FF00000u || (hi == 0x7FF00000u && __double2loint(rms) != 0);
                    const double dec = __dmul_rn(lv, 0.85);
                    lv = rms > dec ? rms : dec;
                    if (poisoned) lv = __longlong_as_double(0x7FF8000000000000ll);
                    if (meter.levels) meter.levels[row * meter.nq + q] = lv;
                }
                meter.state[row] = lv;
            }
        }
        __syncthreads();   // everyone is done reading stage s
        if (threadIdx.x == 0) {
            const unsigned long long tn = t + (unsigned long long)STAGES * gridDim.x;
            if (tn < n_tiles) {
                mbar_expect_tx(&bars[s], tile_bytes);
                bulk_g2s(reinterpret_cast<unsigned char *>(tiles) + (size_t)s * tile_stride_bytes,
                         reinterpret_cast<const unsigned char *>(in) + tn * (unsigned long long)tile_bytes, tile_bytes, &bars[s]);
            }
        }
    }
}

// ---- kernel 2c: frames + level meter, padded-row tiles (the fast meter path) ----------------------
// Same job as kernel 2b with METER (frames out, level meter of App.tsx:267-271 fused in), rebuilt
// around what ncu showed about 2b: it was ISSUE-bound (0.84 warp-instructions per input sample, 76 %
// of the issue slots, HBM at 58 %), and more than half of those instructions were the meter loop's
// bookkeeping -- lane skew against shared-memory bank conflicts (G+7 steps for G of work, one
// address select per step) -- and its partly filled warps.  Here:
//   * rows sit in shared memory at a pitch of an ODD number of 16-byte units (one bulk copy per row
//     instead of one per tile).  Lane l of a warp takes quantum q of row (l mod K): the 8 lanes of an
//     LDS.128 phase read the same granule of 8 consecutive rows, which now fall into 8 different
//     bank groups.  No skew, no selects: the inner loop is LDS.128 + 4 x (2 IMAD, LEA.HI, VIADDMNMX,
//     DFMA), fully unrolled for the 128-sample quantum;
//   * the pad slot after every row holds a NaN, and table entries that point outside the input are
//     redirected to it when the table is staged, so the gather has no per-pick range test;
//   * integer ratios are compile-time (RT), so the 8 picks of a chunk are immediate offsets from one
//     address; row / chunk splits use a host-made reciprocal instead of a division;
//   * a partial last tile is handled in the kernel (no remainder launches).
// The arithmetic (widen_abs_fast and its two escape hatches, rms_of, the Math.max chain) is
// unchanged from kernel 2b, so the results stay bit-identical to the oracle.
// Requirements (launcher): n_in % 4 == 0, quantum % 4 == 0, dense rows (pitch_in == n_in, pitch_out ==
// n_out), in / out 16-byte aligned, K rows + table fit in shared memory.
struct TileMeterGeom {
    uint32_t n_in, n_out;
    uint32_t K;              // rows per tile
    uint32_t rows;           // rows of the batch (the last tile may be partial)
    uint32_t pitch;          // shared-memory row pitch in floats: n_in + 4 or n_in + 8, (pitch / 4) odd
    uint32_t stride;         // integer ratio when RT == 0 and no table
    uint32_t magic_cpr;      // ceil(2^32 / (n_out / 8))  (n_out % 8 == 0 only); 0 stands for a divisor of 1
    uint32_t magic_nout;     // ceil(2^32 / n_out)
    uint32_t magic_K;        // ceil(2^32 / K)
    uint32_t tile_stride_bytes;
};

// x / d for the small x of one tile, d folded into magic = ceil(2^32 / d) on the host (exact while
// x * d < 2^32, which the launcher guarantees); magic == 0 encodes d == 1.
__device__ __forceinline__ uint32_t div_magic(uint32_t x, uint32_t magic) { return magic ? __umulhi(x, magic) : x; }

template <int RT, int STAGES>
__global__ void __launch_bounds__(256)
k_frames_meter(const float *__restrict__ in, int16_t *__restrict__ out, TileMeterGeom geo, unsigned long long n_tiles,
               const uint32_t *__restrict__ lut, int hint, MeterArgs meter)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [STAGES][tile_stride_bytes] | table u32[n_out] (only if lut) | mbarriers | rms double[K * nq]
    const uint32_t n_in = geo.n_in, n_out = geo.n_out, K = geo.K, P = geo.pitch;
    uint32_t *tab = reinterpret_cast<uint32_t *>(smem_raw + (size_t)STAGES * geo.tile_stride_bytes);
    const uint32_t tab_bytes = lut ? ((n_out * 4u + 15u) & ~15u) : 0u;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem_raw + (size_t)STAGES * geo.tile_stride_bytes + tab_bytes);
    double *rms_s = reinterpret_cast<double *>(bars + ((STAGES + 1) & ~1));
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t stride = RT > 0 ? (uint32_t)RT : geo.stride;
    const bool rows_aligned = (n_out & 7u) == 0u;
    const uint32_t cpr = n_out >> 3;
    const uint32_t nq = meter.nq, Q = meter.quantum;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (lut)   // picks outside the input (JS reads `undefined` -> NaN -> 0) go to the row's NaN pad slot
        for (uint32_t i = threadIdx.x; i < n_out; i += blockDim.x) {
            const uint32_t v = __ldg(lut + i);
            tab[i] = v < n_in ? v : n_in;
        }
    for (uint32_t i = threadIdx.x; i < K * STAGES; i += blockDim.x) {   // the pad slots (bulk copies never touch them)
        float *row = reinterpret_cast<float *>(smem_raw + (size_t)(i / K) * geo.tile_stride_bytes) + (size_t)(i % K) * P;
        row[n_in] = __int_as_float(0x7FC00000);
    }
    __syncthreads();

    // one bulk copy per row; a warp issues them, lane r takes rows r, r+32, ...
    auto issue = [&](unsigned long long t, int s) {
        const unsigned long long row0 = t * K;
        const uint32_t nr = (uint32_t)((geo.rows - row0 < K) ? geo.rows - row0 : K);
        if (lane == 0) mbar_expect_tx(&bars[s], nr * n_in * 4u);
        __syncwarp();
        float *dst = reinterpret_cast<float *>(smem_raw + (size_t)s * geo.tile_stride_bytes);
        for (uint32_t r = lane; r < nr; r += 32u)
            bulk_g2s(dst + (size_t)r * P, in + (row0 + r) * n_in, n_in * 4u, &bars[s]);
    };
    if (warp == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; s++) {
            const unsigned long long t = blockIdx.x + (unsigned long long)s * gridDim.x;
            if (t < n_tiles) issue(t, s);
        }
    }

    uint32_t it = 0;
    for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x, it++) {
        const int s = it % STAGES;
        const unsigned long long row0 = t * K;
        const uint32_t nr = (uint32_t)((geo.rows - row0 < K) ? geo.rows - row0 : K);
        const float *tile = reinterpret_cast<const float *>(smem_raw + (size_t)s * geo.tile_stride_bytes);
        // session levels are needed only after the meter: fetch them now, the latency hides behind it
        double lv_in = 0.0;
        if (threadIdx.x < nr) lv_in = meter.state[row0 + threadIdx.x];
        mbar_wait(&bars[s], (it / STAGES) & 1u);

        // ---- frames: pick + encode + store -------------------------------------------------------
        const unsigned long long first = row0 * n_out;
        const uint32_t outs = nr * n_out;
        if (rows_aligned) {
            // rows are whole 8-sample chunks and start 16-byte aligned: chunk g = (row rr, chunk c)
            const uint32_t n_chunks = nr * cpr;
            for (uint32_t g = threadIdx.x; g < n_chunks; g += blockDim.x) {
                const uint32_t rr = div_magic(g, geo.magic_cpr), c = g - rr * cpr;
                const float *row = tile + rr * P;
                int e[8];
                if (lut) {
                    const uint4 a = *reinterpret_cast<const uint4 *>(tab + 8u * c);
                    const uint4 b = *reinterpret_cast<const uint4 *>(tab + 8u * c + 4u);
                    e[0] = enc_s32(row[a.x]); e[1] = enc_s32(row[a.y]); e[2] = enc_s32(row[a.z]); e[3] = enc_s32(row[a.w]);
                    e[4] = enc_s32(row[b.x]); e[5] = enc_s32(row[b.y]); e[6] = enc_s32(row[b.z]); e[7] = enc_s32(row[b.w]);
                } else {
                    const float *p = row + 8u * c * stride;
#pragma unroll
                    for (int j = 0; j < 8; j++) e[j] = enc_s32(p[j * stride]);
                }
                uint4 o4;
                o4.x = pack_s16x2_sat(e[0], e[1]);
                o4.y = pack_s16x2_sat(e[2], e[3]);
                o4.z = pack_s16x2_sat(e[4], e[5]);
                o4.w = pack_s16x2_sat(e[6], e[7]);
                st_128(out + first + (size_t)g * 8u, o4, hint);
            }
        } else {
            // the tile's output is one flat run of `outs` samples starting anywhere: an aligned body of
            // 8-sample chunks (which may straddle frames) plus < 8 samples before and after it
            const unsigned long long body0 = (first + 7ull) & ~7ull;
            const unsigned long long body1 = (first + outs) & ~7ull;
            const uint32_t n_body = body1 > body0 ? (uint32_t)((body1 - body0) >> 3) : 0u;
            const uint32_t head = n_body ? (uint32_t)(body0 - first) : outs;
            auto pick = [&](uint32_t rr, uint32_t o) -> float {
                const uint32_t i = lut ? tab[o] : o * stride;
                return tile[rr * P + i];
            };
            for (uint32_t g = threadIdx.x; g < n_body; g += blockDim.x) {
                const uint32_t f = head + 8u * g;
                uint32_t rr = div_magic(f, geo.magic_nout), o = f - rr * n_out;
                int e[8];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    e[j] = enc_s32(pick(rr, o));
                    if (++o == n_out) { o = 0; rr++; }
                }
                uint4 o4;
                o4.x = pack_s16x2_sat(e[0], e[1]);
                o4.y = pack_s16x2_sat(e[2], e[3]);
                o4.z = pack_s16x2_sat(e[4], e[5]);
                o4.w = pack_s16x2_sat(e[6], e[7]);
                st_128(out + body0 + (size_t)g * 8u, o4, hint);
            }
            const uint32_t tail0 = head + 8u * n_body, n_edge = head + (outs - tail0);
            for (uint32_t k = threadIdx.x; k < n_edge; k += blockDim.x) {
                const uint32_t f = k < head ? k : tail0 + (k - head);
                const uint32_t rr = div_magic(f, geo.magic_nout), o = f - rr * n_out;
                out[first + f] = enc_s16(pick(rr, o));
            }
        }

        // ---- level meter: lane <-> (quantum q, row r), w = q * K + r -----------------------------------
        // The meter takes ceil(items / 32) of the block's 8 warps.  Which ones rotates with the tile
        // index: a warp's scheduler is fixed by its index (warp % 4), so always using warps 0, 1, ...
        // would load the same schedulers in every co-resident block and leave the others idle.
        const uint32_t items = K * nq;
        const uint32_t m_warps = (items + 31u) >> 5;
        const uint32_t rot = m_warps < 8u ? ((uint32_t)t * m_warps) & 7u : 0u;
        for (uint32_t w = (((warp + 8u - rot) & 7u) << 5) + lane; w < ((items + 31u) & ~31u); w += blockDim.x) {
            const uint32_t q = div_magic(w, geo.magic_K), r = w - q * K;
            if (w >= items || r >= nr) continue;            // partial tile / partly filled last warp
            rms_s[w] = quantum_rms(reinterpret_cast<const float4 *>(tile + r * P + q * Q), Q);
        }
        __syncthreads();
        for (uint32_t rr = threadIdx.x; rr < nr; rr += blockDim.x) {
            const unsigned long long row = row0 + rr;
            double lv = rr == threadIdx.x ? lv_in : meter.state[row];
            // Math.max(rms, level * 0.85): NaN if either is NaN, and a NaN level stays NaN
            bool poisoned = lv != lv;
            for (uint32_t q = 0; q < nq; q++) {
                const double rms = rms_s[q * K + rr];
                const uint32_t hi = (uint32_t)__double2hiint(rms) & 0x7FFFFFFFu;
                poisoned |= hi > 0x7FF00000u || (hi == 0x7FF00000u && __double2loint(rms) != 0);
                const double dec = __dmul_rn(lv, 0.85);
                lv = rms > dec ? rms : dec;
                if (poisoned) lv = __longlong_as_double(0x7FF8000000000000ll);
                if (meter.levels) meter.levels[row * nq + q] = lv;
            }
            meter.state[row] = lv;
        }
        __syncthreads();   // everyone is done with stage s (tile and rms_s)
        if (warp == 0) {
            const unsigned long long tn = t + (unsigned long long)STAGES * gridDim.x;
            if (tn < n_tiles) issue(tn, s);
        }
    }
}

// ---- kernel 2d: frames + level meter, warp-pair pipelines ---------------------------------------------
// Kernel 2c cut the instruction count in half and stopped being issue-bound -- and did not get faster:
// ncu showed 37 % issue utilisation, 61 % DRAM, warps parked on the tile's mbarrier and on the block
// barrier.  A block there does load -> pick -> meter -> level chain strictly in sequence, the meter
// alone is a 128-long dependent DFMA chain per lane (>= 1 000 clocks whatever the tile size), and only
// three 62 KB tiles fit per SM, so the time a tile occupies shared memory caps the bytes in flight.
// Shared memory is the buffer between HBM and the compute: throughput = bytes resident / time a byte
// stays resident.  This kernel keeps that time short and the buffer full:
//   * one CTA per SM, persistent; the CTA is NP independent warp PAIRS, each with its own `depth` small
//     stages (a stage = Ks rows holding at most 32 quanta: 4 frames of 960 samples, ONE 2944-sample frame);
//   * in a pair, warp A is the METER: one lane per quantum, then the level chain (all lanes at once in its
//     unrolled form when only the final level is wanted, else one lane per row);
//     warp B does the FRAMES (pick + encode + store) at the same time and is also the pair's loader:
//     when both warps are done with a stage (done[s], two arrivals) it issues the bulk copies of the
//     pair's next tile into it (completing on full[s]).  With depth 2 the other stage is loading while
//     this one is computed on; measured, more pairs with ONE stage each beat fewer, deeper pairs (the
//     other pairs cover a pair's load latency), so the default is depth 1 and 12-16 pairs.  Nothing is
//     shared between pairs, so there is no block-wide barrier after
//     the prologue, no producer warp to become the bottleneck (a single dedicated producer measured ~660
//     clocks + ~130 per copy per stage: 3 samples/clk/SM against the 4.3 the HBM share needs), and no
//     head-of-line blocking between stages;
//   * a row is copied in chunks of C quanta (one bulk copy each, 16 bytes of padding after every chunk)
//     so that the quanta of a row do not all start in the same bank group; the meter lanes are mapped
//     so that the 8 lanes of an LDS.128 phase fall into as many different chunks / rows as there are.
//     One bulk copy PER QUANTUM would be conflict-free but is far too fine for the copy engine
//     (measured ~80 clocks per 512-byte request: 6 B/clk/SM against the 22 B/clk/SM HBM share);
//   * picks go through a shared-memory table of ROW-RELATIVE shared-memory offsets (built once per CTA
//     from the pick table or the stride), so the padded layout costs the gather nothing; a pick outside
//     the input (JS `undefined` -> NaN -> 0) points at a NaN kept in the first chunk's padding.
// Every mbarrier is waited on by the same warp in every one of its phases (a parity wait only means
// something one phase ahead of what the waiter has already seen).
// Arithmetic (quantum_rms, level_step, enc_s32) is shared with the other kernels: bit-identical results.
// Requirements (launcher): n_in % 4 == 0, quantum % 4 == 0, dense rows, in / out 16-byte aligned.
struct WsGeom {
    uint32_t n_in, n_out;
    uint32_t Ks;             // rows per stage
    uint32_t rows;           // rows of the batch (the last tile may be partial)
    uint32_t chunk;          // samples per chunk copy (C * quantum, or n_in for whole rows)
    uint32_t ncopies;        // chunk copies per row = ceil(n_in / chunk)
    uint32_t row_floats;     // shared-memory floats per row = n_in + 4 * ncopies
    uint32_t stride;         // integer ratio when there is no table
    uint32_t magic_cpr, magic_nout, magic_chunk, magic_Ks, magic_ncopies;
    uint32_t stage_bytes;    // Ks * row_floats * 4 rounded up to 128
    unsigned char lane_q[32];   // meter: item w -> row w % Ks, quantum lane_q[w / Ks] (spreads a phase over the chunks)
    uint32_t pairs;          // warp pairs per CTA
    uint32_t depth;          // stages per pair
    uint32_t stg_samples;    // int16 per pair in the output staging buffer: Ks * n_out + 8, rounded up to 8
};

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// CVT: the meter lanes widen with the hardware conversion (quantum_rms_cvt) instead of the integer re-packing.  It
// pays when frames leave meter lanes idle (23 quanta in a 32-lane warp: the kernel is then bound by that warp's issue
// slots, 0.95 -> 0.97 of peak at 2944-sample frames) and costs when all lanes are busy (960-sample frames: the XU pipe
// becomes the limit, 1.03 -> 1.02); the launcher picks by frame length.
template <bool CVT>
__global__ void __launch_bounds__(1024, 1)
k_frames_meter_ws(const float *__restrict__ in, int16_t *__restrict__ out, WsGeom geo, unsigned long long n_tiles,
                  const uint32_t *__restrict__ lut, int hint, MeterArgs meter)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [pairs * depth][stage_bytes] | offset table u32[n_out] | full[NS] | done[NS] | rms [pairs][32] | out staging [pairs][stg_samples]
    const uint32_t n_in = geo.n_in, n_out = geo.n_out, Ks = geo.Ks, NS = geo.pairs * geo.depth;
    uint32_t *tab = reinterpret_cast<uint32_t *>(smem_raw + (size_t)NS * geo.stage_bytes);
    const uint32_t tab_bytes = (n_out * 4u + 15u) & ~15u;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + (size_t)NS * geo.stage_bytes + tab_bytes);
    uint64_t *done = full + NS;
    double *rms_all = reinterpret_cast<double *>(done + NS);
    const uint32_t nq = meter.nq, Q = meter.quantum;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t row_floats = geo.row_floats;               // shared-memory floats per row (chunks + their padding)
    const uint32_t nan_off = min(geo.chunk, n_in);            // the padding right after a row's first chunk holds a NaN

    if (threadIdx.x == 0) {
        for (uint32_t s = 0; s < NS; s++) { mbar_init(&full[s], 1); mbar_init(&done[s], 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (uint32_t o = threadIdx.x; o < n_out; o += blockDim.x) {
        const uint32_t i = lut ? __ldg(lut + o) : o * geo.stride;
        // sample i sits 4 floats further for every chunk boundary before it
        tab[o] = i < n_in ? i + 4u * div_magic(i, geo.magic_chunk) : nan_off;
    }
    for (uint32_t i = threadIdx.x; i < Ks * NS; i += blockDim.x) {   // one NaN per row of every stage
        float *row = reinterpret_cast<float *>(smem_raw + (size_t)(i / Ks) * geo.stage_bytes) + (size_t)(i % Ks) * row_floats;
        row[nan_off] = __int_as_float(0x7FC00000);
    }
    __syncthreads();

    // this CTA's steps: step `it` is tile blockIdx.x + it * gridDim.x; pair p takes steps p, p + pairs, ...
    const uint32_t n_it = blockIdx.x < n_tiles ? (uint32_t)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0u;
    // A warp's scheduler is fixed by its index (warp % 4).  The meter warp issues about twice as many
    // instructions as the frames warp, so the roles alternate from pair to pair in a pattern that puts the
    // meter warps of consecutive pairs on schedulers 0, 2, 1, 3, ... instead of always on 0 and 2.
    const uint32_t pair = warp >> 1, role = ((warp ^ (pair >> 1)) & 1u);
    if (pair >= geo.pairs) return;
    const uint32_t my_steps = pair < n_it ? (n_it - pair + geo.pairs - 1) / geo.pairs : 0u;   // k = 0 .. my_steps-1 -> it = pair + k * pairs
    const uint32_t s0 = pair * geo.depth;
    auto stage_of = [&](uint32_t k) { return s0 + k % geo.depth; };
    auto rows_of = [&](uint32_t k, unsigned long long &row0) {
        row0 = ((unsigned long long)blockIdx.x + (unsigned long long)(pair + k * geo.pairs) * gridDim.x) * Ks;
        return (uint32_t)((geo.rows - row0 < Ks) ? geo.rows - row0 : Ks);
    };

    if (role == 1u) {
        // ---- warp B: loader + frames -----------------------------------------------------------------
        auto load = [&](uint32_t k) {
            unsigned long long row0;
            const uint32_t nr = rows_of(k, row0), s = stage_of(k);
            if (lane == 0) mbar_expect_tx(&full[s], nr * n_in * 4u);
            __syncwarp();
            float *dst = reinterpret_cast<float *>(smem_raw + (size_t)s * geo.stage_bytes);
            for (uint32_t c = lane; c < nr * geo.ncopies; c += 32u) {
                const uint32_t r = div_magic(c, geo.magic_ncopies), j = c - r * geo.ncopies;
                const uint32_t len = min(geo.chunk, n_in - j * geo.chunk);
                bulk_g2s(dst + (size_t)r * row_floats + (size_t)j * (geo.chunk + 4u), in + (row0 + r) * n_in + (size_t)j * geo.chunk,
                         len * 4u, &full[s]);
            }
        };
        for (uint32_t k = 0; k < geo.depth && k < my_steps; k++) load(k);
        // Picks with consecutive lanes on consecutive OUTPUTS: the table reads are then consecutive words and
        // the picks are ratio-strided floats (conflict-free for odd ratios) -- the earlier layout, a lane per
        // 8-output chunk, made every table read and every pick an 8-way bank conflict and put shared memory at
        // 79 % of its bandwidth on 2944-sample frames.  The encoded samples go through a small per-pair
        // staging buffer and leave as 128-bit stores; staging index i holds global output (first - shift) + i,
        // shift = first mod 8, so 16-byte chunks line up in both.
        int16_t *stg = reinterpret_cast<int16_t *>(reinterpret_cast<unsigned char *>(rms_all + (size_t)geo.pairs * 32u)) +
                       (size_t)pair * geo.stg_samples;
        for (uint32_t k = 0; k < my_steps; k++) {
            const uint32_t s = stage_of(k), par = (k / geo.depth) & 1u;
            unsigned long long row0;
            const uint32_t nr = rows_of(k, row0);
            const float *tile = reinterpret_cast<const float *>(smem_raw + (size_t)s * geo.stage_bytes);
            mbar_wait(&full[s], par);
            const unsigned long long first = row0 * n_out;
            const uint32_t outs = nr * n_out, shift = (uint32_t)(first & 7ull);
            if ((n_out & 7u) == 0u) {
                // frames that are whole 8-sample chunks (960 -> 320): a lane per chunk, table read as two
                // 128-bit loads, straight to a 128-bit store.  The picks conflict here too, but these short
                // frames leave shared memory far from saturated and the 8 independent picks per lane are
                // worth more (measured: 1.02 of peak this way, 0.87 through the staging buffer).
                const uint32_t cpr = n_out >> 3, n_chunks = nr * cpr;
                for (uint32_t c8 = lane; c8 < n_chunks; c8 += 32u) {
                    const uint32_t rr = div_magic(c8, geo.magic_cpr), c = c8 - rr * cpr;
                    const float *row = tile + rr * row_floats;
                    const uint4 a = *reinterpret_cast<const uint4 *>(tab + 8u * c);
                    const uint4 b = *reinterpret_cast<const uint4 *>(tab + 8u * c + 4u);
                    int e[8];
                    e[0] = enc_s32(row[a.x]); e[1] = enc_s32(row[a.y]); e[2] = enc_s32(row[a.z]); e[3] = enc_s32(row[a.w]);
                    e[4] = enc_s32(row[b.x]); e[5] = enc_s32(row[b.y]); e[6] = enc_s32(row[b.z]); e[7] = enc_s32(row[b.w]);
                    uint4 o4;
                    o4.x = pack_s16x2_sat(e[0], e[1]);
                    o4.y = pack_s16x2_sat(e[2], e[3]);
                    o4.z = pack_s16x2_sat(e[4], e[5]);
                    o4.w = pack_s16x2_sat(e[6], e[7]);
                    st_128(out + first + (size_t)c8 * 8u, o4, hint);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&done[s]);
                if (k + geo.depth < my_steps) {
                    mbar_wait(&done[s], par);
                    load(k + geo.depth);
                }
                continue;
            }
            for (uint32_t rr = 0; rr < nr; rr++) {
                const float *row = tile + rr * row_floats;
                int16_t *dst = stg + shift + rr * n_out;
                uint32_t o = lane;
                for (; o + 224u < n_out; o += 256u) {              // 8 independent picks per lane in flight
                    uint32_t off[8];
                    float x[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) off[j] = tab[o + 32u * j];
#pragma unroll
                    for (int j = 0; j < 8; j++) x[j] = row[off[j]];
#pragma unroll
                    for (int j = 0; j < 8; j++) dst[o + 32u * j] = enc_s16(x[j]);
                }
                for (; o < n_out; o += 32u) dst[o] = enc_s16(row[tab[o]]);
            }
            __syncwarp();
            // [shift, shift + outs) of the staging buffer is global [first, first + outs): whole 8-sample chunks
            // of it as 128-bit stores, the < 8 samples before and after one by one
            const uint32_t lo = shift ? 8u : 0u, hi = (shift + outs) & ~7u;       // staging indices of the aligned body
            int16_t *gbase = out + (first - shift);
            if (hi > lo) {
                for (uint32_t i = lo + 8u * lane; i < hi; i += 256u) st_128(gbase + i, *reinterpret_cast<const uint4 *>(stg + i), hint);
                if (lane < lo - shift && shift) gbase[shift + lane] = stg[shift + lane];
                if (lane < shift + outs - hi) gbase[hi + lane] = stg[hi + lane];
            } else {
                for (uint32_t i = shift + lane; i < shift + outs; i += 32u) gbase[i] = stg[i];
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&done[s]);
            if (k + geo.depth < my_steps) {                    // refill this stage once the meter warp has left it too
                mbar_wait(&done[s], par);
                load(k + geo.depth);
            }
        }
        return;
    }

    // ---- warp A: level meter ---------------------------------------------------------------------------
    double *rms_s = rms_all + (size_t)pair * 32u;
    for (uint32_t k = 0; k < my_steps; k++) {
        const uint32_t s = stage_of(k), par = (k / geo.depth) & 1u;
        unsigned long long row0;
        const uint32_t nr = rows_of(k, row0);
        const float *tile = reinterpret_cast<const float *>(smem_raw + (size_t)s * geo.stage_bytes);
        double lv_in = 0.0;
        if (lane < nr) lv_in = meter.state[row0 + lane];       // needed after the meter: the latency hides behind it
        mbar_wait(&full[s], par);
        // lane w <-> (row w % Ks, quantum lane_q[w / Ks]) (Ks * nq <= 32)
        if (lane < Ks * nq) {
            const uint32_t kq = div_magic(lane, geo.magic_Ks), r = lane - kq * Ks, q = geo.lane_q[kq];
            const uint32_t i0 = q * Q;
            if (r < nr) {
                const float4 *qsrc = reinterpret_cast<const float4 *>(tile + r * row_floats + i0 + 4u * div_magic(i0, geo.magic_chunk));
                rms_s[r * nq + q] = CVT ? quantum_rms_cvt(qsrc, Q) : quantum_rms(qsrc, Q);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&done[s]);                  // the samples are no longer needed: the stage may be refilled
        if (meter.levels) {
            // per-quantum trace wanted: the plain sequential chain, one lane per row
            if (lane < nr) {
                const unsigned long long row = row0 + lane;
                double lv = lv_in;
                bool poisoned = lv != lv;
                for (uint32_t q = 0; q < nq; q++) {
                    level_step(lv, poisoned, rms_s[lane * nq + q]);
                    meter.levels[row * nq + q] = lv;
                }
                meter.state[row] = lv;
            }
        } else {
            // Only the level after the frame's last quantum is wanted.  level_q = max(rms_q, f(level_{q-1}))
            // with f(x) = fl(x * 0.85); f is monotone, so f(max(a, b)) = max(f(a), f(b)) EXACTLY and the
            // chain unrolls to  level_last = max( f^nq(level_in), max_q f^(nq-1-q)(rms_q) ):  every lane
            // decays its own quantum's rms the right number of times -- all lanes at once instead of one
            // lane per row walking the quanta -- and a row's lanes are then max-reduced.  Math.max's NaN
            // rule (any NaN in, NaN out, and it sticks) is a flag carried beside the values.
            const uint32_t kq = div_magic(lane, geo.magic_Ks), r = lane - kq * Ks;      // this lane's row and position
            const bool mine = lane < Ks * nq && r < nr;
            const uint32_t q = mine ? geo.lane_q[kq] : 0u;
            double v = mine ? rms_s[r * nq + q] : 0.0;
            bool nan = v != v;
            for (uint32_t t = q + 1u; t < nq; t++) v = __dmul_rn(v, 0.85);             // nq-1-q decays (0 for the idle lanes: q = 0, v = 0)
            // the row's incoming level travels with the lane that holds the row's first slot (kq == 0)
            double lvv = __shfl_sync(0xFFFFFFFFu, lv_in, r);                            // lv_in sits in lane r
            if (mine && kq == 0u) {
                nan |= lvv != lvv;
                for (uint32_t t = 0; t < nq; t++) lvv = __dmul_rn(lvv, 0.85);
                v = lvv > v ? lvv : v;
            }
            if (!mine) v = 0.0;
            // lanes of one row are Ks apart (lane = kq * Ks + r, Ks a power of two): xor-reduce over kq
            uint32_t any_nan = nan ? 1u : 0u;
            for (uint32_t d = Ks; d < 32u; d <<= 1) {
                const double o = __shfl_xor_sync(0xFFFFFFFFu, v, d);
                any_nan |= __shfl_xor_sync(0xFFFFFFFFu, any_nan, d);
                v = o > v ? o : v;
            }
            if (lane < nr) meter.state[row0 + lane] = any_nan ? __longlong_as_double(0x7FF8000000000000ll) : v;
        }
        __syncwarp();                                         // rms_s is reused by the next step
    }
}

// ---- kernel 3: resampleTo16k alone (Float32 -> Float32 picks; App.tsx:18-32) ------------------------
// Only for signature parity with the reference's stand-alone function; the batched product path is
// the fused kernels above.  Bit patterns are moved as 32-bit words so NaN payloads survive.
__global__ void __launch_bounds__(256)
k_pick_f32(const uint32_t *__restrict__ in, size_t pitch_in, uint32_t *__restrict__ out, size_t pitch_out,
           uint32_t n_in, uint32_t n_out, const uint32_t *__restrict__ lut, uint32_t stride,
           unsigned long long total)
{
    for (unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned long long r = t / n_out;
        const uint32_t o = (uint32_t)(t - r * n_out);
        const uint32_t i = lut ? __ldg(lut + o) : o * stride;
        out[(size_t)r * pitch_out + o] = (i < n_in) ? __ldg(in + (size_t)r * pitch_in + i) : 0x7FC00000u;
    }
}

// (The device-side aggregator's kernels -- count / scan / scatter / walk / convert / levels / append /
// commit -- are in agg_kernels.cuh.)

// ---- kernel 4: level meter alone (any layout; one thread per session, sequential) ----------------------
// Fallback for batches kernel 2b cannot take (pitched rows, long frames).  Correct, not fast.
__global__ void __launch_bounds__(128)
k_level_meter(const float *__restrict__ in, size_t pitch_in, uint32_t rows, MeterArgs meter)
{
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x) {
        const float *row = in + (size_t)r * pitch_in;
        double lv = meter.state[r];
        for (uint32_t q = 0; q < meter.nq; q++) {
            double sum = 0.0;
            for (uint32_t k = 0; k < meter.quantum; k++) {
                const double x = widen_abs(__ldg(row + (size_t)q * meter.quantum + k));
                sum = __fma_rn(x, x, sum);
            }
            lv = js_max(rms_of(sum, meter.quantum), __dmul_rn(lv, 0.85));
            if (meter.levels) meter.levels[(size_t)r * meter.nq + q] = lv;
        }
        meter.state[r] = lv;
    }
}

// ---- synthetic audio for the benchmarks -----------------------------------------------
__device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
k_synth(float *__restrict__ out, size_t pitch, uint32_t rows, uint32_t n, float inv_rate,
        uint64_t seed, uint64_t row0)
{
    const unsigned long long total = (unsigned long long)rows * n;
    for (unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (unsigned long long)gridDim.x * blockDim.x) {
        const uint32_t r = (uint32_t)(t / n);
        const uint32_t k = (uint32_t)(t - (unsigned long long)r * n);
        const uint64_t session = row0 + r;
        const float f = 110.0f + 7.0f * (float)(session & 511u);
        const float amp = ((session & 7u) == 7u) ? 1.2f : 0.5f;
        const uint64_t h = mix64(mix64(seed ^ (session * 0xD1342543DE82EF95ull)) + k);
        const float u = (float)(uint32_t)(h >> 40) * (1.0f / 8388608.0f) - 1.0f;   // U[-1, 1)
        out[(size_t)r * pitch + k] = amp * sinpif(2.0f * f * (float)k * inv_rate) + 0.05f * u;
    }
}

}  // namespace pcmb200
