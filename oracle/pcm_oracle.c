/*
 * pcm_oracle.c -- CPU ORACLE for the PCM front-end path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is a plain-C restatement of the reference's browser-side audio
 * helpers, written from the ECMAScript semantics of the reference source:
 *
 *   floatTo16BitPCM   /root/reference/apps/web/src/App.tsx:7-16
 *   resampleTo16k     /root/reference/apps/web/src/App.tsx:18-32
 *   frame flush       /root/reference/apps/web/src/App.tsx:273-288  (per-frame
 *                     independence: phase restarts at 0, tail dropped)
 *   level meter       /root/reference/apps/web/src/App.tsx:267-271
 *   keep-alive frame  /root/reference/apps/web/src/App.tsx:291-296
 *
 * PARITY: pinned to outputs of the reference's own source text, executed in the
 * build container by tests/golden/mini_js.py -- a small interpreter for the
 * JavaScript / TypeScript subset these functions are written in, made for this
 * purpose because no JavaScript engine (node, deno, bun, QuickJS, V8 ...) exists in
 * the image or can be installed, and the reference's own test-suite holds no golden
 * vector for these functions.  tests/golden/make_ref_exec.py
 * slices the functions out of App.tsx at run time, runs that text on the committed
 * input vectors and writes tests/golden/ref_exec.npz; tests/test_ref_exec.py requires
 * this file to reproduce every output (encoder, frames, the onmessage handler's flush
 * decisions / frames / levels, the keep-alive timer).  NOT YET CONFIRMED BY A
 * PRODUCTION ENGINE: the language semantics under that run are mini_js.py's reading
 * of ECMA-262 (unit-tested rule by rule in the same test file);
 * `node tools/pin_with_node.mjs` is the one-command check that remains.
 * Beside that pin: (1) the ECMA-262 definitions of Number arithmetic (IEEE-754
 * binary64), Math.floor, Float32Array reads (exact widening) and Int16Array writes
 * (ToInt16: non-finite -> 0, truncate toward zero, modulo 2^16); (2) a second
 * restatement in pure Python floats and (3) a third in integers only
 * (tests/golden/make_golden.py) whose committed outputs this file must reproduce --
 * and which the reference text, run as above, reproduces too.
 *
 * Who may use this file: tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs, as the CHECKER or the timed CPU
 * baseline.  The product (libpcm_b200.so, the N-API addon, the Python mirror)
 * never links, imports or calls it.
 *
 * Build: see oracle/Makefile.  Flags matter: -O2 -fno-fast-math -ffp-contract=off
 * (no reassociation, no fused multiply-add: JS performs one correctly-rounded
 * binary64 operation per operator).
 */
#include <math.h>
#include <pthread.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ECMA-262 7.1.8 ToInt16(v): what `int16Array[i] = v` does to a Number. */
ORACLE_API int16_t oracle_to_int16(double v)
{
    if (!isfinite(v)) return 0;                    /* NaN, +Inf, -Inf -> +0 */
    if (v > -2147483648.0 && v < 2147483648.0) {
        /* common case, what a JS engine's typed-array store does: truncate to int32, keep the
         * low 16 bits (two's complement) */
        int32_t t = (int32_t)v;                    /* C conversion truncates toward zero */
        return (int16_t)(uint16_t)((uint32_t)t & 0xFFFFu);
    }
    double t = trunc(v);                           /* sign(v) * floor(abs(v)) */
    double m = fmod(t, 65536.0);                   /* C fmod keeps the sign of t */
    if (m < 0) m += 65536.0;                       /* int16bit = int modulo 2^16 */
    if (m >= 32768.0) m -= 65536.0;
    return (int16_t)m;
}

/* App.tsx:10-13 for one sample: widen, clamp to [-1, 1] with the reference's
 * comparison order (so NaN falls through both tests), scale by 0x8000 on the
 * negative side and 0x7fff otherwise, in binary64. */
ORACLE_API double oracle_scale_sample(float f)
{
    double s = (double)f;                          /* Float32Array read: exact */
    if (s > 1) s = 1;
    else if (s < -1) s = -1;
    return s < 0 ? s * 32768.0 /* 0x8000 */ : s * 32767.0 /* 0x7fff */;
}

/* App.tsx:7-16 floatTo16BitPCM: out has n elements. */
ORACLE_API void oracle_float_to_16bit_pcm(const float *in, size_t n, int16_t *out)
{
    for (size_t i = 0; i < n; i++)
        out[i] = oracle_to_int16(oracle_scale_sample(in[i]));
}

/* App.tsx:22-24: output length of resampleTo16k for an n-sample input.
 * Returns 0 and sets *status when the JS would not produce an array:
 *   status 0: ok;  1: `new Float32Array(newLen)` would throw RangeError
 *   (negative, infinite or > 2^32-1 length).  NaN length -> ToIndex(NaN) = 0. */
ORACLE_API size_t oracle_resample_len(size_t n, double in_rate, int *status)
{
    if (status) *status = 0;
    if (in_rate == 16000.0) return n;              /* :22 identity */
    double ratio = in_rate / 16000.0;              /* :23 */
    double len = floor((double)n / ratio);         /* :24 */
    if (len != len) return 0;                      /* NaN -> length 0 */
    if (len < 0 || len > 4294967295.0) {           /* RangeError: Invalid typed array length */
        /* -0 is fine (length 0) and compares == 0, so it never lands here */
        if (status) *status = 1;
        return 0;
    }
    return (size_t)len;
}

/* App.tsx:26-30: the pick index of every output sample, produced by the same
 * sequential binary64 running sum.  idx[i] = UINT32_MAX marks a pick outside
 * the input (JS reads `undefined`, stores NaN). */
ORACLE_API void oracle_index_table(size_t n_in, double in_rate, size_t n_out, uint32_t *idx)
{
    double ratio = in_rate / 16000.0;
    double pos = 0;
    for (size_t i = 0; i < n_out; i++) {
        double f = floor(pos);
        idx[i] = (f >= 0 && f < (double)n_in) ? (uint32_t)f : UINT32_MAX;
        pos += ratio;
    }
}

/* App.tsx:18-32 resampleTo16k.  `out` must hold oracle_resample_len() floats.
 * Returns the output length.  At 16 kHz the reference returns the input array
 * itself; here that is a verbatim copy (bit patterns, NaN payloads included). */
ORACLE_API size_t oracle_resample_to_16k(const float *in, size_t n, double in_rate, float *out)
{
    int st;
    size_t n_out = oracle_resample_len(n, in_rate, &st);
    if (st) return 0;
    if (in_rate == 16000.0) {
        if (out != in) memmove(out, in, n * sizeof(float));
        return n;
    }
    double ratio = in_rate / 16000.0;
    double pos = 0;
    for (size_t i = 0; i < n_out; i++) {
        double f = floor(pos);
        if (f >= 0 && f < (double)n) out[i] = in[(size_t)f];
        else out[i] = NAN;                         /* undefined -> NaN in a Float32Array */
        pos += ratio;
    }
    return n_out;
}

/* App.tsx:281-282: one flushed frame = floatTo16BitPCM(resampleTo16k(buf, rate)).
 * Faithful to the reference's allocation behaviour: a fresh Float32Array for the
 * resampled samples and a fresh Int16Array for the result on every call (the
 * result is then copied into the caller's wire buffer, as WebSocket.send does). */
ORACLE_API size_t oracle_frame(const float *in, size_t n_in, double in_rate, int16_t *out)
{
    int st;
    size_t n_out = oracle_resample_len(n_in, in_rate, &st);
    if (st) return 0;
    const float *res = in;
    float *tmp = NULL;
    if (in_rate != 16000.0) {
        tmp = (float *)calloc(n_out ? n_out : 1, sizeof(float));   /* new Float32Array(newLen) */
        oracle_resample_to_16k(in, n_in, in_rate, tmp);
        res = tmp;
    }
    int16_t *pcm = (int16_t *)calloc(n_out ? n_out : 1, sizeof(int16_t)); /* new Int16Array(len) */
    oracle_float_to_16bit_pcm(res, n_out, pcm);
    memcpy(out, pcm, n_out * sizeof(int16_t));
    free(pcm);
    free(tmp);
    return n_out;
}

/* Batched framing restatement: `rows` independent frames of
 * n_in samples, row r at in + r*pitch_in, written to out + r*pitch_out. */
ORACLE_API size_t oracle_frames(const float *in, size_t pitch_in, size_t rows, size_t n_in,
                                double in_rate, int16_t *out, size_t pitch_out)
{
    size_t n_out = 0;
    for (size_t r = 0; r < rows; r++)
        n_out = oracle_frame(in + r * pitch_in, n_in, in_rate, out + r * pitch_out);
    return n_out;
}

/* The same, rows split statically over `threads` POSIX threads (the timed
 * all-cores CPU baseline; sessions are independent so no synchronisation). */
struct frames_job {
    const float *in; size_t pitch_in; size_t r0, r1; size_t n_in; double rate;
    int16_t *out; size_t pitch_out;
};
static void *frames_worker(void *p)
{
    struct frames_job *j = (struct frames_job *)p;
    for (size_t r = j->r0; r < j->r1; r++)
        oracle_frame(j->in + r * j->pitch_in, j->n_in, j->rate, j->out + r * j->pitch_out);
    return NULL;
}
ORACLE_API size_t oracle_frames_mt(const float *in, size_t pitch_in, size_t rows, size_t n_in,
                                   double in_rate, int16_t *out, size_t pitch_out, int threads)
{
    int st;
    size_t n_out = oracle_resample_len(n_in, in_rate, &st);
    if (st) return 0;
    if (threads < 1) threads = 1;
    if ((size_t)threads > rows) threads = rows ? (int)rows : 1;
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    struct frames_job *jobs = (struct frames_job *)malloc(sizeof(struct frames_job) * (size_t)threads);
    for (int t = 0; t < threads; t++) {
        jobs[t] = (struct frames_job){ in, pitch_in, rows * (size_t)t / (size_t)threads,
                                       rows * (size_t)(t + 1) / (size_t)threads, n_in, in_rate,
                                       out, pitch_out };
        pthread_create(&tid[t], NULL, frames_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(tid[t], NULL);
    free(jobs);
    free(tid);
    return n_out;
}

/* App.tsx:267-271 level meter over consecutive quanta of one session:
 *   sum += x*x (binary64), rms = sqrt(sum / len), level = max(rms, prev * 0.85).
 * `levels[q]` receives the state after quantum q; returns the final state.
 * Math.max(NaN, x) is NaN in JS, and NaN then sticks (NaN * 0.85 = NaN). */
ORACLE_API double oracle_level_meter(const float *in, size_t n_quanta, size_t quantum,
                                     double prev, double *levels)
{
    for (size_t q = 0; q < n_quanta; q++) {
        const float *x = in + q * quantum;
        double sum = 0;
        for (size_t i = 0; i < quantum; i++) sum += (double)x[i] * (double)x[i];
        double rms = sqrt(sum / (double)quantum);
        double dec = prev * 0.85;
        double lv;
        if (rms != rms || dec != dec) lv = NAN;     /* Math.max: any NaN -> NaN */
        else lv = rms > dec ? rms : dec;
        prev = lv;
        if (levels) levels[q] = lv;
    }
    return prev;
}

/* App.tsx:273-288 aggregator with an injected clock.  The reference flushes when
 * performance.now() - lastSend >= desiredMs, tested after each 128-sample
 * quantum is appended; the wall clock is replaced by the caller-supplied arrival
 * time of every quantum (milliseconds, binary64), and `lastSend` is reset to
 * that same arrival time at a flush (the reference re-reads the clock a few
 * microseconds later; the injected clock has no such skew).
 * flush_after[q] = 1 when the frame is flushed after quantum q.  Returns the
 * number of flushes.  *last_send is updated in place. */
ORACLE_API size_t oracle_aggregator_plan(const double *arrival_ms, size_t n_quanta,
                                         double desired_ms, double *last_send,
                                         uint8_t *flush_after)
{
    size_t flushes = 0;
    for (size_t q = 0; q < n_quanta; q++) {
        double elapsed = arrival_ms[q] - *last_send;
        if (elapsed >= desired_ms) {
            flush_after[q] = 1;
            *last_send = arrival_ms[q];
            flushes++;
        } else {
            flush_after[q] = 0;
        }
    }
    return flushes;
}

/* App.tsx:292-296 keep-alive: 1600 zero samples (100 ms at 16 kHz). */
ORACLE_API size_t oracle_keepalive_frame(int16_t *out)
{
    memset(out, 0, 1600 * sizeof(int16_t));
    return 1600;
}

/* ---- synthetic benchmark input on the host (SURVEY.md 8d recipe) --------------------------------
 * The same generator as the CUDA library's pcmb200_synth_fill, restated for the host so that
 * `bench.py --impl reference` can build its 2^20-session tick without touching the GPU library:
 *   x[n] = A*sin(2*pi*f*n/rate) + 0.05*u,  f = 110 + 7*(session mod 512) Hz,
 *   A = 1.2 for every 8th session else 0.5,  u ~ U[-1,1) from a counter hash of (seed, session, n).
 * The hash (and therefore u) is bit-identical to the device's; the sine goes through the host
 * libm instead of the device's sinpif, so samples can differ from the device's in the last bit.
 * Not part of the reference: test / measurement infrastructure only. */
static uint64_t synth_mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
struct synth_job { float *out; size_t pitch, r0, r1, n; double rate; uint64_t seed, row0; };
static void *synth_worker(void *p)
{
    struct synth_job *j = (struct synth_job *)p;
    const float inv_rate = (float)(1.0 / j->rate);
    for (size_t r = j->r0; r < j->r1; r++) {
        const uint64_t session = j->row0 + r;
        const float f = 110.0f + 7.0f * (float)(session & 511u);
        const float amp = ((session & 7u) == 7u) ? 1.2f : 0.5f;
        const uint64_t key = synth_mix64(j->seed ^ (session * 0xD1342543DE82EF95ull));
        float *row = j->out + r * j->pitch;
        for (size_t k = 0; k < j->n; k++) {
            const uint64_t h = synth_mix64(key + k);
            const float u = (float)(uint32_t)(h >> 40) * (1.0f / 8388608.0f) - 1.0f;
            const float ph = 2.0f * f * (float)k * inv_rate;
            row[k] = amp * (float)sin(3.14159265358979323846 * (double)ph) + 0.05f * u;
        }
    }
    return NULL;
}
ORACLE_API void oracle_synth_fill(float *out, size_t pitch, size_t rows, size_t n, double rate,
                                  uint64_t seed, uint64_t row0, int threads)
{
    if (threads < 1) threads = 1;
    if ((size_t)threads > rows) threads = rows ? (int)rows : 1;
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    struct synth_job *jobs = (struct synth_job *)malloc(sizeof(struct synth_job) * (size_t)threads);
    for (int t = 0; t < threads; t++) {
        jobs[t] = (struct synth_job){ out, pitch, rows * (size_t)t / (size_t)threads,
                                      rows * (size_t)(t + 1) / (size_t)threads, n, rate, seed, row0 };
        pthread_create(&tid[t], NULL, synth_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++) pthread_join(tid[t], NULL);
    free(jobs);
    free(tid);
}

/* App.tsx:291-296 keep-alive TIMER of one session on an injected clock, event by event: the timer
 * armed at start_ms fires at start_ms + k * 2000 (k = 1, 2, ...); the host looks at its timers at the
 * times polls[0..n_polls) (ascending).  A poll runs the callback once if at least one firing time has
 * passed since the last poll that ran it (a blocked timer's ticks coalesce), and the callback sends
 * only if the socket is open at that poll.  sent[p] = 1 when poll p sends a keep-alive frame.
 * Returns the number of frames sent.  Written as a simulation of firing times, deliberately not as
 * the closed form the library uses. */
ORACLE_API size_t oracle_keepalive_timer(double start_ms, double period_ms, const double *polls, size_t n_polls,
                                         const uint8_t *socket_open, uint8_t *sent)
{
    size_t frames = 0;
    unsigned long long next_k = 1;                  /* the next firing not yet consumed */
    for (size_t p = 0; p < n_polls; p++) {
        int fired = 0;
        /* consume every firing time that is <= this poll.  A firing is "k periods after the start"; whether
         * it has passed is decided on the elapsed time, the way a timer counts intervals */
        while ((polls[p] - start_ms) / period_ms >= (double)next_k) { next_k++; fired = 1; }
        sent[p] = 0;
        if (fired && socket_open[p]) { sent[p] = 1; frames++; }
    }
    return frames;
}
