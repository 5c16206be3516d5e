/*
 * pcm_b200.h -- C ABI of libpcm_b200.so, the B200 (sm_100a) PCM front-end.
 *
 * Drop-in scope: the mic-audio front-end of the reference's browser client
 *   floatTo16BitPCM(float32: Float32Array): Int16Array          apps/web/src/App.tsx:7-16
 *   resampleTo16k(input: Float32Array, rate: number): Float32Array   apps/web/src/App.tsx:18-32
 *   pcm16 = floatTo16BitPCM(resampleTo16k(buf, rate))           apps/web/src/App.tsx:281-282
 * batched over many concurrent sessions.  The reference has no FFI of its own;
 * these entry points are what an N-API addon for that path binds (see
 * INTEGRATION.md, voice-enabled-browser-automation_b200/node/).  The bytes they
 * produce go, unchanged, to DeepgramClient.sendAudio (apps/voice/src/deepgram.ts:16,55-58).
 *
 * Conventions
 *   - every function returns a pcmb200_status (0 = ok); nothing throws across the ABI;
 *   - the caller owns every buffer it passes; the context owns its own staging memory;
 *   - a context is bound to one CUDA device and is NOT thread-safe (one per device per thread);
 *   - `mem` says where the caller's in/out buffers live:
 *       PCMB200_MEM_HOST    host memory (pinned is faster).  The call copies in, computes,
 *                           copies out and returns when the output is complete (synchronous,
 *                           like the reference's functions).
 *       PCMB200_MEM_DEVICE  device memory of the context's device.  The call enqueues work on
 *                           the context's stream and returns; use pcmb200_sync().
 *   - sample layout of a batch: row-major [row][sample]; a row is one frame of one session
 *     (rows = sessions x ticks).  pitch_* are in ELEMENTS, not bytes.
 *   - there is no CPU fallback: without a usable CUDA device pcmb200_ctx_create fails.
 */
#ifndef PCM_B200_H
#define PCM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define PCMB200_API __declspec(dllexport)
#else
#define PCMB200_API __attribute__((visibility("default")))
#endif

typedef struct pcmb200_ctx pcmb200_ctx;

typedef enum pcmb200_status {
    PCMB200_OK = 0,
    PCMB200_ERR_INVALID = 1,    /* bad argument (null pointer, bad mem kind, misuse) */
    PCMB200_ERR_RANGE = 2,      /* the reference would throw RangeError (invalid typed-array length) */
    PCMB200_ERR_CUDA = 3,       /* a CUDA call failed; see pcmb200_last_error */
    PCMB200_ERR_NO_DEVICE = 4,  /* no usable CUDA device: the library has no CPU fallback */
    PCMB200_ERR_NOMEM = 5,      /* host or device allocation failed */
    PCMB200_ERR_CAPACITY = 6    /* caller's output buffer is too small */
} pcmb200_status;

typedef enum pcmb200_mem {
    PCMB200_MEM_HOST = 0,
    PCMB200_MEM_DEVICE = 1
} pcmb200_mem;

/* ---- library / context ---------------------------------------------------------------- */

/* ABI version: major * 1000 + minor. */
PCMB200_API int pcmb200_version(void);
PCMB200_API const char *pcmb200_status_string(int status);
/* Number of CUDA devices visible to this process (0 and PCMB200_ERR_NO_DEVICE if none). */
PCMB200_API int pcmb200_device_count(int *count);
/* Create a context on `device` (creates a stream, events and an empty staging pool). */
PCMB200_API int pcmb200_ctx_create(int device, pcmb200_ctx **out);
PCMB200_API int pcmb200_ctx_destroy(pcmb200_ctx *ctx);
/* Message of the last failure on this context (ctx may be NULL: last failure of this thread). */
PCMB200_API const char *pcmb200_last_error(const pcmb200_ctx *ctx);
/* Wait for everything enqueued on the context's stream. */
PCMB200_API int pcmb200_sync(pcmb200_ctx *ctx);
/* Number of this library's kernels launched through the context so far. */
PCMB200_API int pcmb200_launch_count(const pcmb200_ctx *ctx, uint64_t *launches);
/* Capacity up front (ctx_create(device, max_sessions, max_frame_in, ...)).  Sizes the three
 * staging slots of the host-resident pipeline for batches of up to max_rows frames of up to max_n_in
 * input / max_n_out output samples, and sets aside max_rate_classes (<= 32) pick-index tables of
 * max_n_out entries (one per distinct non-integer (rate, frame length) in use, e.g. the 44.1 kHz
 * family).  After it, pcmb200_frames_encode / _encode_f32_s16 / _resample_to_16k within those limits
 * call no cudaMalloc, cudaFree or device-wide synchronisation: larger batches are cut into more
 * slices instead of growing the slots.  Without it the library grows on demand (the first large
 * batch of a live server then pays an allocation stall).  pcmb200_alloc_stats reports how often the
 * library allocated / freed device memory and synchronised a stream on its own behalf. */
PCMB200_API int pcmb200_ctx_reserve(pcmb200_ctx *ctx, uint32_t max_rows, uint32_t max_n_in, uint32_t max_n_out,
                                    uint32_t max_rate_classes);
PCMB200_API int pcmb200_alloc_stats(const pcmb200_ctx *ctx, uint64_t *device_mallocs, uint64_t *device_frees,
                                    uint64_t *stream_syncs);
/* Name, SM count and total memory of the context's device. */
PCMB200_API int pcmb200_device_info(const pcmb200_ctx *ctx, char *name, size_t name_cap,
                                    int *sm_count, size_t *total_mem);

/* ---- frame arithmetic (host side, data independent) --------------------------------------- */

/* Output length of resampleTo16k for an n_in-sample input: n_in at 16 kHz (App.tsx:22), else
 * floor(n_in / (in_rate / 16000)) in binary64 (App.tsx:23-24).  PCMB200_ERR_RANGE where the
 * reference's `new Float32Array(newLen)` throws (rate <= 0 ...); a NaN length gives 0. */
PCMB200_API int pcmb200_resample_len(size_t n_in, double in_rate, size_t *n_out);
/* The pick index of each of n_out output samples, from the reference's sequential binary64
 * running sum `idx += ratio` (App.tsx:26-30).  0xFFFFFFFF marks a pick outside the input. */
PCMB200_API int pcmb200_index_table(size_t n_in, double in_rate, size_t n_out, uint32_t *idx);

/* ---- memory --------------------------------------------------------------------------------- */

PCMB200_API int pcmb200_device_alloc(pcmb200_ctx *ctx, size_t bytes, void **ptr);
PCMB200_API int pcmb200_device_free(pcmb200_ctx *ctx, void *ptr);
/* Page-locked host memory: H2D/D2H of pinned buffers runs at PCIe speed and overlaps compute. */
PCMB200_API int pcmb200_pinned_alloc(pcmb200_ctx *ctx, size_t bytes, void **ptr);
/* Same, write-combined (cudaHostAllocWriteCombined): for buffers the host only WRITES (input
 * samples on their way to the GPU).  The device reads them without snooping the CPU caches; host
 * reads of such memory are very slow.  Free with pcmb200_pinned_free. */
PCMB200_API int pcmb200_pinned_alloc_write_combined(pcmb200_ctx *ctx, size_t bytes, void **ptr);
/* ctx may be NULL here (page-locked host memory is not tied to a device). */
PCMB200_API int pcmb200_pinned_free(pcmb200_ctx *ctx, void *ptr);
/* Blocking copies on the context's stream. */
PCMB200_API int pcmb200_memcpy_h2d(pcmb200_ctx *ctx, void *dst_device, const void *src_host, size_t bytes);
PCMB200_API int pcmb200_memcpy_d2h(pcmb200_ctx *ctx, void *dst_host, const void *src_device, size_t bytes);
PCMB200_API int pcmb200_memset_device(pcmb200_ctx *ctx, void *dst_device, int value, size_t bytes);

/* ---- the hot path ---------------------------------------------------------------------------- */

/* floatTo16BitPCM (App.tsx:7-16) over n samples: clamp to [-1,1], scale by 0x8000 (negative) or
 * 0x7fff, truncate toward zero; NaN -> 0.  Bit-exact with the reference's binary64 arithmetic. */
PCMB200_API int pcmb200_encode_f32_s16(pcmb200_ctx *ctx, const float *in, int16_t *out, size_t n, int mem);

/* resampleTo16k (App.tsx:18-32): nearest-sample pick with the reference's running-sum indices,
 * Float32 in, Float32 out (bit patterns copied unmodified).  out_cap is in elements; *n_out is
 * set to the produced length.  At 16 kHz the reference returns its input; here that is a copy. */
PCMB200_API int pcmb200_resample_to_16k(pcmb200_ctx *ctx, const float *in, size_t n_in, double in_rate,
                                        float *out, size_t out_cap, size_t *n_out, int mem);

/* The fused frame path (App.tsx:281-282) over a batch: for each of `rows` frames of n_in samples
 * at in_rate, out_row = floatTo16BitPCM(resampleTo16k(in_row, in_rate)); no intermediate Float32
 * array is materialised.  Every frame restarts the pick phase at 0 and drops its tail, as the
 * reference does at each flush (App.tsx:26, :286-287).  *n_out (may be NULL) receives the
 * per-row output length; pitch_out must be >= that length. */
PCMB200_API int pcmb200_frames_encode(pcmb200_ctx *ctx, const float *in, size_t pitch_in,
                                      int16_t *out, size_t pitch_out, uint32_t rows, uint32_t n_in,
                                      double in_rate, int mem, size_t *n_out);

/* ---- frames + level meter ----------------------------------------------------------------------- */

/* pcmb200_frames_encode plus the reference's audio level meter (App.tsx:267-271), fused so the
 * samples are read from HBM once: for every whole `quantum` (128 = the AudioWorklet render quantum,
 * App.tsx:46-55) of every frame, rms = Math.sqrt(sum(x*x) / quantum) in binary64 in sample order,
 * and per session level = Math.max(rms, level * 0.85) (setLevel, App.tsx:271).  Device-resident
 * buffers; asynchronous on the context's stream.
 *   level_state  [rows] doubles, in/out: each session's level before / after this frame
 *   levels       [rows][n_in / quantum] doubles (level after every quantum) or NULL
 * The trailing n_in % quantum samples of a frame are not metered (the reference's frames are whole
 * multiples of the quantum). */
PCMB200_API int pcmb200_frames_encode_meter(pcmb200_ctx *ctx, const float *in_device, size_t pitch_in,
                                            int16_t *out_device, size_t pitch_out, uint32_t rows, uint32_t n_in,
                                            double in_rate, uint32_t quantum, double *level_state_device,
                                            double *levels_device, size_t *n_out);

/* ---- capture aggregator -------------------------------------------------------------------------- */

/* Device-side restatement of the reference's capture aggregator (App.tsx:148-157, :273-288): each
 * session appends 128-sample quanta; the quantum that arrives with now - lastSend >= desired_ms
 * flushes the session's WHOLE buffer as one frame, pcm16 = floatTo16BitPCM(resampleTo16k(buf, rate))
 * (:281-282), and resets buffer and lastSend (:286-287).  The wall clock is injected: arrival_ms[i]
 * is performance.now() at quantum i.  Sessions of one aggregator share the capture rate.
 * Session state, buffers and the flush plan of every push live on the GPU (no host threads).
 * A session's waiting quanta are kept CONVERTED (the int16 outputs each quantum owns do not depend on
 * the frame's final length, App.tsx:26-30), so device memory per session is about
 * 2 * N_out(max_frame_quanta * quantum) + 8 * max_frame_quanta + 28 bytes -- 3 KB for 32 quanta at 48 kHz
 * (a buffer of samples would be 16 KB).
 *   quantum          samples per quantum, a multiple of 4, at most 876 (the AudioWorklet quantum is 128)
 *   max_frame_quanta  longest frame, in quanta (a push that would exceed it is rejected whole); one frame
 *                     must fit on chip: max_frame_quanta * (quantum + 4) * 4 bytes <= 200 KB
 *   max_push_quanta   most quanta in one push;   max_out_samples_per_push  output arena (host path)
 * A rejected push (unknown session id, over-long buffer, more frames than max_frames, output too small)
 * returns an error and leaves every session exactly as it was. */
typedef struct pcmb200_aggregator pcmb200_aggregator;
PCMB200_API int pcmb200_aggregator_create(pcmb200_ctx *ctx, uint32_t max_sessions, double in_rate, uint32_t quantum,
                                          double desired_ms, uint32_t max_frame_quanta, uint32_t max_push_quanta,
                                          size_t max_out_samples_per_push, pcmb200_aggregator **agg);
PCMB200_API int pcmb200_aggregator_destroy(pcmb200_ctx *ctx, pcmb200_aggregator *agg);
/* Initial state / stop() of the reference (App.tsx:148-157, :325-330): empty buffer, lastSend = now.
 * Enqueued on the context's stream (ordered with the pushes); _range resets `count` sessions at once. */
PCMB200_API int pcmb200_aggregator_reset_session(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t session, double now_ms);
PCMB200_API int pcmb200_aggregator_reset_range(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t first_session,
                                               uint32_t count, double now_ms);
/* Level meter of the aggregated sessions (App.tsx:267-271): when enabled (before the first push), the
 * quanta of every flushed frame go through rms = sqrt(sum(x*x)/quantum), level = Math.max(rms, level * 0.85)
 * in arrival order (binary64, bit-identical to the reference's arithmetic).  get_levels copies the
 * current level of `count` sessions to the host (waits for the pushes enqueued so far).  The level
 * lags the reference's by the quanta still waiting in a session's buffer. */
PCMB200_API int pcmb200_aggregator_set_meter(pcmb200_ctx *ctx, pcmb200_aggregator *agg, int enabled);
PCMB200_API int pcmb200_aggregator_get_levels(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t first_session,
                                              uint32_t count, double *levels_host);
/* Append n_quanta quanta, in arrival order (quantum i belongs to sessions[i], samples + i*quantum),
 * and convert every frame that flushes.  Frame f (in flush order) is frame_len[f] int16 samples of
 * session frame_session[f] at out + frame_offset[f]; offsets are multiples of 8 samples, so
 * out_capacity must hold every frame rounded up to a multiple of 8 samples.
 * sessions / arrival_ms / the frame arrays are HOST arrays here; samples and out are host or device
 * memory (samples_mem / out_mem).  Returns when the frames and the frame arrays are complete. */
PCMB200_API int pcmb200_aggregator_push(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t n_quanta,
                                        const uint32_t *sessions, const double *arrival_ms, const float *samples,
                                        int samples_mem, int16_t *out, size_t out_capacity, int out_mem,
                                        uint32_t *frame_session, uint32_t *frame_len, unsigned long long *frame_offset,
                                        uint32_t max_frames, uint32_t *n_frames);
/* The same with EVERYTHING in device memory (a producer on the GPU: a decoder, a mixer): nothing
 * crosses PCIe but the 24-byte outcome.  Asynchronous on the context's stream; the frame arrays may
 * be NULL.  pcmb200_aggregator_wait waits for the push, returns its status and the number of frames /
 * output samples (padded total); one push may be pending per aggregator. */
PCMB200_API int pcmb200_aggregator_push_device(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t n_quanta,
                                               const uint32_t *sessions_device, const double *arrival_ms_device,
                                               const float *samples_device, int16_t *out_device, size_t out_capacity,
                                               uint32_t *frame_session_device, uint32_t *frame_len_device,
                                               unsigned long long *frame_offset_device, uint32_t max_frames);
PCMB200_API int pcmb200_aggregator_wait(pcmb200_ctx *ctx, pcmb200_aggregator *agg, uint32_t *n_frames, size_t *out_samples);

/* ---- wire egress helpers ------------------------------------------------------------------------ */

/* The reference's keep-alive (App.tsx:291-296): every PCMB200_KEEPALIVE_PERIOD_MS while the socket
 * is open it sends `new Int16Array(1600)` -- 100 ms of silence at 16 kHz.  Writes the 1600 zero
 * samples to `out` (host memory, capacity in samples) and sets *n; PCMB200_ERR_CAPACITY if too small.
 * The wire format of every frame this library produces is the reference's: raw native-endian
 * (little-endian on x86-64 / aarch64) int16 mono 16 kHz, no header (App.tsx:284, deepgram.ts:36-45);
 * pcmb200_wire_is_little_endian() returns 1 on such hosts. */
#define PCMB200_KEEPALIVE_SAMPLES 1600u
#define PCMB200_KEEPALIVE_PERIOD_MS 2000u
PCMB200_API int pcmb200_keepalive_frame(int16_t *out, size_t capacity, size_t *n);
PCMB200_API int pcmb200_wire_is_little_endian(void);
/* Per-session wire payloads over a converted batch: the reference sends one WebSocket binary message
 * per frame, payload = pcm16.buffer (App.tsx:284), relayed untouched (apps/voice/src/server.ts:260-265)
 * into sendAudio (deepgram.ts:55-58).  Row r of a [rows][pitch_out] int16 block is the payload
 * (byte_offset[r], byte_length[r] = 2 * n_out) of that block: hand out views, not copies.
 * _frames does the same for an aggregator push (frame_offset / frame_len in samples). */
PCMB200_API int pcmb200_egress_table(uint32_t rows, size_t n_out, size_t pitch_out, unsigned long long *byte_offset,
                                     uint32_t *byte_length);
PCMB200_API int pcmb200_egress_table_frames(uint32_t n_frames, const unsigned long long *frame_offset, const uint32_t *frame_len,
                                            unsigned long long *byte_offset, uint32_t *byte_length);
/* Keep-alive schedule of a fleet (App.tsx:291-296: setInterval(send 1600 zeros if the socket is open, 2000),
 * armed by start(), cleared by stop() :309-312) on an injected clock.  start_ms[s]: when session s armed
 * its timer (NaN = not armed); ticks_done[s] (in/out): timer ticks already served; socket_open[s]: the
 * readyState === OPEN test.  Writes the sessions that must be sent a keep-alive frame now (each at most
 * once per poll; elapsed ticks are marked done also for closed sockets, whose callback returns early).
 * PCMB200_ERR_CAPACITY (and no change) if more than max_due are due.  period_ms is
 * PCMB200_KEEPALIVE_PERIOD_MS in the reference. */
PCMB200_API int pcmb200_keepalive_due(double now_ms, uint32_t n_sessions, const double *start_ms,
                                      unsigned long long *ticks_done, const uint8_t *socket_open, double period_ms,
                                      uint32_t *due_sessions, uint32_t max_due, uint32_t *n_due);

/* ---- per-tick replay --------------------------------------------------------------------------- */

/* (The graph owns a private copy of its pick table, so it stays valid whatever the context's table
 * cache does afterwards.)
 * A tick whose buffers, shape and rate do not change (the steady state of a fleet: same sessions,
 * same frame length every 20 ms) can be recorded once as a CUDA graph and replayed, which removes
 * the per-launch CPU cost that dominates small ticks (config 2: 1.97 MB per tick is 0.3 us of HBM
 * time).  Device-resident buffers only.  The graph is replayed on the context's stream. */
typedef struct pcmb200_tick_graph pcmb200_tick_graph;
PCMB200_API int pcmb200_tick_graph_create(pcmb200_ctx *ctx, const float *in_device, size_t pitch_in,
                                          int16_t *out_device, size_t pitch_out, uint32_t rows, uint32_t n_in,
                                          double in_rate, pcmb200_tick_graph **graph);
PCMB200_API int pcmb200_tick_graph_launch(pcmb200_ctx *ctx, pcmb200_tick_graph *graph);
PCMB200_API int pcmb200_tick_graph_destroy(pcmb200_ctx *ctx, pcmb200_tick_graph *graph);

/* ---- measurement helpers (used by bench.py and the tests; not part of the reference surface) -- */

/* Fill a device matrix with the benchmark's synthetic audio: per row
 * x[n] = A*sin(2*pi*f*n/rate) + 0.05*u, f = 110 + 7*(session mod 512) Hz, A = 1.2 for every 8th
 * session else 0.5, u ~ U[-1,1) from a counter-based hash of (seed, session, n);
 * session = row0 + row. */
PCMB200_API int pcmb200_synth_fill(pcmb200_ctx *ctx, float *out_device, size_t pitch, uint32_t rows,
                                   uint32_t n, double rate, uint64_t seed, uint64_t row0);
/* Bare copy probe: `reps` rounds of one pinned H2D copy (h2d_bytes) and one pinned D2H copy (d2h_bytes)
 * in flight together, no kernels: the platform ceiling of the host-resident path (tools/pcie_probe.py
 * runs it from N processes at once).  *milliseconds = device time of all `reps` rounds. */
PCMB200_API int pcmb200_pcie_probe(pcmb200_ctx *ctx, size_t h2d_bytes, size_t d2h_bytes, int reps, float *milliseconds);
/* CUDA-event stopwatch on the context's stream: everything the context enqueues between start and
 * stop (the pipelined host path included) is inside the measured interval. */
PCMB200_API int pcmb200_timer_start(pcmb200_ctx *ctx);
PCMB200_API int pcmb200_timer_stop(pcmb200_ctx *ctx, float *milliseconds);
/* Tuning knobs for sweeps (defaults are the measured best): "block" (threads per
 * block), "blocks_per_sm" (0 = one tile per block; > 0 = capped, looping grid), "unroll", "path"
 * (0 auto, 1 general kernel only, 2 general path with the tile-staged kernel), "coalesced", "tile",
 * "tile_kb", "stages", "meter_tile_kb", "meter_stages", "occupancy", "store_hint",
 * "slice_bytes" (host pipeline slice), "meter_kernel", "ws_smem_kb", "ws_pairs", "ws_depth", "ws_rows", "ws_chunk",
 * "ws_cvt_quanta" (kernel 2d: frames of at least this many quanta widen float -> double in hardware; 0 = never);
 * for the aggregator "agg_block" / "agg_stage_kb" (convert kernel: threads per CTA, staging memory per step; read
 * when an aggregator is created), "agg_persistent" (convert grid = the CTAs resident at once), "agg_pdl" (the
 * kernels of a push are chained by programmatic dependent launch; 0 = plain launches), "agg_walk_regs" (flush-rule
 * kernel: index lists of up to 32 or 4 quanta in registers; 0 = chosen from the previous push; 4 also selects the other
 * kernels built for one-quantum-per-session pushes) and "agg_warp_frames" (0: those pushes keep the CTA-per-frame
 * convert kernel).
 * "overlap_ticks" = 1 changes behaviour, not just speed: consecutive device-resident dense integer-ratio
 * calls on this context may then OVERLAP on the GPU (the next tick's blocks start while the previous
 * tick's last wave drains; programmatic dependent launch).  Only for callers whose consecutive calls
 * are independent (different frames in different buffers); copies, pcmb200_sync and the stopwatch still
 * wait for everything. */
PCMB200_API int pcmb200_tune(pcmb200_ctx *ctx, const char *key, long long value);

/* ---- allocation-free variant ------------------------------------------------------------------ */

/* No counterpart in the reference (which allocates fresh typed arrays on every call); added for
 * callers that must not allocate device memory while they run: a real-time tick thread that cannot
 * take the device-wide stall of cudaMalloc/cudaFree, or a process whose GPU memory is already owned
 * by another framework's allocator, where a late cudaMalloc fails.
 *
 * Same result as pcmb200_frames_encode(..., PCMB200_MEM_DEVICE) for device-resident `in` / `out`,
 * computed by the general pick kernel with its index table in a fixed 64 KiB array inside the
 * library (held by every GPU that loads the library, until the process exits).  No context is
 * needed and nothing is allocated, on the device or the host.
 *   - limit: at most PCMB200_NOALLOC_MAX_OUT output samples per frame (about 1 s of 16 kHz audio);
 *     longer frames return PCMB200_ERR_CAPACITY (message via pcmb200_last_error(NULL));
 *   - the device holding in/out must be the calling thread's current device; `cuda_stream` is a
 *     cudaStream_t (NULL = default stream); wait != 0 blocks until the output is complete;
 *   - the first call on a device resolves the array's address: make it at start-up, before any
 *     stream capture;
 *   - all calls on a device share the array and it keeps the last call's table: calls whose GPU work
 *     may overlap (issued on different streams) race and must be serialised by the caller ON THE GPU:
 *     the later stream has to wait (cudaStreamWaitEvent) on an event recorded in the earlier stream
 *     after the earlier call -- calling later on the host is not enough, even for the same frame shape;
 *     calls on one stream are ordered by the stream and are safe. */
#define PCMB200_NOALLOC_MAX_OUT 16384u
PCMB200_API int pcmb200_frames_encode_noalloc(const float *in_device, size_t pitch_in, int16_t *out_device,
                                              size_t pitch_out, uint32_t rows, uint32_t n_in, double in_rate,
                                              void *cuda_stream, int wait, size_t *n_out);

#ifdef __cplusplus
}
#endif
#endif /* PCM_B200_H */
