// addon.cc -- thin N-API shim over the C ABI of libpcm_b200.so (include/pcm_b200.h).
//
// Exports the reference's two signatures unchanged (apps/web/src/App.tsx:7, :18-21) plus the batched
// frame call, so `floatTo16BitPCM(resampleTo16k(buf, rate)).buffer` (App.tsx:281-284) keeps working
// and its bytes go to the unchanged DeepgramClient.sendAudio (apps/voice/src/deepgram.ts:55-58).
// All logic lives in the C ABI library; this file only validates arguments, moves pointers and maps
// status codes to JS exceptions.  No CPU fallback: if no CUDA device is usable, module load throws.
#include "napi_min.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/pcm_b200.h"

namespace {

pcmb200_ctx *g_ctx = nullptr;
std::mutex g_ctx_mutex;   // a context is not thread-safe: the async worker and the main thread take turns

bool throw_status(napi_env env, int status)
{
    if (status == PCMB200_OK) return false;
    const char *msg = pcmb200_last_error(g_ctx);
    if (!msg || !*msg) msg = pcmb200_last_error(nullptr);
    if (!msg || !*msg) msg = pcmb200_status_string(status);
    char code[32];
    std::snprintf(code, sizeof code, "PCMB200_%d", status);
    if (status == PCMB200_ERR_RANGE) napi_throw_range_error(env, code, msg);   // `new Float32Array(newLen)` would throw
    else napi_throw_error(env, code, msg);
    return true;
}

// Returns the Float32Array's data pointer / length, or throws TypeError (an extension: the
// reference does not validate, TypeScript does it at compile time).
bool get_f32(napi_env env, napi_value v, const float **data, size_t *len)
{
    bool is_ta = false;
    napi_typedarray_type t;
    void *p = nullptr;
    if (napi_is_typedarray(env, v, &is_ta) != napi_ok || !is_ta ||
        napi_get_typedarray_info(env, v, &t, len, &p, nullptr, nullptr) != napi_ok || t != napi_float32_array) {
        napi_throw_type_error(env, "PCMB200_TYPE", "expected a Float32Array");
        return false;
    }
    *data = static_cast<const float *>(p);
    return true;
}

napi_value make_typed(napi_env env, napi_typedarray_type t, size_t len, size_t elem, void **data)
{
    napi_value ab = nullptr, ta = nullptr;
    if (napi_create_arraybuffer(env, len * elem, data, &ab) != napi_ok) return nullptr;
    if (napi_create_typedarray(env, t, len, ab, 0, &ta) != napi_ok) return nullptr;
    return ta;
}

// floatTo16BitPCM(float32: Float32Array): Int16Array            App.tsx:7-16
napi_value FloatTo16BitPCM(napi_env env, napi_callback_info info)
{
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    const float *in;
    size_t n;
    if (argc < 1 || !get_f32(env, argv[0], &in, &n)) return nullptr;
    void *out = nullptr;
    napi_value res = make_typed(env, napi_int16_array, n, 2, &out);   // a fresh array, like `new Int16Array(len)`
    if (!res) return nullptr;
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    if (throw_status(env, pcmb200_encode_f32_s16(g_ctx, in, static_cast<int16_t *>(out), n, PCMB200_MEM_HOST))) return nullptr;
    return res;
}

// resampleTo16k(input: Float32Array, inputSampleRate: number): Float32Array     App.tsx:18-32
napi_value ResampleTo16k(napi_env env, napi_callback_info info)
{
    size_t argc = 2;
    napi_value argv[2];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    const float *in;
    size_t n;
    double rate = 0;
    if (argc < 2 || !get_f32(env, argv[0], &in, &n)) return nullptr;
    if (napi_get_value_double(env, argv[1], &rate) != napi_ok) {
        napi_throw_type_error(env, "PCMB200_TYPE", "inputSampleRate must be a number");
        return nullptr;
    }
    if (rate == 16000.0) return argv[0];                               // App.tsx:22 returns the same object
    size_t n_out = 0;
    if (throw_status(env, pcmb200_resample_len(n, rate, &n_out))) return nullptr;
    void *out = nullptr;
    napi_value res = make_typed(env, napi_float32_array, n_out, 4, &out);
    if (!res) return nullptr;
    size_t got = 0;
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    if (throw_status(env, pcmb200_resample_to_16k(g_ctx, in, n, rate, static_cast<float *>(out), n_out, &got, PCMB200_MEM_HOST)))
        return nullptr;
    return res;
}

// framesEncode(batch: Float32Array, rows: number, nIn: number, rate: number, out?: Int16Array): Int16Array
// One fused floatTo16BitPCM(resampleTo16k(row, rate)) per row (App.tsx:281-282), rows dense.
napi_value FramesEncode(napi_env env, napi_callback_info info)
{
    size_t argc = 5;
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    const float *in;
    size_t n;
    uint32_t rows = 0, n_in = 0;
    double rate = 0;
    if (argc < 4 || !get_f32(env, argv[0], &in, &n)) return nullptr;
    if (napi_get_value_uint32(env, argv[1], &rows) != napi_ok || napi_get_value_uint32(env, argv[2], &n_in) != napi_ok ||
        napi_get_value_double(env, argv[3], &rate) != napi_ok || (size_t)rows * n_in > n) {
        napi_throw_type_error(env, "PCMB200_TYPE", "framesEncode(batch, rows, nIn, rate): rows * nIn exceeds batch.length");
        return nullptr;
    }
    size_t n_out = 0;
    if (throw_status(env, pcmb200_resample_len(n_in, rate, &n_out))) return nullptr;
    void *out = nullptr;
    napi_value res = nullptr;
    if (argc >= 5) {   // caller-provided (e.g. pinned) output
        bool is_ta = false;
        napi_typedarray_type t;
        size_t len = 0;
        if (napi_is_typedarray(env, argv[4], &is_ta) != napi_ok || !is_ta ||
            napi_get_typedarray_info(env, argv[4], &t, &len, &out, nullptr, nullptr) != napi_ok ||
            t != napi_int16_array || len < (size_t)rows * n_out) {
            napi_throw_type_error(env, "PCMB200_TYPE", "out must be an Int16Array of at least rows * resampleLen(nIn, rate)");
            return nullptr;
        }
        res = argv[4];
    } else {
        res = make_typed(env, napi_int16_array, (size_t)rows * n_out, 2, &out);
        if (!res) return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    if (throw_status(env, pcmb200_frames_encode(g_ctx, in, n_in, static_cast<int16_t *>(out), n_out, rows, n_in, rate,
                                                PCMB200_MEM_HOST, nullptr)))
        return nullptr;
    return res;
}

// framesEncodeAsync(batch, rows, nIn, rate, out): Promise<Int16Array>
// The many-session path: the conversion (H2D, kernel, D2H and the wait for them) runs on a libuv
// worker thread, so the event loop that owns the WebSockets is never blocked on a CUDA sync.
// `batch` and `out` must stay untouched until the promise settles (use the pinned allocators).
struct AsyncJob {
    napi_async_work work = nullptr;
    napi_deferred deferred = nullptr;
    napi_ref batch_ref = nullptr, out_ref = nullptr;
    const float *in = nullptr;
    int16_t *out = nullptr;
    uint32_t rows = 0, n_in = 0;
    size_t n_out = 0;
    double rate = 0;
    int status = PCMB200_OK;
    std::string error;
};

void async_execute(napi_env, void *data)
{
    AsyncJob *j = static_cast<AsyncJob *>(data);
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    if (!g_ctx) {   // the environment is shutting down (cleanup ran first): fail the job, touch nothing
        j->status = PCMB200_ERR_INVALID;
        j->error = "pcm_b200: the addon's context was destroyed (environment shutting down)";
        return;
    }
    j->status = pcmb200_frames_encode(g_ctx, j->in, j->n_in, j->out, j->n_out, j->rows, j->n_in, j->rate, PCMB200_MEM_HOST, nullptr);
    if (j->status != PCMB200_OK) j->error = pcmb200_last_error(g_ctx);
}

void async_complete(napi_env env, napi_status, void *data)
{
    AsyncJob *j = static_cast<AsyncJob *>(data);
    napi_value out_arr = nullptr;
    napi_get_reference_value(env, j->out_ref, &out_arr);
    if (j->status == PCMB200_OK) {
        napi_resolve_deferred(env, j->deferred, out_arr);
    } else {
        napi_value msg = nullptr, err = nullptr;
        napi_create_string_utf8(env, j->error.c_str(), NAPI_AUTO_LENGTH, &msg);
        if (j->status == PCMB200_ERR_RANGE) napi_create_range_error(env, nullptr, msg, &err);
        else napi_create_error(env, nullptr, msg, &err);
        napi_reject_deferred(env, j->deferred, err);
    }
    napi_delete_reference(env, j->batch_ref);
    napi_delete_reference(env, j->out_ref);
    napi_delete_async_work(env, j->work);
    delete j;
}

napi_value FramesEncodeAsync(napi_env env, napi_callback_info info)
{
    size_t argc = 5;
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    const float *in;
    size_t n;
    uint32_t rows = 0, n_in = 0;
    double rate = 0;
    if (argc < 5 || !get_f32(env, argv[0], &in, &n)) {
        if (argc < 5) napi_throw_type_error(env, "PCMB200_TYPE", "framesEncodeAsync(batch, rows, nIn, rate, out)");
        return nullptr;
    }
    if (napi_get_value_uint32(env, argv[1], &rows) != napi_ok || napi_get_value_uint32(env, argv[2], &n_in) != napi_ok ||
        napi_get_value_double(env, argv[3], &rate) != napi_ok || (size_t)rows * n_in > n) {
        napi_throw_type_error(env, "PCMB200_TYPE", "framesEncodeAsync: rows * nIn exceeds batch.length");
        return nullptr;
    }
    size_t n_out = 0;
    if (throw_status(env, pcmb200_resample_len(n_in, rate, &n_out))) return nullptr;
    bool is_ta = false;
    napi_typedarray_type t;
    size_t len = 0;
    void *out = nullptr;
    if (napi_is_typedarray(env, argv[4], &is_ta) != napi_ok || !is_ta ||
        napi_get_typedarray_info(env, argv[4], &t, &len, &out, nullptr, nullptr) != napi_ok || t != napi_int16_array ||
        len < (size_t)rows * n_out) {
        napi_throw_type_error(env, "PCMB200_TYPE", "out must be an Int16Array of at least rows * resampleLen(nIn, rate)");
        return nullptr;
    }
    AsyncJob *j = new AsyncJob();
    j->in = in; j->out = static_cast<int16_t *>(out); j->rows = rows; j->n_in = n_in; j->n_out = n_out; j->rate = rate;
    napi_value promise = nullptr, name = nullptr;
    napi_create_promise(env, &j->deferred, &promise);
    napi_create_reference(env, argv[0], 1, &j->batch_ref);   // keep both arrays alive until the job is done
    napi_create_reference(env, argv[4], 1, &j->out_ref);
    napi_create_string_utf8(env, "pcm_b200.framesEncodeAsync", NAPI_AUTO_LENGTH, &name);
    napi_create_async_work(env, nullptr, name, async_execute, async_complete, j, &j->work);
    napi_queue_async_work(env, j->work);
    return promise;
}

// resampleLen(n: number, rate: number): number                   App.tsx:23-24
napi_value ResampleLen(napi_env env, napi_callback_info info)
{
    size_t argc = 2;
    napi_value argv[2];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    double n = 0, rate = 0;
    if (argc < 2 || napi_get_value_double(env, argv[0], &n) != napi_ok || napi_get_value_double(env, argv[1], &rate) != napi_ok || n < 0) {
        napi_throw_type_error(env, "PCMB200_TYPE", "resampleLen(n, rate)");
        return nullptr;
    }
    size_t n_out = 0;
    if (throw_status(env, pcmb200_resample_len((size_t)n, rate, &n_out))) return nullptr;
    napi_value r = nullptr;
    napi_create_double(env, (double)n_out, &r);
    return r;
}

// GC finalizer of a pinned ArrayBuffer.  It can run on the main thread while a worker is inside the
// context (which records errors in it), and after cleanup() has destroyed the context: take the same
// lock as every other user, and let the library free without a context then (ctx == NULL is allowed
// for pcmb200_pinned_free: page-locked host memory does not belong to a device).
void free_pinned(napi_env, void *data, void *)
{
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    pcmb200_pinned_free(g_ctx, data);
}

// allocPinnedFloat32(n) / allocPinnedInt16(n): typed arrays over page-locked memory, so JS fills
// H2D-ready buffers and receives D2H results without a staging copy.
napi_value alloc_pinned(napi_env env, napi_callback_info info, napi_typedarray_type t, size_t elem)
{
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    double n = 0;
    if (argc < 1 || napi_get_value_double(env, argv[0], &n) != napi_ok || n < 0) {
        napi_throw_type_error(env, "PCMB200_TYPE", "length must be a non-negative number");
        return nullptr;
    }
    void *p = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_ctx_mutex);
        if (throw_status(env, pcmb200_pinned_alloc(g_ctx, (size_t)n * elem, &p))) return nullptr;
    }
    napi_value ab = nullptr, ta = nullptr;
    if (napi_create_external_arraybuffer(env, p, (size_t)n * elem, free_pinned, nullptr, &ab) != napi_ok ||
        napi_create_typedarray(env, t, (size_t)n, ab, 0, &ta) != napi_ok) {
        free_pinned(env, p, nullptr);
        napi_throw_error(env, "PCMB200_NAPI", "could not wrap pinned memory");
        return nullptr;
    }
    return ta;
}
napi_value AllocPinnedFloat32(napi_env env, napi_callback_info info) { return alloc_pinned(env, info, napi_float32_array, 4); }
napi_value AllocPinnedInt16(napi_env env, napi_callback_info info) { return alloc_pinned(env, info, napi_int16_array, 2); }

// ---- the capture aggregator (App.tsx:148-157, :273-288) for many connections, on the GPU ---------------------------
// aggregatorCreate(maxSessions, rate, maxFrameQuanta, maxPushQuanta, maxOutSamples): number      -> a small handle
// aggregatorReset(h, slot, nowMs)                              buffer empty, lastSend = nowMs (a new connection)
// aggregatorPush(h, n, sessions: Uint32Array, timesMs: Float64Array, samples: Float32Array, out: Int16Array,
//                frameSession: Uint32Array, frameLen: Uint32Array, frameOffset: Float64Array): number of frames
//                (frame f = out.subarray(frameOffset[f], frameOffset[f] + frameLen[f]) of session frameSession[f])
// aggregatorDestroy(h)
constexpr int kMaxAggregators = 16;
struct AggSlot { pcmb200_aggregator *agg = nullptr; unsigned long long *offsets = nullptr; uint32_t max_push = 0; };
AggSlot g_aggs[kMaxAggregators];

bool get_ta(napi_env env, napi_value v, napi_typedarray_type want, void **data, size_t *len, const char *what)
{
    bool is_ta = false;
    napi_typedarray_type t;
    if (napi_is_typedarray(env, v, &is_ta) != napi_ok || !is_ta ||
        napi_get_typedarray_info(env, v, &t, len, data, nullptr, nullptr) != napi_ok || t != want) {
        napi_throw_type_error(env, "PCMB200_TYPE", what);
        return false;
    }
    return true;
}

AggSlot *agg_of(napi_env env, napi_value v)
{
    uint32_t h = 0;
    if (napi_get_value_uint32(env, v, &h) != napi_ok || h >= (uint32_t)kMaxAggregators || !g_aggs[h].agg) {
        napi_throw_type_error(env, "PCMB200_TYPE", "not an aggregator handle");
        return nullptr;
    }
    return &g_aggs[h];
}

napi_value AggregatorCreate(napi_env env, napi_callback_info info)
{
    size_t argc = 5;
    napi_value argv[5];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    uint32_t max_sessions = 0, max_frame_quanta = 0, max_push = 0;
    double rate = 0, max_out = 0;
    if (argc < 5 || napi_get_value_uint32(env, argv[0], &max_sessions) != napi_ok || napi_get_value_double(env, argv[1], &rate) != napi_ok ||
        napi_get_value_uint32(env, argv[2], &max_frame_quanta) != napi_ok || napi_get_value_uint32(env, argv[3], &max_push) != napi_ok ||
        napi_get_value_double(env, argv[4], &max_out) != napi_ok || !(max_out >= 1)) {
        napi_throw_type_error(env, "PCMB200_TYPE", "aggregatorCreate(maxSessions, rate, maxFrameQuanta, maxPushQuanta, maxOutSamples)");
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    int h = 0;
    while (h < kMaxAggregators && g_aggs[h].agg) h++;
    if (h == kMaxAggregators) { napi_throw_error(env, "PCMB200_6", "too many aggregators (16)"); return nullptr; }
    unsigned long long *offs = static_cast<unsigned long long *>(std::malloc(sizeof(unsigned long long) * (max_push ? max_push : 1)));
    if (!offs) { napi_throw_error(env, "PCMB200_5", "out of host memory"); return nullptr; }
    pcmb200_aggregator *a = nullptr;
    // 128-sample quanta and the reference's 60 ms (App.tsx:156)
    if (throw_status(env, pcmb200_aggregator_create(g_ctx, max_sessions, rate, 128, 60.0, max_frame_quanta, max_push, (size_t)max_out, &a))) {
        std::free(offs);
        return nullptr;
    }
    g_aggs[h] = AggSlot{a, offs, max_push};
    napi_value res = nullptr;
    napi_create_double(env, (double)h, &res);
    return res;
}

napi_value AggregatorReset(napi_env env, napi_callback_info info)
{
    size_t argc = 3;
    napi_value argv[3];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    uint32_t slot = 0;
    double now = 0;
    AggSlot *s = argc >= 3 ? agg_of(env, argv[0]) : nullptr;
    if (!s) return nullptr;
    if (napi_get_value_uint32(env, argv[1], &slot) != napi_ok || napi_get_value_double(env, argv[2], &now) != napi_ok) {
        napi_throw_type_error(env, "PCMB200_TYPE", "aggregatorReset(handle, slot, nowMs)");
        return nullptr;
    }
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    throw_status(env, pcmb200_aggregator_reset_session(g_ctx, s->agg, slot, now));
    return nullptr;
}

napi_value AggregatorPush(napi_env env, napi_callback_info info)
{
    size_t argc = 9;
    napi_value argv[9];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    AggSlot *s = argc >= 9 ? agg_of(env, argv[0]) : nullptr;
    if (!s) return nullptr;
    uint32_t n = 0;
    void *sess, *times, *samples, *out, *fs, *fl, *fo;
    size_t n_sess, n_times, n_samples, n_out, n_fs, n_fl, n_fo;
    if (napi_get_value_uint32(env, argv[1], &n) != napi_ok ||
        !get_ta(env, argv[2], napi_uint32_array, &sess, &n_sess, "sessions must be a Uint32Array") ||
        !get_ta(env, argv[3], napi_float64_array, &times, &n_times, "timesMs must be a Float64Array") ||
        !get_ta(env, argv[4], napi_float32_array, &samples, &n_samples, "samples must be a Float32Array") ||
        !get_ta(env, argv[5], napi_int16_array, &out, &n_out, "out must be an Int16Array") ||
        !get_ta(env, argv[6], napi_uint32_array, &fs, &n_fs, "frameSession must be a Uint32Array") ||
        !get_ta(env, argv[7], napi_uint32_array, &fl, &n_fl, "frameLen must be a Uint32Array") ||
        !get_ta(env, argv[8], napi_float64_array, &fo, &n_fo, "frameOffset must be a Float64Array"))
        return nullptr;
    size_t max_frames = n_fs < n_fl ? n_fs : n_fl;
    if (n_fo < max_frames) max_frames = n_fo;
    if (max_frames > s->max_push) max_frames = s->max_push;
    if (n > n_sess || n > n_times || (size_t)n * 128 > n_samples) {
        napi_throw_type_error(env, "PCMB200_TYPE", "aggregatorPush: n exceeds sessions / timesMs / samples (128 samples per quantum)");
        return nullptr;
    }
    uint32_t nf = 0;
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    if (throw_status(env, pcmb200_aggregator_push(g_ctx, s->agg, n, static_cast<const uint32_t *>(sess), static_cast<const double *>(times),
                                                  static_cast<const float *>(samples), PCMB200_MEM_HOST, static_cast<int16_t *>(out), n_out,
                                                  PCMB200_MEM_HOST, static_cast<uint32_t *>(fs), static_cast<uint32_t *>(fl), s->offsets,
                                                  (uint32_t)max_frames, &nf)))
        return nullptr;
    for (uint32_t f = 0; f < nf; f++) static_cast<double *>(fo)[f] = (double)s->offsets[f];   // sample offsets: far below 2^53
    napi_value res = nullptr;
    napi_create_double(env, (double)nf, &res);
    return res;
}

napi_value AggregatorDestroy(napi_env env, napi_callback_info info)
{
    size_t argc = 1;
    napi_value argv[1];
    napi_get_cb_info(env, info, &argc, argv, nullptr, nullptr);
    AggSlot *s = argc >= 1 ? agg_of(env, argv[0]) : nullptr;
    if (!s) return nullptr;
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    pcmb200_aggregator_destroy(g_ctx, s->agg);
    std::free(s->offsets);
    *s = AggSlot{};
    return nullptr;
}

// Environment teardown.  Waits (on the lock) for a conversion that is still running on a worker
// thread, then destroys the context; jobs that start later see g_ctx == NULL and reject.
void cleanup(void *)
{
    std::lock_guard<std::mutex> lock(g_ctx_mutex);
    for (AggSlot &s : g_aggs) {
        if (s.agg) pcmb200_aggregator_destroy(g_ctx, s.agg);
        std::free(s.offsets);
        s = AggSlot{};
    }
    pcmb200_ctx_destroy(g_ctx);
    g_ctx = nullptr;
}

napi_value Init(napi_env env, napi_value exports)
{
    int device = 0;
    if (const char *e = std::getenv("PCMB200_DEVICE")) device = std::atoi(e);
    if (throw_status(env, pcmb200_ctx_create(device, &g_ctx))) return nullptr;   // loud failure, no CPU path
    napi_add_env_cleanup_hook(env, cleanup, nullptr);
    const napi_property_descriptor props[] = {
        {"floatTo16BitPCM", nullptr, FloatTo16BitPCM, nullptr, nullptr, nullptr, 0, nullptr},
        {"resampleTo16k", nullptr, ResampleTo16k, nullptr, nullptr, nullptr, 0, nullptr},
        {"framesEncode", nullptr, FramesEncode, nullptr, nullptr, nullptr, 0, nullptr},
        {"framesEncodeAsync", nullptr, FramesEncodeAsync, nullptr, nullptr, nullptr, 0, nullptr},
        {"resampleLen", nullptr, ResampleLen, nullptr, nullptr, nullptr, 0, nullptr},
        {"allocPinnedFloat32", nullptr, AllocPinnedFloat32, nullptr, nullptr, nullptr, 0, nullptr},
        {"allocPinnedInt16", nullptr, AllocPinnedInt16, nullptr, nullptr, nullptr, 0, nullptr},
        {"aggregatorCreate", nullptr, AggregatorCreate, nullptr, nullptr, nullptr, 0, nullptr},
        {"aggregatorReset", nullptr, AggregatorReset, nullptr, nullptr, nullptr, 0, nullptr},
        {"aggregatorPush", nullptr, AggregatorPush, nullptr, nullptr, nullptr, 0, nullptr},
        {"aggregatorDestroy", nullptr, AggregatorDestroy, nullptr, nullptr, nullptr, 0, nullptr},
    };
    napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
    return exports;
}

napi_module g_module = {1, 0, __FILE__, Init, "pcm_b200_napi", nullptr, {nullptr, nullptr, nullptr, nullptr}};

}  // namespace

// Node resolves this symbol (Node-API module version 1) when it dlopens the .node file.
extern "C" __attribute__((visibility("default"))) napi_value napi_register_module_v1(napi_env env, napi_value exports)
{
    return Init(env, exports);
}
extern "C" __attribute__((constructor)) void pcmb200_napi_register(void)
{
    // legacy registration path; harmless when napi_register_module_v1 is used
    (void)g_module;
}
