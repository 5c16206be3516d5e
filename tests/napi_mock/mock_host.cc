// mock_host.cc -- a minimal stand-in for the Node-API runtime, test infrastructure only.
//
// Node is not installed on this image, so the N-API shim (voice-enabled-browser-automation_b200/
// node/addon.cc) cannot be loaded by a real `node`.  This program implements the ~27 napi_*
// functions the shim uses with just enough of their documented semantics (values, typed arrays over
// array buffers, pending exceptions, promises, async work), exports them (-rdynamic), dlopens the
// built pcm_b200_napi.node exactly as Node would, calls napi_register_module_v1 and then drives the
// exported functions the way JavaScript would.  Exit code 0 = every check passed.
//
//   g++ -std=c++17 -rdynamic tests/napi_mock/mock_host.cc -ldl -lpthread -o mock_host
//   ./mock_host voice-enabled-browser-automation_b200/lib/pcm_b200_napi.node [--expect-no-device]
#include <dlfcn.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../voice-enabled-browser-automation_b200/node/napi_min.h"

// ---- the mock object model --------------------------------------------------------------------------
struct napi_value__ {
    enum Kind { Undefined, Number, String, ArrayBuffer, TypedArray, Object, Function, Error, Promise } kind = Undefined;
    double num = 0;
    std::string str;                 // String value / Error message
    std::string err_class;           // "Error" | "TypeError" | "RangeError"
    std::string code;
    std::vector<uint8_t> owned;      // ArrayBuffer storage (when not external)
    void *data = nullptr;            // ArrayBuffer / TypedArray data pointer
    size_t byte_len = 0;
    napi_finalize finalize = nullptr;
    napi_typedarray_type ta_type = napi_int8_array;
    size_t length = 0, byte_offset = 0;
    napi_value__ *buffer = nullptr;
    std::map<std::string, napi_value__ *> props;
    napi_callback cb = nullptr;
    int promise_state = 0;           // 0 pending, 1 resolved, 2 rejected
    napi_value__ *promise_value = nullptr;
};
struct napi_env__ {
    napi_value__ *pending = nullptr;
    std::vector<std::function<void()>> cleanup;
    std::vector<std::function<void()>> completions;   // async work finished, to run on the "main thread"
    std::vector<std::thread> workers;
};
struct napi_callback_info__ { std::vector<napi_value> args; };
struct napi_deferred__ { napi_value__ *promise; };
struct napi_async_work__ { napi_async_execute_callback exec; napi_async_complete_callback done; void *data; };
struct napi_ref__ { napi_value__ *v; };

static napi_value mk(napi_value__::Kind k) { napi_value v = new napi_value__(); v->kind = k; return v; }
static size_t elem_size(napi_typedarray_type t)
{
    switch (t) {
    case napi_int8_array: case napi_uint8_array: case napi_uint8_clamped_array: return 1;
    case napi_int16_array: case napi_uint16_array: return 2;
    case napi_int32_array: case napi_uint32_array: case napi_float32_array: return 4;
    default: return 8;
    }
}
static napi_status throw_err(napi_env env, const char *cls, const char *code, const char *msg)
{
    napi_value e = mk(napi_value__::Error);
    e->err_class = cls; e->code = code ? code : ""; e->str = msg ? msg : "";
    env->pending = e;
    return napi_ok;
}

extern "C" {
napi_status napi_get_cb_info(napi_env, napi_callback_info info, size_t *argc, napi_value *argv, napi_value *, void **)
{
    const size_t cap = *argc;
    for (size_t i = 0; i < cap; i++) argv[i] = i < info->args.size() ? info->args[i] : mk(napi_value__::Undefined);
    *argc = info->args.size();
    return napi_ok;
}
napi_status napi_is_typedarray(napi_env, napi_value v, bool *r) { *r = v && v->kind == napi_value__::TypedArray; return napi_ok; }
napi_status napi_get_typedarray_info(napi_env, napi_value v, napi_typedarray_type *type, size_t *length, void **data,
                                     napi_value *ab, size_t *off)
{
    if (!v || v->kind != napi_value__::TypedArray) return napi_invalid_arg;
    if (type) *type = v->ta_type;
    if (length) *length = v->length;
    if (data) *data = static_cast<uint8_t *>(v->buffer->data) + v->byte_offset;
    if (ab) *ab = v->buffer;
    if (off) *off = v->byte_offset;
    return napi_ok;
}
napi_status napi_create_arraybuffer(napi_env, size_t n, void **data, napi_value *result)
{
    napi_value v = mk(napi_value__::ArrayBuffer);
    v->owned.assign(n ? n : 1, 0);
    v->data = v->owned.data(); v->byte_len = n;
    if (data) *data = v->data;
    *result = v;
    return napi_ok;
}
napi_status napi_create_external_arraybuffer(napi_env, void *ext, size_t n, napi_finalize fin, void *, napi_value *result)
{
    napi_value v = mk(napi_value__::ArrayBuffer);
    v->data = ext; v->byte_len = n; v->finalize = fin;
    *result = v;
    return napi_ok;
}
napi_status napi_create_typedarray(napi_env, napi_typedarray_type t, size_t len, napi_value ab, size_t off, napi_value *result)
{
    if (!ab || ab->kind != napi_value__::ArrayBuffer || off + len * elem_size(t) > ab->byte_len) return napi_invalid_arg;
    napi_value v = mk(napi_value__::TypedArray);
    v->ta_type = t; v->length = len; v->byte_offset = off; v->buffer = ab;
    *result = v;
    return napi_ok;
}
napi_status napi_get_value_double(napi_env, napi_value v, double *r)
{
    if (!v || v->kind != napi_value__::Number) return napi_number_expected;
    *r = v->num;
    return napi_ok;
}
napi_status napi_get_value_uint32(napi_env, napi_value v, uint32_t *r)
{
    if (!v || v->kind != napi_value__::Number) return napi_number_expected;
    *r = (uint32_t)(long long)v->num;   // ToUint32 for the in-range values used here
    return napi_ok;
}
napi_status napi_create_double(napi_env, double d, napi_value *r) { *r = mk(napi_value__::Number); (*r)->num = d; return napi_ok; }
napi_status napi_create_string_utf8(napi_env, const char *s, size_t n, napi_value *r)
{
    *r = mk(napi_value__::String);
    (*r)->str = n == NAPI_AUTO_LENGTH ? std::string(s) : std::string(s, n);
    return napi_ok;
}
static napi_status mk_error(const char *cls, napi_value code, napi_value msg, napi_value *r)
{
    *r = mk(napi_value__::Error);
    (*r)->err_class = cls; (*r)->str = msg ? msg->str : ""; (*r)->code = code ? code->str : "";
    return napi_ok;
}
napi_status napi_create_error(napi_env, napi_value c, napi_value m, napi_value *r) { return mk_error("Error", c, m, r); }
napi_status napi_create_range_error(napi_env, napi_value c, napi_value m, napi_value *r) { return mk_error("RangeError", c, m, r); }
napi_status napi_throw(napi_env env, napi_value e) { env->pending = e; return napi_ok; }
napi_status napi_throw_error(napi_env env, const char *c, const char *m) { return throw_err(env, "Error", c, m); }
napi_status napi_throw_type_error(napi_env env, const char *c, const char *m) { return throw_err(env, "TypeError", c, m); }
napi_status napi_throw_range_error(napi_env env, const char *c, const char *m) { return throw_err(env, "RangeError", c, m); }
napi_status napi_define_properties(napi_env, napi_value obj, size_t n, const napi_property_descriptor *p)
{
    for (size_t i = 0; i < n; i++) {
        napi_value f = mk(napi_value__::Function);
        f->cb = p[i].method;
        obj->props[p[i].utf8name] = f;
    }
    return napi_ok;
}
napi_status napi_add_env_cleanup_hook(napi_env env, void (*fun)(void *), void *arg)
{
    env->cleanup.push_back([=] { fun(arg); });
    return napi_ok;
}
napi_status napi_create_promise(napi_env, napi_deferred *d, napi_value *p)
{
    *p = mk(napi_value__::Promise);
    *d = new napi_deferred__{*p};
    return napi_ok;
}
napi_status napi_resolve_deferred(napi_env, napi_deferred d, napi_value v) { d->promise->promise_state = 1; d->promise->promise_value = v; delete d; return napi_ok; }
napi_status napi_reject_deferred(napi_env, napi_deferred d, napi_value v) { d->promise->promise_state = 2; d->promise->promise_value = v; delete d; return napi_ok; }
napi_status napi_create_async_work(napi_env, napi_value, napi_value, napi_async_execute_callback ex, napi_async_complete_callback done,
                                   void *data, napi_async_work *r)
{
    *r = new napi_async_work__{ex, done, data};
    return napi_ok;
}
napi_status napi_queue_async_work(napi_env env, napi_async_work w)
{
    // execute on another thread (like libuv's pool); completion runs later on the main thread
    env->workers.emplace_back([env, w] { w->exec(env, w->data); });
    env->completions.push_back([env, w] { w->done(env, napi_ok, w->data); });
    return napi_ok;
}
napi_status napi_delete_async_work(napi_env, napi_async_work w) { delete w; return napi_ok; }
napi_status napi_create_reference(napi_env, napi_value v, uint32_t, napi_ref *r) { *r = new napi_ref__{v}; return napi_ok; }
napi_status napi_delete_reference(napi_env, napi_ref r) { delete r; return napi_ok; }
napi_status napi_get_reference_value(napi_env, napi_ref r, napi_value *v) { *v = r->v; return napi_ok; }
void napi_module_register(napi_module *) {}
}

// ---- the "JavaScript" side -----------------------------------------------------------------------------
static int g_failed = 0;
#define CHECK(cond, what)                                                        \
    do {                                                                         \
        if (cond) std::printf("ok   %s\n", what);                                \
        else { std::printf("FAIL %s\n", what); g_failed++; }                     \
    } while (0)

static napi_value call(napi_env env, napi_value exports, const char *name, std::vector<napi_value> args)
{
    env->pending = nullptr;
    napi_callback_info__ info{std::move(args)};
    return exports->props.at(name)->cb(env, &info);
}
static napi_value num(double d) { napi_value v = mk(napi_value__::Number); v->num = d; return v; }
template <typename T>
static napi_value typed(napi_env env, napi_typedarray_type t, const std::vector<T> &vals)
{
    void *data; napi_value ab, ta;
    napi_create_arraybuffer(env, vals.size() * sizeof(T), &data, &ab);
    std::memcpy(data, vals.data(), vals.size() * sizeof(T));
    napi_create_typedarray(env, t, vals.size(), ab, 0, &ta);
    return ta;
}
template <typename T> static T *ptr(napi_value ta) { return reinterpret_cast<T *>(static_cast<uint8_t *>(ta->buffer->data) + ta->byte_offset); }

int main(int argc, char **argv)
{
    if (argc < 2) { std::fprintf(stderr, "usage: mock_host <addon.node> [--expect-no-device]\n"); return 2; }
    const bool expect_no_device = argc > 2 && std::string(argv[2]) == "--expect-no-device";
    void *h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { std::fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto reg = reinterpret_cast<napi_value (*)(napi_env, napi_value)>(dlsym(h, "napi_register_module_v1"));
    CHECK(reg != nullptr, "addon exports napi_register_module_v1");
    if (!reg) return 1;
    napi_env__ env_obj;
    napi_env env = &env_obj;
    napi_value exports = mk(napi_value__::Object);
    napi_value ret = reg(env, exports);

    if (expect_no_device) {   // no GPU: loading the module must throw, not fall back to a CPU path
        CHECK(env->pending != nullptr && ret == nullptr, "module load throws without a CUDA device");
        CHECK(env->pending && env->pending->str.find("no usable CUDA device") != std::string::npos, "the message says why");
        CHECK(exports->props.empty(), "nothing was exported");
        return g_failed ? 1 : 0;
    }
    CHECK(env->pending == nullptr && ret == exports, "module load succeeds");
    for (const char *n : {"floatTo16BitPCM", "resampleTo16k", "framesEncode", "framesEncodeAsync", "resampleLen",
                          "allocPinnedFloat32", "allocPinnedInt16"})
        CHECK(exports->props.count(n) == 1, n);

    // floatTo16BitPCM(Float32Array) -> fresh Int16Array                        App.tsx:7-16
    napi_value f = typed<float>(env, napi_float32_array, {1.f, -1.f, 0.5f, -0.5f, NAN, INFINITY, -INFINITY, 0.99999994f, 1.5f});
    napi_value r = call(env, exports, "floatTo16BitPCM", {f});
    const int16_t want[] = {32767, -32768, 16383, -16384, 0, 32767, -32768, 32766, 32767};
    CHECK(r && r->kind == napi_value__::TypedArray && r->ta_type == napi_int16_array && r->length == 9 && r->buffer != f->buffer,
          "floatTo16BitPCM returns a fresh Int16Array of the same length");
    CHECK(r && std::memcmp(ptr<int16_t>(r), want, sizeof want) == 0, "floatTo16BitPCM values (clamp, truncation, NaN -> 0)");
    r = call(env, exports, "floatTo16BitPCM", {typed<int16_t>(env, napi_int16_array, {1, 2})});
    CHECK(r == nullptr && env->pending && env->pending->err_class == "TypeError", "floatTo16BitPCM(Int16Array) throws TypeError");
    r = call(env, exports, "floatTo16BitPCM", {typed<float>(env, napi_float32_array, {})});
    CHECK(r && r->length == 0 && !env->pending, "floatTo16BitPCM(empty) -> empty");

    // resampleTo16k(Float32Array, rate)                                          App.tsx:18-32
    napi_value six = typed<float>(env, napi_float32_array, {0.f, 1.f, 2.f, 3.f, 4.f, 5.f});
    r = call(env, exports, "resampleTo16k", {six, num(16000)});
    CHECK(r == six, "resampleTo16k at 16 kHz returns the same object");
    r = call(env, exports, "resampleTo16k", {six, num(48000)});
    CHECK(r && r->ta_type == napi_float32_array && r->length == 2 && ptr<float>(r)[0] == 0.f && ptr<float>(r)[1] == 3.f,
          "resampleTo16k(48 kHz) picks samples 0 and 3");
    r = call(env, exports, "resampleTo16k", {six, num(8000)});
    CHECK(r && r->length == 12 && ptr<float>(r)[1] == 0.f && ptr<float>(r)[2] == 1.f && ptr<float>(r)[11] == 5.f,
          "resampleTo16k(8 kHz) repeats samples");
    r = call(env, exports, "resampleTo16k", {six, num(0)});
    CHECK(r == nullptr && env->pending && env->pending->err_class == "RangeError", "resampleTo16k(rate 0) throws RangeError");
    r = call(env, exports, "resampleLen", {num(2944), num(48000)});
    CHECK(r && r->num == 981, "resampleLen(2944, 48000) == 981");

    // framesEncode(batch, rows, nIn, rate[, out])                               App.tsx:281-282 batched
    std::vector<float> batch(4 * 960);
    for (size_t i = 0; i < batch.size(); i++) batch[i] = 0.9f * std::sin(0.01f * (float)i);
    napi_value b = typed<float>(env, napi_float32_array, batch);
    r = call(env, exports, "framesEncode", {b, num(4), num(960), num(48000)});
    bool same = r && r->length == 4 * 320;
    for (size_t k = 0; same && k < 4 * 320; k++) {
        const double s = batch[3 * k];
        const double v = s < 0 ? s * 32768.0 : s * 32767.0;
        same = ptr<int16_t>(r)[k] == (int16_t)std::trunc(v);
    }
    CHECK(same, "framesEncode(4 x 960 @ 48 kHz) equals trunc(scale(pick every third sample))");
    r = call(env, exports, "framesEncode", {b, num(5), num(960), num(48000)});
    CHECK(r == nullptr && env->pending && env->pending->err_class == "TypeError", "framesEncode rejects rows * nIn > batch.length");

    // pinned allocators: external array buffers with a finalizer
    napi_value pin_in = call(env, exports, "allocPinnedFloat32", {num(2 * 960)});
    napi_value pin_out = call(env, exports, "allocPinnedInt16", {num(2 * 320)});
    CHECK(pin_in && pin_in->ta_type == napi_float32_array && pin_in->length == 1920 && pin_in->buffer->finalize != nullptr,
          "allocPinnedFloat32 returns an external Float32Array");
    CHECK(pin_out && pin_out->ta_type == napi_int16_array && pin_out->length == 640, "allocPinnedInt16");
    std::memcpy(ptr<float>(pin_in), batch.data(), 1920 * sizeof(float));
    r = call(env, exports, "framesEncode", {pin_in, num(2), num(960), num(48000), pin_out});
    CHECK(r == pin_out && ptr<int16_t>(pin_out)[1] == (int16_t)std::trunc(batch[3] * 32767.0), "framesEncode into a caller-provided pinned Int16Array");

    // framesEncodeAsync -> Promise resolved with `out` after the worker finished
    std::memset(ptr<int16_t>(pin_out), 0, 640 * 2);
    napi_value p = call(env, exports, "framesEncodeAsync", {pin_in, num(2), num(960), num(48000), pin_out});
    CHECK(p && p->kind == napi_value__::Promise && p->promise_state == 0, "framesEncodeAsync returns a pending Promise");
    for (auto &w : env->workers) w.join();           // "event loop": wait for the pool, then run completions
    env->workers.clear();
    for (auto &c : env->completions) c();
    env->completions.clear();
    CHECK(p && p->promise_state == 1 && p->promise_value == pin_out, "the Promise resolves with the out array");
    CHECK(ptr<int16_t>(pin_out)[2] == (int16_t)std::trunc(batch[6] * (batch[6] < 0 ? 32768.0 : 32767.0)), "async result is correct");
    p = call(env, exports, "framesEncodeAsync", {pin_in, num(2), num(960), num(-5), pin_out});
    CHECK(p == nullptr && env->pending && env->pending->err_class == "RangeError", "framesEncodeAsync(rate < 0) throws RangeError");

    // the capture aggregator (App.tsx:273-288) for two connections: 23 quanta each, 128e3/48000 ms apart; the 23rd
    // arrives 61.3 ms after lastSend = 0 and flushes 2944 samples -> 981 outputs per connection
    {
        napi_value h = call(env, exports, "aggregatorCreate", {num(4), num(48000), num(32), num(64), num(4096)});
        CHECK(h && h->kind == napi_value__::Number && !env->pending, "aggregatorCreate returns a handle");
        call(env, exports, "aggregatorReset", {h, num(0), num(0)});
        call(env, exports, "aggregatorReset", {h, num(1), num(0)});
        CHECK(!env->pending, "aggregatorReset");
        const int nq = 23, n = 2 * nq;
        std::vector<uint32_t> sess(n);
        std::vector<double> times(n);
        std::vector<float> x((size_t)n * 128);
        for (int q = 0; q < nq; q++)
            for (int c = 0; c < 2; c++) {
                const int i = q * 2 + c;
                sess[i] = (uint32_t)c;
                times[i] = (q + 1) * (128e3 / 48000);
                for (int k = 0; k < 128; k++) x[(size_t)i * 128 + k] = (c ? -0.5f : 0.25f) + 0.001f * (float)((q * 128 + k) % 97);
            }
        napi_value a_sess = typed<uint32_t>(env, napi_uint32_array, sess), a_t = typed<double>(env, napi_float64_array, times);
        napi_value a_x = typed<float>(env, napi_float32_array, x);
        napi_value a_out = typed<int16_t>(env, napi_int16_array, std::vector<int16_t>(4096, 0x1111));
        napi_value a_fs = typed<uint32_t>(env, napi_uint32_array, std::vector<uint32_t>(8, 9)), a_fl = typed<uint32_t>(env, napi_uint32_array, std::vector<uint32_t>(8, 9));
        napi_value a_fo = typed<double>(env, napi_float64_array, std::vector<double>(8, -1.0));
        r = call(env, exports, "aggregatorPush", {h, num(n), a_sess, a_t, a_x, a_out, a_fs, a_fl, a_fo});
        CHECK(r && !env->pending && r->num == 2.0, "aggregatorPush: both connections flush one frame");
        bool ok = r && ptr<uint32_t>(a_fl)[0] == 981 && ptr<uint32_t>(a_fl)[1] == 981 && ptr<uint32_t>(a_fs)[0] == 0 && ptr<uint32_t>(a_fs)[1] == 1 &&
                  ptr<double>(a_fo)[0] == 0.0 && ptr<double>(a_fo)[1] == 984.0;
        CHECK(ok, "frames: session order, 981 samples each, second frame starts at the next multiple of 8");
        // output o of connection c is sample 3 * o of its 2944-sample buffer: quantum q = 3o / 128, k = 3o % 128
        for (int c = 0; c < 2 && ok; c++)
            for (int o = 0; o < 981; o += 140) {
                const int q = (3 * o) / 128, k = (3 * o) % 128;
                const float v = x[(size_t)(q * 2 + c) * 128 + k];
                const int16_t want = (int16_t)std::trunc((double)v * (v < 0 ? 32768.0 : 32767.0));
                ok = ok && ptr<int16_t>(a_out)[c * 984 + o] == want;
            }
        CHECK(ok && ptr<int16_t>(a_out)[2 * 984] == 0x1111, "frame samples are the reference's (pick every 3rd, encode); nothing written past the frames");
        r = call(env, exports, "aggregatorPush", {h, num(n), a_sess, a_t, a_x, a_fs, a_fs, a_fl, a_fo});
        CHECK(r == nullptr && env->pending && env->pending->err_class == "TypeError", "aggregatorPush with a wrong array type throws TypeError");
        call(env, exports, "aggregatorDestroy", {h});
        CHECK(!env->pending, "aggregatorDestroy");
        r = call(env, exports, "aggregatorReset", {h, num(0), num(0)});
        CHECK(env->pending && env->pending->err_class == "TypeError", "a destroyed handle is refused");
    }

    // garbage collection of the pinned buffers, then environment teardown
    pin_in->buffer->finalize(env, pin_in->buffer->data, nullptr);
    pin_out->buffer->finalize(env, pin_out->buffer->data, nullptr);
    for (auto &c : env->cleanup) c();
    CHECK(true, "finalizers and cleanup hook ran");
    std::printf("%s\n", g_failed ? "FAILED" : "ALL OK");
    return g_failed ? 1 : 0;
}
