// napi_min.h -- hand-declared subset of Node-API (N-API) used by addon.cc.
// This image has neither Node nor node_api.h.  N-API is a stable C ABI, so the prototypes below
// (names, argument order and enum values as documented for Node-API version 6) are all the shim
// needs; the symbols stay undefined in the .node file and bind to the Node process at load time.
// When building inside a Node toolchain, include <node_api.h> instead (-DPCMB200_HAVE_NODE_API_H).
#pragma once
#ifdef PCMB200_HAVE_NODE_API_H
#include <node_api.h>
#else
#include <stddef.h>
#include <stdint.h>
extern "C" {
typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_callback_info__ *napi_callback_info;
typedef enum { napi_ok = 0, napi_invalid_arg, napi_object_expected, napi_string_expected, napi_name_expected,
               napi_function_expected, napi_number_expected, napi_boolean_expected, napi_array_expected,
               napi_generic_failure, napi_pending_exception } napi_status;
typedef enum { napi_int8_array, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array, napi_uint16_array,
               napi_int32_array, napi_uint32_array, napi_float32_array, napi_float64_array,
               napi_bigint64_array, napi_biguint64_array } napi_typedarray_type;
typedef napi_value (*napi_callback)(napi_env env, napi_callback_info info);
typedef void (*napi_finalize)(napi_env env, void *finalize_data, void *finalize_hint);
typedef struct { const char *utf8name; napi_value name; napi_callback method; napi_callback getter;
                 napi_callback setter; napi_value value; int attributes; void *data; } napi_property_descriptor;
typedef struct napi_module { int nm_version; unsigned int nm_flags; const char *nm_filename;
                             napi_value (*nm_register_func)(napi_env, napi_value); const char *nm_modname;
                             void *nm_priv; void *reserved[4]; } napi_module;

napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t *argc, napi_value *argv, napi_value *this_arg, void **data);
napi_status napi_is_typedarray(napi_env, napi_value, bool *result);
napi_status napi_get_typedarray_info(napi_env, napi_value, napi_typedarray_type *type, size_t *length, void **data,
                                     napi_value *arraybuffer, size_t *byte_offset);
napi_status napi_create_arraybuffer(napi_env, size_t byte_length, void **data, napi_value *result);
napi_status napi_create_external_arraybuffer(napi_env, void *external_data, size_t byte_length, napi_finalize finalize_cb,
                                             void *finalize_hint, napi_value *result);
napi_status napi_create_typedarray(napi_env, napi_typedarray_type type, size_t length, napi_value arraybuffer,
                                   size_t byte_offset, napi_value *result);
napi_status napi_get_value_double(napi_env, napi_value, double *result);
napi_status napi_get_value_uint32(napi_env, napi_value, uint32_t *result);
napi_status napi_create_double(napi_env, double value, napi_value *result);
napi_status napi_create_string_utf8(napi_env, const char *str, size_t length, napi_value *result);
napi_status napi_create_error(napi_env, napi_value code, napi_value msg, napi_value *result);
napi_status napi_create_range_error(napi_env, napi_value code, napi_value msg, napi_value *result);
napi_status napi_throw(napi_env, napi_value error);
napi_status napi_throw_error(napi_env, const char *code, const char *msg);
napi_status napi_throw_type_error(napi_env, const char *code, const char *msg);
napi_status napi_throw_range_error(napi_env, const char *code, const char *msg);
napi_status napi_define_properties(napi_env, napi_value object, size_t property_count, const napi_property_descriptor *properties);
napi_status napi_add_env_cleanup_hook(napi_env, void (*fun)(void *arg), void *arg);
typedef struct napi_deferred__ *napi_deferred;
typedef struct napi_async_work__ *napi_async_work;
typedef struct napi_ref__ *napi_ref;
typedef void (*napi_async_execute_callback)(napi_env env, void *data);
typedef void (*napi_async_complete_callback)(napi_env env, napi_status status, void *data);
napi_status napi_create_promise(napi_env, napi_deferred *deferred, napi_value *promise);
napi_status napi_resolve_deferred(napi_env, napi_deferred deferred, napi_value resolution);
napi_status napi_reject_deferred(napi_env, napi_deferred deferred, napi_value rejection);
napi_status napi_create_async_work(napi_env, napi_value async_resource, napi_value async_resource_name,
                                   napi_async_execute_callback execute, napi_async_complete_callback complete, void *data,
                                   napi_async_work *result);
napi_status napi_queue_async_work(napi_env, napi_async_work work);
napi_status napi_delete_async_work(napi_env, napi_async_work work);
napi_status napi_create_reference(napi_env, napi_value value, uint32_t initial_refcount, napi_ref *result);
napi_status napi_delete_reference(napi_env, napi_ref ref);
napi_status napi_get_reference_value(napi_env, napi_ref ref, napi_value *result);
void napi_module_register(napi_module *mod);
}
#define NAPI_AUTO_LENGTH ((size_t)-1)
#endif
