/*
 * addon/symrelay_napi.c -- thin N-API shim over include/symrelay.h (SOURCE ONLY).
 *
 * Not built or run in this repository's image: there is no Node.js, npm or
 * node_api.h here.  N-API is a stable C ABI, so the
 * handful of prototypes used are declared by hand below; on a machine with
 * Node >= 18 build with
 *     gcc -shared -fPIC -I../include symrelay_napi.c -L../symmetry_b200 -lsymrelay -o symrelay.node
 * It binds exactly the entry points that replace sodium-native's
 * crypto_secretstream_xchacha20poly1305_{init_push,init_pull,push,pull} as
 * reached from /root/reference/src/provider.ts:248 (`peer.write(chunk)`) and
 * :174 (`peer.on('data')`): host code stays TypeScript (addon/symrelay.ts).
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include "symrelay.h"

/* ---- minimal N-API surface (from node_api.h / js_native_api.h) ---------- */
typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_callback_info__ *napi_callback_info;
typedef enum { napi_ok = 0 } napi_status;
typedef napi_value (*napi_callback)(napi_env, napi_callback_info);
typedef void (*napi_finalize)(napi_env, void *data, void *hint);
typedef struct { const char *utf8name; napi_value name; napi_callback method; napi_callback getter;
                 napi_callback setter; napi_value value; int attributes; void *data; } napi_property_descriptor;
extern napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t *argc, napi_value *argv, napi_value *self, void **data);
extern napi_status napi_get_value_uint32(napi_env, napi_value, uint32_t *);
extern napi_status napi_get_value_int64(napi_env, napi_value, int64_t *);
extern napi_status napi_get_value_external(napi_env, napi_value, void **);
extern napi_status napi_create_external(napi_env, void *, napi_finalize, void *, napi_value *);
extern napi_status napi_create_external_arraybuffer(napi_env, void *, size_t, napi_finalize, void *, napi_value *);
extern napi_status napi_get_typedarray_info(napi_env, napi_value, int *type, size_t *len, void **data, napi_value *ab, size_t *off);
extern napi_status napi_create_uint32(napi_env, uint32_t, napi_value *);
extern napi_status napi_create_int64(napi_env, int64_t, napi_value *);
extern napi_status napi_throw_error(napi_env, const char *code, const char *msg);
extern napi_status napi_define_properties(napi_env, napi_value obj, size_t n, const napi_property_descriptor *);

static symrelay_ctx *ctx_of(napi_env env, napi_value v) { void *p = NULL; napi_get_value_external(env, v, &p); return (symrelay_ctx *)p; }
static void *bytes_of(napi_env env, napi_value v, size_t *len) { void *d = NULL; size_t n = 0; napi_get_typedarray_info(env, v, NULL, &n, &d, NULL, NULL); if (len) *len = n; return d; }
static napi_value fail(napi_env env, symrelay_ctx *c) { napi_throw_error(env, "ESYMRELAY", symrelay_last_error(c)); return NULL; }
static void free_ctx(napi_env env, void *data, void *hint) { (void)env; (void)hint; symrelay_destroy((symrelay_ctx *)data); }

/* create(gpuIds: Int32Array) -> external */
static napi_value js_create(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value argv[1]; napi_get_cb_info(env, info, &argc, argv, NULL, NULL);
    size_t n = 0; const int *ids = (const int *)bytes_of(env, argv[0], &n);
    symrelay_ctx *c = symrelay_create(ids, (int)n, NULL);
    if (!c) return fail(env, NULL);
    napi_value out; napi_create_external(env, c, free_ctx, NULL, &out); return out;
}
/* streamOpen(ctx, txKey, txHeaderOrState, rxKey, rxHeaderOrState, providerKind) -> slot
 * call site: hyperswarm 'connection' (provider.ts:55-58), with peer._encrypt.{key,state} and
 * peer._decrypt.{key,state} (types.ts:168-177) */
static napi_value js_stream_open(napi_env env, napi_callback_info info) {
    size_t argc = 6; napi_value a[6]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    symrelay_ctx *c = ctx_of(env, a[0]); size_t tl = 0, rl = 0; uint32_t kind = 0, slot = 0;
    const uint8_t *tk = bytes_of(env, a[1], NULL), *ti = bytes_of(env, a[2], &tl);
    const uint8_t *rk = bytes_of(env, a[3], NULL), *ri = bytes_of(env, a[4], &rl);
    napi_get_value_uint32(env, a[5], &kind);
    if (symrelay_stream_open(c, tk, ti, (int)tl, rk, ri, (int)rl, (int)kind, &slot)) return fail(env, c);
    napi_value out; napi_create_uint32(env, slot, &out); return out;
}
static napi_value js_stream_close(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0; napi_get_value_uint32(env, a[1], &slot);
    if (symrelay_stream_close(ctx_of(env, a[0]), slot)) return fail(env, ctx_of(env, a[0]));
    return NULL;
}
/* streamExport(ctx, slot, tx52: Uint8Array, rx52: Uint8Array): hand the state back to sodium-native */
static napi_value js_stream_export(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0; napi_get_value_uint32(env, a[1], &slot);
    if (symrelay_stream_export(ctx_of(env, a[0]), slot, bytes_of(env, a[2], NULL), bytes_of(env, a[3], NULL))) return fail(env, ctx_of(env, a[0]));
    return NULL;
}
/* allocPinned(ctx, bytes) -> ArrayBuffer over pinned memory: chunks are copied straight into it */
static void free_pinned(napi_env env, void *data, void *hint) { (void)env; symrelay_free_pinned((symrelay_ctx *)hint, data); }
static napi_value js_alloc_pinned(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    int64_t n = 0; napi_get_value_int64(env, a[1], &n);
    void *p = symrelay_alloc_pinned(ctx_of(env, a[0]), (uint64_t)n);
    if (!p) return fail(env, ctx_of(env, a[0]));
    napi_value out; napi_create_external_arraybuffer(env, p, (size_t)n, free_pinned, ctx_of(env, a[0]), &out); return out;
}
/* batchSubmit(ctx, desc: Uint32Array(4n), inArena, outArena, status: Uint32Array(n)) -> ticket */
static napi_value js_batch_submit(napi_env env, napi_callback_info info) {
    size_t argc = 5; napi_value a[5]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    symrelay_ctx *c = ctx_of(env, a[0]); size_t dl = 0, il = 0, ol = 0;
    const symrelay_desc *d = (const symrelay_desc *)bytes_of(env, a[1], &dl);
    const uint8_t *in = bytes_of(env, a[2], &il); uint8_t *out = bytes_of(env, a[3], &ol);
    uint32_t *st = bytes_of(env, a[4], NULL);
    int64_t t = symrelay_batch_submit(c, d, (uint32_t)(dl / 4), in, il, out, ol, st, NULL, NULL, 0);
    if (t < 0) return fail(env, c);
    napi_value r; napi_create_int64(env, t, &r); return r;
}
/* batchWait(ctx, ticket): in production call it from a napi_async_work / threadsafe-function worker so the
 * event loop never blocks on the GPU (SURVEY 8(b) "Threading") */
static napi_value js_batch_wait(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    int64_t t = 0; napi_get_value_int64(env, a[1], &t);
    if (symrelay_batch_wait(ctx_of(env, a[0]), t)) return fail(env, ctx_of(env, a[0]));
    return NULL;
}

napi_value napi_register_module_v1(napi_env env, napi_value exports) {
    const napi_property_descriptor props[] = {
        {"create", 0, js_create, 0, 0, 0, 0, 0}, {"streamOpen", 0, js_stream_open, 0, 0, 0, 0, 0},
        {"streamClose", 0, js_stream_close, 0, 0, 0, 0, 0}, {"streamExport", 0, js_stream_export, 0, 0, 0, 0, 0},
        {"allocPinned", 0, js_alloc_pinned, 0, 0, 0, 0, 0}, {"batchSubmit", 0, js_batch_submit, 0, 0, 0, 0, 0},
        {"batchWait", 0, js_batch_wait, 0, 0, 0, 0, 0},
    };
    napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
    return exports;
}
