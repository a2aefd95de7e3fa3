/*
 * addon/symrelay_napi.c -- thin N-API shim over include/symrelay.h (SOURCE ONLY).
 *
 * Not built or run in this repository's image: there is no Node.js, npm, node_api.h or uv.h here
 * (SURVEY.md section 0 fact 6).  N-API and libuv are stable C ABIs, so the prototypes used are declared
 * by hand below (tests/test_abi_symbols.py runs `gcc -fsyntax-only` over this file); on a machine with
 * Node >= 18 build with
 *     gcc -shared -fPIC -I../include symrelay_napi.c -L../symmetry_b200 -lsymrelay -o symrelay.node
 * It binds the entry points that replace sodium-native's
 * crypto_secretstream_xchacha20poly1305_{init_push,init_pull,push,pull} as reached from
 * /root/reference/src/provider.ts:248 (`peer.write(chunk)`) and :174 (`peer.on('data')`): host code stays
 * TypeScript (addon/symrelay.ts).  The event loop never blocks on the GPU: completions arrive through the
 * library's eventfd (symrelay_event_fd) watched by a uv_poll handle on Node's own loop.
 */
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "symrelay.h"

/* ---- minimal N-API surface (from node_api.h / js_native_api.h) ---------- */
typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_ref__ *napi_ref;
typedef struct napi_handle_scope__ *napi_handle_scope;
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
extern napi_status napi_create_buffer_copy(napi_env, size_t, const void *, void **, napi_value *);
extern napi_status napi_get_typedarray_info(napi_env, napi_value, int *type, size_t *len, void **data, napi_value *ab, size_t *off);
extern napi_status napi_create_uint32(napi_env, uint32_t, napi_value *);
extern napi_status napi_create_int32(napi_env, int32_t, napi_value *);
extern napi_status napi_create_int64(napi_env, int64_t, napi_value *);
extern napi_status napi_create_double(napi_env, double, napi_value *);
extern napi_status napi_create_object(napi_env, napi_value *);
extern napi_status napi_create_array_with_length(napi_env, size_t, napi_value *);
extern napi_status napi_set_element(napi_env, napi_value, uint32_t, napi_value);
extern napi_status napi_set_named_property(napi_env, napi_value, const char *, napi_value);
extern napi_status napi_get_null(napi_env, napi_value *);
extern napi_status napi_get_undefined(napi_env, napi_value *);
extern napi_status napi_throw_error(napi_env, const char *code, const char *msg);
extern napi_status napi_define_properties(napi_env, napi_value obj, size_t n, const napi_property_descriptor *);
extern napi_status napi_create_reference(napi_env, napi_value, uint32_t, napi_ref *);
extern napi_status napi_get_reference_value(napi_env, napi_ref, napi_value *);
extern napi_status napi_delete_reference(napi_env, napi_ref);
extern napi_status napi_open_handle_scope(napi_env, napi_handle_scope *);
extern napi_status napi_close_handle_scope(napi_env, napi_handle_scope);
extern napi_status napi_call_function(napi_env, napi_value recv, napi_value fn, size_t argc, const napi_value *argv, napi_value *result);
struct uv_loop_s;
extern napi_status napi_get_uv_event_loop(napi_env, struct uv_loop_s **);
/* ---- minimal libuv surface (uv.h) ---- */
typedef struct uv_poll_s uv_poll_t;
typedef void (*uv_poll_cb)(uv_poll_t *, int status, int events);
extern size_t uv_handle_size(int type);       /* UV_POLL == 8 in uv_handle_type */
extern int uv_poll_init(struct uv_loop_s *, uv_poll_t *, int fd);
extern int uv_poll_start(uv_poll_t *, int events, uv_poll_cb); /* UV_READABLE == 1 */
extern int uv_poll_stop(uv_poll_t *);
extern void uv_handle_set_data(void *handle, void *data);
extern void *uv_handle_get_data(const void *handle);
extern long read(int, void *, unsigned long);

typedef struct {            /* one per context: what the uv_poll callback needs */
    symrelay_ctx *ctx;
    napi_env env;
    napi_ref on_ready;      /* JS function called when completions are waiting */
    uv_poll_t *poll;
} shim_ctx;

static shim_ctx *shim_of(napi_env env, napi_value v) { void *p = NULL; napi_get_value_external(env, v, &p); return (shim_ctx *)p; }
static symrelay_ctx *ctx_of(napi_env env, napi_value v) { shim_ctx *s = shim_of(env, v); return s ? s->ctx : NULL; }
static void *bytes_of(napi_env env, napi_value v, size_t *len) { void *d = NULL; size_t n = 0; napi_get_typedarray_info(env, v, NULL, &n, &d, NULL, NULL); if (len) *len = n; return d; }
static napi_value fail(napi_env env, symrelay_ctx *c) { napi_throw_error(env, "ESYMRELAY", symrelay_last_error(c)); return NULL; }
static napi_value i64(napi_env env, int64_t v) { napi_value r; napi_create_int64(env, v, &r); return r; }
static void free_shim(napi_env env, void *data, void *hint) {
    shim_ctx *s = (shim_ctx *)data; (void)hint;
    if (s->poll) uv_poll_stop(s->poll);
    if (s->on_ready) napi_delete_reference(env, s->on_ready);
    symrelay_destroy(s->ctx);
    free(s);
}

/* create(gpuIds: Int32Array, flushMicros, maxStreams) -> external */
static napi_value js_create(napi_env env, napi_callback_info info) {
    size_t argc = 3; napi_value a[3]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    size_t n = 0; const int *ids = (const int *)bytes_of(env, a[0], &n);
    symrelay_opts o; memset(&o, 0, sizeof o);
    if (argc > 1) napi_get_value_uint32(env, a[1], &o.sched_flush_us);
    if (argc > 2) napi_get_value_uint32(env, a[2], &o.max_streams);
    symrelay_ctx *c = symrelay_create(ids, (int)n, &o);
    if (!c) return fail(env, NULL);
    shim_ctx *s = (shim_ctx *)calloc(1, sizeof *s);
    s->ctx = c; s->env = env;
    napi_value out; napi_create_external(env, s, free_shim, NULL, &out); return out;
}

/* onReady(ctx, fn): fn() is called on the event loop whenever finished frames are waiting (eventfd -> uv_poll) */
static void on_event(uv_poll_t *h, int status, int events) {
    shim_ctx *s = (shim_ctx *)uv_handle_get_data(h); (void)status; (void)events;
    uint64_t count; if (read(symrelay_event_fd(s->ctx), &count, 8) < 0) return;   /* drain the counter */
    napi_handle_scope scope; napi_open_handle_scope(s->env, &scope);
    napi_value fn, undef; napi_get_reference_value(s->env, s->on_ready, &fn); napi_get_undefined(s->env, &undef);
    napi_call_function(s->env, undef, fn, 0, NULL, NULL);
    napi_close_handle_scope(s->env, scope);
}
static napi_value js_on_ready(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    shim_ctx *s = shim_of(env, a[0]);
    const int fd = symrelay_event_fd(s->ctx);            /* starts the library's watcher thread */
    if (fd < 0) return fail(env, s->ctx);
    napi_create_reference(env, a[1], 1, &s->on_ready);
    struct uv_loop_s *loop = NULL; napi_get_uv_event_loop(env, &loop);
    s->poll = (uv_poll_t *)malloc(uv_handle_size(8 /* UV_POLL */));
    uv_poll_init(loop, s->poll, fd);
    uv_handle_set_data(s->poll, s);
    uv_poll_start(s->poll, 1 /* UV_READABLE */, on_event);
    return NULL;
}

/* streamOpen(ctx, txKey, txHeaderOrState, rxKey|null, rxHeaderOrState|null, providerKind) -> slot
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
    napi_value r; napi_create_int32(env, symrelay_stream_close(ctx_of(env, a[0]), slot), &r); return r;   /* SYMRELAY_EAGAIN: frames still queued */
}
/* streamExport(ctx, slot, tx52: Uint8Array, rx52: Uint8Array): hand the state back to sodium-native */
static napi_value js_stream_export(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0; napi_get_value_uint32(env, a[1], &slot);
    if (symrelay_stream_export(ctx_of(env, a[0]), slot, bytes_of(env, a[2], NULL), bytes_of(env, a[3], NULL))) return fail(env, ctx_of(env, a[0]));
    return NULL;
}
/* rxExpectHeader(ctx, slot, rxKey, expectedStreamId|null); setKeepAlive(ctx, slot, on) */
static napi_value js_rx_expect_header(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0; napi_get_value_uint32(env, a[1], &slot);
    if (symrelay_rx_expect_header(ctx_of(env, a[0]), slot, bytes_of(env, a[2], NULL), bytes_of(env, a[3], NULL))) return fail(env, ctx_of(env, a[0]));
    return NULL;
}
static napi_value js_set_keepalive(napi_env env, napi_callback_info info) {
    size_t argc = 3; napi_value a[3]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0, on = 0; napi_get_value_uint32(env, a[1], &slot); napi_get_value_uint32(env, a[2], &on);
    if (symrelay_stream_set_keepalive(ctx_of(env, a[0]), slot, (int)on)) return fail(env, ctx_of(env, a[0]));
    return NULL;
}
/* connectBatch(ctx, gpu, hashes: Uint8Array(64n), initiator: Uint8Array(n), nsInitiator, nsResponder,
 *              localIds: Uint8Array(32n), remoteIds: Uint8Array(32n), headers: Uint8Array(24n), firstMessages: Uint8Array(59n)) */
static napi_value js_connect_batch(napi_env env, napi_callback_info info) {
    size_t argc = 10; napi_value a[10]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t gpu = 0; size_t n = 0; napi_get_value_uint32(env, a[1], &gpu);
    const uint8_t *ini = bytes_of(env, a[3], &n);
    if (symrelay_connect_batch(ctx_of(env, a[0]), (int)gpu, (uint32_t)n, bytes_of(env, a[2], NULL), ini, bytes_of(env, a[4], NULL),
                               bytes_of(env, a[5], NULL), NULL, bytes_of(env, a[6], NULL), bytes_of(env, a[7], NULL),
                               bytes_of(env, a[8], NULL), bytes_of(env, a[9], NULL))) return fail(env, ctx_of(env, a[0]));
    return NULL;
}

/* submitSeal(ctx, slot, chunk: Uint8Array, flags) -> ticket | SYMRELAY_EAGAIN  (`peer.write(chunk)`, provider.ts:248)
 * The chunk is copied into the library's pinned ring before the call returns: V8 may move or free it afterwards. */
static napi_value js_submit_seal(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0, flags = 0; size_t len = 0; napi_get_value_uint32(env, a[1], &slot); napi_get_value_uint32(env, a[3], &flags);
    const uint8_t *p = bytes_of(env, a[2], &len);
    const int64_t t = symrelay_submit_seal(ctx_of(env, a[0]), slot, p, (uint32_t)len, flags);
    if (t < 0 && t != SYMRELAY_EAGAIN) return fail(env, ctx_of(env, a[0]));
    return i64(env, t);
}
/* rxFeed(ctx, slot, segment: Uint8Array, tickets: BigInt64Array|Float64Array view as Int64) -> {frames, consumed}
 * secret-stream `_onrawdata` (types.ts:156-159): raw transport bytes, cut anywhere */
static napi_value js_rx_feed(napi_env env, napi_callback_info info) {
    size_t argc = 4; napi_value a[4]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    uint32_t slot = 0, used = 0; size_t len = 0, tl = 0; napi_get_value_uint32(env, a[1], &slot);
    const uint8_t *p = bytes_of(env, a[2], &len);
    int64_t *tickets = (int64_t *)bytes_of(env, a[3], &tl);
    const int n = symrelay_rx_feed(ctx_of(env, a[0]), slot, p, (uint32_t)len, &used, tickets, (int)tl);
    if (n < 0 && n != SYMRELAY_EAGAIN) return fail(env, ctx_of(env, a[0]));       /* SYMRELAY_EPROTO: destroy the stream */
    napi_value o, v; napi_create_object(env, &o);
    napi_create_int32(env, n, &v); napi_set_named_property(env, o, "frames", v);
    napi_create_uint32(env, used, &v); napi_set_named_property(env, o, "consumed", v);
    return o;
}
static napi_value js_flush(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    napi_value r; napi_create_int32(env, symrelay_flush(ctx_of(env, a[0])), &r); return r;
}
/* poll(ctx, max) -> [{ticket, slot, status, tag, open, data: Buffer, delta: Buffer, latencyNs}]
 * Never blocks.  data/delta are COPIED out of the pinned ring and the completion is released here, so JS owns plain
 * Buffers (an external ArrayBuffer over the ring + release in its finalizer saves the copy; kept simple here). */
static napi_value js_poll(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    symrelay_ctx *c = ctx_of(env, a[0]); uint32_t max = 256; if (argc > 1) napi_get_value_uint32(env, a[1], &max);
    if (max > 1024) max = 1024;
    symrelay_completion comp[1024];
    const int n = symrelay_poll(c, comp, (int)max);
    if (n < 0) return fail(env, c);
    napi_value arr; napi_create_array_with_length(env, (size_t)n, &arr);
    for (int i = 0; i < n; i++) {
        napi_value o, v; void *copy; napi_create_object(env, &o);
        napi_set_named_property(env, o, "ticket", i64(env, comp[i].ticket));
        napi_create_uint32(env, comp[i].slot, &v); napi_set_named_property(env, o, "slot", v);
        napi_create_uint32(env, SYMRELAY_STATUS_CODE(comp[i].status), &v); napi_set_named_property(env, o, "status", v);
        napi_create_uint32(env, SYMRELAY_STATUS_TAG(comp[i].status), &v); napi_set_named_property(env, o, "tag", v);
        napi_create_uint32(env, (comp[i].flags & SYMRELAY_F_OPEN) ? 1u : 0u, &v); napi_set_named_property(env, o, "open", v);
        napi_create_buffer_copy(env, comp[i].len, comp[i].data ? (const void *)comp[i].data : (const void *)"", &copy, &v);
        napi_set_named_property(env, o, "data", v);
        napi_create_buffer_copy(env, comp[i].delta_len, comp[i].delta ? (const void *)comp[i].delta : (const void *)"", &copy, &v);
        napi_set_named_property(env, o, "delta", v);
        napi_create_double(env, (double)comp[i].latency_ns, &v); napi_set_named_property(env, o, "latencyNs", v);
        napi_set_element(env, arr, (uint32_t)i, o);
        symrelay_release(c, comp[i].ticket);
    }
    return arr;
}
/* schedStats(ctx) -> {frames, batches, p50Us, p99Us, meanBatchFill} */
static napi_value js_sched_stats(napi_env env, napi_callback_info info) {
    size_t argc = 1; napi_value a[1]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    symrelay_sched_stats_t st; if (symrelay_sched_stats(ctx_of(env, a[0]), &st)) return fail(env, ctx_of(env, a[0]));
    napi_value o, v; napi_create_object(env, &o);
    napi_create_double(env, (double)st.frames, &v); napi_set_named_property(env, o, "frames", v);
    napi_create_double(env, (double)st.batches, &v); napi_set_named_property(env, o, "batches", v);
    napi_create_double(env, st.p50_us, &v); napi_set_named_property(env, o, "p50Us", v);
    napi_create_double(env, st.p99_us, &v); napi_set_named_property(env, o, "p99Us", v);
    napi_create_double(env, st.mean_batch_fill, &v); napi_set_named_property(env, o, "meanBatchFill", v);
    return o;
}

/* ---- the batch calls, for callers that build their own batches (benchmarks) ---- */
static void free_pinned(napi_env env, void *data, void *hint) { (void)env; symrelay_free_pinned((symrelay_ctx *)hint, data); }
static napi_value js_alloc_pinned(napi_env env, napi_callback_info info) {
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    int64_t n = 0; napi_get_value_int64(env, a[1], &n);
    void *p = symrelay_alloc_pinned(ctx_of(env, a[0]), (uint64_t)n);
    if (!p) return fail(env, ctx_of(env, a[0]));
    napi_value out; napi_create_external_arraybuffer(env, p, (size_t)n, free_pinned, ctx_of(env, a[0]), &out); return out;
}
/* batchSubmit(ctx, desc: Uint32Array(4n), inArena, outArena, status: Uint32Array(n)) -> ticket; every buffer must stay
 * untouched until batchWait(ticket) -- use one arena pair per batch in flight */
static napi_value js_batch_submit(napi_env env, napi_callback_info info) {
    size_t argc = 5; napi_value a[5]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    symrelay_ctx *c = ctx_of(env, a[0]); size_t dl = 0, il = 0, ol = 0;
    const symrelay_desc *d = (const symrelay_desc *)bytes_of(env, a[1], &dl);
    const uint8_t *in = bytes_of(env, a[2], &il); uint8_t *out = bytes_of(env, a[3], &ol);
    uint32_t *st = bytes_of(env, a[4], NULL);
    int64_t t = symrelay_batch_submit(c, d, (uint32_t)(dl / 4), in, il, out, ol, st, NULL, NULL, 0);
    if (t < 0) return fail(env, c);
    return i64(env, t);
}
static napi_value js_batch_wait(napi_env env, napi_callback_info info) {   /* blocks: benchmarks only, never on a serving loop */
    size_t argc = 2; napi_value a[2]; napi_get_cb_info(env, info, &argc, a, NULL, NULL);
    int64_t t = 0; napi_get_value_int64(env, a[1], &t);
    if (symrelay_batch_wait(ctx_of(env, a[0]), t)) return fail(env, ctx_of(env, a[0]));
    return NULL;
}

napi_value napi_register_module_v1(napi_env env, napi_value exports) {
    const napi_property_descriptor props[] = {
        {"create", 0, js_create, 0, 0, 0, 0, 0}, {"onReady", 0, js_on_ready, 0, 0, 0, 0, 0},
        {"streamOpen", 0, js_stream_open, 0, 0, 0, 0, 0}, {"streamClose", 0, js_stream_close, 0, 0, 0, 0, 0},
        {"streamExport", 0, js_stream_export, 0, 0, 0, 0, 0}, {"rxExpectHeader", 0, js_rx_expect_header, 0, 0, 0, 0, 0},
        {"setKeepAlive", 0, js_set_keepalive, 0, 0, 0, 0, 0}, {"connectBatch", 0, js_connect_batch, 0, 0, 0, 0, 0},
        {"submitSeal", 0, js_submit_seal, 0, 0, 0, 0, 0}, {"rxFeed", 0, js_rx_feed, 0, 0, 0, 0, 0},
        {"flush", 0, js_flush, 0, 0, 0, 0, 0}, {"poll", 0, js_poll, 0, 0, 0, 0, 0},
        {"schedStats", 0, js_sched_stats, 0, 0, 0, 0, 0}, {"allocPinned", 0, js_alloc_pinned, 0, 0, 0, 0, 0},
        {"batchSubmit", 0, js_batch_submit, 0, 0, 0, 0, 0}, {"batchWait", 0, js_batch_wait, 0, 0, 0, 0, 0},
    };
    napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
    return exports;
}
