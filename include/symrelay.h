/*
 * symrelay.h -- C ABI of the B200-native batched secretstream relay.
 *
 * Drop-in boundary for the one data-parallel path of Symmetry's provider:
 * the per-chunk body of the relay loop, /root/reference/src/provider.ts:240-257
 * (parse -> extract delta -> `peer.write(chunk)`), and its mirror on receive,
 * provider.ts:174 (`peer.on('data')`).  In the reference every `peer.write`
 * ends in ONE synchronous libsodium call per frame four packages down
 * (@hyperswarm/secret-stream `_write` -> sodium-secretstream `Push.next` ->
 * sodium-native `crypto_secretstream_xchacha20poly1305_push`;
 * package-lock.json:763-777, 5725-5758; state shape src/types.ts:168-177).
 * This library does the same arithmetic, bit-exact, for the frames of MANY
 * connections per call on a B200.  Plain C types only; no exceptions; every
 * function returns 0 / a non-negative value on success or a negative
 * SYMRELAY_E* code.  There is no CPU fallback: without a CUDA device
 * symrelay_create() fails.
 *
 * Wire format (unchanged, so an unmodified peer decrypts it):
 *     uint24-LE(mlen+17) ‖ tag^ks ‖ c[mlen] ‖ mac[16]      (mlen + 20 bytes)
 */
#ifndef SYMRELAY_H
#define SYMRELAY_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SYMRELAY_API __attribute__((visibility("default")))
#else
#define SYMRELAY_API
#endif

#define SYMRELAY_ABI_VERSION 1

#define SYMRELAY_KEYBYTES 32
#define SYMRELAY_HEADERBYTES 24
#define SYMRELAY_STATEBYTES 52  /* libsodium crypto_secretstream_xchacha20poly1305_state; src/types.ts:170,175 */
#define SYMRELAY_WIRE_OVERHEAD 20 /* 3-byte length prefix + 17 (ABYTES) */
#define SYMRELAY_MAX_FRAME ((1u << 24) - 1u - 17u) /* uint24 prefix caps mlen */

/* error codes (negative return values) */
#define SYMRELAY_OK 0
#define SYMRELAY_EINVAL (-22)
#define SYMRELAY_EAGAIN (-11) /* ring full: maps to `peer.write() === false`, provider.ts:248-252 */
#define SYMRELAY_ENOMEM (-12)
#define SYMRELAY_ENOENT (-2)
#define SYMRELAY_EPROTO (-71) /* the peer broke the framing protocol (bad first message): secret-stream destroys the stream */
#define SYMRELAY_ECUDA (-1000) /* CUDA failure: fatal for the batch, see symrelay_last_error() */

/* per-frame status word written by the kernels: bits 0..7 code, bits 8..15 opened tag */
#define SYMRELAY_ST_OK 0u
#define SYMRELAY_ST_BAD_MAC 1u            /* pull returned -1: that slot's state is untouched; no plaintext is released
                                             (the frame's output bytes are left untouched or zeroed; the host batch calls
                                             never let unauthenticated plaintext reach the caller's memory, the
                                             device-resident entry points may write a big frame's plaintext to the
                                             caller's HBM buffer and wipe it before the kernel ends) */
#define SYMRELAY_ST_TOO_SHORT 2u          /* wire frame shorter than 17 bytes after the prefix */
#define SYMRELAY_ST_EXTRACT_THROW 3u      /* the reference would throw a TypeError here (utils.ts:38-50) */
#define SYMRELAY_ST_EXTRACT_UNSUPPORTED 4u /* value shape outside the GPU extractor's domain: use the JS helper */
/* (3 and 4: the chunk is NOT sealed -- that slot's state is untouched and its wire bytes are left untouched or zeroed:
 *  the seal kernel may run beside the extractor and takes back what it sealed ahead of such a verdict) */
#define SYMRELAY_ST_RETRY 5u /* frame of a run behind a failed or REKEY frame: not processed, resubmit it */
#define SYMRELAY_ST_FAILED 6u /* per-frame front end only: the batch this frame sat in was refused as a whole
                                (symrelay_last_error() says why); nothing was sealed or opened, state untouched */
#define SYMRELAY_ST_PENDING 0xffu
#define SYMRELAY_STATUS_CODE(s) ((s) & 0xffu)
#define SYMRELAY_STATUS_TAG(s) (((s) >> 8) & 0xffu)

/* descriptor flags (upper byte of len_flags) */
#define SYMRELAY_F_TAG_MASK 0x03000000u /* secretstream tag for seal: MESSAGE 0, PUSH 1, REKEY 2, FINAL 3 */
#define SYMRELAY_F_TAG_SHIFT 24
#define SYMRELAY_F_OPEN 0x04000000u     /* frame is an inbound wire frame to open (else: seal) */
#define SYMRELAY_F_EXTRACT 0x08000000u  /* also run the SSE `data:` test + delta extraction on the plaintext */
#define SYMRELAY_F_RUN 0x10000000u      /* open frame that belongs to a run: consecutive descriptors of one slot, all
                                          F_OPEN|F_RUN, are frames queued behind each other on that stream (at most
                                          64); they are opened in parallel from the MACs on the wire */
#define SYMRELAY_LEN_MASK 0x00ffffffu

/* provider kinds: src/constants.ts:22-29 (lmstudio / oobabooga / unknown take the default branch) */
enum symrelay_provider {
    SYMRELAY_PROVIDER_DEFAULT = 0, /* litellm, lmstudio, oobabooga, anything else: utils.ts:45-50 */
    SYMRELAY_PROVIDER_OLLAMA = 1,  /* ollama, openwebui: utils.ts:38-42 */
    SYMRELAY_PROVIDER_LLAMACPP = 2 /* utils.ts:43-44 */
};

/*
 * One frame of a batch ("key, nonce/state, offset, length" of the north star):
 * the key and nonce live in the device-resident state table, addressed by slot.
 *   seal: plaintext  = in  + in_off  (in_off  % 16 == 0),  mlen = len
 *         wire frame -> out + out_off (out_off % 16 == 12), mlen + 20 bytes
 *         (in place allowed with in == out and in_off == out_off + 4: the layout
 *          secret-stream's `_write` uses, `wrapped.set(data, 4)`)
 *   open: wire frame = in  + in_off  (in_off  % 16 == 12), mlen + 20 bytes, len = mlen
 *         plaintext  -> out + out_off (out_off % 16 == 0)
 * A batch holds at most ONE frame per (slot, direction): frames of one stream
 * are MAC-chained and go into consecutive batches
 * -- except inbound runs marked SYMRELAY_F_RUN, whose chain is on the wire
 * (at most 64 frames per run).  The host-buffer entry points check both rules
 * and refuse the batch (SYMRELAY_EINVAL); the device entry points cannot read
 * the descriptors, there a frame beyond 64 in its run, or a SYMRELAY_F_RUN frame
 * given to symrelay_batch_device (no scratch), completes with SYMRELAY_ST_RETRY.
 * Arenas must be readable up to the next 16-byte boundary past each frame.
 */
typedef struct symrelay_desc {
    uint32_t slot;
    uint32_t in_off;
    uint32_t out_off;
    uint32_t len_flags; /* len (24 bits) | SYMRELAY_F_* */
} symrelay_desc;

/* Work tile produced by symrelay_plan(): a warp's worth of frames. */
typedef struct symrelay_tile {
    uint32_t first;   /* index of the tile's first descriptor */
    uint32_t n_kind;  /* bits 0..15 frames in the tile, bit 16: bulk tile (one frame, > 32 ChaCha blocks) */
} symrelay_tile;

/* Result of the fused extractor for one frame (see symrelay_batch_* `extract` argument). */
typedef struct symrelay_delta {
    uint32_t off; /* byte offset into the delta arena */
    uint32_t len; /* WTF-8 bytes appended to `completion` (provider.ts:243) */
} symrelay_delta;

#define SYMRELAY_OPT_NO_ZERO_COPY 1u /* opts.flags: always stage host batches through device buffers */

typedef struct symrelay_opts {
    uint32_t max_streams;     /* capacity of the slot table (default 65536) */
    uint32_t pipeline_depth;  /* host-batch pipeline stages (default 4) */
    uint64_t stage_bytes;     /* device staging bytes per stage and direction (default 64 MiB) */
    uint32_t max_batch_frames; /* descriptors per host batch (default 131072) */
    uint32_t flags;
    /* per-frame front end (symrelay_submit_* / poll): */
    uint32_t sched_flush_us;     /* deadline after a batch's first frame; 0 = flush on size / request only */
    uint32_t sched_batches;      /* batch buffers per GPU (default pipeline_depth + 4) */
    uint32_t sched_batch_frames; /* frames per batch (default 16384) */
    uint32_t reserved;
    uint64_t sched_arena_bytes;  /* pinned bytes per batch and direction (default 4 MiB) */
} symrelay_opts;

/* One finished frame, delivered by symrelay_poll in per-stream submit order.  `data` points into a
 * pinned library arena that stays valid until symrelay_release(ticket): the wire frame (len = mlen+20)
 * for a seal, the plaintext (len = mlen) for an open; NULL/0 when status is not OK. */
typedef struct symrelay_completion {
    int64_t ticket;
    uint32_t slot;
    uint32_t status;      /* SYMRELAY_ST_* | opened tag << 8 */
    const uint8_t *data;
    uint32_t len;
    uint32_t flags;       /* SYMRELAY_F_OPEN / _EXTRACT as submitted */
    const uint8_t *delta; /* extracted delta (F_EXTRACT seals), WTF-8 */
    uint32_t delta_len;
    uint32_t pad;
    uint64_t latency_ns;  /* delivery time - submit time: the relay latency this library adds */
} symrelay_completion;

typedef struct symrelay_sched_stats_t {
    uint64_t frames, batches, zero_copy_batches, staged_batches;
    double mean_batch_fill, p50_us, p99_us, p999_us;
} symrelay_sched_stats_t;

typedef struct symrelay_ctx symrelay_ctx;

/* ---- lifetime ------------------------------------------------------------ */
SYMRELAY_API int symrelay_abi_version(void);
/* Creates a context on gpu_ids[0..n_gpus) (n_gpus >= 1).  Slots are sharded
 * slot % n_gpus, no state ever crosses GPUs.  Hook:
 * once per process, before hyperswarm's 'connection' events (provider.ts:55-58). */
SYMRELAY_API symrelay_ctx *symrelay_create(const int *gpu_ids, int n_gpus, const symrelay_opts *opts);
SYMRELAY_API void symrelay_destroy(symrelay_ctx *ctx);
SYMRELAY_API const char *symrelay_last_error(symrelay_ctx *ctx);
SYMRELAY_API const char *symrelay_strerror(int code);
/* kernels launched by this context so far (bench.py's gpu_launches) */
SYMRELAY_API uint64_t symrelay_launch_count(symrelay_ctx *ctx);

/* ---- streams: replaces sodium-secretstream's Push/Pull construction -------
 * (`new Push(tx)` / `Pull.init(header)` in secret-stream `_setupSecretStream`;
 * the objects typed at src/types.ts:168-177).  `*_init_len` is 24 (a header:
 * the sub-key is derived with HChaCha20 on the GPU, libsodium init_push/_pull)
 * or 52 (an already initialised libsodium state to adopt), or 0 with a NULL
 * key to leave that direction closed. */
SYMRELAY_API int symrelay_stream_open(symrelay_ctx *ctx, const uint8_t *tx_key, const uint8_t *tx_init, int tx_init_len,
                         const uint8_t *rx_key, const uint8_t *rx_init, int rx_init_len, int provider_kind,
                         uint32_t *slot_out);
/* Batched variant for connection churn: n streams, keys
 * 32 B each, inits `init_len` bytes each (24 or 52), slots_out[n]. */
SYMRELAY_API int symrelay_stream_open_batch(symrelay_ctx *ctx, uint32_t n, const uint8_t *tx_keys, const uint8_t *tx_inits,
                               const uint8_t *rx_keys, const uint8_t *rx_inits, int init_len, int provider_kind,
                               uint32_t *slots_out);
/* Connection set-up in batch: what secret-stream computes per connection before it sends its
 * first message, for n new connections at once on GPU `gpu_index`.  Per connection i:
 *     local id   = BLAKE2b-256(key = handshake_hashes[i] (64 B), message = the namespace of OUR role)
 *     remote id  = the same with the peer's role's namespace: what symrelay_rx_expect_header should be given
 *     header     = 24 fresh bytes: the secretstream header of the outbound direction (libsodium init_push draws them
 *                  from its RNG; here ChaCha20 of a 32-byte seed -- seed32 NULL: read from the OS generator)
 * i.e. libsodium crypto_generichash(out32, ns, 32, hash, 64) twice and randombytes_buf(24).  initiator[i] != 0 says which
 * of the two 32-byte namespaces (hypercore-crypto `namespace('hyperswarm/secret-stream', 3)`, supplied by the caller)
 * is ours.  Outputs (each may be NULL): local_ids n x 32, remote_ids n x 32, headers n x 24, first_messages n x 59
 * (uint24-LE(56) ‖ local id ‖ header, ready for the socket).  Then symrelay_stream_open_batch(tx_inits = headers, 24).
 * Hook: hyperswarm's 'connection' event, src/provider.ts:55-58. */
SYMRELAY_API int symrelay_connect_batch(symrelay_ctx *ctx, int gpu_index, uint32_t n, const uint8_t *handshake_hashes,
                                        const uint8_t *initiator, const uint8_t ns_initiator[32], const uint8_t ns_responder[32],
                                        const uint8_t *seed32, uint8_t *local_ids, uint8_t *remote_ids, uint8_t *headers,
                                        uint8_t *first_messages);
/* Wipes the slot's keys on the device, forgets its half-received inbound frame (symrelay_rx_feed) and frees
 * the slot.  Close only after every frame submitted for the slot has completed: a frame still in flight would
 * be sealed with the wiped state.  The per-frame front end enforces that: while one of the slot's frames sits
 * in a batch that is filling or in flight the call returns SYMRELAY_EAGAIN (flush, poll, close again).  Host
 * batches that name a closed slot are refused (SYMRELAY_ENOENT). */
SYMRELAY_API int symrelay_stream_close(symrelay_ctx *ctx, uint32_t slot);
/* Copies the live libsodium-compatible 52-byte states back (either may be NULL). */
SYMRELAY_API int symrelay_stream_export(symrelay_ctx *ctx, uint32_t slot, uint8_t tx_state[SYMRELAY_STATEBYTES],
                           uint8_t rx_state[SYMRELAY_STATEBYTES]);

/* ---- planning: pack descriptors into warp tiles (host, no GPU) ------------
 * Writes at most max_tiles tiles -- packed tiles (several small frames that
 * are consecutive in `desc`) first, then *n_bulk_out bulk tiles (one frame of
 * more than 30 payload blocks each) -- and returns the total number needed
 * (call with tiles == NULL to size), or SYMRELAY_EINVAL for a bad descriptor. */
SYMRELAY_API int64_t symrelay_plan(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                      uint32_t *n_bulk_out);
/* The same for callers of the device entry points whose frames arrive in any order: writes the
 * descriptors SORTED by ChaCha20 block count to sorted_out[n] (a batch holds at most one frame per
 * (slot, direction), so any order is legal), tiles each size class in closed form and returns the
 * permutation: sorted_out[k] == desc[perm_out[k]], and the status word of sorted descriptor k belongs
 * to the caller's frame perm_out[k].  Upload sorted_out, not desc.  Packs BASELINE config 3 into
 * 9,054 tiles instead of 19,441.  Returns the tile count (tiles == NULL sizes and writes nothing),
 * SYMRELAY_EINVAL for a bad descriptor or a batch with SYMRELAY_F_RUN frames (their order is the
 * chain: use symrelay_plan).  No counterpart in the reference (one frame per call there). */
SYMRELAY_API int64_t symrelay_plan_sorted(const symrelay_desc *desc, uint32_t n, symrelay_tile *tiles, uint32_t max_tiles,
                             uint32_t *n_bulk_out, symrelay_desc *sorted_out, uint32_t *perm_out);
/* Validates one descriptor's alignment and length rules; 0 or SYMRELAY_EINVAL. */
SYMRELAY_API int symrelay_desc_check(const symrelay_desc *d);

/* ---- synchronous batch: replaces N calls of push/pull ----------------------
 * (sodium.crypto_secretstream_xchacha20poly1305_push / _pull, one frame each,
 * reached from provider.ts:248 / :174).  Mixed seal and open frames allowed. */
/* All pointers are DEVICE memory on the context's GPU `gpu_index`; the launch
 * goes to `cuda_stream` (a cudaStream_t, NULL = default stream) and returns
 * without synchronising.  d_delta / d_delta_arena may be NULL when no
 * descriptor carries SYMRELAY_F_EXTRACT. */
SYMRELAY_API int symrelay_batch_device(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc, uint32_t n,
                          const symrelay_tile *d_tiles, uint32_t n_tiles, uint32_t n_bulk_tiles,
                          const uint8_t *d_in, uint8_t *d_out, uint32_t *d_status, symrelay_delta *d_delta,
                          uint8_t *d_delta_arena, void *cuda_stream);
/* symrelay_batch_device for batches that contain SYMRELAY_F_RUN frames: the caller also passes
 * SCRATCH device memory (16-byte aligned, 16 bytes per descriptor) that the nonce pre-pass fills
 * and the tile kernels read; its contents are meaningless afterwards.  Replaces a loop of
 * `Pull.next` calls over the frames secret-stream's `_onrawdata` split out of one UDX segment. */
SYMRELAY_API int symrelay_batch_device_runs(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc, uint32_t n,
                                            const symrelay_tile *d_tiles, uint32_t n_tiles, uint32_t n_bulk_tiles,
                                            const uint8_t *d_in, uint8_t *d_out, uint32_t *d_status, void *d_scratch,
                                            uint64_t scratch_bytes, void *cuda_stream);

/* Same work as symrelay_batch_device_runs, for callers that must not allocate.  No counterpart in
 * the reference (it has no GPU and no scratch).  A provider normally shares its GPU with the model
 * server it relays for, which owns almost all of HBM by the time bulk segments arrive, so scratch
 * requested then may not exist; and the caller is an event-loop thread that cannot afford the
 * device-wide synchronisation of cudaMalloc/cudaFree.  Here the nonce scratch is a fixed array inside
 * the library: 1 MiB of HBM on every GPU that loads it, kept until the process exits, at most
 * SYMRELAY_STATIC_RUN_MAX (65,536) descriptors per call -- larger batches are refused with
 * SYMRELAY_EINVAL and a message naming the limit.  The call itself performs no cudaMalloc, cudaFree
 * or other allocating call; all inputs and outputs are still the caller's device buffers.
 * The array's device address is looked up once per GPU inside symrelay_create (with that GPU
 * current), so no call of this function does set-up work.  All calls on a GPU share the one array
 * and it keeps the previous call's contents: nothing in it needs clearing (each call's pre-pass
 * rewrites every entry its kernels read), but two calls whose kernels overlap on the GPU (different
 * streams) race on it -- issue them on one stream or otherwise order them. */
#define SYMRELAY_STATIC_RUN_MAX 65536u
SYMRELAY_API int symrelay_batch_device_runs_static(symrelay_ctx *ctx, int gpu_index, const symrelay_desc *d_desc,
                                                   uint32_t n, const symrelay_tile *d_tiles, uint32_t n_tiles,
                                                   uint32_t n_bulk_tiles, const uint8_t *d_in, uint8_t *d_out,
                                                   uint32_t *d_status, void *cuda_stream);

/* HOST buffers (pinned from symrelay_alloc_pinned for full speed, pageable
 * accepted): plans, copies in, launches, copies out, waits.  Batches of up to
 * 1 MiB (in + out) whose arenas came from symrelay_alloc_pinned skip the copies:
 * the kernels read and write the pinned host arenas directly over PCIe.  status[n] and, if
 * any descriptor has F_EXTRACT, delta[n] + delta_arena (delta_arena_bytes) are
 * filled.  in_bytes/out_bytes are the arena sizes (multiples of 16). */
SYMRELAY_API int symrelay_batch_host(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, const uint8_t *in,
                        uint64_t in_bytes, uint8_t *out, uint64_t out_bytes, uint32_t *status,
                        symrelay_delta *delta, uint8_t *delta_arena, uint64_t delta_arena_bytes);
/* Pipelined form of the same call: returns a ticket (>= 0) at once; buffers
 * must stay valid until symrelay_batch_wait(ticket) returns.  Batches execute
 * in submission order, so consecutive batches may carry consecutive frames of
 * the same streams.  The out arena, status and delta are valid only after
 * symrelay_batch_wait(ticket) has returned: a pipelined small batch keeps its
 * results on the device until the next batch's kernel (or the wait itself)
 * posts them to the host, and the whole out arena is written, gaps included. */
SYMRELAY_API int64_t symrelay_batch_submit(symrelay_ctx *ctx, const symrelay_desc *desc, uint32_t n, const uint8_t *in,
                              uint64_t in_bytes, uint8_t *out, uint64_t out_bytes, uint32_t *status,
                              symrelay_delta *delta, uint8_t *delta_arena, uint64_t delta_arena_bytes);
SYMRELAY_API int symrelay_batch_wait(symrelay_ctx *ctx, int64_t ticket);

/* ---- per-frame asynchronous front end: the shape of the reference's own calls ----
 * symrelay_submit_seal  <-  `peer.write(chunk)` (src/provider.ts:97,103,125,234,248,260): copies the
 *     chunk into the pinned ring and returns a ticket at once; SYMRELAY_EAGAIN when the ring is full
 *     (= `write()` returned false: retry after a completion, the reference's 'drain', :250-252).
 *     flags: SYMRELAY_F_EXTRACT and/or a tag in SYMRELAY_F_TAG_MASK.
 * symrelay_submit_open  <-  secret-stream `_incoming` feeding `peer.on('data')` (:110,:174): `wire` is
 *     one whole frame including its uint24-LE prefix; frames shorter than 17 bytes after the prefix or
 *     with a wrong prefix complete with SYMRELAY_ST_TOO_SHORT (secret-stream destroys the stream).
 * symrelay_flush closes the batch being filled; batches also close on size and on opts.sched_flush_us.
 * symrelay_poll never blocks; symrelay_poll_wait blocks until the oldest in-flight batch is done.
 * Completions of one stream arrive in submit order.  Every delivered completion must be released. */
SYMRELAY_API int64_t symrelay_submit_seal(symrelay_ctx *ctx, uint32_t slot, const uint8_t *plain, uint32_t len,
                                          uint32_t flags);
SYMRELAY_API int64_t symrelay_submit_open(symrelay_ctx *ctx, uint32_t slot, const uint8_t *wire, uint32_t wire_len);
/* symrelay_rx_feed  <-  secret-stream `_onrawdata` (the state typed at src/types.ts:156-159: _state, _len,
 *     _tmp, _message): feed raw transport bytes of one connection in arrival order, cut anywhere (inside
 *     the 3-byte length too).  Every frame completed by this segment is submitted as an open frame;
 *     several frames of one segment join one batch as a run (SYMRELAY_F_RUN) and are opened in parallel.
 *     Returns the number of frames submitted (their tickets in tickets[0..)), *consumed = bytes of `seg`
 *     taken; when the ring is full it returns what it could (or SYMRELAY_EAGAIN if nothing) and the caller
 *     feeds the rest of the segment after draining completions.  A completion with SYMRELAY_ST_RETRY (only
 *     behind a REKEY-tagged frame or a counter wrap inside a run) must be resubmitted with symrelay_submit_open, in
 *     the order the completions arrive.  Frames of that stream that were already submitted behind the RETRY frames
 *     (a later segment, a later batch) were opened against a state that lags behind them: they are delivered as
 *     SYMRELAY_ST_RETRY as well, not as SYMRELAY_ST_BAD_MAC -- resubmit them too, after the earlier ones.  Only a
 *     SYMRELAY_ST_BAD_MAC means what it means in the reference: the stream is dead. */
SYMRELAY_API int symrelay_rx_feed(symrelay_ctx *ctx, uint32_t slot, const uint8_t *seg, uint32_t n, uint32_t *consumed,
                                  int64_t *tickets, int max_tickets);
/* symrelay_rx_expect_header  <-  secret-stream's handling of the connection's FIRST message: before any encrypted
 *     frame each side sends uint24-LE(56) ‖ streamId[32] ‖ header[24] in the clear (`_onrawdata` with `_state` 0, the
 *     fields typed at src/types.ts:156-159).  After this call the next frame symrelay_rx_feed completes on the slot is
 *     taken as that message: its stream id is compared with expected_stream_id (NULL: not checked), its header and
 *     rx_key initialise the slot's receive state on the GPU (HChaCha20 sub-key, libsodium init_pull), and the frames
 *     after it are opened as usual.  A first message of another size or with another id makes symrelay_rx_feed return
 *     SYMRELAY_EPROTO: the caller destroys the connection, as secret-stream does.  Open the slot with rx_key NULL.
 * symrelay_stream_set_keepalive  <-  secret-stream `setKeepAlive` / the `_incoming` rule "an empty message on a
 *     stream with keep-alive on is swallowed": zero-length inbound frames are opened (the chain advances) but never
 *     delivered by symrelay_poll. */
SYMRELAY_API int symrelay_rx_expect_header(symrelay_ctx *ctx, uint32_t slot, const uint8_t rx_key[SYMRELAY_KEYBYTES],
                                           const uint8_t *expected_stream_id);
SYMRELAY_API int symrelay_stream_set_keepalive(symrelay_ctx *ctx, uint32_t slot, int on);
SYMRELAY_API int symrelay_flush(symrelay_ctx *ctx);
SYMRELAY_API int symrelay_poll(symrelay_ctx *ctx, symrelay_completion *out, int max);
SYMRELAY_API int symrelay_poll_wait(symrelay_ctx *ctx, symrelay_completion *out, int max);
SYMRELAY_API int symrelay_release(symrelay_ctx *ctx, int64_t ticket);
/* Completion wake-up for an event loop.  The reference is woken by libuv when its sockets have something for it
 * (`peer.on('data')`, 'drain': src/provider.ts:174, :250-252) and never polls; SURVEY.md 8(b) asks the same of this
 * boundary.  Returns an eventfd (owned by the context, closed by symrelay_destroy) that becomes readable whenever a
 * batch of the per-frame front end has finished: read the 8-byte counter, then call symrelay_poll, which delivers
 * the completions without waiting.  The first call starts one watcher thread for the context; that thread also
 * performs the deadline flushes (opts.sched_flush_us), so an idle loop need not call symrelay_poll for them.
 * The N-API shim hands the descriptor to uv_poll_init. */
SYMRELAY_API int symrelay_event_fd(symrelay_ctx *ctx);
/* frames/batches handled by the front end and the p50/p99/p99.9 of their added latency */
SYMRELAY_API int symrelay_sched_stats(symrelay_ctx *ctx, symrelay_sched_stats_t *out);
SYMRELAY_API int symrelay_sched_reset_stats(symrelay_ctx *ctx);
/* How the fused extractor fared on the front end's SYMRELAY_F_EXTRACT frames, per provider kind (enum symrelay_provider):
 * counts[4 * kind + j], j = 0 delta extracted, 1 the reference would have thrown (SYMRELAY_ST_EXTRACT_THROW), 2 handed back
 * to the host helper (SYMRELAY_ST_EXTRACT_UNSUPPORTED: each of those costs the stream a second batch), 3 anything else.
 * Tells an operator how often real provider output leaves the GPU extractor's domain.  No counterpart in the reference. */
SYMRELAY_API int symrelay_extract_stats(symrelay_ctx *ctx, uint64_t counts[12]);

/* ---- after the path: sealed frames to datagrams, many per system call ----
 * In the reference each `peer.write` ends in udx-native 1.10.3 (package-lock.json:6241-6253; UDXStream / UDXSocket are
 * typed at src/types.ts:91-122) giving the kernel ONE packet per frame.  symrelay_udp_send_batch takes the completions
 * symrelay_poll delivered (sealed frames only; others are skipped), cuts every wire frame into datagrams of at most `mtu`
 * bytes -- a 16-byte symrelay_udp_hdr plus a slice gathered straight from the pinned ring, no copy -- and hands them to
 * the kernel up to 1,024 per sendmmsg(2).  dst[i] / dst_len[i]: the destination of completion i (NULL on a connected
 * socket); conn_id[i]: what the receiver knows the connection by; stream_off[i]: in/out, the connection's outbound byte
 * count so far (the receiver orders slices by it and feeds them to symrelay_rx_feed).  Returns the number of datagrams
 * sent, SYMRELAY_EAGAIN when the socket buffer is full (stats say how far it got), SYMRELAY_EINVAL otherwise.  It does not
 * speak UDX: acknowledgements, retransmission and congestion control stay in udx-native; this is the send primitive.
 * SYMRELAY_UDP_ONE_PER_CALL reproduces the reference's one system call per packet, for comparison. */
#define SYMRELAY_UDP_MAGIC 0x554d5953u /* "SYMU" */
#define SYMRELAY_UDP_HDR 16u
#define SYMRELAY_UDP_ONE_PER_CALL 1u
typedef struct symrelay_udp_hdr {
    uint32_t magic, conn;
    uint64_t offset; /* of this slice in the connection's outbound byte stream */
} symrelay_udp_hdr;
typedef struct symrelay_udp_stats {
    uint64_t frames, datagrams, syscalls, payload_bytes;
} symrelay_udp_stats;
SYMRELAY_API int symrelay_udp_send_batch(int fd, const symrelay_completion *comp, int n, const void *const *dst,
                                         const uint32_t *dst_len, const uint32_t *conn_id, uint64_t *stream_off, uint32_t mtu,
                                         uint32_t flags, symrelay_udp_stats *stats);

/* Measurement aid (bench.py): sustained integer throughput of this GPU in
 * Tera-ops/s for ChaCha20's ADD:XOR:ROTATE 1:1:1 mix and for 32x32->64
 * multiply-adds -- the integer roofline denominators.
 * No counterpart in the reference. */
SYMRELAY_API int symrelay_probe_int_peak(symrelay_ctx *ctx, int gpu_index, double *arx_tops, double *imad_tops);

/* pinned host memory for arenas (the N-API shim exposes it as an external ArrayBuffer) */
SYMRELAY_API void *symrelay_alloc_pinned(symrelay_ctx *ctx, uint64_t bytes);
SYMRELAY_API void symrelay_free_pinned(symrelay_ctx *ctx, void *p);

#ifdef __cplusplus
}
#endif
#endif /* SYMRELAY_H */
