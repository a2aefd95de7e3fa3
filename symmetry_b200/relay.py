"""Host-side front end of the C ABI (ctypes), in the reference's vocabulary.

The reference's seal/open path is `peer.write(chunk)` / `peer.on('data')`
(/root/reference/src/provider.ts:248, :174) on a hyperswarm secret-stream whose
crypto state is typed at /root/reference/src/types.ts:168-177.  `SymRelay` owns
the device-resident table of those states (one slot per connection) and runs
batches of frames from many connections through the CUDA kernels.  No Node.js
exists in this image, so this module is also the only runnable caller of the
library; the N-API shim a maintainer would add is in INTEGRATION.md.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

DESC_DTYPE = np.dtype([("slot", "<u4"), ("in_off", "<u4"), ("out_off", "<u4"), ("len_flags", "<u4")])
TILE_DTYPE = np.dtype([("first", "<u4"), ("n_kind", "<u4")])
DELTA_DTYPE = np.dtype([("off", "<u4"), ("len", "<u4")])

F_OPEN = 0x04000000
F_EXTRACT = 0x08000000
F_RUN = 0x10000000
LEN_MASK = 0x00FFFFFF
ST_OK, ST_BAD_MAC, ST_TOO_SHORT, ST_EXTRACT_THROW, ST_EXTRACT_UNSUPPORTED, ST_RETRY, ST_FAILED = 0, 1, 2, 3, 4, 5, 6
EAGAIN = -11
EPROTO = -71
PROVIDER_DEFAULT, PROVIDER_OLLAMA, PROVIDER_LLAMACPP = 0, 1, 2
STATEBYTES, HEADERBYTES, KEYBYTES, WIRE_OVERHEAD = 52, 24, 32, 20

# src/constants.ts:22-29 -> extractor branch (utils.ts:36-50)
PROVIDER_KIND = {"ollama": PROVIDER_OLLAMA, "openwebui": PROVIDER_OLLAMA, "llamacpp": PROVIDER_LLAMACPP,
                 "litellm": PROVIDER_DEFAULT, "lmstudio": PROVIDER_DEFAULT, "oobabooga": PROVIDER_DEFAULT}


def provider_kind(api_provider: str) -> int:
    return PROVIDER_KIND.get(api_provider, PROVIDER_DEFAULT)


class RelayError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"symrelay error {code}: {msg}")
        self.code = code


def _ptr(a) -> int:
    if a is None:
        return 0
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return int(a)


def _bytes_ptr(b):
    return None if b is None else bytes(b)  # ctypes passes bytes to a void* parameter as-is


class SymRelay:
    """One context = one state table per GPU + the batch pipeline."""

    def __init__(self, gpu_ids=(0,), max_streams: int = 65536, pipeline_depth: int = 4,
                 stage_bytes: int = 64 << 20, max_batch_frames: int = 131072, zero_copy: bool = True,
                 flush_us: int = 0, sched_batches: int = 0, sched_batch_frames: int = 0, sched_arena_bytes: int = 0):
        self.lib = _lib.load()
        ids = (C.c_int * len(gpu_ids))(*gpu_ids)
        opts = _lib.SymrelayOpts(max_streams, pipeline_depth, stage_bytes, max_batch_frames, 0 if zero_copy else 1,
                                 flush_us, sched_batches, sched_batch_frames, 0, sched_arena_bytes)
        self._cbuf = (_lib.SymrelayCompletion * 4096)()
        self.n_gpus = len(gpu_ids)
        self.ctx = self.lib.symrelay_create(ids, len(gpu_ids), C.byref(opts))
        if not self.ctx:
            raise RelayError(-1000, self.lib.symrelay_last_error(None).decode())
        self._pinned = []

    # -- lifetime ----------------------------------------------------------------
    def close(self):
        if getattr(self, "ctx", None):
            for p in self._pinned:
                self.lib.symrelay_free_pinned(self.ctx, p)
            self._pinned = []
            self.lib.symrelay_destroy(self.ctx)
            self.ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int) -> int:
        if rc < 0:
            raise RelayError(rc, self.lib.symrelay_last_error(self.ctx).decode())
        return rc

    @property
    def launches(self) -> int:
        return int(self.lib.symrelay_launch_count(self.ctx))

    # -- streams -----------------------------------------------------------------
    def open_stream(self, tx_key=None, tx_init=None, rx_key=None, rx_init=None, provider: int = 0) -> int:
        """`new Push(tx)` + `Pull.init(header)` for one connection; init is a 24-byte header or a
        52-byte libsodium state.  Returns the slot."""
        slot = C.c_uint32(0)
        self._check(self.lib.symrelay_stream_open(
            self.ctx, _bytes_ptr(tx_key), _bytes_ptr(tx_init), len(tx_init) if tx_init else 0,
            _bytes_ptr(rx_key), _bytes_ptr(rx_init), len(rx_init) if rx_init else 0, provider, C.byref(slot)))
        return slot.value

    def open_streams(self, tx_keys, tx_inits, rx_keys, rx_inits, provider: int = 0) -> np.ndarray:
        """Batched open: arrays [n,32] and [n,24|52] (uint8); either direction may be None."""
        ref = tx_keys if tx_keys is not None else rx_keys
        n = len(ref)
        inits = tx_inits if tx_inits is not None else rx_inits
        init_len = inits.shape[1]
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.uint8)
                for a in (tx_keys, tx_inits, rx_keys, rx_inits)]
        slots = np.zeros(n, dtype=np.uint32)
        self._check(self.lib.symrelay_stream_open_batch(self.ctx, n, _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2]),
                                                        _ptr(arrs[3]), init_len, provider, _ptr(slots)))
        return slots

    def close_stream(self, slot: int) -> None:
        self._check(self.lib.symrelay_stream_close(self.ctx, int(slot)))

    def export(self, slot: int):
        """(tx_state, rx_state): the live 52-byte libsodium states of a slot."""
        tx = C.create_string_buffer(STATEBYTES)
        rx = C.create_string_buffer(STATEBYTES)
        self._check(self.lib.symrelay_stream_export(self.ctx, int(slot), tx, rx))
        return tx.raw, rx.raw

    # -- batches -----------------------------------------------------------------
    def plan(self, descs: np.ndarray):
        nb = C.c_uint32(0)
        need = self._check(self.lib.symrelay_plan(_ptr(descs), len(descs), None, 0, C.byref(nb)))
        tiles = np.zeros(max(1, need), dtype=TILE_DTYPE)
        self._check(self.lib.symrelay_plan(_ptr(descs), len(descs), _ptr(tiles), need, C.byref(nb)))
        return tiles[:need], nb.value

    def plan_sorted(self, descs: np.ndarray):
        """(tiles, n_bulk, sorted descriptors, perm): descriptors grouped by size for the device entry points;
        the status of sorted descriptor k belongs to descs[perm[k]]."""
        nb = C.c_uint32(0)
        need = self._check(self.lib.symrelay_plan_sorted(_ptr(descs), len(descs), None, 0, C.byref(nb), None, None))
        tiles = np.zeros(max(1, need), dtype=TILE_DTYPE)
        sorted_descs = np.zeros(len(descs), dtype=descs.dtype)
        perm = np.zeros(len(descs), dtype=np.uint32)
        self._check(self.lib.symrelay_plan_sorted(_ptr(descs), len(descs), _ptr(tiles), need, C.byref(nb),
                                                  _ptr(sorted_descs), _ptr(perm)))
        return tiles[:need], nb.value, sorted_descs, perm

    def batch_host(self, descs: np.ndarray, arena_in: np.ndarray, arena_out: np.ndarray,
                   status: np.ndarray | None = None, delta: np.ndarray | None = None,
                   delta_arena: np.ndarray | None = None) -> np.ndarray:
        """Synchronous mixed seal/open batch on HOST arenas (N x push/pull in one call).  Frames with
        F_EXTRACT also fill delta[i] = (off, len) and the delta bytes in delta_arena."""
        if status is None:
            status = np.full(len(descs), 0xFF, dtype=np.uint32)
        self._check(self.lib.symrelay_batch_host(
            self.ctx, _ptr(descs), len(descs), _ptr(arena_in), arena_in.nbytes, _ptr(arena_out), arena_out.nbytes,
            _ptr(status), _ptr(delta) or None, _ptr(delta_arena) or None,
            0 if delta_arena is None else delta_arena.nbytes))
        return status

    def submit(self, descs: np.ndarray, arena_in, in_bytes: int, arena_out, out_bytes: int, status,
               delta=None, delta_arena=None, delta_arena_bytes: int = 0) -> int:
        """Pipelined batch; buffers must stay alive until wait(ticket)."""
        return self._check(self.lib.symrelay_batch_submit(
            self.ctx, _ptr(descs), len(descs), _ptr(arena_in), in_bytes, _ptr(arena_out), out_bytes, _ptr(status),
            _ptr(delta) or None, _ptr(delta_arena) or None, delta_arena_bytes))

    def wait(self, ticket: int) -> None:
        self._check(self.lib.symrelay_batch_wait(self.ctx, ticket))

    def batch_device(self, d_desc: int, n: int, d_tiles: int, n_tiles: int, n_bulk: int, d_in: int, d_out: int,
                     d_status: int, stream: int = 0, gpu_index: int = 0, d_delta: int = 0,
                     d_delta_arena: int = 0) -> None:
        """One launch per tile class (plus the extractor when d_delta is given) on device-resident
        buffers (raw device pointers), asynchronous on `stream`."""
        self._check(self.lib.symrelay_batch_device(self.ctx, gpu_index, d_desc, n, d_tiles, n_tiles, n_bulk, d_in,
                                                   d_out, d_status, d_delta or None, d_delta_arena or None, stream))

    # -- per-frame front end (the shape of peer.write / peer.on('data')) ----------
    def submit_seal(self, slot: int, chunk: bytes, extract: bool = False, tag: int = 0) -> int:
        """`peer.write(chunk)`: returns a ticket, or raises RelayError(EAGAIN) when the ring is full."""
        flags = (F_EXTRACT if extract else 0) | ((tag & 3) << 24)
        return self._check(self.lib.symrelay_submit_seal(self.ctx, int(slot), bytes(chunk), len(chunk), flags))

    def submit_open(self, slot: int, wire: bytes) -> int:
        """One whole inbound wire frame (uint24-LE prefix included)."""
        return self._check(self.lib.symrelay_submit_open(self.ctx, int(slot), bytes(wire), len(wire)))

    def rx_feed(self, slot: int, segment: bytes):
        """secret-stream `_onrawdata`: raw transport bytes of one connection, cut anywhere.  Returns
        (tickets of the frames this segment completed, bytes consumed)."""
        tick = (C.c_int64 * 256)()
        used = C.c_uint32(0)
        n = self.lib.symrelay_rx_feed(self.ctx, int(slot), bytes(segment), len(segment), C.byref(used), tick, 256)
        if n < 0 and n != EAGAIN:
            self._check(n)
        return [tick[i] for i in range(max(n, 0))], used.value

    def connect_batch(self, handshake_hashes: np.ndarray, initiator, ns_initiator: bytes, ns_responder: bytes,
                      seed: bytes | None = None, gpu_index: int = 0):
        """Connection set-up for n new connections at once: (local ids [n,32], remote ids [n,32], headers [n,24],
        first messages [n,59]) -- the BLAKE2b stream ids and fresh secretstream headers secret-stream computes per
        connection before its first message."""
        hh = np.ascontiguousarray(handshake_hashes, dtype=np.uint8)
        n = len(hh)
        ini = np.ascontiguousarray(initiator, dtype=np.uint8)
        lid, rid = np.zeros((n, 32), np.uint8), np.zeros((n, 32), np.uint8)
        hdr, fm = np.zeros((n, 24), np.uint8), np.zeros((n, 59), np.uint8)
        self._check(self.lib.symrelay_connect_batch(self.ctx, gpu_index, n, _ptr(hh), _ptr(ini), bytes(ns_initiator),
                                                    bytes(ns_responder), _bytes_ptr(seed), _ptr(lid), _ptr(rid), _ptr(hdr), _ptr(fm)))
        return lid, rid, hdr, fm

    def rx_expect_header(self, slot: int, rx_key: bytes, expected_stream_id: bytes | None = None) -> None:
        """The next frame fed on this slot is the peer's first message (uint24(56) ‖ streamId ‖ header)."""
        self._check(self.lib.symrelay_rx_expect_header(self.ctx, int(slot), bytes(rx_key), _bytes_ptr(expected_stream_id)))

    def set_keepalive(self, slot: int, on: bool = True) -> None:
        self._check(self.lib.symrelay_stream_set_keepalive(self.ctx, int(slot), 1 if on else 0))

    def event_fd(self) -> int:
        """eventfd that becomes readable when a batch of the per-frame front end has finished (what the N-API shim
        gives to uv_poll); starts the context's watcher thread on first use."""
        return self._check(self.lib.symrelay_event_fd(self.ctx))

    def flush(self) -> None:
        self._check(self.lib.symrelay_flush(self.ctx))

    def poll(self, wait: bool = False, release: bool = True):
        """Finished frames as dicts (data/delta copied out); releases them unless told otherwise."""
        fn = self.lib.symrelay_poll_wait if wait else self.lib.symrelay_poll
        n = self._check(fn(self.ctx, self._cbuf, len(self._cbuf)))
        out = []
        for i in range(n):
            c = self._cbuf[i]
            out.append({"ticket": c.ticket, "slot": c.slot, "status": c.status & 0xFF, "tag": (c.status >> 8) & 0xFF,
                        "open": bool(c.flags & F_OPEN),
                        "data": C.string_at(c.data, c.len) if c.data else b"",
                        "delta": C.string_at(c.delta, c.delta_len) if c.delta else b"",
                        "latency_ns": c.latency_ns})
            if release:
                self._check(self.lib.symrelay_release(self.ctx, c.ticket))
        return out

    def release(self, ticket: int) -> None:
        self._check(self.lib.symrelay_release(self.ctx, ticket))

    def sched_stats(self) -> dict:
        st = _lib.SymrelaySchedStats()
        self._check(self.lib.symrelay_sched_stats(self.ctx, C.byref(st)))
        return {f: getattr(st, f) for f, _ in st._fields_}

    def extract_stats(self) -> dict:
        """{provider kind: {"ok", "throw", "unsupported", "other"}} for the front end's F_EXTRACT frames."""
        c = (C.c_uint64 * 12)()
        self._check(self.lib.symrelay_extract_stats(self.ctx, c))
        names = ("ok", "throw", "unsupported", "other")
        return {k: {nm: int(c[4 * k + j]) for j, nm in enumerate(names)} for k in range(3)}

    def probe_int_peak(self, gpu_index: int = 0):
        """(arx_tops, imad_tops): measured integer throughput, the ALU roofline denominators."""
        a, m = C.c_double(0), C.c_double(0)
        self._check(self.lib.symrelay_probe_int_peak(self.ctx, gpu_index, C.byref(a), C.byref(m)))
        return a.value, m.value

    def batch_device_runs(self, d_desc: int, n: int, d_tiles: int, n_tiles: int, n_bulk: int, d_in: int, d_out: int,
                          d_status: int, d_scratch: int, scratch_bytes: int, stream: int = 0, gpu_index: int = 0):
        """batch_device for batches with F_RUN frames; d_scratch: 16 bytes per descriptor of device scratch."""
        self._check(self.lib.symrelay_batch_device_runs(self.ctx, gpu_index, d_desc, n, d_tiles, n_tiles, n_bulk, d_in,
                                                        d_out, d_status, d_scratch, scratch_bytes, stream))

    def batch_device_runs_static(self, d_desc: int, n: int, d_tiles: int, n_tiles: int, n_bulk: int, d_in: int,
                                 d_out: int, d_status: int, stream: int = 0, gpu_index: int = 0):
        """batch_device_runs for callers that cannot allocate (the GPU is full of model weights, or the
        caller is an event-loop thread that must not trigger cudaMalloc's device-wide sync): the nonce
        scratch is a fixed 1 MiB array inside the library, so n is capped at 65,536 descriptors.  Calls
        share the array, so keep them ordered on one stream."""
        self._check(self.lib.symrelay_batch_device_runs_static(self.ctx, gpu_index, d_desc, n, d_tiles, n_tiles, n_bulk,
                                                               d_in, d_out, d_status, stream))

    def alloc_pinned(self, nbytes: int) -> np.ndarray:
        """Pinned host arena as a uint8 numpy view (freed with the context)."""
        p = self.lib.symrelay_alloc_pinned(self.ctx, nbytes)
        if not p:
            raise RelayError(-12, self.lib.symrelay_last_error(self.ctx).decode())
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p)
        return np.frombuffer(buf, dtype=np.uint8)
