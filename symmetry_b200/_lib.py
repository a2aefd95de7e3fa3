"""ctypes binding of libsymrelay.so (include/symrelay.h).

This is the only way the package computes anything: if the CUDA library is
missing the import fails loudly -- there is no CPU fallback (BASELINE.json
north_star)."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# SYMRELAY_SO lets A/B experiments load another build of the same library (never a different implementation)
SO_PATH = os.environ.get("SYMRELAY_SO") or os.path.join(HERE, "libsymrelay.so")

# every symbol include/symrelay.h declares (tests/test_abi_symbols.py cross-checks the header)
EXPORTS = (
    "symrelay_abi_version", "symrelay_create", "symrelay_destroy", "symrelay_last_error", "symrelay_strerror",
    "symrelay_launch_count", "symrelay_stream_open", "symrelay_stream_open_batch", "symrelay_stream_close",
    "symrelay_stream_export", "symrelay_plan", "symrelay_plan_sorted", "symrelay_desc_check", "symrelay_batch_device",
    "symrelay_batch_host", "symrelay_batch_submit", "symrelay_batch_wait", "symrelay_alloc_pinned",
    "symrelay_free_pinned", "symrelay_probe_int_peak", "symrelay_submit_seal", "symrelay_submit_open",
    "symrelay_flush", "symrelay_poll", "symrelay_poll_wait", "symrelay_release", "symrelay_sched_stats",
    "symrelay_sched_reset_stats", "symrelay_batch_device_runs", "symrelay_batch_device_runs_static",
    "symrelay_rx_feed", "symrelay_event_fd", "symrelay_connect_batch", "symrelay_rx_expect_header",
    "symrelay_stream_set_keepalive", "symrelay_extract_stats", "symrelay_udp_send_batch",
)


class SymrelayOpts(C.Structure):
    _fields_ = [("max_streams", C.c_uint32), ("pipeline_depth", C.c_uint32), ("stage_bytes", C.c_uint64),
                ("max_batch_frames", C.c_uint32), ("flags", C.c_uint32), ("sched_flush_us", C.c_uint32),
                ("sched_batches", C.c_uint32), ("sched_batch_frames", C.c_uint32), ("reserved", C.c_uint32),
                ("sched_arena_bytes", C.c_uint64)]


class SymrelayCompletion(C.Structure):
    _fields_ = [("ticket", C.c_int64), ("slot", C.c_uint32), ("status", C.c_uint32), ("data", C.c_void_p),
                ("len", C.c_uint32), ("flags", C.c_uint32), ("delta", C.c_void_p), ("delta_len", C.c_uint32),
                ("pad", C.c_uint32), ("latency_ns", C.c_uint64)]


class SymrelayUdpStats(C.Structure):
    _fields_ = [("frames", C.c_uint64), ("datagrams", C.c_uint64), ("syscalls", C.c_uint64), ("payload_bytes", C.c_uint64)]


class SymrelaySchedStats(C.Structure):
    _fields_ = [("frames", C.c_uint64), ("batches", C.c_uint64), ("zero_copy_batches", C.c_uint64),
                ("staged_batches", C.c_uint64), ("mean_batch_fill", C.c_double), ("p50_us", C.c_double),
                ("p99_us", C.c_double), ("p999_us", C.c_double)]


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: build it with `python -m symmetry_b200.build` (needs nvcc). "
            "symmetry_b200 has no CPU fallback.")
    L = C.CDLL(SO_PATH)
    vp, u32, u64, i64, ci = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int64, C.c_int
    L.symrelay_abi_version.restype = ci
    L.symrelay_create.restype = vp
    L.symrelay_create.argtypes = [C.POINTER(ci), ci, C.POINTER(SymrelayOpts)]
    L.symrelay_destroy.argtypes = [vp]
    L.symrelay_destroy.restype = None
    L.symrelay_last_error.restype = C.c_char_p
    L.symrelay_last_error.argtypes = [vp]
    L.symrelay_strerror.restype = C.c_char_p
    L.symrelay_strerror.argtypes = [ci]
    L.symrelay_launch_count.restype = u64
    L.symrelay_launch_count.argtypes = [vp]
    L.symrelay_stream_open.restype = ci
    L.symrelay_stream_open.argtypes = [vp, vp, vp, ci, vp, vp, ci, ci, C.POINTER(u32)]
    L.symrelay_stream_open_batch.restype = ci
    L.symrelay_stream_open_batch.argtypes = [vp, u32, vp, vp, vp, vp, ci, ci, vp]
    L.symrelay_stream_close.restype = ci
    L.symrelay_stream_close.argtypes = [vp, u32]
    L.symrelay_stream_export.restype = ci
    L.symrelay_stream_export.argtypes = [vp, u32, vp, vp]
    L.symrelay_plan.restype = i64
    L.symrelay_plan.argtypes = [vp, u32, vp, u32, C.POINTER(u32)]
    L.symrelay_plan_sorted.restype = i64
    L.symrelay_plan_sorted.argtypes = [vp, u32, vp, u32, C.POINTER(u32), vp, vp]
    L.symrelay_desc_check.restype = ci
    L.symrelay_desc_check.argtypes = [vp]
    L.symrelay_batch_device.restype = ci
    L.symrelay_batch_device.argtypes = [vp, ci, vp, u32, vp, u32, u32, vp, vp, vp, vp, vp, vp]
    L.symrelay_batch_host.restype = ci
    L.symrelay_batch_host.argtypes = [vp, vp, u32, vp, u64, vp, u64, vp, vp, vp, u64]
    L.symrelay_batch_submit.restype = i64
    L.symrelay_batch_submit.argtypes = [vp, vp, u32, vp, u64, vp, u64, vp, vp, vp, u64]
    L.symrelay_batch_wait.restype = ci
    L.symrelay_batch_wait.argtypes = [vp, i64]
    L.symrelay_alloc_pinned.restype = vp
    L.symrelay_alloc_pinned.argtypes = [vp, u64]
    L.symrelay_free_pinned.restype = None
    L.symrelay_free_pinned.argtypes = [vp, vp]
    L.symrelay_probe_int_peak.restype = ci
    L.symrelay_probe_int_peak.argtypes = [vp, ci, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.symrelay_batch_device_runs.restype = ci
    L.symrelay_batch_device_runs.argtypes = [vp, ci, vp, u32, vp, u32, u32, vp, vp, vp, vp, u64, vp]
    L.symrelay_batch_device_runs_static.restype = ci
    L.symrelay_batch_device_runs_static.argtypes = [vp, ci, vp, u32, vp, u32, u32, vp, vp, vp, vp]
    L.symrelay_submit_seal.restype = i64
    L.symrelay_submit_seal.argtypes = [vp, u32, vp, u32, u32]
    L.symrelay_submit_open.restype = i64
    L.symrelay_submit_open.argtypes = [vp, u32, vp, u32]
    L.symrelay_rx_feed.restype = ci
    L.symrelay_rx_feed.argtypes = [vp, u32, vp, u32, C.POINTER(u32), C.POINTER(i64), ci]
    L.symrelay_connect_batch.restype = ci
    L.symrelay_connect_batch.argtypes = [vp, ci, u32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.symrelay_rx_expect_header.restype = ci
    L.symrelay_rx_expect_header.argtypes = [vp, u32, vp, vp]
    L.symrelay_stream_set_keepalive.restype = ci
    L.symrelay_stream_set_keepalive.argtypes = [vp, u32, ci]
    L.symrelay_udp_send_batch.restype = ci
    L.symrelay_udp_send_batch.argtypes = [ci, C.POINTER(SymrelayCompletion), ci, vp, vp, vp, vp, u32, u32, C.POINTER(SymrelayUdpStats)]
    L.symrelay_extract_stats.restype = ci
    L.symrelay_extract_stats.argtypes = [vp, C.POINTER(u64)]
    L.symrelay_event_fd.restype = ci
    L.symrelay_event_fd.argtypes = [vp]
    L.symrelay_flush.restype = ci
    L.symrelay_flush.argtypes = [vp]
    for name in ("symrelay_poll", "symrelay_poll_wait"):
        getattr(L, name).restype = ci
        getattr(L, name).argtypes = [vp, C.POINTER(SymrelayCompletion), ci]
    L.symrelay_release.restype = ci
    L.symrelay_release.argtypes = [vp, i64]
    L.symrelay_sched_stats.restype = ci
    L.symrelay_sched_stats.argtypes = [vp, C.POINTER(SymrelaySchedStats)]
    L.symrelay_sched_reset_stats.restype = ci
    L.symrelay_sched_reset_stats.argtypes = [vp]
    if L.symrelay_abi_version() != 1:
        raise ImportError("libsymrelay.so ABI version mismatch")
    _lib = L
    return L
