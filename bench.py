#!/usr/bin/env python3
"""bench.py -- batched secretstream seal+open throughput on B200(s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4] [--impl ours|reference]

A "step" is one pass of the relay hot path over one batch: every stream seals
one outbound frame AND opens one inbound frame (one mixed batch = one kernel
launch per tile class).  Default workload = BASELINE.json configs[1]: 1,024
concurrent streams x 64 B token-delta frames, batched seal+open on 1 B200.
`value` is device-resident (inputs already in HBM); `e2e` goes through the
C ABI's host-buffer batch call with pinned host arenas, H2D and D2H inside the
timed region.  `--impl reference` times the reference's CPU path for the same
workload: real libsodium (the C library under sodium-native) driven by the C
loop in oracle/cpu_bench.c on all host cores.  Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

L2_BYTES = 126 << 20
METRIC = "sealed_frames_per_sec"


# ------------------------------------------------------------------ workload --
def workload(name: str, n_gpus: int):
    from symmetry_b200 import synth

    if name == "c2":
        lens = np.full(1024, 64, dtype=np.int64)
        desc = "configs[1]: 1,024 concurrent streams x 64 B token-delta frames, batched seal+open"
        scaling = "weak"
    elif name == "c3":
        total = 65536
        lens = synth.power_law_lens(total)[: total // n_gpus]
        desc = "configs[2]: 65,536 streams, power-law frames 32 B-16 KiB (alpha 1.2), sharded by slot, frames in arrival order"
        if os.environ.get("BENCH_C3_SORTED", "0") not in ("", "0"):
            lens = np.sort(lens)     # what a gather step that buckets frames by size would hand over
            desc = desc.replace("frames in arrival order", "frames bucketed by size")
        scaling = "strong"
    elif name == "c4":
        lens = np.full(16384, 65536, dtype=np.int64)
        desc = "configs[3]: bulk 64 KiB frames, 16,384 streams per round (1 GiB), seal+open"
        scaling = "weak"
    else:
        raise SystemExit(f"unknown workload {name}")
    return lens, desc, scaling


class Layout:
    """Ring-slot layout [plain P | wire Wb | recovered P]; step i reads slot i (plain_i ‖ wire_{i-D})
    and writes wire_i ‖ recovered_{i-D} into slot i+D at offset P."""

    def __init__(self, lens):
        import batch_util as bu

        self.lens = [int(x) for x in lens]
        self.p_offs, self.P = bu.plain_layout(self.lens)
        self.w_offs, self.Wb = bu.wire_layout(self.lens)
        self.in_bytes = self.P + self.Wb
        self.out_bytes = self.Wb + self.P
        self.slot_bytes = self.P + self.Wb + self.P

    def descs(self, slots, with_open=True):
        import batch_util as bu

        seal = bu.make_descs(slots, self.lens, self.p_offs, self.w_offs)
        if not with_open:
            return seal
        opn = bu.make_descs(slots, self.lens, [self.P + o for o in self.w_offs], [self.Wb + o for o in self.p_offs],
                            open_=True)
        return np.concatenate([seal, opn])


def algorithmic_bytes(lens) -> int:
    """2*mlen + 92 per frame (seal and open alike)."""
    return int(sum(2 * int(L) + 92 for L in lens))


def int_ops(lens) -> int:
    """992 * nC + 70 * nP per frame."""
    t = 0
    for L in lens:
        L = int(L)
        t += 992 * (2 + (L + 63) // 64) + 70 * ((80 + L + L % 16 + 15) // 16)
    return t


# -------------------------------------------------------------------- clocks --
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.perf_counter(), line.strip()))

    def stop(self, windows):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        all_sm = []
        for t, line in self.samples:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                clk, mx = float(f[0]), float(f[1])
            except ValueError:
                continue
            all_sm.append(clk)
            if any(a - 0.03 <= t <= b + 0.03 for a, b in windows):
                sm.append(clk)
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        use = sm or all_sm
        return {"sm_mhz": float(np.median(use)) if use else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples_under_load": len(sm)}


# --------------------------------------------------------------- CPU baseline --
def cpu_run(lens, steps, warmup, threads, prefer="reference"):
    """Times seal+open of one frame per stream per step on host cores.  Returns (seconds for `steps`, kind)."""
    from oracle import secretstream as ss_port
    from symmetry_b200 import synth

    lib = ss_port.lib()
    kind = "port"
    push = C.cast(lib.cpub_port_push, C.c_void_p).value
    pull = C.cast(lib.cpub_port_pull, C.c_void_p).value
    if prefer == "reference":
        try:
            from oracle import libsodium as ls

            push = ls.fn_ptr("crypto_secretstream_xchacha20poly1305_push")
            pull = ls.fn_ptr("crypto_secretstream_xchacha20poly1305_pull")
            kind = "reference"
        except Exception:
            pass
    n = len(lens)
    keys, hdrs = synth.stream_keys(n)
    tx = np.zeros((n, 52), dtype=np.uint8)
    for i in range(n):
        tx[i] = np.frombuffer(ss_port.State.init(bytes(keys[i]), bytes(hdrs[i])).raw, dtype=np.uint8)
    rx = tx.copy()
    lay = Layout(lens)
    ring = 4
    rng = np.random.default_rng(7)
    plain = rng.integers(0, 256, ring * lay.P, dtype=np.uint8)
    wire = np.zeros(ring * lay.Wb + 64, dtype=np.uint8)
    p_off = np.asarray(lay.p_offs, dtype=np.uint64)
    w_off = np.asarray(lay.w_offs, dtype=np.uint64)
    lens32 = np.asarray(lens, dtype=np.uint32)

    def go(k):
        return lib.cpub_run(push, pull, tx.ctypes.data, rx.ctypes.data, n, plain.ctypes.data, p_off.ctypes.data,
                            lens32.ctypes.data, lay.P, wire.ctypes.data, w_off.ctypes.data, lay.Wb, ring, k, 1, 1,
                            threads)

    if warmup:
        assert go(warmup) >= 0
    sec = go(steps)
    assert sec >= 0, "CPU path reported a failed pull"
    assert (tx == rx).all(), "CPU seal/open chains diverged"
    return sec, kind


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model() -> str:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


# ------------------------------------------------------------------ reference --
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lens, desc, scaling = workload(args.workload, 1)
    if scaling == "weak" and args.gpus > 1:
        lens = np.tile(lens, args.gpus)      # same whole-job batch as our arm at N GPUs (per-GPU work is fixed)
    threads = host_threads()
    # bound the run: a step is the same batch, but never let warmup+steps exceed ~2 minutes
    probe, kind = cpu_run(lens, 1, 0, threads)
    max_steps = max(1, int(120.0 / max(probe, 1e-6)))
    steps = min(args.steps, max_steps)
    warm = min(args.warmup, max(1, max_steps // 10))
    sec, kind = cpu_run(lens, steps, warm, threads)
    frames = len(lens) * steps
    val = frames / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": sec / steps * 1e3, "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "streams": len(lens), "frames_per_step": 2 * len(lens),
                   "plaintext_bytes_per_step": int(sum(lens))},
        "gb_per_s": float(sum(lens)) * steps / sec / 1e9,
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": f"{steps} steps of the full batch (seal+open, {len(lens)} streams), "
                                   f"libsodium 1.0.20 (PyNaCl build) via oracle/cpu_bench.c, {cpu_model()}"},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------- ours --
def run_ours(args):
    import torch
    import torch.distributed as dist

    import symmetry_b200 as sb
    from oracle import libsodium as ls
    from symmetry_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    n_gpus = max(world, 1)
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")

    lens, desc, scaling = workload(args.workload, n_gpus)
    n = len(lens)
    lay = Layout(lens)
    K, W = args.steps, args.warmup
    D_DEV, D_E2E = 1, int(os.environ.get("BENCH_E2E_DEPTH", "6"))
    stage_bytes = max(64 << 20, (lay.in_bytes + 4095) // 4096 * 4096)
    relay = sb.SymRelay(gpu_ids=(local,), max_streams=max(4 * n, 4096), pipeline_depth=D_E2E,
                        stage_bytes=stage_bytes, max_batch_frames=max(2 * n, 4096))
    keys, hdrs = synth.stream_keys(n, first=rank * n)
    sample = sorted(set([0, 1, n // 2, n - 1] + [int(i) for i in np.argsort(lens)[-2:]]))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident run: `value` ---------------------------
    slots = relay.open_streams(keys, hdrs, keys, hdrs)
    R = int(min(2048, max(4, math.ceil(2 * L2_BYTES / lay.slot_bytes))))
    ring = torch.empty(R * lay.slot_bytes, dtype=torch.uint8, device=dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    ring.random_(0, 256, generator=gen)
    flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device=dev)
    d_full = lay.descs(slots, True)
    d_seal = lay.descs(slots, False)
    # Ragged batches are planned with the library's sorting planner (descriptors grouped by size class, statuses
    # come back through a permutation); BENCH_PLAN=caller keeps the frames in arrival order.
    plan_mode = os.environ.get("BENCH_PLAN", "sorted" if args.workload == "c3" else "caller")
    if plan_mode == "sorted":
        tiles_full, nb_full, d_full, _perm_full = relay.plan_sorted(d_full)
        tiles_seal, nb_seal, d_seal, _perm_seal = relay.plan_sorted(d_seal)
    else:
        tiles_full, nb_full = relay.plan(d_full)
        tiles_seal, nb_seal = relay.plan(d_seal)
    t_dfull = torch.from_numpy(d_full.view(np.uint8).copy()).to(dev)
    t_dseal = torch.from_numpy(d_seal.view(np.uint8).copy()).to(dev)
    t_tfull = torch.from_numpy(tiles_full.view(np.uint8).copy()).to(dev)
    t_tseal = torch.from_numpy(tiles_seal.view(np.uint8).copy()).to(dev)
    t_status = torch.full((2 * n,), 0xFF, dtype=torch.int32, device=dev)
    base = ring.data_ptr()
    launches_per_step = (1 if len(tiles_full) - nb_full else 0) + (1 if nb_full else 0)

    def dev_step(i, full=True):
        stream = torch.cuda.current_stream().cuda_stream
        src = base + (i % R) * lay.slot_bytes
        dst = base + ((i + D_DEV) % R) * lay.slot_bytes + lay.P
        if full:
            relay.batch_device(t_dfull.data_ptr(), 2 * n, t_tfull.data_ptr(), len(tiles_full), nb_full, src, dst,
                               t_status.data_ptr(), stream=stream)
        else:
            relay.batch_device(t_dseal.data_ptr(), n, t_tseal.data_ptr(), len(tiles_seal), nb_seal, src, dst,
                               t_status.data_ptr(), stream=stream)

    # keep a host copy of the sample streams' plaintext per ring slot for the CPU replay
    ring_plain_host = {}
    ring_view = ring.view(R, lay.slot_bytes)
    for s in sample:
        o, L = lay.p_offs[s], lay.lens[s]
        ring_plain_host[s] = ring_view[:, o:o + L].cpu().numpy().copy()

    step_no = 0
    for _ in range(D_DEV):                      # priming: seal only (no inbound frames exist yet)
        dev_step(step_no, full=False)
        step_no += 1
    for _ in range(W):
        dev_step(step_no)
        step_no += 1
    # The inner loop is launch-bound, so R consecutive steps (one lap of the ring) are captured in a
    # CUDA graph and replayed; the remaining K % R steps are launched directly.  Every step is still
    # its own kernel launch (graph node).
    n_replays = 0 if args.no_graph else K // R
    n_rem = K - n_replays * R
    graph = graph_rem = None
    launches_per_lap = launches_rem = 0
    if not args.no_graph:
        while step_no % R:                       # extra untimed warm-up so a lap starts at slot 0
            dev_step(step_no)
            step_no += 1
        torch.cuda.synchronize()
        if n_replays:
            l0 = relay.launches
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                for j in range(R):
                    dev_step(step_no + j)
            launches_per_lap = relay.launches - l0
        if n_rem:                                # the last K % R steps: a second, shorter graph
            l0 = relay.launches
            graph_rem = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph_rem):
                for j in range(n_rem):
                    dev_step(step_no + j)
            launches_rem = relay.launches - l0
        # One untimed replay of each graph: the first launch of an instantiated graph also uploads it to
        # the device (about a millisecond for a lap of 1,152 nodes), which is set-up, not a step.  Both are
        # whole-lap neutral: after the shorter graph the lap is completed with direct launches.
        if graph is not None:
            graph.replay()
            step_no += R
        if graph_rem is not None:
            graph_rem.replay()
            step_no += n_rem
            while step_no % R:
                dev_step(step_no)
                step_no += 1
        torch.cuda.synchronize()
    flush.zero_()                                # evict whatever set-up left in L2
    clocks = ClockSampler(local)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = relay.launches
    t_w0 = time.perf_counter()
    e0.record()
    for _ in range(n_replays):
        graph.replay()
        step_no += R
    if graph_rem is not None:
        graph_rem.replay()
        step_no += n_rem
    elif n_rem:
        for _ in range(n_rem):
            dev_step(step_no)
            step_no += 1
    e1.record()
    barrier()
    t_w1 = time.perf_counter()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    gpu_launches = relay.launches - launches0 + n_replays * launches_per_lap + (launches_rem if graph_rem is not None else 0)
    total_steps = step_no

    # parity spot-check of the timed run against libsodium (tx chain, rx chain, last wire, recovered plaintext)
    st = t_status.cpu().numpy()
    assert (st == 0).all(), "non-OK frame status in the last timed step"
    last = total_steps - 1
    wire_slot = ring_view[(last + D_DEV) % R].cpu().numpy()
    for s in sample:
        ref = ls.State(bytes(keys[s]), bytes(hdrs[s]))
        rx_expect = None
        w = None
        for i in range(total_steps):
            if i == total_steps - D_DEV:
                rx_expect = ref.raw
            w = ref.seal_frame(bytes(ring_plain_host[s][i % R]))
        tx_state, rx_state = relay.export(int(slots[s]))
        assert tx_state == ref.raw, f"tx chain of stream {s} differs from libsodium after {total_steps} frames"
        assert rx_state == rx_expect, f"rx chain of stream {s} differs from libsodium"
        o = lay.P + lay.w_offs[s]
        assert bytes(wire_slot[o:o + len(w)]) == w, f"last wire frame of stream {s} differs from libsodium"
        ro = lay.P + lay.Wb + lay.p_offs[s]
        assert bytes(wire_slot[ro:ro + lay.lens[s]]) == bytes(ring_plain_host[s][(last - D_DEV) % R]), "recovered plaintext"

    # integer roofline denominators, measured on this GPU right now
    arx_tops, imad_tops = relay.probe_int_peak(0)

    # ---------------- end-to-end through host arenas: `e2e` -------------------
    del ring, flush
    torch.cuda.empty_cache()
    slots2 = relay.open_streams(keys, hdrs, keys, hdrs)
    R2 = max(2 * D_E2E + 2, 8)
    if args.no_e2e:
        R2 = 1
    hring = relay.alloc_pinned(R2 * lay.slot_bytes)
    hring[:] = np.random.default_rng(99 + rank).integers(0, 256, hring.size, dtype=np.uint8)
    hplain = {s: [bytes(hring[j * lay.slot_bytes + lay.p_offs[s]: j * lay.slot_bytes + lay.p_offs[s] + lay.lens[s]])
                  for j in range(R2)] for s in sample}
    e_full = lay.descs(slots2, True)
    e_seal = lay.descs(slots2, False)
    hstat = [np.full(2 * n, 0xFF, dtype=np.uint32) for _ in range(R2)]
    hbase = hring.ctypes.data
    tickets = {}

    def e2e_step(i, full=True):
        if i - D_E2E in tickets:
            relay.wait(tickets.pop(i - D_E2E))
        d = e_full if full else e_seal
        tickets[i] = relay.submit(d, hbase + (i % R2) * lay.slot_bytes, lay.in_bytes,
                                  hbase + ((i + D_E2E) % R2) * lay.slot_bytes + lay.P, lay.out_bytes, hstat[i % R2])

    def e2e_drain():
        for k in sorted(tickets):
            relay.wait(tickets.pop(k))

    K2 = K if args.e2e_steps <= 0 else args.e2e_steps
    if args.no_e2e:
        K2 = 0
    i2 = 0
    for _ in range(D_E2E if K2 else 0):
        e2e_step(i2, full=False)
        i2 += 1
    for _ in range(W if K2 else 0):
        e2e_step(i2)
        i2 += 1
    e2e_drain()
    barrier()
    # The timed loop is driven from C (tools/e2e_driver.c, a plain caller of the public C ABI, as the
    # N-API addon would be): per batch it calls symrelay_batch_wait(ticket[i-4]) then
    # symrelay_batch_submit(...) on the same ring -- the calls e2e_step() makes, minus the interpreter.
    e2e_driver = "python"
    t0 = time.perf_counter()
    if K2 and not args.py_e2e:
        from symmetry_b200 import build as sbuild

        drv = C.CDLL(sbuild.build_driver())
        drv.e2e_ring_run.restype = C.c_double
        drv.e2e_ring_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                                     C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        stat_all = np.full(R2 * 2 * n, 0xFF, dtype=np.uint32)
        secs = drv.e2e_ring_run(relay.ctx, e_full.ctypes.data, 2 * n, hbase, lay.slot_bytes, lay.in_bytes, lay.P,
                                lay.out_bytes, R2, D_E2E, i2, K2, stat_all.ctypes.data, None)
        assert secs > 0, f"e2e driver failed: {secs}"
        for i in range(i2, i2 + K2):          # every ring slot a timed step wrote must be all-OK
            j = i % R2
            hstat[j][:] = stat_all[j * 2 * n:(j + 1) * 2 * n]
        i2 += K2
        e2e_driver = "tools/e2e_driver.c"
        t1 = t0 + secs
    else:
        for _ in range(K2):
            e2e_step(i2)
            i2 += 1
        e2e_drain()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
    e2e_sec = max_over_ranks(t1 - t0)
    clk = clocks.stop([(t_w0, t_w1), (t0, t1)])
    for j in sorted({i % R2 for i in range(max(0, i2 - K2), i2)} if K2 else set()):
        assert (hstat[j] == 0).all(), "non-OK status in the e2e run"
    for s in (sample if K2 else []):
        ref = ls.State(bytes(keys[s]), bytes(hdrs[s]))
        for i in range(i2):
            w = ref.seal_frame(hplain[s][i % R2])
        assert relay.export(int(slots2[s]))[0] == ref.raw, "e2e tx chain differs from libsodium"
        o = ((i2 - 1 + D_E2E) % R2) * lay.slot_bytes + lay.P + lay.w_offs[s]
        assert bytes(hring[o:o + len(w)]) == w, "e2e last wire frame differs from libsodium"

    if rank != 0:
        relay.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- CPU baseline on this box's host cores (rank 0, N=1) -----
    cpu = None
    if world == 1 and not args.no_cpu:
        threads = host_threads()
        probe, kind = cpu_run(lens, 2, 0, threads)
        per_step_cpu_s = probe / 2 * threads           # CPU-seconds of work per step
        cpu_steps = int(min(max(4, 15.0 / max(per_step_cpu_s, 1e-9)), 200000))
        sec, kind = cpu_run(lens, cpu_steps, max(1, cpu_steps // 20), threads)
        sec1, _ = cpu_run(lens, max(1, cpu_steps // threads), 1, 1)
        cpu = {"value": n * cpu_steps / sec, "unit": "frames/s", "cores": threads, "kind": kind,
               "sample": f"{cpu_steps} steps of the same batch (seal+open per stream), ~{per_step_cpu_s * cpu_steps:.0f} "
                         f"CPU-s; libsodium 1.0.20 (PyNaCl build, the C library under sodium-native) via "
                         f"oracle/cpu_bench.c; {cpu_model()}",
               "single_core_value": n * max(1, cpu_steps // threads) / sec1}

    # ---------------- p50 added relay latency (the metric's third figure), rank 0 at N=1 ----------
    latency = None
    if world == 1 and not args.no_latency:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import latency_bench

            relay.close()
            relay = None
            latency = latency_bench.run(streams=65536, rate=2e6, nbytes=200, seconds=1.0, flush_us=100)
        except Exception as e:  # the headline numbers do not depend on it
            latency = {"error": repr(e)}

    frames_per_step = n                                  # sealed frames; each step also opens n frames
    dev_s = dev_ms * 1e-3
    value = n_gpus * frames_per_step * K / dev_s
    alg_bytes = algorithmic_bytes(lens) * 2              # seal + open
    ops = int_ops(lens) * 2
    hbm_peak, hbm_src = hbm_peak_gbs()
    per_launch_ms = dev_ms / K
    achieved_gbs = alg_bytes / (per_launch_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(args.workload)
        except Exception:
            traffic = None
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
        "ms_per_step": per_launch_ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": desc, "streams_per_gpu": n, "plan": ("symrelay_plan_sorted (library groups the frames by size class)" if plan_mode == "sorted" else "symrelay_plan (caller's descriptor order)"), "frames_per_step_per_gpu": 2 * n,
                   "plaintext_bytes_per_step_per_gpu": int(sum(lens)),
                   "l2": f"inputs larger than L2: ring of {R} distinct batch slots, {R * lay.slot_bytes >> 20} MiB, "
                         f"visited round-robin; L2 flushed once before timing",
                   "step": "every stream seals one frame and opens one frame (mixed batch)",
                   "launch": ("direct launches" if args.no_graph else
                              f"{n_replays} replays of a CUDA graph holding {R} steps + one graph of the last {n_rem} steps, "
                              "each replayed once untimed first (graph upload)")},
        "gb_per_s": n_gpus * float(sum(lens)) * K / dev_s / 1e9,
        "opened_frames_per_sec": value,
        "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved_gbs / hbm_peak, "traffic": traffic,
                     "peak_source": hbm_src,
                     "kernel": "ss_packed_kernel" if nb_full == 0 else "ss_bulk_kernel",
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "note": "the path is integer-ALU bound (SURVEY 8(d)); see alu_roofline"},
        "alu_roofline": {"bound": "int_alu", "achieved": ops / (per_launch_ms * 1e-3) / 1e12,
                         "peak": arx_tops, "unit": "Tint-op/s", "frac": ops / (per_launch_ms * 1e-3) / 1e12 / arx_tops,
                         "int_ops_per_launch": ops, "peak_source": "symrelay_probe_int_peak (ChaCha ARX mix, measured now)",
                         "imad_wide_tops": imad_tops},
        "e2e": {"value": (n_gpus * frames_per_step * K2 / e2e_sec) if K2 else None, "unit": "frames/s",
                "h2d_bytes_per_step": lay.in_bytes + 2 * n * 16 + len(tiles_full) * 8,
                "d2h_bytes_per_step": lay.out_bytes + 2 * n * 4, "steps": K2,
                "ms_per_step": (e2e_sec / K2 * 1e3) if K2 else None,
                "path": "symrelay_batch_submit/wait on pinned host arenas, pipeline depth " + str(D_E2E) + ", host clock; loop driver: " + e2e_driver},
        "gpu_launches": int(gpu_launches),
        "launches_per_step": launches_per_step,
        "clocks": clk,
        "parity": "timed run spot-checked against libsodium: tx/rx chains, last wire frames, recovered plaintext",
    }
    if cpu:
        line["cpu_baseline"] = cpu
    if latency:
        line["latency"] = latency
    print(json.dumps(line), flush=True)
    if relay is not None:
        relay.close()
    if world > 1:
        dist.destroy_process_group()


def hbm_peak_gbs():
    """(GB/s, where it came from): the driver-written measured copy bandwidth, else the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured)"
    except Exception:
        return 6650.0, "fallback 6650"


# ------------------------------------------------------------------------- c5 --
def run_c5(args):
    """configs[4]: fused relay -- per batch one synthetic chat-completion chunk per stream:
    prefix test + JSON validation + delta extraction + seal (device-resident)."""
    import torch

    import batch_util as bu
    import symmetry_b200 as sb
    from oracle import extract as ex
    from oracle import libsodium as ls
    from symmetry_b200 import synth

    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    n, n_kinds = int(os.environ.get("BENCH_C5_STREAMS", "16384")), 16   # configs[4] says 16k; the env knob is for the batch-size note in DESIGN.md
    K, W = args.steps, args.warmup
    relay = sb.SymRelay(gpu_ids=(0,), max_streams=2 * n)
    keys, hdrs = synth.stream_keys(n)
    slots = relay.open_streams(keys, hdrs, None, None, provider=sb.PROVIDER_OLLAMA)
    # 64 distinct completions cycled over the streams; ring slot j holds chunk j of each stream
    base_streams = [synth.chat_chunks(n_kinds, seed=5, stream_id=i) for i in range(64)]
    R = n_kinds + 2
    slot_chunks = [[base_streams[s % 64][j] for s in range(n)] for j in range(R)]
    lays = []
    for j in range(R):
        lens = [len(c) for c in slot_chunks[j]]
        p_offs, P = bu.plain_layout(lens)
        w_offs, Wb = bu.wire_layout(lens)
        lays.append((lens, p_offs, P, w_offs, Wb))
    Pmax = max(l[2] for l in lays)
    Wmax = max(l[4] for l in lays)
    t_in = torch.zeros(R, Pmax, dtype=torch.uint8, device=dev)
    t_out = torch.zeros(R, Wmax, dtype=torch.uint8, device=dev)
    t_darena = torch.zeros(R, Pmax, dtype=torch.uint8, device=dev)
    t_delta = torch.zeros(R, n, 2, dtype=torch.int32, device=dev)
    t_status = torch.zeros(R, n, dtype=torch.int32, device=dev)
    t_desc, t_tiles, meta = [], [], []
    for j, (lens, p_offs, P, w_offs, Wb) in enumerate(lays):
        host = np.zeros(Pmax, dtype=np.uint8)
        for c, o in zip(slot_chunks[j], p_offs):
            host[o:o + len(c)] = np.frombuffer(c, dtype=np.uint8)
        t_in[j].copy_(torch.from_numpy(host))
        d = bu.make_descs(slots, lens, p_offs, w_offs, extract=True)
        tiles, nb = relay.plan(d)
        t_desc.append(torch.from_numpy(d.view(np.uint8).copy()).to(dev))
        t_tiles.append(torch.from_numpy(tiles.view(np.uint8).copy()).to(dev))
        meta.append((len(tiles), nb))

    def step(i):
        j = i % R
        relay.batch_device(t_desc[j].data_ptr(), n, t_tiles[j].data_ptr(), meta[j][0], meta[j][1], t_in[j].data_ptr(),
                           t_out[j].data_ptr(), t_status[j].data_ptr(), stream=torch.cuda.current_stream().cuda_stream,
                           d_delta=t_delta[j].data_ptr(), d_delta_arena=t_darena[j].data_ptr())

    i = 0
    for _ in range(W):
        step(i); i += 1
    while i % R:
        step(i); i += 1
    torch.cuda.synchronize()
    l0 = relay.launches
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for j in range(R):
            step(i + j)
    per_lap = relay.launches - l0
    g.replay()                                   # untimed: the first launch uploads the graph
    i += R
    reps = max(1, K // R)
    clocks = ClockSampler(0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.perf_counter()
    e0.record()
    for _ in range(reps):
        g.replay()
        i += R
    e1.record()
    torch.cuda.synchronize()
    tw1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    steps = reps * R
    # parity: statuses OK, deltas == oracle for the last lap, chains == libsodium on sampled streams
    assert (t_status.cpu().numpy() == 0).all()
    dl = t_delta.cpu().numpy()
    da = t_darena.cpu().numpy()
    for j in range(R):
        for s in range(0, n, 1021):
            off, ln = int(dl[j, s, 0]), int(dl[j, s, 1])
            code, want = ex.extract("ollama", slot_chunks[j][s])
            assert code == ex.OK and bytes(da[j, off:off + ln]) == want, (j, s)
    for s in (0, 1, n - 1):
        ref = ls.State(bytes(keys[s]), bytes(hdrs[s]))
        for k in range(i):
            w = ref.seal_frame(slot_chunks[k % R][s])
        assert relay.export(int(slots[s]))[0] == ref.raw, "c5 tx chain differs from libsodium"
        jl = (i - 1) % R
        o = lays[jl][3][s]
        assert bytes(t_out[jl, o:o + len(w)].cpu().numpy()) == w
    arx, imad = relay.probe_int_peak(0)
    clk = clocks.stop([(tw0, tw1)])
    all_lens = [x for l in lays for x in l[0]]
    mean_len = sum(all_lens) / len(all_lens)
    ops = sum(int_ops(l[0]) + 4 * sum(l[0]) for l in lays) / R          # +4 ops/byte for the extractor
    per_step = ms / steps
    alg_bytes = sum(2 * sum(l[0]) + 92 * len(l[0]) for l in lays) / R
    hbm_peak, hbm_src = hbm_peak_gbs()
    line = {
        "metric": METRIC, "value": n * steps / (ms * 1e-3), "unit": "frames/s", "n_gpus": 1, "steps": steps,
        "warmup": W, "ms_per_step": per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"configs[4]: fused SSE data: test + JSON validate + delta extract + seal, {n:,} streams "
                               "of synthetic llama3:8b chat completions, apiProvider ollama",
                   "streams_per_gpu": n, "mean_chunk_bytes": mean_len, "launch": f"{reps} replays of a CUDA graph of {R} steps",
                   "l2": f"ring of {R} distinct batches ({(R * (2 * Pmax + Wmax)) >> 20} MiB) visited round-robin"},
        "gb_per_s": mean_len * n * steps / (ms * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "achieved": alg_bytes / (per_step * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": alg_bytes / (per_step * 1e-3) / 1e9 / hbm_peak, "traffic": None, "peak_source": hbm_src,
                     "kernel": "ss_extract_kernel + ss_packed_kernel (whole step)",
                     "algorithmic_bytes_per_step": alg_bytes,
                     "note": "2*mlen+92 bytes per frame (SURVEY 8(d)); the step is bound by the extractor's serial "
                             "per-chunk instruction stream, see alu_roofline and DESIGN.md 4.1"},
        "alu_roofline": {"bound": "int_alu", "achieved": ops / (per_step * 1e-3) / 1e12, "peak": arx, "unit": "Tint-op/s",
                         "frac": ops / (per_step * 1e-3) / 1e12 / arx},
        "gpu_launches": reps * per_lap, "launches_per_step": per_lap / R, "clocks": clk,
        "parity": "deltas == oracle/extract.py on sampled frames of every ring slot; chains == libsodium",
    }
    print(json.dumps(line), flush=True)
    relay.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="default: 20000 for c2, 200 for c3, 64 for c4, 360 for c5")
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = same as --steps")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-latency", action="store_true", help="skip the p50 added-latency leg (64k streams, Poisson)")
    ap.add_argument("--py-e2e", action="store_true", help="drive the e2e loop from Python instead of tools/e2e_driver.c")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (huge workloads)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step directly (no CUDA graph)")
    args = ap.parse_args()
    if args.steps is None:   # a step of c3/c4/c5 is 10-400x a step of c2
        args.steps = {"c2": 20000, "c3": 200, "c4": 64, "c5": 360}.get(args.workload, 20000)
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "c5":
        run_c5(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
