#!/usr/bin/env python3
"""bench.py -- batched secretstream seal+open throughput on B200(s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c1|c3|c4|c5|c4sweep]
                    [--impl ours|reference] [--sub c3,c4,c5|none]

A "step" is one pass of the relay hot path over one batch: every stream seals
one outbound frame AND opens one inbound frame (one mixed batch = one kernel
launch per tile class).  Default workload = BASELINE.json configs[1]: 1,024
concurrent streams x 64 B token-delta frames, batched seal+open on 1 B200.
`value` is device-resident (inputs already in HBM); `e2e` goes through the
C ABI's host-buffer batch call with pinned host arenas, H2D and D2H inside the
timed region.  The default line also carries `c4` and `c3` sub-objects (the
bulk roofline probe and the ragged 65,536-stream batch sharded over the N
ranks), measured in the same run, device-resident.  `--impl reference` times
the reference's CPU path for the same workload: real libsodium (the C library
under sodium-native) driven by the C loop in oracle/cpu_bench.c on all host
cores.  Prints ONE JSON line.
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
GATE_CYCLES = 400_000                       # ~200 us at 1.965 GHz: the spin kernel ahead of the device-timed region
DEFAULT_STEPS = {"c2": 20000, "c3": 200, "c4": 64, "c5": 360, "c1": 20, "c4sweep": 8}
SUB_STEPS = {"c4": 64, "c3": 200, "c5": 360}          # steps of the sub-lines riding on the default line


# ------------------------------------------------------------------ workload --
def workload(name: str, n_gpus: int, rank: int = 0):
    """(lens of this rank's streams, description, scaling, first stream id of this rank)."""
    from symmetry_b200 import shard, synth

    if name == "c2":
        lens = np.full(1024, 64, dtype=np.int64)
        desc = "configs[1]: 1,024 concurrent streams x 64 B token-delta frames, batched seal+open"
        return lens, desc, "weak", rank * 1024
    if name == "c3":
        total = 65536
        # BENCH_C3_SHARD_OF=8 on one GPU: the share rank 0 of 8 gets (a one-GPU look at the strong-scaling step)
        of = int(os.environ.get("BENCH_C3_SHARD_OF", "0") or 0)
        lo, hi = shard.shard_streams(total, of if of > 1 and n_gpus == 1 else n_gpus, rank)      # connections shard by key; no data-path collective
        lens = synth.power_law_lens(total)[lo:hi]
        desc = "configs[2]: 65,536 streams, power-law frames 32 B-16 KiB (alpha 1.2), sharded by connection over the GPUs, frames in arrival order"
        if of > 1 and n_gpus == 1:
            desc += f" [BENCH_C3_SHARD_OF={of}: only the {hi - lo} streams rank 0 of {of} would get]"
        if os.environ.get("BENCH_C3_SORTED", "0") not in ("", "0"):
            lens = np.sort(lens)     # what a gather step that buckets frames by size would hand over
            desc = desc.replace("frames in arrival order", "frames bucketed by size")
        return lens, desc, "strong", lo
    if name == "c4":
        lens = np.full(16384, 65536, dtype=np.int64)
        desc = "configs[3]: bulk 64 KiB frames, 16,384 streams per round (1 GiB), seal+open"
        return lens, desc, "weak", rank * 16384
    raise SystemExit(f"unknown workload {name}")


class Layout:
    """Ring-slot layout [plain P | wire Wb | recovered P]; step i reads slot i (plain_i ‖ wire_{i-D})
    and writes wire_i ‖ recovered_{i-D} into slot i+D at offset P."""

    def __init__(self, lens):
        lens = np.asarray(lens, dtype=np.int64)
        self.lens = lens
        pw = np.maximum(16, (lens + 15) & ~15)
        ww = (12 + lens + 20 + 15) & ~15
        self.p_offs = np.concatenate([[0], np.cumsum(pw)[:-1]]).astype(np.int64) if len(lens) else np.zeros(0, np.int64)
        self.w_offs = (np.concatenate([[0], np.cumsum(ww)[:-1]]) + 12).astype(np.int64) if len(lens) else np.zeros(0, np.int64)
        self.P = int(pw.sum())
        self.Wb = int(ww.sum())
        self.in_bytes = self.P + self.Wb
        self.out_bytes = self.Wb + self.P
        self.slot_bytes = self.P + self.Wb + self.P

    def descs(self, slots, with_open=True):
        import batch_util as bu

        seal = bu.make_descs(slots, self.lens, self.p_offs, self.w_offs)
        if not with_open:
            return seal
        opn = bu.make_descs(slots, self.lens, self.P + self.w_offs, self.Wb + self.p_offs, open_=True)
        return np.concatenate([seal, opn])


def algorithmic_bytes(lens) -> int:
    """2*mlen + 92 per frame (seal and open alike)."""
    lens = np.asarray(lens, dtype=np.int64)
    return int((2 * lens + 92).sum())


def int_ops(lens) -> int:
    """992 * nC + 70 * nP per frame."""
    L = np.asarray(lens, dtype=np.int64)
    return int((992 * (2 + (L + 63) // 64) + 70 * ((80 + L + L % 16 + 15) // 16)).sum())


def ring_slots(lay: Layout) -> int:
    return int(min(2048, max(4, math.ceil(2 * L2_BYTES / max(lay.slot_bytes, 1)))))


def make_config(name: str, n_gpus: int, K: int, no_graph: bool, plan_mode: str) -> dict:
    """The ONE description of the workload both arms print (the driver compares the two dicts)."""
    lens, desc, scaling, _ = workload(name, n_gpus, 0)
    lay = Layout(lens)
    R = ring_slots(lay)
    n_replays = 0 if no_graph else K // R
    return {
        "workload": desc, "n_gpus": n_gpus, "streams_per_gpu": int(len(lens)),
        "frames_per_step_per_gpu": 2 * int(len(lens)), "plaintext_bytes_per_step_per_gpu": int(lens.sum()),
        "step": "every stream seals one frame and opens one frame (mixed batch); both chains run through all steps",
        "plan": ("symrelay_plan_sorted (library groups the frames by size class)" if plan_mode == "sorted"
                 else "symrelay_plan (caller's descriptor order)"),
        "l2": f"GPU arm: inputs larger than L2 -- ring of {R} distinct batch slots, {R * lay.slot_bytes >> 20} MiB, visited "
              f"round-robin, L2 flushed once before timing; CPU arm: ring of 4 batch slots in host memory",
        "launch": ("GPU arm: direct launches" if no_graph else
                   f"GPU arm: {n_replays} replays of a CUDA graph holding {R} steps + one graph of the last {K - n_replays * R} "
                   "steps, each replayed once untimed first (graph upload)")
                  + "; CPU arm: a pool of host threads started and warmed before the timed steps",
        "timing": "GPU arm: CUDA events on the launching stream around exactly K steps, max over ranks; a ~200 us spin kernel "
                  "ahead of the start event lets the host queue the timed launches first, so host launch latency is outside "
                  "the window; CPU arm: each worker thread's own clock, latest end minus earliest start",
    }


def plan_mode_for(name: str) -> str:
    return os.environ.get("BENCH_PLAN", "sorted" if name == "c3" else "caller")


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
    """Times seal+open of one frame per stream per step on host cores: `warmup` untimed steps and then
    `steps` timed ones on the same pool of threads.  Returns (seconds for `steps`, kind)."""
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
    sec = lib.cpub_run(push, pull, tx.ctypes.data, rx.ctypes.data, n, plain.ctypes.data, p_off.ctypes.data,
                       lens32.ctypes.data, lay.P, wire.ctypes.data, w_off.ctypes.data, lay.Wb, ring, steps, warmup, 1, 1,
                       threads)
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
    name = args.workload if args.workload in ("c2", "c3", "c4") else "c2"
    lens, desc, scaling, _ = workload(name, 1)
    if scaling == "weak" and args.gpus > 1:
        lens = np.tile(lens, args.gpus)      # same whole-job batch as our arm at N GPUs (per-GPU work is fixed)
    threads = host_threads()
    # bound the run: a step is the same batch, but never let warmup+steps exceed ~2 minutes
    probe, kind = cpu_run(lens, 2, 1, threads)
    max_steps = max(1, int(120.0 / max(probe / 2, 1e-6)))
    steps = min(args.steps, max_steps)
    # The pool warms up on the same threads before the timed steps; a short timed region (the driver's
    # --steps 20 is about a millisecond of CPU work) needs more warm-up than --warmup to reach the steady
    # state the long cpu_baseline sample of the other arm measures (idle cores take tens of milliseconds to reach
    # their working clocks), so warm for at least ~0.3 s.
    warm = max(args.warmup, min(max_steps, int(0.3 / max(probe / 2, 1e-9))))
    sec, kind = cpu_run(lens, steps, warm, threads)
    frames = len(lens) * steps
    val = frames / sec
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": sec / steps * 1e3, "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": make_config(name, args.gpus, args.steps, args.no_graph, plan_mode_for(name)),
        "gb_per_s": float(sum(lens)) * steps / sec / 1e9,
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": f"{steps} steps of the whole-job batch (seal+open, {len(lens)} streams) after {warm} "
                                   f"warm-up steps on the same thread pool, libsodium 1.0.20 (PyNaCl build) via "
                                   f"oracle/cpu_bench.c, {cpu_model()}"},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------- ours --
class Job:
    """torch.distributed plumbing shared by the legs of one run."""

    def __init__(self):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{self.local}"))
            self.host_group = dist.new_group(backend="gloo")   # barriers that must not synchronise the device
        self.n_gpus = max(self.world, 1)
        torch.cuda.set_device(self.local)
        self.dev = torch.device(f"cuda:{self.local}")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def host_barrier(self):
        """Ranks line up without a device synchronisation (the resident relay kernel stays alive)."""
        if self.world > 1:
            self.dist.barrier(group=self.host_group)

    def max_over_ranks(self, x: float) -> float:
        from symmetry_b200 import shard

        return shard.max_over_ranks(x, self.dist if self.world > 1 else None, self.dev)

    def sum_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def measure_device(job: Job, name: str, K: int, W: int, no_graph: bool, relay=None, keep_relay=False, custom=None):
    """Device-resident leg of one workload: K timed steps (CUDA events on the launching stream, barrier +
    synchronize on both sides, max over ranks), then the spot check against libsodium.  Returns a dict."""
    import symmetry_b200 as sb
    from oracle import libsodium as ls
    from symmetry_b200 import synth

    torch = job.torch
    dev = job.dev
    if custom is not None:          # (lens, description, scaling, keys, headers): the size sweep brings its own
        lens, desc, scaling, keys, hdrs = custom
    else:
        lens, desc, scaling, first = workload(name, job.n_gpus, job.rank)
        keys, hdrs = synth.stream_keys(len(lens), first=first)
    n = len(lens)
    lay = Layout(lens)
    D = 1
    own_relay = relay is None
    if own_relay:
        relay = sb.SymRelay(gpu_ids=(job.local,), max_streams=max(2 * n, 4096), pipeline_depth=1, stage_bytes=1 << 20,
                            max_batch_frames=4096)
    sample = sorted(set([0, 1, n // 2, n - 1] + [int(i) for i in np.argsort(lens)[-2:]]))
    slots = relay.open_streams(keys, hdrs, keys, hdrs)
    R = ring_slots(lay)
    ring = torch.empty(R * lay.slot_bytes, dtype=torch.uint8, device=dev)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + job.rank)
    ring.random_(0, 256, generator=gen)
    flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device=dev)
    d_full = lay.descs(slots, True)
    d_seal = lay.descs(slots, False)
    # Ragged batches are planned with the library's sorting planner (descriptors grouped by size class, statuses
    # come back through a permutation); BENCH_PLAN=caller keeps the frames in arrival order.
    plan_mode = plan_mode_for(name)
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
        dst = base + ((i + D) % R) * lay.slot_bytes + lay.P
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
        o, L = int(lay.p_offs[s]), int(lay.lens[s])
        ring_plain_host[s] = ring_view[:, o:o + L].cpu().numpy().copy()

    step_no = 0
    for _ in range(D):                      # priming: seal only (no inbound frames exist yet)
        dev_step(step_no, full=False)
        step_no += 1
    for _ in range(W):
        dev_step(step_no)
        step_no += 1
    # The inner loop is launch-bound, so R consecutive steps (one lap of the ring) are captured in a
    # CUDA graph and replayed; the remaining K % R steps are launched directly.  Every step is still
    # its own kernel launch (graph node).
    n_replays = 0 if no_graph else K // R
    n_rem = K - n_replays * R
    graph = graph_rem = None
    launches_per_lap = launches_rem = 0
    if not no_graph:
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
    job.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = relay.launches
    t_w0 = time.perf_counter()
    # A spin kernel (no memory traffic) keeps the stream busy while the host queues the start event, the timed launches
    # and the end event: the two events then bracket device time only.  Without it the driver's 20-step run (one graph
    # launch of ~130 us of work) also times the host's cudaGraphLaunch call, which with eight ranks sharing the host cores
    # was the whole "scaling loss" of the device-resident number.
    if hasattr(torch.cuda, "_sleep"):
        torch.cuda._sleep(GATE_CYCLES)
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
    job.barrier()
    t_w1 = time.perf_counter()
    dev_ms = job.max_over_ranks(e0.elapsed_time(e1))
    gpu_launches = relay.launches - launches0 + n_replays * launches_per_lap + (launches_rem if graph_rem is not None else 0)
    total_steps = step_no

    # parity spot-check of the timed run against libsodium (tx chain, rx chain, last wire, recovered plaintext)
    st = t_status.cpu().numpy()
    assert (st == 0).all(), f"{name}: non-OK frame status in the last timed step"
    last = total_steps - 1
    wire_slot = ring_view[(last + D) % R].cpu().numpy()
    for s in sample:
        ref = ls.State(bytes(keys[s]), bytes(hdrs[s]))
        rx_expect = None
        w = None
        for i in range(total_steps):
            if i == total_steps - D:
                rx_expect = ref.raw
            w = ref.seal_frame(bytes(ring_plain_host[s][i % R]))
        tx_state, rx_state = relay.export(int(slots[s]))
        assert tx_state == ref.raw, f"{name}: tx chain of stream {s} differs from libsodium after {total_steps} frames"
        assert rx_state == rx_expect, f"{name}: rx chain of stream {s} differs from libsodium"
        o = lay.P + int(lay.w_offs[s])
        assert bytes(wire_slot[o:o + len(w)]) == w, f"{name}: last wire frame of stream {s} differs from libsodium"
        ro = lay.P + lay.Wb + int(lay.p_offs[s])
        assert bytes(wire_slot[ro:ro + int(lay.lens[s])]) == bytes(ring_plain_host[s][(last - D) % R]), "recovered plaintext"

    arx_tops, imad_tops = relay.probe_int_peak(0)    # integer roofline denominators, measured on this GPU right now
    del ring, flush, ring_view
    torch.cuda.empty_cache()
    if own_relay and not keep_relay:
        relay.close()
        relay = None

    # whole-job figures: every rank's units over the slowest rank's time
    dev_s = dev_ms * 1e-3
    frames_all = job.sum_over_ranks(float(n)) * K
    plain_all = job.sum_over_ranks(float(lens.sum())) * K
    alg_bytes = algorithmic_bytes(lens) * 2              # this rank's launch: seal + open
    ops = int_ops(lens) * 2
    hbm_peak, hbm_src = hbm_peak_gbs()
    per_step_ms = dev_ms / K
    achieved_gbs = alg_bytes / (per_step_ms * 1e-3) / 1e9
    achieved_tops = ops / (per_step_ms * 1e-3) / 1e12
    res = {
        "workload": desc, "scaling": scaling, "steps": K, "ms_per_step": per_step_ms,
        "value": frames_all / dev_s, "unit": "frames/s", "gb_per_s": plain_all / dev_s / 1e9,
        "streams_per_gpu": n,
        "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved_gbs / hbm_peak, "traffic": None, "peak_source": hbm_src,
                     "kernel": kernel_label(tiles_full, nb_full),
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "note": "the path is integer-ALU bound (SURVEY 8(d)); see alu_roofline.  traffic: measured only "
                             "under ncu, see traffic_ncu"},
        "alu_roofline": {"bound": "int_alu", "achieved": achieved_tops, "peak": arx_tops, "unit": "Tint-op/s",
                         "frac": achieved_tops / arx_tops, "int_ops_per_launch": ops,
                         "peak_source": "symrelay_probe_int_peak (ChaCha ARX mix, measured now)",
                         "imad_wide_tops": imad_tops},
        "gpu_launches": int(gpu_launches), "launches_per_step": launches_per_step,
        "parity": "timed run spot-checked against libsodium: tx/rx chains, last wire frames, recovered plaintext: ok",
    }
    tn = traffic_ncu(name)
    if tn:
        res["roofline"]["traffic_ncu"] = tn
    extra = {"lens": lens, "lay": lay, "keys": keys, "hdrs": hdrs, "sample": sample, "window": (t_w0, t_w1),
             "relay": relay, "n_tiles_full": len(tiles_full), "plan_mode": plan_mode, "R": R}
    return res, extra


def kernel_label(tiles, n_bulk) -> str:
    """Which of the library's tile kernels a batch with these tiles launches (see launch_tiles in csrc/symrelay.cu)."""
    kinds = tiles["n_kind"]
    small, bulk = kinds[: len(tiles) - n_bulk], kinds[len(tiles) - n_bulk:]
    names = []
    if len(small) > 148 * 8:
        if ((small >> 17) & 1).any():
            names.append("ss_thread_kernel")
        if (((small >> 17) & 1) == 0).any():
            names.append("ss_packed_only_kernel")
    elif len(small):
        names.append("ss_packed_kernel")
    if len(bulk) and (((bulk >> 24) & 63) != 0).any():
        names.append("ss_strided_kernel")
    if len(bulk) and (((bulk >> 24) & 63) == 0).any():
        names.append("ss_bulk_kernel")
    return " + ".join(names) + (" (concurrent)" if len(names) > 1 else "")


def traffic_ncu(name: str):
    """DRAM bytes per launch of the dominant kernel from a separate `ncu --set full` capture (never measured in a
    timed run): {"bytes": ..., "source": file under profiles/}; None when no capture of this round is committed."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(name)
        return t if isinstance(t, dict) else None
    except Exception:
        return None


def pcie_peaks(job: Job):
    """(h2d GB/s, d2h GB/s): pinned 64 MiB cudaMemcpyAsync each way, best of 5, CUDA events."""
    torch = job.torch
    nbytes = 64 << 20
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d = torch.empty(nbytes, dtype=torch.uint8, device=job.dev)
    best = [1e30, 1e30]
    for rep in range(6):
        for k in (0, 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if k == 0:
                d.copy_(h, non_blocking=True)
            else:
                h.copy_(d, non_blocking=True)
            e1.record()
            torch.cuda.synchronize()
            if rep:
                best[k] = min(best[k], e0.elapsed_time(e1))
    return nbytes / (best[0] * 1e-3) / 1e9, nbytes / (best[1] * 1e-3) / 1e9


def pcie_concurrent(job: Job):
    """What the box's host-memory/PCIe path gives when EVERY rank copies in BOTH directions at once (copy engines, pinned
    64 MiB buffers, two streams per GPU, ranks released together): per-rank GB/s each way and the sum over ranks.  This is
    the ceiling the N-GPU end-to-end number lives under; the one-at-a-time peaks of pcie_peaks() are not."""
    torch = job.torch
    nbytes, reps = 64 << 20, 4
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(nbytes, dtype=torch.uint8, device=job.dev)
    d_out = torch.empty(nbytes, dtype=torch.uint8, device=job.dev)
    s_up, s_down = torch.cuda.Stream(device=job.dev), torch.cuda.Stream(device=job.dev)
    for timed in (False, True):
        job.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        with torch.cuda.stream(s_up):
            ev[0].record()
            for _ in range(reps):
                d_in.copy_(h_in, non_blocking=True)
            ev[1].record()
        with torch.cuda.stream(s_down):
            ev[2].record()
            for _ in range(reps):
                h_out.copy_(d_out, non_blocking=True)
            ev[3].record()
        torch.cuda.synchronize()
    up = nbytes * reps / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9
    down = nbytes * reps / (ev[2].elapsed_time(ev[3]) * 1e-3) / 1e9
    return {"h2d_gbs_this_rank": up, "d2h_gbs_this_rank": down,
            "h2d_gbs_all_ranks": job.sum_over_ranks(up), "d2h_gbs_all_ranks": job.sum_over_ranks(down),
            "what": "pinned 64 MiB cudaMemcpyAsync in both directions at once on every rank, ranks released together"}


def run_ours(args):
    import symmetry_b200 as sb
    from oracle import libsodium as ls

    job = Job()
    torch = job.torch
    name = args.workload
    K, W = args.steps, args.warmup
    D_E2E = int(os.environ.get("BENCH_E2E_DEPTH", "12"))
    lens, desc, scaling, first = workload(name, job.n_gpus, job.rank)
    n = len(lens)
    lay0 = Layout(lens)
    stage_bytes = max(64 << 20, (lay0.in_bytes + 4095) // 4096 * 4096)
    relay = sb.SymRelay(gpu_ids=(job.local,), max_streams=max(4 * n, 4096), pipeline_depth=D_E2E,
                        stage_bytes=stage_bytes, max_batch_frames=max(2 * n, 4096))
    clocks = ClockSampler(job.local)
    res, x = measure_device(job, name, K, W, args.no_graph, relay=relay)
    lay, keys, hdrs, sample = x["lay"], x["keys"], x["hdrs"], x["sample"]
    windows = [x["window"]]

    # ---------------- end-to-end through host arenas: `e2e` -------------------
    slots2 = relay.open_streams(keys, hdrs, keys, hdrs)
    R2 = max(2 * D_E2E + 2, 8)
    if args.no_e2e:
        R2 = 1
    hring = relay.alloc_pinned(R2 * lay.slot_bytes)
    hring[:] = np.random.default_rng(99 + job.rank).integers(0, 256, hring.size, dtype=np.uint8)
    hplain = {s: [bytes(hring[j * lay.slot_bytes + int(lay.p_offs[s]): j * lay.slot_bytes + int(lay.p_offs[s]) + int(lay.lens[s])])
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
    # The timed loop is driven from C (tools/e2e_driver.c, a plain caller of the public C ABI, as the
    # N-API addon would be): per batch it calls symrelay_batch_wait(ticket[i-depth]) then
    # symrelay_batch_submit(...) on the same ring -- the calls e2e_step() makes, minus the interpreter.
    # The W warm-up steps go through the same driver and the timed call follows them at once (ranks line up at a
    # host barrier, not a device synchronisation): the relay is measured as it runs under load, resident kernel
    # alive, not as it starts from idle.
    e2e_driver = "python"
    batch_lat = None
    drv = stat_all = None
    if K2 and not args.py_e2e:
        from symmetry_b200 import build as sbuild

        drv = C.CDLL(sbuild.build_driver())
        drv.e2e_ring_run.restype = C.c_double
        drv.e2e_ring_run.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                                     C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        stat_all = np.full(R2 * 2 * n, 0xFF, dtype=np.uint32)
        e2e_args = (relay.ctx, e_full.ctypes.data, 2 * n, hbase, lay.slot_bytes, lay.in_bytes, lay.P, lay.out_bytes, R2, D_E2E)
        e2e_drain()
        assert drv.e2e_ring_run(*e2e_args, i2, W, stat_all.ctypes.data, None) > 0
        i2 += W
        job.host_barrier()
    else:
        for _ in range(W if K2 else 0):
            e2e_step(i2)
            i2 += 1
        e2e_drain()
        job.barrier()
    t0 = time.perf_counter()
    if K2 and not args.py_e2e:
        secs = drv.e2e_ring_run(*e2e_args, i2, K2, stat_all.ctypes.data, None)
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
    e2e_sec = job.max_over_ranks(t1 - t0)
    windows.append((t0, t1))
    # The same loop with descriptors that differ from batch to batch (the same frames, listed in the opposite order on
    # odd steps): every batch takes the library's general path -- scan, plan, per-stream stamps, descriptor upload --
    # instead of the repeating-batch fast path the steady token-streaming shape gets.
    general = None
    if K2 and not args.py_e2e and name == "c2":
        drv.e2e_ring_run2.restype = C.c_double
        drv.e2e_ring_run2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                                      C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        e_alt = np.ascontiguousarray(e_full[::-1])
        Kg = min(max(K2, 200), 2000)
        # the status words of a reversed batch come back in its own order: all must be OK either way
        secs_g = drv.e2e_ring_run2(relay.ctx, e_full.ctypes.data, e_alt.ctypes.data, 2 * n, hbase, lay.slot_bytes, lay.in_bytes,
                                   lay.P, lay.out_bytes, R2, D_E2E, i2, Kg, stat_all.ctypes.data, None)
        assert secs_g > 0, f"e2e driver (general path) failed: {secs_g}"
        for i in range(i2, i2 + Kg):
            j = i % R2
            hstat[j][:] = stat_all[j * 2 * n:(j + 1) * 2 * n]
        i2 += Kg
        secs_g = job.max_over_ranks(secs_g)
        general = {"value": job.n_gpus * n * Kg / secs_g, "unit": "frames/s", "steps": Kg, "ms_per_step": secs_g / Kg * 1e3,
                   "what": "no two consecutive batches have the same descriptor array: the library's general host path every time"}
    clk = clocks.stop(windows)
    for j in sorted({i % R2 for i in range(max(0, i2 - K2), i2)} if K2 else set()):
        assert (hstat[j] == 0).all(), "non-OK status in the e2e run"
    # per-batch latency of the synchronous call over a 1,000-deep chain ("batch latency p50/p99")
    if K2 and not args.py_e2e and job.rank == 0 and name == "c2" and os.environ.get("BENCH_SYNC_LAT", "1") != "0":
        drv.e2e_sync_latency.restype = C.c_int
        drv.e2e_sync_latency.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                                         C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        nlat = 1000
        lat_ns = np.zeros(nlat, dtype=np.uint64)
        stat_all[:] = 0xFF
        rc = drv.e2e_sync_latency(relay.ctx, e_full.ctypes.data, 2 * n, hbase, lay.slot_bytes, lay.in_bytes, lay.P,
                                  lay.out_bytes, R2, D_E2E, i2, nlat, stat_all.ctypes.data, lat_ns.ctypes.data)
        assert rc == 0, f"sync latency driver failed: {rc}"
        i2 += nlat
        batch_lat = {"batches": nlat, "p50_us": float(np.percentile(lat_ns, 50)) / 1e3,
                     "p99_us": float(np.percentile(lat_ns, 99)) / 1e3, "mean_us": float(lat_ns.mean()) / 1e3,
                     "what": "symrelay_batch_host round trip (host arenas in, results visible to the host) of one "
                             "2,048-frame batch, 1,000 consecutive batches on the same chains, one in flight"}
    for s in (sample if K2 else []):
        ref = ls.State(bytes(keys[s]), bytes(hdrs[s]))
        for i in range(i2):
            w = ref.seal_frame(hplain[s][i % R2])
        assert relay.export(int(slots2[s]))[0] == ref.raw, "e2e tx chain differs from libsodium"
        o = ((i2 - 1 + D_E2E) % R2) * lay.slot_bytes + lay.P + int(lay.w_offs[s])
        assert bytes(hring[o:o + len(w)]) == w, "e2e last wire frame differs from libsodium"
    relay.close()
    relay = None
    h2d_peak, d2h_peak = pcie_peaks(job)
    pcie_all = pcie_concurrent(job)

    # ---------------- sub-lines: the other single-kernel claims, measured in this same run ----------
    subs = {}
    for sub in ([] if args.sub in ("", "none") else args.sub.split(",")):
        if sub == name or sub == "c5":   # (c5 is one GPU's figure: rank 0 measures it below)
            continue
        r, _ = measure_device(job, sub, SUB_STEPS.get(sub, 32), max(3, min(W, 8)), args.no_graph)
        r["config"] = make_config(sub, job.n_gpus, r["steps"], args.no_graph, plan_mode_for(sub))
        subs[sub] = r

    if job.rank != 0:
        job.close()
        return

    # ---------------- CPU baseline on this box's host cores (rank 0, N=1) -----
    cpu = None
    if job.world == 1 and not args.no_cpu:
        threads = host_threads()
        probe, kind = cpu_run(lens, 4, 2, threads)
        per_step_cpu_s = probe / 4 * threads           # CPU-seconds of work per step
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
    if job.world == 1 and not args.no_latency:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import latency_bench

            latency = latency_bench.run(streams=65536, rate=2e6, nbytes=200, seconds=1.0, flush_us=100)
        except Exception as e:  # the headline numbers do not depend on it
            latency = {"error": repr(e)}

    # ---------------- the step after the path: what sending the sealed frames costs the host ----------
    udp = None
    if job.world == 1 and not args.no_latency:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import udp_bench

            udp = udp_bench.run(frames=2 * n if name == "c2" else 2048, nbytes=84, batches=60)
        except Exception as e:
            udp = {"error": repr(e)}

    # ---------------- BASELINE config 5 (extract + seal fused), one GPU's figure, rank 0 ----------
    if "c5" in ([] if args.sub in ("", "none") else args.sub.split(",")) and name != "c5":
        try:
            subs["c5"] = c5_line(SUB_STEPS["c5"], 5)
        except Exception as e:
            subs["c5"] = {"error": repr(e)}

    h2d = lay.in_bytes + 2 * n * 16 + x["n_tiles_full"] * 8
    d2h = lay.out_bytes + 2 * n * 4
    e2e_ms = (e2e_sec / K2 * 1e3) if K2 else None
    e2e = {"value": (job.n_gpus * n * K2 / e2e_sec) if K2 else None, "unit": "frames/s",
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": K2, "ms_per_step": e2e_ms,
           "path": "symrelay_batch_submit/wait on pinned host arenas, pipeline depth " + str(D_E2E) +
                   ", host clock; loop driver: " + e2e_driver}
    if K2:
        # PCIe is full duplex: the bound for a step is the slower direction at that direction's measured peak
        t_bound = max(h2d / (h2d_peak * 1e9), d2h / (d2h_peak * 1e9))
        e2e["roofline"] = {"bound": "pcie", "achieved": (h2d + d2h) / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s",
                           "peak": (h2d + d2h) / t_bound / 1e9, "frac": t_bound / (e2e_ms * 1e-3),
                           "h2d_peak_gbs": h2d_peak, "d2h_peak_gbs": d2h_peak,
                           "peak_source": "pinned 64 MiB cudaMemcpyAsync each way, measured in this run; peak = both "
                                          "directions' bytes over the slower direction's time at its peak (full duplex)"}
    if K2:
        # the same traffic against what the box gives when all N ranks copy both ways at once (the N-GPU ceiling)
        t_all = max(h2d / (pcie_all["h2d_gbs_this_rank"] * 1e9), d2h / (pcie_all["d2h_gbs_this_rank"] * 1e9))
        e2e["roofline"]["all_ranks_copying"] = dict(pcie_all, frac=t_all / (e2e_ms * 1e-3))
    if batch_lat:
        e2e["batch_latency"] = batch_lat
    if general:
        e2e["general_path"] = general
    line = {
        "metric": METRIC, "value": res["value"], "unit": "frames/s", "n_gpus": job.n_gpus, "steps": K, "warmup": W,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": make_config(name, job.n_gpus, K, args.no_graph, x["plan_mode"]),
        "gb_per_s": res["gb_per_s"], "opened_frames_per_sec": res["value"],
        "roofline": res["roofline"], "alu_roofline": res["alu_roofline"], "e2e": e2e,
        "gpu_launches": res["gpu_launches"], "launches_per_step": res["launches_per_step"],
        "clocks": clk, "parity": res["parity"],
    }
    for k, v in subs.items():
        line[k] = v
    if cpu:
        line["cpu_baseline"] = cpu
    if latency:
        line["latency"] = latency
    if udp:
        line["udp_send"] = udp
    print(json.dumps(line), flush=True)
    job.close()


def hbm_peak_gbs():
    """(GB/s, where it came from): the driver-written measured copy bandwidth, else the profiling guide's fallback."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured)"
    except Exception:
        return 6650.0, "fallback 6650"


# ------------------------------------------------------------------------- c5 --
def run_c5(args):
    print(json.dumps(c5_line(args.steps, args.warmup)), flush=True)


def c5_line(K: int, W: int) -> dict:
    """configs[4]: fused relay -- per batch one synthetic chat-completion chunk per stream:
    prefix test + JSON validation + delta extraction + seal (device-resident).  One GPU (the current device's figure)."""
    import torch

    import batch_util as bu
    import symmetry_b200 as sb
    from oracle import extract as ex
    from oracle import libsodium as ls
    from symmetry_b200 import synth

    torch.cuda.set_device(0)
    dev = torch.device("cuda:0")
    n, n_kinds = int(os.environ.get("BENCH_C5_STREAMS", "16384")), 16   # configs[4] says 16k; the env knob is for the batch-size note in DESIGN.md
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
                     "kernel": ("ss_extract_seal_kernel (extractor and seal in one kernel)" if per_lap == R else
                                "ss_extract_kernel + ss_packed_kernel (whole step)"),
                     "algorithmic_bytes_per_step": alg_bytes,
                     "note": "2*mlen+92 bytes per frame (SURVEY 8(d)); the step is bound by the extractor's serial "
                             "per-chunk instruction streams, see alu_roofline and DESIGN.md 4.6"},
        "alu_roofline": {"bound": "int_alu", "achieved": ops / (per_step * 1e-3) / 1e12, "peak": arx, "unit": "Tint-op/s",
                         "frac": ops / (per_step * 1e-3) / 1e12 / arx},
        "gpu_launches": reps * per_lap, "launches_per_step": per_lap / R, "clocks": clk,
        "parity": "deltas == oracle/extract.py on sampled frames of every ring slot; chains == libsodium",
    }
    relay.close()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="default: 20000 for c2, 200 for c3, 64 for c4, 360 for c5")
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--sub", default=None, help="sub-lines carried by the default (c2) line: 'c3,c4,c5' (default) or 'none'")
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = same as --steps")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-latency", action="store_true", help="skip the p50 added-latency leg (64k streams, Poisson)")
    ap.add_argument("--py-e2e", action="store_true", help="drive the e2e loop from Python instead of tools/e2e_driver.c")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (huge workloads)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step directly (no CUDA graph)")
    ap.add_argument("--single-process", action="store_true",
                    help="with --gpus N (no torchrun): one process, one thread, one context over N GPUs; end-to-end leg only")
    args = ap.parse_args()
    if args.steps is None:   # a step of c3/c4/c5 is 10-400x a step of c2
        args.steps = DEFAULT_STEPS.get(args.workload, 20000)
    if args.sub is None:
        # (c3 before c4: measured right behind c4 in one process, c3's four concurrent kernels take 63 us instead of the
        # 58-59 of its stand-alone run -- neither 0.3 s of idling nor an emptied allocator cache changes that; c4 and c5
        # read the same in any order)
        args.sub = "c3,c4,c5" if args.workload == "c2" else "none"
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    elif args.single_process:
        import bench_extra

        bench_extra.run_single_process(args)
    elif args.workload == "c5":
        run_c5(args)
    elif args.workload == "c1":
        import bench_extra

        bench_extra.run_c1(args)
    elif args.workload == "c4sweep":
        import bench_extra

        bench_extra.run_c4_sweep(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
