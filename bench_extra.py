"""bench_extra.py -- the two workloads of SURVEY.md 8(d) that are not batches of one shape:

  --workload c1       BASELINE configs[0] restated: ONE connection, one llama3:8b-style completion (preamble,
                      256 delta chunks, stop chunk, `data: [DONE]`, trailer: /root/reference/src/provider.ts:234-262),
                      its frames strictly in sequence through symrelay_submit_seal(F_EXTRACT) on one host core,
                      with libsodium push+pull of the same frames on one core beside it.  This is the worst case
                      for a batching relay (a batch of one) and the per-core unit of the CPU baseline.
  --workload c4sweep  configs[3]'s bandwidth sweep: frame size 64 B -> 64 KiB in doublings at 1 GiB of plaintext
                      per round, device-resident, each point against the ALU and HBM rooflines.

Each prints ONE JSON line in bench.py's format."""
from __future__ import annotations

import ctypes as C
import json
import os
import time

import numpy as np

import bench as B


def c1_frames(key: str = "c1-emitter-key", n_chunks: int = 256, seed: int = 1):
    """(frames, extract flags) of one inference as the relay loop writes them."""
    from symmetry_b200 import synth

    body = synth.chat_chunks(n_chunks, seed=seed, stream_id=0)
    pre = json.dumps({"symmetryEmitterKey": key}, separators=(",", ":")).encode()          # provider.ts:234-238
    post = json.dumps({"key": "inferenceEnded", "data": key}, separators=(",", ":")).encode()  # :260-262
    return [pre] + body + [post], [0] + [1] * len(body) + [0]


def run_c1(args):
    import torch

    import symmetry_b200 as sb
    from oracle import extract as ex
    from oracle import libsodium as ls
    from oracle import secretstream as ss_port
    from symmetry_b200 import build as sbuild
    from symmetry_b200 import synth

    torch.cuda.set_device(0)
    frames, flags = c1_frames()
    n = len(frames)
    blob = np.frombuffer(b"".join(frames), dtype=np.uint8).copy()
    off = np.zeros(n + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(f) for f in frames])
    fl = np.asarray(flags, dtype=np.uint8)
    keys, hdrs = synth.stream_keys(1)
    key, hdr = bytes(keys[0]), bytes(hdrs[0])
    K, W = args.steps, max(args.warmup, 3)
    drv = C.CDLL(sbuild.build_driver())
    drv.c1_run.restype = C.c_double
    drv.c1_run.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                           C.POINTER(C.c_uint64)]
    clocks = B.ClockSampler(0)
    with sb.SymRelay(gpu_ids=(0,), max_streams=64, sched_batch_frames=64, sched_arena_bytes=1 << 16) as relay:
        slot = relay.open_stream(key, hdr, None, None, provider=sb.PROVIDER_OLLAMA)
        # parity pass through the Python binding: every wire frame vs libsodium, every delta vs the extractor oracle
        ref = ls.State(key, hdr)
        completion = b""
        for f, x in zip(frames, flags):
            t = relay.submit_seal(slot, f, extract=bool(x))
            relay.flush()
            (c,) = relay.poll(wait=True)
            assert c["ticket"] == t and c["status"] == 0 and c["data"] == ref.seal_frame(f)
            if x:
                code, want = ex.extract("ollama", f)
                assert code == ex.OK and c["delta"] == want
                completion += c["delta"]
        db = C.c_uint64(0)
        assert drv.c1_run(relay.ctx, slot, blob.ctypes.data, off.ctypes.data, fl.ctypes.data, n, W, C.byref(db)) > 0
        l0 = relay.launches
        tw0 = time.perf_counter()
        secs = drv.c1_run(relay.ctx, slot, blob.ctypes.data, off.ctypes.data, fl.ctypes.data, n, K, C.byref(db))
        tw1 = time.perf_counter()
        assert secs > 0, f"c1 driver failed: {secs} ({relay.lib.symrelay_last_error(relay.ctx)})"
        launches = relay.launches - l0
        assert db.value == K * len(completion), "delta bytes of the timed run differ from the parity pass"
        for r in range(W + K):
            for f in frames:
                ref.seal_frame(f)
        assert relay.export(slot)[0] == ref.raw, "c1 tx chain differs from libsodium after the timed run"
        arx, imad = relay.probe_int_peak(0)
    clk = clocks.stop([(tw0, tw1)])
    # the same connection on one host core: libsodium push + pull per frame (secret-stream `_write` / `_incoming`)
    lib = ss_port.lib()
    tx = np.frombuffer(ls.State(key, hdr).raw, dtype=np.uint8).copy()
    rx = tx.copy()
    scratch = np.zeros(int(max(len(f) for f in frames)) + 64, dtype=np.uint8)
    push = ls.fn_ptr("crypto_secretstream_xchacha20poly1305_push")
    pull = ls.fn_ptr("crypto_secretstream_xchacha20poly1305_pull")
    cpu_reps = 2000
    lib.cpub_chain_run(push, pull, tx.ctypes.data, rx.ctypes.data, blob.ctypes.data, off.ctypes.data, n, 50, scratch.ctypes.data)
    csec = lib.cpub_chain_run(push, pull, tx.ctypes.data, rx.ctypes.data, blob.ctypes.data, off.ctypes.data, n, cpu_reps,
                              scratch.ctypes.data)
    assert csec > 0 and (tx == rx).all()
    t0 = time.perf_counter()
    for f, x in zip(frames, flags):
        if x:
            ex.extract("ollama", f)
    py_extract_us = (time.perf_counter() - t0) / sum(flags) * 1e6
    lens = np.asarray([len(f) for f in frames])
    per_frame_us = secs / (K * n) * 1e6
    hbm_peak, hbm_src = B.hbm_peak_gbs()
    alg = B.algorithmic_bytes(lens) / n
    ops = (B.int_ops(lens) + 4 * int(lens.sum())) / n
    val = K * n / secs
    line = {
        "metric": B.METRIC, "value": val, "unit": "frames/s", "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": secs / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": "configs[0] restated (no Node, no hyperdht here): 1 provider connection, one llama3:8b-style "
                               "completion per step = preamble + 256 delta chunks + stop chunk + data: [DONE] + trailer "
                               f"({n} frames, {int(lens.mean())} B mean), frames strictly in sequence: symrelay_submit_seal"
                               "(F_EXTRACT) -> flush -> completion, one host core, apiProvider ollama",
                   "frames_per_step": n, "l2": "not applicable: one frame in flight, the loop is round-trip bound"},
        "ns_per_frame": per_frame_us * 1e3, "gb_per_s": float(lens.sum()) * K / secs / 1e9,
        "roofline": {"bound": "hbm", "achieved": alg / (per_frame_us * 1e-6) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": alg / (per_frame_us * 1e-6) / 1e9 / hbm_peak, "traffic": None, "peak_source": hbm_src,
                     "kernel": "ss_extract_kernel + ss_packed_kernel, one frame per launch",
                     "note": "a batch of one frame: the time is launch + PCIe round trip, not the kernel"},
        "alu_roofline": {"bound": "int_alu", "achieved": ops / (per_frame_us * 1e-6) / 1e12, "peak": arx,
                         "unit": "Tint-op/s", "frac": ops / (per_frame_us * 1e-6) / 1e12 / arx},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": int(lens.sum()) + 24 * n,
                "d2h_bytes_per_step": int(lens.sum()) + 24 * n + int(lens.sum()), "steps": K,
                "path": "the per-frame ABI on host buffers IS the measured path (tools/e2e_driver.c c1_run)"},
        "cpu_baseline": {"value": cpu_reps * n / csec, "unit": "frames/s", "cores": 1, "kind": "reference",
                         "ns_per_frame": csec / (cpu_reps * n) * 1e9,
                         "sample": f"{cpu_reps} passes over the same {n} frames, libsodium 1.0.20 push + pull per frame on one "
                                   f"core via oracle/cpu_bench.c cpub_chain_run; {B.cpu_model()}; the reference's JS extractor "
                                   f"is not runnable here (its Python restatement takes {py_extract_us:.1f} us per chunk, "
                                   "information only)"},
        "gpu_launches": int(launches), "launches_per_step": launches / K, "clocks": clk,
        "parity": "every wire frame == libsodium and every delta == oracle/extract.py in a pass before timing; tx chain "
                  "== libsodium after it",
    }
    print(json.dumps(line), flush=True)


def run_c4_sweep(args):
    job = B.Job()
    assert job.world == 1, "the size sweep is a single-GPU roofline probe"
    round_bytes = int(os.environ.get("BENCH_SWEEP_BYTES", str(1 << 30)))
    sizes = [64 << k for k in range(11)]
    if os.environ.get("BENCH_SWEEP_SIZES"):
        sizes = [int(x) for x in os.environ["BENCH_SWEEP_SIZES"].split(",")]
    K, W = args.steps, max(3, min(args.warmup, 4))
    clocks = B.ClockSampler(job.local)
    points, windows = [], []
    rng = np.random.default_rng(0x5EED0001)
    for size in sizes:
        n = max(1, round_bytes // size)
        lens = np.full(n, size, dtype=np.int64)
        keys = rng.integers(0, 256, (n, 32), dtype=np.uint8)      # 16.7 M BLAKE2b derivations in Python would take minutes
        hdrs = rng.integers(0, 256, (n, 24), dtype=np.uint8)
        desc = f"configs[3] sweep: {n:,} streams x {size} B frames ({round_bytes >> 20} MiB of plaintext per round), seal+open"
        r, x = B.measure_device(job, "c4sweep", K, W, args.no_graph, custom=(lens, desc, "weak", keys, hdrs))
        windows.append(x["window"])
        points.append({"frame_bytes": size, "streams": n, "ms_per_step": r["ms_per_step"], "frames_per_s": r["value"],
                       "plaintext_gb_per_s": r["gb_per_s"], "alu_frac": r["alu_roofline"]["frac"],
                       "hbm_frac": r["roofline"]["frac"], "kernel": r["roofline"]["kernel"],
                       "launches_per_step": r["launches_per_step"]})
        last = r
    clk = clocks.stop(windows)
    line = {
        "metric": B.METRIC, "value": last["value"], "unit": "frames/s", "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": last["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": f"configs[3]: seal/open bandwidth sweep, frame size {sizes[0]} B -> {sizes[-1]} B, "
                               f"{round_bytes >> 20} MiB of plaintext per round, device-resident; value = the last size",
                   "keys": "random per stream (numpy default_rng), sub-keys derived on the GPU",
                   "l2": "every round's input exceeds L2 several times over; ring of >= 4 distinct batch slots"},
        "gb_per_s": last["gb_per_s"], "roofline": last["roofline"], "alu_roofline": last["alu_roofline"],
        "sweep": points, "gpu_launches": sum(int(p["launches_per_step"]) * K for p in points), "clocks": clk,
        "parity": "each point spot-checked against libsodium (tx/rx chains, last wire frames, recovered plaintext)",
    }
    print(json.dumps(line), flush=True)
    job.close()


def run_single_process(args):
    """--single-process --gpus N: ONE process, ONE thread, one context over N GPUs -- the reference's own shape (one
    Node process, src/provider.ts:37-58).  Streams are sharded slot % N inside the context; the
    thread visits the GPUs round-robin through symrelay_batch_submit/_wait (tools/e2e_driver.c).  Prints the e2e
    figure as `value` too: there is no device-resident leg in this mode."""
    import torch

    import symmetry_b200 as sb
    from oracle import libsodium as ls
    from symmetry_b200 import build as sbuild
    from symmetry_b200 import synth

    N = args.gpus
    assert torch.cuda.device_count() >= N, f"needs {N} GPUs"
    name = args.workload if args.workload in ("c2", "c3", "c4") else "c2"
    K, W = args.steps, max(args.warmup, 3)
    D = int(os.environ.get("BENCH_E2E_DEPTH", "12"))
    per = []
    total = 0
    for g in range(N):
        lens, desc_txt, scaling, first = B.workload(name, N, g)
        per.append({"lens": lens, "lay": B.Layout(lens), "first": first})
        total += len(lens)
    stage_bytes = max(64 << 20, max((p["lay"].in_bytes + 4095) // 4096 * 4096 for p in per))
    relay = sb.SymRelay(gpu_ids=tuple(range(N)), max_streams=max(4 * total, 4096), pipeline_depth=D, stage_bytes=stage_bytes,
                        max_batch_frames=max(2 * max(len(p["lens"]) for p in per), 4096))
    # open all streams, then give GPU g the slots that landed on it (slot % N == g)
    keys, hdrs = synth.stream_keys(total)
    slots = relay.open_streams(keys, hdrs, keys, hdrs)
    by_gpu = [[i for i in range(total) if int(slots[i]) % N == g] for g in range(N)]
    R2 = max(2 * D + 2, 8)
    descs, rings, stats, samples = [], [], [], []
    for g in range(N):
        idx = by_gpu[g][: len(per[g]["lens"])]
        assert len(idx) == len(per[g]["lens"]), "slots did not spread evenly over the GPUs"
        lay = per[g]["lay"]
        descs.append(np.ascontiguousarray(lay.descs(slots[idx], True)))
        ring = relay.alloc_pinned(R2 * lay.slot_bytes)
        ring[:] = np.random.default_rng(500 + g).integers(0, 256, ring.size, dtype=np.uint8)
        rings.append(ring)
        stats.append(np.full(R2 * 2 * len(idx), 0xFF, dtype=np.uint32))
        samples.append([(0, idx[0]), (len(idx) - 1, idx[-1])])
        per[g]["idx"] = idx
        per[g]["plain"] = {j: [bytes(ring[r * lay.slot_bytes + int(lay.p_offs[j]): r * lay.slot_bytes + int(lay.p_offs[j]) + int(lay.lens[j])])
                               for r in range(R2)] for j, _ in samples[g]}
    drv = C.CDLL(sbuild.build_driver())
    drv.e2e_ring_run_multi.restype = C.c_double
    PP, U32, U64 = C.c_void_p * N, C.c_uint32 * N, C.c_uint64 * N
    drv.e2e_ring_run_multi.argtypes = [C.c_void_p, C.c_int, PP, U32, PP, U64, U64, U64, U64, C.c_uint32, C.c_uint32, C.c_uint64,
                                       C.c_uint32, PP]

    def go(first_step, steps):
        return drv.e2e_ring_run_multi(
            relay.ctx, N, PP(*[d.ctypes.data for d in descs]), U32(*[len(d) for d in descs]), PP(*[r.ctypes.data for r in rings]),
            U64(*[p["lay"].slot_bytes for p in per]), U64(*[p["lay"].in_bytes for p in per]), U64(*[p["lay"].P for p in per]),
            U64(*[p["lay"].out_bytes for p in per]), R2, D, first_step, steps, PP(*[s.ctypes.data for s in stats]))

    # priming: the first D steps seal only (no inbound frames exist yet), as in bench.py
    seal_descs = [np.ascontiguousarray(p["lay"].descs(slots[p["idx"]], False)) for p in per]
    full = descs
    descs = seal_descs
    for s in stats:
        s[:] = 0xFF
    step = 0
    assert go(step, D) > 0
    step += D
    descs = full
    assert go(step, W) > 0
    step += W
    clocks = B.ClockSampler(0)
    tw0 = time.perf_counter()
    secs = go(step, K)
    tw1 = time.perf_counter()
    assert secs > 0, f"driver failed: {secs} {relay.lib.symrelay_last_error(relay.ctx)}"
    step += K
    clk = clocks.stop([(tw0, tw1)])
    for g in range(N):
        n_g = len(per[g]["idx"])
        st = stats[g].reshape(R2, 2 * n_g)
        for i in range(step - K, step):
            assert (st[i % R2] == 0).all(), f"non-OK status on GPU {g}"
        lay = per[g]["lay"]
        for j, gi in samples[g]:
            ref = ls.State(bytes(keys[gi]), bytes(hdrs[gi]))
            for i in range(step):
                w = ref.seal_frame(per[g]["plain"][j][i % R2])
            assert relay.export(int(slots[gi]))[0] == ref.raw, f"tx chain of a stream on GPU {g} differs from libsodium"
            o = ((step - 1 + D) % R2) * lay.slot_bytes + lay.P + int(lay.w_offs[j])
            assert bytes(rings[g][o:o + len(w)]) == w
    frames = sum(len(p["lens"]) for p in per)
    val = frames * K / secs
    lay0 = per[0]["lay"]
    line = {
        "metric": B.METRIC, "value": val, "unit": "frames/s", "n_gpus": N, "steps": K, "warmup": W, "ms_per_step": secs / K * 1e3,
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": dict(B.make_config(name, N, K, True, B.plan_mode_for(name)),
                       mode="single process, single thread, one context over all GPUs (tools/e2e_driver.c e2e_ring_run_multi)"),
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": int(sum(p["lay"].in_bytes + 2 * len(p["lens"]) * 16 for p in per)),
                "d2h_bytes_per_step": int(sum(p["lay"].out_bytes + 2 * len(p["lens"]) * 4 for p in per)), "steps": K,
                "ms_per_step": secs / K * 1e3,
                "path": f"symrelay_batch_submit/wait on pinned host arenas, {N} GPUs visited round-robin by one thread, depth {D}"},
        "note": "value repeats e2e: this mode has no device-resident leg (run without --single-process for it)",
        "gpu_launches": relay.launches, "clocks": clk,
        "parity": "statuses all OK; sampled tx chains and last wire frames == libsodium on every GPU",
    }
    print(json.dumps(line), flush=True)
    relay.close()
