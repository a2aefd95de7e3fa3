"""The C-ABI library loads on a machine without a GPU and exports every symbol
include/symrelay.h declares; computing entry points refuse to run without a device."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "symrelay.h")) as f:
        text = f.read()
    return sorted(set(re.findall(r"SYMRELAY_API[^;(]*?\b(symrelay_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from symmetry_b200 import _lib, build

    build.build()
    L = _lib.load()
    decl = declared_symbols()
    assert len(decl) >= 18
    for name in decl:
        assert hasattr(L, name), f"{name} declared in include/symrelay.h but not exported"
    assert sorted(_lib.EXPORTS) == decl, "symmetry_b200/_lib.py EXPORTS out of sync with the header"
    assert L.symrelay_abi_version() == 1
    assert L.symrelay_strerror(-22) == b"invalid argument"


def test_planner_and_desc_check_are_host_only():
    import numpy as np

    import batch_util as bu
    from symmetry_b200 import _lib

    L = _lib.load()
    lens = [64] * 20 + [5000]
    p, _ = bu.plain_layout(lens)
    w, _ = bu.wire_layout(lens)
    d = bu.make_descs(np.arange(len(lens)), lens, p, w)
    import ctypes as C

    nb = C.c_uint32(0)
    assert L.symrelay_plan(d.ctypes.data, len(d), None, 0, C.byref(nb)) == 3 and nb.value == 1
    assert L.symrelay_desc_check(d[:1].ctypes.data) == 0
    d["out_off"][0] = 16
    assert L.symrelay_desc_check(d[:1].ctypes.data) == -22


def test_planner_marks_wide_tiles_only_when_big_frames_are_few():
    """Frames of 128 blocks and more (> 8,000 payload bytes) get a whole 128-thread block (tile bit 18) when the batch
    holds at most 2,048 of them, fewer than 262,144 ChaCha20 blocks in all and at most 32,768 frames (beyond that it is
    bound by total work, not by its longest frame); a batch of many big frames keeps one warp per frame.  Both planners agree."""
    import ctypes as C

    import numpy as np

    import batch_util as bu
    from symmetry_b200 import _lib

    L = _lib.load()

    def plan(lens, sorted_):
        p, _ = bu.plain_layout(lens)
        w, _ = bu.wire_layout(lens)
        d = bu.make_descs(np.arange(len(lens)), lens, p, w)
        tiles = np.zeros((len(lens), 2), dtype=np.uint32)
        nb = C.c_uint32(0)
        if sorted_:
            sd = np.zeros_like(d)
            perm = np.zeros(len(lens), dtype=np.uint32)
            nt = L.symrelay_plan_sorted(d.ctypes.data, len(d), tiles.ctypes.data, len(tiles), C.byref(nb), sd.ctypes.data,
                                        perm.ctypes.data)
            d = sd
        else:
            nt = L.symrelay_plan(d.ctypes.data, len(d), tiles.ctypes.data, len(tiles), C.byref(nb))
        assert nt > 0
        return tiles[:nt], d

    for sorted_ in (False, True):
        lens = [64] * 40 + [700] * 9 + [8000, 8001, 16384, 3000, 70000]
        tiles, d = plan(lens, sorted_)
        wide = [int(d["len_flags"][t[0]]) & 0xFFFFFF for t in tiles if t[1] & (1 << 18)]
        assert sorted(wide) == [8001, 16384, 70000], (sorted_, wide)
        assert all((t[1] & 0xFFFF) == 1 and (t[1] & (1 << 16)) and not (t[1] >> 24) for t in tiles if t[1] & (1 << 18))
        tiles, d = plan([64] * 10 + [8001] * 2049, sorted_)
        assert not any(t[1] & (1 << 18) for t in tiles), sorted_
        tiles, d = plan([64] * 10 + [8001] * 2047, sorted_)           # 2,047 x 128 + 10 x 3 = 262,046 blocks
        assert sum(1 for t in tiles if t[1] & (1 << 18)) == 2047, sorted_
        tiles, d = plan([64] * 100 + [8001] * 2047, sorted_)          # 262,316 blocks: bound by total work
        assert not any(t[1] & (1 << 18) for t in tiles), sorted_
        tiles, d = plan([64] * 32767 + [9000], sorted_)
        assert sum(1 for t in tiles if t[1] & (1 << 18)) == 1, sorted_
        tiles, d = plan([64] * 32768 + [9000], sorted_)               # 32,769 frames
        assert not any(t[1] & (1 << 18) for t in tiles), sorted_


def test_no_cpu_fallback_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import symmetry_b200 as sb

    with pytest.raises(sb.RelayError, match="no CPU path"):
        sb.SymRelay()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "symmetry_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn
                assert "ssref_" not in src and "libss_emu" not in src and "nacl" not in src, fn


def test_avx2_descriptor_scan_matches_the_scalar_loop(tmp_path):
    """The host planner's first pass has an AVX2 form (symmetry_b200/csrc/plan_scan_avx2.cpp) that the library
    picks at run time; tests/emu/scan_check.cpp runs it against the scalar loop of plan.h on 20,000 random
    batches with misaligned, oversized, out-of-bounds and overlapping descriptors mixed in."""
    import subprocess

    exe = str(tmp_path / "scan_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "emu", "scan_check.cpp"),
                    os.path.join(ROOT, "symmetry_b200", "csrc", "plan_scan_avx2.cpp")], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    if res.returncode == 77:
        pytest.skip("this CPU has no AVX2")
    assert res.returncode == 0, res.stdout + res.stderr


def test_bucketed_planner_covers_every_descriptor_once(tmp_path):
    """Large host batches are planned bucketed (descriptors sorted by block count, closed-form tiles, a
    permutation for the statuses): tests/emu/bucket_check.cpp validates the plan of 131,072 power-law frames."""
    import subprocess

    exe = str(tmp_path / "bucket_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "emu", "bucket_check.cpp"),
                    os.path.join(ROOT, "symmetry_b200", "csrc", "plan_scan_avx2.cpp")], check=True)
    res = subprocess.run([exe], capture_output=True, text=True)
    assert res.returncode == 0 and "valid=1" in res.stdout, res.stdout + res.stderr
    # batches this large are planned on four threads (the caller and three workers); without them -- kill switch, a busy
    # pool, a forked child -- the caller runs the four shares itself and must arrive at the very same plan
    env = dict(os.environ, SYMRELAY_PLAN_THREADS="0")
    res1 = subprocess.run([exe], capture_output=True, text=True, env=env)
    assert res1.returncode == 0 and "valid=1" in res1.stdout, res1.stdout + res1.stderr
    plan = re.search(r"plan=([0-9a-f]{16})", res.stdout).group(1)
    assert plan == re.search(r"plan=([0-9a-f]{16})", res1.stdout).group(1), (res.stdout, res1.stdout)


def test_planner_threads_survive_fork_and_exit():
    """The library's planner starts three worker threads with its first large batch.  The process must still exit
    cleanly, and a forked child (which has no such threads) must plan by itself instead of waiting for them."""
    import subprocess
    import sys

    code = r'''
import os, sys, ctypes as C
import numpy as np
sys.path.insert(0, os.path.join(%r, "tests"))
sys.path.insert(0, %r)
import batch_util as bu
from symmetry_b200 import _lib
L = _lib.load()
n = 70000
lens = [64 + (i * 37) %% 900 for i in range(n)]
p, _ = bu.plain_layout(lens); w, _ = bu.wire_layout(lens)
d = bu.make_descs(np.arange(n), lens, p, w)
def plan():
    tiles = np.zeros((n, 2), dtype=np.uint32); sd = np.zeros_like(d); perm = np.zeros(n, dtype=np.uint32); nb = C.c_uint32(0)
    nt = L.symrelay_plan_sorted(d.ctypes.data, n, tiles.ctypes.data, n, C.byref(nb), sd.ctypes.data, perm.ctypes.data)
    assert nt > 0 and sorted(perm.tolist()) == list(range(n))
    return tiles[:nt].tobytes() + sd.tobytes() + perm.tobytes()
a = plan()
pid = os.fork()
if pid == 0:
    sys.exit(0 if plan() == a else 3)      # an ordinary exit: the child runs the library's exit handlers too
_, st = os.waitpid(pid, 0)
assert os.WEXITSTATUS(st) == 0, st
assert plan() == a
print("ok")
''' % (ROOT, ROOT)
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0 and "ok" in res.stdout, res.stdout + res.stderr


def test_napi_shim_is_valid_c_against_the_header():
    """addon/symrelay_napi.c cannot be built here (no node_api.h, no Node), but it must at least be valid C that agrees
    with include/symrelay.h: every ABI call it makes is type-checked against the header's prototypes."""
    import subprocess

    res = subprocess.run(["gcc", "-fsyntax-only", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                          os.path.join(ROOT, "addon", "symrelay_napi.c")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    src = open(os.path.join(ROOT, "addon", "symrelay_napi.c")).read()
    for fn in ("symrelay_submit_seal", "symrelay_rx_feed", "symrelay_poll", "symrelay_release", "symrelay_event_fd",
               "symrelay_rx_expect_header", "symrelay_stream_set_keepalive", "symrelay_connect_batch"):
        assert fn + "(" in src, f"the shim does not bind {fn}"
