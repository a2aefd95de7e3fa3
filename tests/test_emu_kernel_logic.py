"""CPU unit test of the PRODUCT's device code (symmetry_b200/csrc/ss_core.cuh) and
planner (plan.h), compiled with g++ under tests/emu/'s 32-thread warp emulator.
It checks lane mapping, window building, the lane-parallel Poly1305 and the
nonce chain against libsodium on machines without a GPU.  The GPU parity tests
(-m gpu) are the ones that count for the product; this one finds logic bugs early."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

import batch_util as bu
import parity_cases as pc

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "emu")


class EmuBackend:
    """State table in a numpy array (device layout), batches run by the warp emulator."""

    def __init__(self, lib, max_slots=4096):
        self.lib = lib
        self.table = bu.aligned_zeros(max_slots * 2 * 48).reshape(max_slots, 2, 48)
        self.next = 0

    def open_streams(self, tx_states, rx_states):
        slots = []
        for tx, rx in zip(tx_states, rx_states):
            s = self.next
            self.next += 1
            self.table[s, 0, :44] = np.frombuffer(tx[:44], dtype=np.uint8)
            self.table[s, 1, :44] = np.frombuffer(rx[:44], dtype=np.uint8)
            slots.append(s)
        return slots

    def plan(self, descs):
        nb = C.c_uint32(0)
        need = self.lib.emu_plan(descs.ctypes.data, len(descs), None, 0, C.byref(nb))
        assert need >= 0
        tiles = np.zeros(max(1, need), dtype=bu.TILE_DTYPE)
        assert self.lib.emu_plan(descs.ctypes.data, len(descs), tiles.ctypes.data, need, C.byref(nb)) == need
        return tiles[:need], nb.value

    def run(self, descs, arena_in, arena_out):
        tiles, n_bulk = self.plan(descs)
        status = np.full(len(descs), 0xFF, dtype=np.uint32)
        self.lib.emu_run_batch(descs.ctypes.data, tiles.ctypes.data, len(tiles), n_bulk, arena_in.ctypes.data,
                               arena_out.ctypes.data, self.table.ctypes.data, status.ctypes.data)
        return status

    def export(self, slot):
        return bu.unpack_state(self.table, slot, 0), bu.unpack_state(self.table, slot, 1)


@pytest.fixture(scope="module")
def emu_lib():
    subprocess.run(["make", "-C", EMU_DIR, "-s"], check=True)
    L = C.CDLL(os.path.join(EMU_DIR, "_build", "libss_emu.so"))
    L.emu_plan.restype = C.c_int64
    L.emu_plan.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
    L.emu_run_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p]
    return L


@pytest.fixture(params=[0, 2, 6], ids=["direct", "staged_input", "staged_input_output"])
def be(emu_lib, request):
    """Every scenario runs twice: plain per-lane loads, and with each packed tile's input window pulled
    through shared memory first (SS_ARG_STAGE_INPUT, what zero-copy host batches use)."""
    emu_lib.emu_set_arg_flags(request.param)
    b = EmuBackend(emu_lib)
    b.preserves_gaps = not (request.param & 4)   # staged output rewrites whole 16-byte rows of its window
    yield b
    emu_lib.emu_set_arg_flags(0)


def test_planner_packs_and_splits(be):
    lens = [64] * 25 + [0, 1, 1920, 1921, 5000, 30, 200, 200, 200, 200, 200, 200]
    p, _ = bu.plain_layout(lens)
    w, _ = bu.wire_layout(lens)
    d = bu.make_descs(np.arange(len(lens)), lens, p, w)
    tiles, n_bulk = be.plan(d)
    assert n_bulk == 2
    covered = []
    for t in tiles[: len(tiles) - n_bulk]:
        nfr = int(t["n_kind"]) & 0xFFFF
        assert 1 <= nfr <= 16 and not (int(t["n_kind"]) >> 16)
        blocks = sum(2 + (lens[i] + 63) // 64 for i in range(int(t["first"]), int(t["first"]) + nfr))
        assert blocks <= 32
        covered += list(range(int(t["first"]), int(t["first"]) + nfr))
    for t in tiles[len(tiles) - n_bulk :]:
        assert int(t["n_kind"]) == (1 | (1 << 16)) and lens[int(t["first"])] > 1920
        covered.append(int(t["first"]))
    assert sorted(covered) == list(range(len(lens)))
    assert (int(tiles[0]["n_kind"]) & 0xFFFF) == 10  # ten 64-byte frames (3 blocks each) share one warp
    d["in_off"][3] += 4  # misaligned descriptor is rejected
    assert be.lib.emu_plan(d.ctypes.data, len(d), None, 0, None) == -22


def test_packed_every_small_length(be):
    lens = list(range(0, 131))
    pc.seal_open_roundtrip(be, lens, [0] * len(lens), seed=1, chain=2)


def test_packed_tags_and_rekey(be):
    rng = random.Random(5)
    lens = [rng.choice([0, 1, 15, 16, 17, 63, 64, 65, 100, 200, 500, 1000, 1919, 1920]) for _ in range(40)]
    pc.seal_open_roundtrip(be, lens, [i % 4 for i in range(40)], seed=2, chain=3)


def test_bulk_lengths(be):
    lens = [1921, 1984, 2048, 4095, 4096, 4097, 16384, 65536, 65537, 64 * 1022, 64 * 1023 + 1]
    pc.seal_open_roundtrip(be, lens, [0, 1, 2, 3, 0, 0, 0, 0, 2, 0, 0], seed=3, chain=2)


def test_mixed_ragged_with_forgeries(be):
    lens = pc.power_law_lens(random.Random(11), 96) + [3000, 0, 0, 64]
    pc.seal_open_roundtrip(be, lens, [0] * len(lens), seed=4, chain=2, corrupt=(0, 5, 17, 96, 97, 99))


def test_golden_chains_through_emulated_kernel(be, ss_vectors):
    assert pc.golden_chains(be, ss_vectors[:3], max_frames=40) >= 80


def test_runs_of_queued_frames(be):
    n, n_retry = pc.run_open_scenario(be, seed=3)
    assert n > 60 and n_retry > 0


def test_thread_per_frame_tiles(be):
    """The emulator library is built with the thread-tile threshold lowered to 64 frames so this path
    (one lane = one whole frame, serial blocks, plain Horner) is covered on CPU."""
    lens = [64] * 70 + [0] * 40 + [200] * 33 + [256] * 32 + [1] * 32 + [17, 300, 64, 64]
    p, _ = bu.plain_layout(lens)
    w, _ = bu.wire_layout(lens)
    tiles, n_bulk = be.plan(bu.make_descs(np.arange(len(lens)), lens, p, w))
    kinds = [int(t["n_kind"]) >> 16 for t in tiles]
    assert kinds.count(2) >= 5 and kinds.count(0) >= 2 and n_bulk == 0   # both small-tile kinds in one batch
    tags = [(i % 4) if i % 7 == 0 else 0 for i in range(len(lens))]
    pc.seal_open_roundtrip(be, lens, tags, seed=8, chain=3, corrupt=(0, 5, 71, 111, 150, 180))


def test_thread_tiles_in_runs(be):
    n, n_retry = pc.run_open_scenario(be, seed=9, n_streams=40)
    assert n > 64


@pytest.mark.parametrize("length", [385, 449, 700, 1024, 1025, 1857, 1920])
def test_strided_tiles_mid_size_frames(be, length):
    """Uniform batches of 7..32-block frames are cut into strided tiles (k = 2, 4, 8 lanes per frame, 32/k frames per
    warp, r^(4k) between a lane's windows): all lane counts, partial tail tiles, tags, forged frames."""
    n = 19
    lens = [length] * n
    p, _ = bu.plain_layout(lens)
    w, _ = bu.wire_layout(lens)
    tiles, n_bulk = be.plan(bu.make_descs(np.arange(n), lens, p, w))
    assert n_bulk == len(tiles) and all((int(t["n_kind"]) >> 24) & 63 in (2, 4, 8) for t in tiles)
    pc.seal_open_roundtrip(be, lens, [i % 4 for i in range(n)], seed=500 + length, chain=2, corrupt=(0, 7, 18),
                           speculative_min_len=385)


def test_stream_id_blake2b_matches_libsodium(emu_lib):
    """secret-stream's stream id = crypto_generichash(out32, namespace, key = handshake hash): the product's BLAKE2b
    (ss_blake2b.cuh, run here on the CPU) against libsodium's, random keys and messages."""
    import random

    import nacl.bindings as nb

    rng = random.Random(77)
    emu_lib.emu_blake2b256_keyed.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
    for _ in range(200):
        key, msg = rng.randbytes(64), rng.randbytes(32)
        out = C.create_string_buffer(32)
        emu_lib.emu_blake2b256_keyed(key, msg, out)
        assert out.raw == nb.crypto_generichash_blake2b_salt_personal(msg, digest_size=32, key=key)
