"""Batch construction shared by the emulator test (CPU) and the GPU parity tests.

Layouts follow include/symrelay.h: seal plaintext at 16-aligned offsets, wire
frames at offsets == 12 (mod 16) so the ciphertext is 16-aligned."""
from __future__ import annotations

import numpy as np

F_OPEN = 0x04000000
F_EXTRACT = 0x08000000
F_RUN = 0x10000000

DESC_DTYPE = np.dtype([("slot", "<u4"), ("in_off", "<u4"), ("out_off", "<u4"), ("len_flags", "<u4")])
TILE_DTYPE = np.dtype([("first", "<u4"), ("n_kind", "<u4")])


def align16(n: int) -> int:
    return (n + 15) & ~15


def aligned_zeros(nbytes: int) -> np.ndarray:
    """uint8 array whose data pointer is 64-byte aligned."""
    raw = np.zeros(nbytes + 64, dtype=np.uint8)
    off = (-raw.ctypes.data) % 64
    return raw[off : off + nbytes]


def plain_layout(lens):
    """16-aligned plaintext offsets; returns (offsets, arena_bytes)."""
    offs, cur = [], 0
    for n in lens:
        offs.append(cur)
        cur += max(16, align16(n))
    return offs, cur


def wire_layout(lens):
    """Wire-frame offsets == 12 mod 16; returns (offsets, arena_bytes)."""
    offs, cur = [], 0
    for n in lens:
        offs.append(cur + 12)
        cur += align16(12 + n + 20)
    return offs, cur


def make_descs(slots, lens, in_offs, out_offs, *, open_=False, tags=None, extract=False, run=False) -> np.ndarray:
    d = np.zeros(len(lens), dtype=DESC_DTYPE)
    d["slot"] = slots
    d["in_off"] = in_offs
    d["out_off"] = out_offs
    lf = np.asarray(lens, dtype=np.uint32).copy()
    if tags is not None:
        lf |= (np.asarray(tags, dtype=np.uint32) & 3) << 24
    if open_:
        lf |= F_OPEN
    if extract:
        lf |= F_EXTRACT
    if run:
        lf |= F_RUN
    d["len_flags"] = lf
    return d


def pack_state_table(n_slots: int, tx_states, rx_states) -> np.ndarray:
    """[slot][dir] -> 48 bytes: k[32] | nonce[12] | meta[4] (device layout of the emulator/kernels)."""
    tab = aligned_zeros(n_slots * 2 * 48).reshape(n_slots, 2, 48)
    for s in range(n_slots):
        for d, states in ((0, tx_states), (1, rx_states)):
            if states is not None and states[s] is not None:
                tab[s, d, :44] = np.frombuffer(states[s][:44], dtype=np.uint8)
    return tab


def unpack_state(tab: np.ndarray, slot: int, d: int) -> bytes:
    """Back to libsodium's 52-byte form (pad bytes are always zero)."""
    return bytes(tab[slot, d, :44]) + bytes(8)
