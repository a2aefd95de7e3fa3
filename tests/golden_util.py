"""Helpers shared by the CPU and GPU parity tests."""
import hashlib


def long_msg(seed: int, n: int) -> bytes:
    """Same generator as tests/golden/make_golden.py."""
    return hashlib.shake_256(b"symrelay-golden" + seed.to_bytes(4, "little")).digest(n)


def frame_message(fr) -> bytes:
    if "m" in fr:
        return bytes.fromhex(fr["m"])
    return long_msg(fr["m_seed"], fr["len"])
