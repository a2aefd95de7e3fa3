"""Synthetic workloads of BASELINE.json's configs.

Keys and headers: key_s = BLAKE2b-256("symrelay-key" ‖ LE32(seed) ‖ LE32(s)),
header_s = first 24 B of BLAKE2b-256("symrelay-hdr" ‖ LE32(seed) ‖ LE32(s));
payload bytes from numpy's default_rng(seed).  The chat-completion generator
restates what an OpenAI-compatible endpoint streams to the reference's relay
loop (/root/reference/src/provider.ts:240-257): one SSE event per chunk."""
from __future__ import annotations

import hashlib
import struct

import numpy as np

SEED = 0x5EED0001


def stream_keys(n: int, seed: int = SEED, first: int = 0):
    keys = np.empty((n, 32), dtype=np.uint8)
    hdrs = np.empty((n, 24), dtype=np.uint8)
    for i in range(n):
        tail = struct.pack("<II", seed & 0xFFFFFFFF, first + i)
        keys[i] = np.frombuffer(hashlib.blake2b(b"symrelay-key" + tail, digest_size=32).digest(), dtype=np.uint8)
        hdrs[i] = np.frombuffer(hashlib.blake2b(b"symrelay-hdr" + tail, digest_size=32).digest()[:24], dtype=np.uint8)
    return keys, hdrs


def power_law_lens(n: int, seed: int = SEED, alpha: float = 1.2, cap: int = 16384) -> np.ndarray:
    """Config 3: L = min(cap, floor(32 * U^(-1/alpha))), U ~ Uniform(0,1]."""
    u = 1.0 - np.random.default_rng(seed).random(n)
    return np.minimum(cap, np.floor(32.0 * u ** (-1.0 / alpha))).astype(np.int64)


VOCAB = [" the", " a", " of", " and", " to", " in", " is", " that", " it", " for", " as", " with", " on", " be",
         " this", " by", " are", " or", " an", " from", " at", " which", " one", " all", " we", " can", " has",
         " there", " more", " if", " about", " model", " token", " stream", " relay", " GPU", "ing", "ed", "s",
         ".", ",", "!", "?", "\n", " \"quoted\"", " café", " naïve", " 中文", " \U0001F600", " tab\t",
         " back\\slash", " 1", " 42", " 3.14", "A", "B", " I", " Hello", " world", " data", " data:", " {json}"]


def chat_chunks(n_chunks: int, seed: int, stream_id: int = 0):
    """Config 1/5: the body chunks of one llama3:8b-style streamed completion, as bytes:
    n_chunks delta events, a stop event, then `data: [DONE]`."""
    import json

    rng = np.random.default_rng((seed, stream_id))
    out = []
    created = 1730000000 + stream_id
    cid = f"chatcmpl-{stream_id:06d}"
    for i in range(n_chunks):
        tok = VOCAB[int(rng.integers(0, len(VOCAB)))]
        delta = {"role": "assistant", "content": tok}
        ev = {"id": cid, "object": "chat.completion.chunk", "created": created, "model": "llama3:8b",
              "system_fingerprint": "fp_ollama",
              "choices": [{"index": 0, "delta": delta, "finish_reason": None}]}
        out.append(("data: " + json.dumps(ev, separators=(",", ":"), ensure_ascii=bool(i % 3 == 0)) + "\n\n").encode())
    ev = {"id": cid, "object": "chat.completion.chunk", "created": created, "model": "llama3:8b",
          "system_fingerprint": "fp_ollama", "choices": [{"index": 0, "delta": {}, "finish_reason": "stop"}]}
    out.append(("data: " + json.dumps(ev, separators=(",", ":")) + "\n\n").encode())
    out.append(b"data: [DONE]\n\n")
    return out
