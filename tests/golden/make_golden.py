#!/usr/bin/env python3
"""Regenerates the fixtures in this directory.  Run from the repo root:

    python tests/golden/make_golden.py

secretstream_vectors.json   produced by the REAL libsodium 1.0.20 (PyNaCl's
    bundled build, the C library behind sodium-native 4.1.1 which the
    reference reaches from /root/reference/src/provider.ts:248) -- never by
    this repo's own code.  Each chain: key, header, then frames pushed in
    order on one state with the wire bytes (uint24-LE prefix ‖ push output;
    sha256 only when longer than 192 B) and the 52-byte state after each.
extract_cases.json          the delta-extraction quirk matrix of SURVEY.md
    section 8(a) rows A2-A4, evaluated by oracle/extract.py (the restatement of
    /root/reference/src/utils.ts:16-52; no JS engine exists here, so these
    freeze the restatement rather than pin it).
"""
import hashlib
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import extract as ex  # noqa: E402
from oracle import libsodium as ls  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def chain(key, header, frames):
    st = ls.State(key, header)
    out = {"key": key.hex(), "header": header.hex(), "state0": st.raw.hex(), "frames": []}
    for m, tag in frames:
        wire = st.seal_frame(m, tag)
        rec = {"len": len(m), "tag": tag, "state_after": st.raw.hex(),
               "wire_sha256": hashlib.sha256(wire).hexdigest()}
        if len(m) <= 192:
            rec["m"] = m.hex()
            rec["wire"] = wire.hex()
        else:
            rec["m_seed"] = None  # filled by caller for long frames
        out["frames"].append(rec)
    return out


def long_msg(seed: int, n: int) -> bytes:
    """Deterministic long message: SHAKE-256 of the seed (reproducible without numpy)."""
    return hashlib.shake_256(b"symrelay-golden" + seed.to_bytes(4, "little")).digest(n)


def make_secretstream():
    chains = []
    # 1. the chain printed in SURVEY.md 8(c)
    msgs = [b"", b'{"symmetryEmitterKey":"abc"}',
            b'data: {"choices":[{"index":0,"delta":{"content":" Hello"}}]}\n\n',
            bytes(range(64)), b'{"key":"inferenceEnded","data":"abc"}']
    chains.append(chain(bytes(range(32)), bytes(range(0x40, 0x58)), [(m, 0) for m in msgs]))
    # 2. every length 0..130 on one chain (block and pad boundaries), tag 0
    rng = random.Random(0x5EED0001)
    key, hdr = rng.randbytes(32), rng.randbytes(24)
    chains.append(chain(key, hdr, [(rng.randbytes(n), 0) for n in range(0, 131)]))
    # 3. all four tags incl. REKEY / FINAL, mixed lengths
    key, hdr = rng.randbytes(32), rng.randbytes(24)
    frames = []
    for i in range(48):
        n = rng.choice([0, 1, 15, 16, 17, 47, 48, 63, 64, 65, 100, 128, 190])
        frames.append((rng.randbytes(n), i % 4))
    chains.append(chain(key, hdr, frames))
    # 4. long frames (sha256 only): sizes around 64-B / 2-KiB tiles and the bench sizes
    key, hdr = rng.randbytes(32), rng.randbytes(24)
    st = ls.State(key, hdr)
    c = {"key": key.hex(), "header": hdr.hex(), "state0": st.raw.hex(), "frames": []}
    for i, n in enumerate([193, 255, 256, 1000, 1919, 1920, 1921, 1984, 2048, 4095, 4096, 4097, 16383, 16384,
                           65535, 65536, 65537, 131072 + 5]):
        m = long_msg(i, n)
        wire = st.seal_frame(m, 0)
        c["frames"].append({"len": n, "tag": 0, "m_seed": i, "state_after": st.raw.hex(),
                            "wire_sha256": hashlib.sha256(wire).hexdigest(), "wire_head": wire[:20].hex(),
                            "mac": wire[-16:].hex()})
    chains.append(c)
    with open(os.path.join(HERE, "secretstream_vectors.json"), "w") as f:
        json.dump({"source": "libsodium 1.0.20 via PyNaCl 1.6.2 (tests/golden/make_golden.py)",
                   "chains": chains}, f, indent=0)
    return sum(len(c["frames"]) for c in chains)


EXTRACT_CHUNKS = [
    # mainstream OpenAI-format SSE events
    b'data: {"choices":[{"index":0,"delta":{"content":" Hello"}}]}\n\n',
    b'data: {"id":"chatcmpl-1","object":"chat.completion.chunk","created":1730000000,"model":"llama3:8b",'
    b'"system_fingerprint":"fp_ollama","choices":[{"index":0,"delta":{"role":"assistant","content":"tok"},'
    b'"finish_reason":null}]}\n\n',
    b'data: {"choices":[{"index":0,"delta":{},"finish_reason":"stop"}]}\n\n',
    b'data: {"choices":[{"index":0,"delta":{"role":"assistant","content":""},"finish_reason":null}]}\n\n',
    b"data: [DONE]\n\n",
    b'data:{"choices":[{"delta":{"content":"nospace"}}]}',
    b'data: \t\r\n {"choices":[{"delta":{"content":"ws"}}]} \r\n\t ',
    # no prefix -> JSON.parse(whole)
    b'{"choices":[{"delta":{"content":"bare"}}]}',
    b'{"content":"llama.cpp native","stop":false}',
    b'data: {"content":" tok","stop":false,"id_slot":0}\n\n',
    b' data: {"choices":[{"delta":{"content":"leading space"}}]}',
    b'Data: {"choices":[{"delta":{"content":"case"}}]}',
    b'\xef\xbb\xbfdata: {"choices":[{"delta":{"content":"bom"}}]}',
    # multi-event / partial chunks
    b'data: {"choices":[{"delta":{"content":"a"}}]}\n\ndata: {"choices":[{"delta":{"content":"b"}}]}\n\n',
    b'data: {"choices":[{"delta":{"content":"cut',
    b'"}}]}\n\n',
    b"data: ",
    b"data:",
    b"",
    b"\n",
    # `data:` inside the JSON truncates at split()
    b'data: {"choices":[{"delta":{"content":"say data: hi"}}]}\n\n',
    b'{"choices":[{"delta":{"content":"say data: hi"}}]}',
    # TypeError matrix
    b'data: {"choices":[]}\n\n',
    b'data: {"usage":{"prompt_tokens":1}}\n\n',
    b'data: {"choices":[null]}',
    b'data: {"choices":[{}]}',
    b'data: {"choices":[{"delta":null}]}',
    b'data: {"choices":null}',
    b'data: {"choices":[5]}',
    b'data: {"choices":["s"]}',
    b'data: {"choices":[[]]}',
    b'data: {"choices":"abc"}',
    b'data: {"choices":""}',
    b'data: {"choices":{"0":{"delta":{"content":"objidx"}}}}',
    b'data: {"choices":{}}',
    b'data: {"choices":7}',
    b'data: {"choices":true}',
    b"data: 5",
    b"data: null",
    b"data: true",
    b'data: "str"',
    b"data: []",
    b"data: {}",
    b'data: [{"choices":1}]',
    # content of other types / truthiness / coercions
    b'data: {"choices":[{"delta":{"content":null}}]}',
    b'data: {"choices":[{"delta":{"content":0}}]}',
    b'data: {"choices":[{"delta":{"content":-0}}]}',
    b'data: {"choices":[{"delta":{"content":0.0e5}}]}',
    b'data: {"choices":[{"delta":{"content":5}}]}',
    b'data: {"choices":[{"delta":{"content":-12}}]}',
    b'data: {"choices":[{"delta":{"content":1.5}}]}',
    b'data: {"choices":[{"delta":{"content":1e21}}]}',
    b'data: {"choices":[{"delta":{"content":123456789012}}]}',
    b'data: {"choices":[{"delta":{"content":true}}]}',
    b'data: {"choices":[{"delta":{"content":false}}]}',
    b'data: {"choices":[{"delta":{"content":{}}}]}',
    b'data: {"choices":[{"delta":{"content":{"a":1}}}]}',
    b'data: {"choices":[{"delta":{"content":[]}}]}',
    b'data: {"choices":[{"delta":{"content":[1,"a",null,[2,3]]}}]}',
    b'data: {"choices":[{"delta":{"content":"undefined"}}]}',
    b'data: {"choices":[{"delta":"str"}]}',
    b'data: {"choices":[{"delta":5}]}',
    b'data: {"choices":[{"delta":[]}]}',
    b'data: {"choices":[{"delta":["x"]}]}',
    b'{"content":null}', b'{"content":5}', b'{"content":true}', b'{"content":{}}', b'{"content":[1,2]}',
    b'{"content":""}', b'{"nocontent":1}', b'[1]', b'"s"', b"7",
    # duplicate keys: last wins
    b'data: {"choices":[{"delta":{"content":"first","content":"second"}}]}',
    b'data: {"choices":[{"delta":{"content":"x"}}],"choices":[]}',
    b'data: {"choices":[],"choices":[{"delta":{"content":"late"}}]}',
    b'data: {"choices":[{"delta":{"content":"x"},"delta":{}}]}',
    b'data: {"choices":[{"delta":{},"delta":{"content":"y"}}]}',
    b'data: {"choices":[{"delta":{"content":"x","content":null}}]}',
    b'{"content":"a","content":"b"}',
    # second array element ignored, nested look-alikes ignored
    b'data: {"choices":[{"delta":{"content":"one"}},{"delta":{"content":"two"}}]}',
    b'data: {"x":{"choices":[{"delta":{"content":"nested"}}]},"choices":[{"delta":{"content":"top"}}]}',
    b'data: {"choices":[{"delta":{"content":"c","x":{"content":"inner"}}}]}',
    b'data: {"choices":[{"x":{"delta":{"content":"no"}},"delta":{"content":"yes"}}]}',
    # escapes / unicode
    b'data: {"choices":[{"delta":{"content":"q\\"b\\\\s\\/n\\nt\\tr\\rb\\bf\\f"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\u00e9\\u4e2d\\ud83d\\ude00"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\ud83d"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\ude00\\ud83d"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\u0000\\u001f"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\uD83D\\uDE00 \\uABCD"}}]}',
    'data: {"choices":[{"delta":{"content":"é中😀"}}]}'.encode(),
    b'data: {"choices":[{"delta":{"content":"\\u00zz"}}]}',
    b'data: {"choices":[{"delta":{"content":"\\x41"}}]}',
    b'data: {"choices":[{"delta":{"content":"bad\\"}}]}',
    b'data: {"choices":[{"delta":{"content":"raw\ttab"}}]}',
    b'data: {"choices":[{"delta":{"content":"raw\nnl"}}]}',
    b'data: {"choices":[{"delta":{"content":"\x7f del ok"}}]}',
    b'data: {"ch\\u006fices":[{"d\\u0065lta":{"cont\\u0065nt":"escaped keys"}}]}',
    # invalid UTF-8 -> U+FFFD
    b'data: {"choices":[{"delta":{"content":"\xff\xfe"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xe4\xb8"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xe4\xb8x"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xf0\x9f\x98"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xed\xa0\x80"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xc0\xaf"}}]}',
    b'data: {"choices":[{"delta":{"content":"\xf4\x90\x80\x80"}}]}',
    b'data: {"choices":[{"delta":{"content":"\x80"}}]}',
    b'data: {"choices":[{"delta":{"content":"ok"}}]}\xff',
    b'data: {"choices":[{"delta":{"content":"ok"}}]\xc3}',
    # strict-grammar rejections
    b'data: {"choices":[{"delta":{"content":"x"}}],}',
    b'data: {"choices":[{"delta":{"content":"x"}},]}',
    b"data: {'choices':[]}",
    b'data: {"a":NaN}', b'data: {"a":Infinity}', b'data: {"a":-Infinity}', b'data: {"a":01}', b'data: {"a":1.}',
    b'data: {"a":.5}', b'data: {"a":+1}', b'data: {"a":1e}', b'data: {"a":1e+}', b'data: {"a":-}',
    b'data: {"a":0x10}', b'data: {"a":tru}', b'data: {"a":nul}', b'data: {"a":True}',
    b'data: {"a":1 "b":2}', b'data: {"a" 1}', b'data: {a:1}', b'data: {"a":1}}', b'data: {"a":1}]',
    b'data: {"a":[1,2}', b'data: {"a":{"b":1]', b'data: {"a":1} x', b'data: {"a":1}\x0b', b'data: {"a":1}\x0c',
    b'data: {"a":1}\xc2\xa0', b'data: //c\n{}', b'data: {"a":1,"choices":[{"delta":{"content":"n"}}],"z":-1.25E+3}',
    b'data: {"choices":[{"delta":{"content":"deep"},"x":[[[[[[[[[[1]]]]]]]]]]}]}',
    b'data: ' + b"[" * 40 + b"]" * 40,
    b'data: {"choices":[{"delta":{"content":"big"}}],"pad":"' + b"x" * 600 + b'"}',
]


def make_extract():
    cases = []
    for prov in ("ollama", "openwebui", "llamacpp", "litellm", "lmstudio", "oobabooga", "somethingelse"):
        for ch in EXTRACT_CHUNKS:
            st, delta = ex.extract(prov, ch)
            cases.append({"provider": prov, "chunk": ch.hex(), "status": st, "delta": delta.hex()})
    with open(os.path.join(HERE, "extract_cases.json"), "w") as f:
        json.dump({"source": "oracle/extract.py (restatement of src/utils.ts:16-52; parity unpinned)",
                   "cases": cases}, f, indent=0)
    return len(cases)


if __name__ == "__main__":
    print("secretstream frames:", make_secretstream())
    print("extract cases:", make_extract())
