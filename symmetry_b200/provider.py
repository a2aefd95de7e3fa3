"""Host-side mirror of the reference's per-connection stream handler.

`handle_inference_requests` restates the control flow of
/root/reference/src/provider.ts:195-275 (`handleInferenceRequest`) for MANY
connections at once on top of the per-frame front end of the C ABI:

    provider.ts:234-238   peer.write(JSON.stringify({symmetryEmitterKey}))        preamble frame
    provider.ts:240-257   for each upstream chunk: completion += delta; peer.write(chunk)
    provider.ts:260-262   peer.write(createMessage('inferenceEnded', key))         trailer frame
    provider.ts:270-274   a TypeError from the extractor ends that relay: the chunk is NOT written and
                          no trailer follows (other connections are unaffected)

Names follow the reference (src/utils.ts:12-14 createMessage, src/constants.ts:3-20).  The
TypeScript original cannot run in this image (no Node.js); this module is what the parity tests
drive, and addon/symrelay.ts is the glue a maintainer would use in the real provider.
"""
from __future__ import annotations

import json

from . import relay as _r

serverMessageKeys = {"inference": "inference", "inferenceEnded": "inferenceEnded", "newConversation": "newConversation"}


def createMessage(key: str, data=None) -> str:  # src/utils.ts:12-14
    obj = {"key": key}
    if data is not None:
        obj["data"] = data
    return json.dumps(obj, separators=(",", ":"), ensure_ascii=False)


def safeParseJson(data: str):  # src/utils.ts:4-10
    try:
        return json.loads(data)
    except Exception:
        return None


class PeerRelayState:
    """What the loop keeps per connection (provider.ts:230-262)."""

    def __init__(self, slot: int, emitter_key: str, chunks):
        self.slot = slot
        self.emitter_key = emitter_key
        self.frames = [json.dumps({"symmetryEmitterKey": emitter_key}, separators=(",", ":")).encode()]
        self.extract = [False]
        for c in chunks:
            self.frames.append(bytes(c))
            self.extract.append(True)
        self.frames.append(createMessage(serverMessageKeys["inferenceEnded"], emitter_key).encode())
        self.extract.append(False)
        self.next = 0
        self.completion = b""
        self.wire = []          # sealed frames in order, as they go to the socket
        self.aborted = False    # the reference threw a TypeError (provider.ts:270-274)
        self.unsupported = []   # chunk indices the GPU extractor handed back to the host helper
        self.done = False


def handle_inference_requests(relay: "_r.SymRelay", peers, host_extract=None):
    """Relays every peer's stream to completion.  One frame per connection per batch (the MAC
    chain), all connections of a round in one batch.  `host_extract(provider_slot, chunk)` is the
    reference's own JS helper stand-in for chunks the GPU reports as UNSUPPORTED: it must return
    (ok, delta_bytes); without it such a chunk raises."""
    by_ticket = {}
    active = [p for p in peers if not p.done]
    while active:
        for p in active:
            while True:
                try:
                    t = relay.submit_seal(p.slot, p.frames[p.next], extract=p.extract[p.next])
                    break
                except _r.RelayError as e:      # ring full: `write()` returned false -> wait for 'drain'
                    if e.code != _r.EAGAIN:
                        raise
                    _drain(relay, by_ticket, wait=True)
            by_ticket[t] = p
        relay.flush()
        while by_ticket:
            _drain(relay, by_ticket, wait=True, host_extract=host_extract)
        active = [p for p in peers if not p.done]
    return peers


def _drain(relay, by_ticket, wait, host_extract=None):
    for c in relay.poll(wait=wait):
        p = by_ticket.pop(c["ticket"], None)
        if p is None:
            continue
        if c["status"] == _r.ST_OK:
            p.completion += c["delta"]
            p.wire.append(c["data"])
            p.next += 1
        elif c["status"] == _r.ST_EXTRACT_THROW:
            p.aborted = True
            p.done = True
            continue
        elif c["status"] == _r.ST_EXTRACT_UNSUPPORTED:
            if host_extract is None:
                raise RuntimeError("chunk needs the host helper (extractor reported UNSUPPORTED)")
            ok, delta = host_extract(p, p.frames[p.next])
            p.unsupported.append(p.next)
            if not ok:
                p.aborted = True
                p.done = True
                continue
            p.completion += delta
            p.extract[p.next] = False          # resubmit the same chunk for sealing only
        else:
            raise RuntimeError(f"frame failed with status {c['status']}")
        if p.next >= len(p.frames):
            p.done = True


# ---- transcript output: provider.ts:264-269, 277-297 -------------------------
def _json_stringify(value) -> str:
    """JSON.stringify for the shapes saveCompletion writes (arrays of {role, content} objects):
    compact separators, non-ASCII left as is, lone surrogates escaped as \\udxxx (ES2019)."""
    s = json.dumps(value, separators=(",", ":"), ensure_ascii=False)
    if any(0xD800 <= ord(c) <= 0xDFFF for c in s):
        s = "".join(f"\\u{ord(c):04x}" if 0xD800 <= ord(c) <= 0xDFFF else c for c in s)
    return s


def saveCompletion(completion: bytes, peer_public_key: bytes, conversation_index: int, messages, path: str) -> str:
    """Writes `<path>/<peerPubKeyHex>-<conversationIndex>.json` = JSON.stringify([...messages,
    {role:'assistant', content: completion}]) exactly as provider.ts:282-293 does, fed from the
    extractor's accumulated delta bytes (WTF-8) instead of a JS string.  Returns the file name."""
    import os

    text = completion.decode("utf-8", "surrogatepass")
    # halves of a surrogate pair that arrived in different chunks join up in the JS string
    text = text.encode("utf-16", "surrogatepass").decode("utf-16", "surrogatepass")
    body = _json_stringify(list(messages) + [{"role": "assistant", "content": text}])
    name = os.path.join(path, f"{peer_public_key.hex()}-{conversation_index}.json")
    with open(name, "w", encoding="utf-8") as f:
        f.write(body)
    return name
