"""Host-side mirror pieces that need no GPU: message helpers (src/utils.ts:4-14) and the transcript
writer fed from extractor output (src/provider.ts:277-297)."""
import json

from oracle import extract as ex
from symmetry_b200 import provider, synth


def test_create_message_matches_json_stringify_shape():
    assert provider.createMessage("inferenceEnded", "abc") == '{"key":"inferenceEnded","data":"abc"}'
    assert provider.createMessage("ping") == '{"key":"ping"}'          # JSON.stringify drops undefined
    assert provider.safeParseJson('{"key":"inference","data":{"key":"k","messages":[]}}')["key"] == "inference"
    assert provider.safeParseJson("not json") is None


def test_save_completion_from_extracted_deltas(tmp_path):
    chunks = synth.chat_chunks(40, seed=3, stream_id=7)
    completion = b"".join(ex.extract("ollama", c)[1] for c in chunks)
    # a surrogate pair split across two chunks joins in the JS string (`completion +=` is UTF-16)
    completion += ex.extract("ollama", b'data: {"choices":[{"delta":{"content":"\\ud83d"}}]}')[1]
    completion += ex.extract("ollama", b'data: {"choices":[{"delta":{"content":"\\ude00"}}]}')[1]
    completion += ex.extract("ollama", b'data: {"choices":[{"delta":{"content":"\\ud800!"}}]}')[1]   # stays lone
    msgs = [{"role": "user", "content": "hi   there"}]
    name = provider.saveCompletion(completion, bytes(range(32)), 3, msgs, str(tmp_path))
    assert name.endswith("000102030405060708090a0b0c0d0e0f101112131415161718191a1b1c1d1e1f-3.json")
    raw = open(name, encoding="utf-8").read()
    assert raw.startswith('[{"role":"user","content":"hi   there"},{"role":"assistant","content":"')
    assert "\U0001F600" in raw and "\\ud800!" in raw
    back = json.loads(raw)
    assert back[1]["role"] == "assistant" and back[1]["content"].startswith(back[1]["content"][:1])
    want = "".join(ex.extract("ollama", c)[1].decode() for c in chunks)
    assert back[1]["content"].startswith(want)
