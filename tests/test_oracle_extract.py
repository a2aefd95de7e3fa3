"""Table-driven checks of oracle/extract.py (restatement of
/root/reference/src/utils.ts:16-52 + provider.ts:243-246).  The expectations
written out by hand below are derived from ECMAScript semantics (SURVEY.md
section 8(a) rows A2-A4); tests/golden/extract_cases.json then freezes the whole
quirk matrix so the GPU extractor and later edits are held to it."""
import pytest

from oracle import extract as ex

OLL, LCP, DEF = "ollama", "llamacpp", "litellm"

HAND = [
    # (provider, chunk, status, delta)
    (OLL, b'data: {"choices":[{"index":0,"delta":{"content":" Hello"}}]}\n\n', ex.OK, b" Hello"),
    (OLL, b"data: [DONE]\n\n", ex.OK, b""),                       # SyntaxError -> undefined -> ''
    (LCP, b"data: [DONE]\n\n", ex.OK, b"undefined"),              # += undefined
    (DEF, b"data: [DONE]\n\n", ex.OK, b""),
    (OLL, b'data: {"choices":[]}\n\n', ex.THROW, b""),            # choices[0] undefined -> .delta throws
    (OLL, b'data: {"usage":{}}\n\n', ex.THROW, b""),              # choices undefined -> [0] throws
    (OLL, b'data: {"choices":[{}]}', ex.OK, b""),                 # delta?.content
    (DEF, b'data: {"choices":[{}]}', ex.THROW, b""),              # delta.content throws in default branch
    (DEF, b'data: {"choices":[{"delta":{"content":"undefined"}}]}', ex.OK, b""),
    (OLL, b'data: {"choices":[{"delta":{"content":"undefined"}}]}', ex.OK, b"undefined"),
    (OLL, b'data: {"choices":[{"delta":{"content":"a"}}]}\n\ndata: {"choices":[{"delta":{"content":"b"}}]}\n\n',
     ex.OK, b"a"),                                                # only the first event is parsed
    (OLL, b'data: {"choices":[{"delta":{"content":"say data: hi"}}]}', ex.OK, b""),  # split truncates
    (OLL, b'{"choices":[{"delta":{"content":"say data: hi"}}]}', ex.OK, b"say data: hi"),
    (OLL, b'data: {"choices":[{"delta":{"content":"cut', ex.OK, b""),
    (OLL, b' data: {"choices":[{"delta":{"content":"x"}}]}', ex.OK, b""),  # no prefix, JSON.parse(' data:..') fails
    (OLL, b"data: null", ex.OK, b""),
    (OLL, b"data: 5", ex.THROW, b""),
    (LCP, b'{"content":"x"}', ex.OK, b"x"),
    (LCP, b'{"content":null}', ex.OK, b"null"),
    (LCP, b'{"content":5}', ex.OK, b"5"),
    (LCP, b'{"content":1.5}', ex.OK, b"1.5"),
    (LCP, b'{"content":true}', ex.OK, b"true"),
    (LCP, b'{"content":{}}', ex.OK, b"[object Object]"),
    (LCP, b'{"content":[1,"a",null,[2,3]]}', ex.OK, b"1,a,,2,3"),
    (LCP, b"data: null", ex.OK, b"undefined"),
    (OLL, b'data: {"choices":[{"delta":{"content":0}}]}', ex.OK, b""),
    (OLL, b'data: {"choices":[{"delta":{"content":5}}]}', ex.OK, b"5"),
    (OLL, b'data: {"choices":[{"delta":{"content":"first","content":"second"}}]}', ex.OK, b"second"),
    (OLL, b'data: {"choices":[{"delta":{"content":"x"}}],"choices":[]}', ex.THROW, b""),
    (OLL, b'data: {"choices":[{"delta":{"content":"\\ud83d\\ude00"}}]}', ex.OK, "😀".encode()),
    (OLL, b'data: {"choices":[{"delta":{"content":"\\ud83d"}}]}', ex.OK, b"\xed\xa0\xbd"),
    (OLL, b'data: {"choices":[{"delta":{"content":"\xff"}}]}', ex.OK, b"\xef\xbf\xbd"),
    (OLL, b'data: {"choices":[{"delta":{"content":"raw\ttab"}}]}', ex.OK, b""),
    (OLL, b'data: {"a":NaN,"choices":[{"delta":{"content":"x"}}]}', ex.OK, b""),
    (OLL, b'\xef\xbb\xbfdata: {"choices":[{"delta":{"content":"bom"}}]}', ex.OK, b""),
    (OLL, b'data: {"choices":{"0":{"delta":{"content":"objidx"}}}}', ex.OK, b"objidx"),
    (OLL, b'data: {"choices":"abc"}', ex.OK, b""),
    (DEF, b'data: {"choices":"abc"}', ex.THROW, b""),
]


@pytest.mark.parametrize("prov,chunk,status,delta", HAND)
def test_hand_cases(prov, chunk, status, delta):
    assert ex.extract(prov, chunk) == (status, delta)


def test_default_branch_covers_lmstudio_oobabooga():
    ch = b'data: {"choices":[{}]}'
    for p in ("litellm", "lmstudio", "oobabooga", "anything"):
        assert ex.extract(p, ch)[0] == ex.THROW
    for p in ("ollama", "openwebui"):
        assert ex.extract(p, ch) == (ex.OK, b"")


def test_number_to_string():
    for x, s in [(1.0, "1"), (1e21, "1e+21"), (1e20, "1" + "0" * 20), (1e-7, "1e-7"), (0.000001, "0.000001"),
                 (1.5e300, "1.5e+300"), (-2.5, "-2.5"), (5e-324, "5e-324"), (123456789012345680000.0,
                                                                               "123456789012345680000"),
                 (-0.0, "0"), (float("inf"), "Infinity")]:
        assert ex.number_to_string(x) == s


def test_frozen_matrix(extract_cases):
    assert len(extract_cases) >= 900
    for c in extract_cases:
        st, delta = ex.extract(c["provider"], bytes.fromhex(c["chunk"]))
        assert (st, delta.hex()) == (c["status"], c["delta"]), c
