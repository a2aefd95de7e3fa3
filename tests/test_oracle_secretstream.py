"""Pins oracle/secretstream_ref.c: published KATs, the libsodium-generated golden
chains (tests/golden/), and live differential runs against libsodium 1.0.20."""
import hashlib
import random

import pytest

from oracle import libsodium as ls
from oracle import secretstream as ss
from golden_util import frame_message


def test_hchacha20_xchacha_draft_2_2_1():
    out = ss.hchacha20(bytes(range(32)), bytes.fromhex("000000090000004a0000000031415927"))
    assert out.hex() == "82413b4227b27bfed30e42508a877d73a0f9e4d58a74a853c12ec41326d3ecdc"


def test_chacha20_block_rfc8439_2_3_2():
    out = ss.chacha20_block(bytes(range(32)), 1, bytes.fromhex("000000090000004a00000000"))
    assert out.hex() == (
        "10f1e7e4d13b5915500fdd1fa32071c4c7d1f4c733c068030422aa9ac3d46c4e"
        "d2826446079faa0914c2d705d98b02a2b5129cd1de164eb9cbd083e8a2503c4e")


def test_chacha20_encrypt_rfc8439_2_4_2():
    pt = (b"Ladies and Gentlemen of the class of '99: If I could offer you only one tip for the future, "
          b"sunscreen would be it.")
    ct = ss.chacha20_ietf_xor_ic(pt, bytes.fromhex("000000000000004a00000000"), 1, bytes(range(32)))
    assert ct.hex().startswith("6e2e359a2568f98041ba0728dd0d6981")
    assert ct.hex().endswith("0bbf74a35be6b40b8eedf2785e42874d")


def test_poly1305_rfc8439_2_5_2():
    key = bytes.fromhex("85d6be7857556d337f4452fe42d506a80103808afb0db2fd4abff6af4149f51b")
    assert ss.poly1305(b"Cryptographic Forum Research Group", key).hex() == "a8061dc1305136c6c22b8baf0c0127a9"


def test_poly1305_wraparound_cases():
    # h >= p before the final reduction, and carries through every limb
    key = bytes.fromhex("02" + "00" * 31)
    assert ss.poly1305(b"\xff" * 16, key) == ls.poly1305(b"\xff" * 16, key)
    key = bytes([0xFF] * 16) + bytes([0xFF] * 16)
    for n in (1, 15, 16, 17, 32, 48, 255):
        assert ss.poly1305(b"\xff" * n, key) == ls.poly1305(b"\xff" * n, key)


def test_survey_golden_chain():
    st = ss.State.init(bytes(range(32)), bytes(range(0x40, 0x58)))
    assert st.raw.hex() == ("001b38f1bc654a0470f0172049103eccb67d8bb16b11d2a468db66a2dd53d47d"
                            "010000005051525354555657" + "00" * 8)
    w = st.seal_frame(b"")
    assert w.hex() == "1100000dff2db0dcccf8b9cea79016f3d4ae8e52"
    assert st.raw[32:44].hex() == "02000000af7ce28f98adef99"
    w = st.seal_frame(b'{"symmetryEmitterKey":"abc"}')
    assert w.hex() == ("2d000029707c985c38db9aa6387a79d4e10c8c7bd4d8ec17d4f0f5d2b8c93bb4c0706058ee92012bc5ce71a4"
                       "b02da072")
    w = st.seal_frame(b'data: {"choices":[{"index":0,"delta":{"content":" Hello"}}]}\n\n')
    assert w[:4].hex() == "4f0000d7" and w[-16:].hex() == "30027e481be4d86e6e0f69ea806e4fc0"
    assert hashlib.sha256(w[3:]).hexdigest()[:32] == "f49afd2e1cd46c24a751492e9021c3f4"  # sha256 of push output
    w = st.seal_frame(bytes(range(64)))
    assert w[:4].hex() == "51000036" and w[-16:].hex() == "4a7bc4d9e18f1f1e6e735108ed1bcf69"
    assert hashlib.sha256(w[3:]).hexdigest()[:32] == "0298d73f008aaae076794b04eb28f2b2"
    w = st.seal_frame(b'{"key":"inferenceEnded","data":"abc"}')
    assert w[:4].hex() == "36000005" and w[-16:].hex() == "b4831badf79caeedb726b743143215f1"
    assert hashlib.sha256(w[3:]).hexdigest()[:32] == "8c57ae1fd7a9c4be394907c4c35fab5f"
    assert st.raw[32:44].hex() == "06000000a1f623eb7bc8872f"


def test_golden_chains_seal_and_open(ss_vectors):
    n = 0
    for ch in ss_vectors:
        key, hdr = bytes.fromhex(ch["key"]), bytes.fromhex(ch["header"])
        tx, rx = ss.State.init(key, hdr), ss.State.init(key, hdr)
        assert tx.raw.hex() == ch["state0"]
        for fr in ch["frames"]:
            m = frame_message(fr)
            wire = tx.seal_frame(m, fr["tag"])
            assert hashlib.sha256(wire).hexdigest() == fr["wire_sha256"]
            if "wire" in fr:
                assert wire.hex() == fr["wire"]
            assert tx.raw.hex() == fr["state_after"]
            got = rx.open_frame(wire)
            assert got == (m, fr["tag"])
            assert rx.raw.hex() == fr["state_after"]
            n += 1
    assert n >= 200


@pytest.mark.parametrize("seed", range(4))
def test_differential_vs_libsodium(seed):
    rng = random.Random(seed)
    lens = [0, 1, 15, 16, 17, 31, 32, 33, 47, 48, 63, 64, 65, 100, 127, 128, 129, 1000, 1919, 1920, 1921, 5000]
    for _ in range(150):
        key, hdr = rng.randbytes(32), rng.randbytes(24)
        a, b, c = ss.State.init(key, hdr), ls.State(key, hdr), ss.State.init(key, hdr)
        assert a.raw == b.raw
        for _f in range(6):
            m = rng.randbytes(rng.choice(lens))
            tag = rng.choice([0, 0, 0, 1, 2, 3])
            ad = rng.randbytes(rng.choice([0, 0, 1, 16, 17]))
            wa, wb = a.push(m, tag, ad), b.push(m, tag, ad or None)
            assert wa == wb and a.raw == b.raw
            assert c.pull(wb, ad) == (m, tag) and c.raw == b.raw
            bad = bytearray(wb)
            bad[rng.randrange(len(bad))] ^= 1 << rng.randrange(8)
            before = c.raw
            assert c.pull(bytes(bad), ad) is None and c.raw == before
            assert c.pull(wb[:16], ad) is None and c.raw == before  # clen < 17


def test_primitives_vs_libsodium():
    rng = random.Random(7)
    for _ in range(100):
        key, nonce = rng.randbytes(32), rng.randbytes(12)
        data = rng.randbytes(rng.randrange(0, 300))
        ic = rng.choice([0, 1, 2, 0xFFFFFFFE])
        if ic == 0xFFFFFFFE and len(data) > 128:
            data = data[:128]
        assert ss.chacha20_ietf_xor_ic(data, nonce, ic, key) == ls.chacha20_ietf_xor_ic(data, nonce, ic, key)
        assert ss.hchacha20(key, nonce + b"abcd") == ls.hchacha20(key, nonce + b"abcd")
        assert ss.poly1305(data, key) == ls.poly1305(data, key)


def test_counter_wrap_triggers_rekey():
    key, hdr = bytes(range(32)), bytes(24)
    a, b = ss.State.init(key, hdr), ls.State(key, hdr)
    raw = bytearray(a.raw)
    raw[32:36] = (0xFFFFFFFF).to_bytes(4, "little")
    a = ss.State(bytes(raw))
    b.st.statebuf[0:52] = bytes(raw)
    wa, wb = a.push(b"wrap"), b.push(b"wrap")
    assert wa == wb and a.raw == b.raw
    assert a.raw[32:36] == (1).to_bytes(4, "little") and a.raw[:32] != bytes(raw[:32])
