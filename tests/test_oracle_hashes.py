"""Known-answer tests that pin the oracle's standard digests (CPU)."""
import hashlib
import random

import oracle


def test_sha256_fips_vectors():
    # FIPS 180-4 / NIST example vectors
    assert oracle.sha256(b"abc").hex() == "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad"
    assert oracle.sha256(b"").hex() == "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"
    m = b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq"
    assert oracle.sha256(m).hex() == "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1"


def test_sha256_against_hashlib_all_paddings():
    rng = random.Random(7)
    for n in list(range(0, 200)) + [1000, 4095, 4096, 4097, 65536 + 55, 65536 + 56]:
        data = bytes(rng.getrandbits(8) for _ in range(n))
        assert oracle.sha256(data) == hashlib.sha256(data).digest(), n


def test_xxh64_known_answers():
    # xxHash specification / reference implementation sanity values
    assert oracle.xxh64(b"", 0) == 0xEF46DB3751D8E999
    assert oracle.xxh64(b"a", 0) == 0xD24EC4F1A98C6E5B
    assert oracle.xxh64(b"abc", 0) == 0x44BC2CF5AD770999
    assert oracle.xxh64(b"Nobody inspects the spammish repetition", 0) == 0xFBCEA83C8A378BF1


def test_h64_is_length_and_position_sensitive():
    base = bytes(range(256)) * 40  # 10240 bytes: three tiles
    h = oracle.h64(base)
    assert h == oracle.h64(bytes(base))
    seen = {h}
    for pos in (0, 1, 15, 16, 4095, 4096, 8191, 8192, 10239):
        b = bytearray(base)
        b[pos] ^= 1
        seen.add(oracle.h64(bytes(b)))
    assert len(seen) == 10
    # trailing zeros change the digest (length is mixed in), so does swapping two vectors
    assert oracle.h64(base + b"\0") != h
    sw = bytearray(base)
    sw[0:16], sw[16:32] = sw[16:32], sw[0:16]
    assert oracle.h64(bytes(sw)) != h
    sw = bytearray(base)
    sw[0:4096], sw[4096:8192] = sw[4096:8192], sw[0:4096]
    assert bytes(sw) == base or oracle.h64(bytes(sw)) != h
