"""Pins the CPU oracle (oracle/gh_oracle.c) before anything is compared against it.

The reference has no digest fixtures, so the pins are the standards'
known-answer vectors (tests/golden/vectors.json, produced by hashlib/OpenSSL) and a differential
run against hashlib over padding-boundary and random lengths.
"""
import hashlib

import numpy as np
import pytest

import oracle

ALGS = [(oracle.SHA1, "sha1"), (oracle.MD5, "md5"), (oracle.SHA256, "sha256")]


@pytest.mark.parametrize("alg,name", ALGS)
def test_known_answer_vectors(golden, alg, name):
    msgs = {"empty": b"", "abc": b"abc",
            "448bit": b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq",
            "million_a": b"a" * 1000000}
    for row in golden["kat"]:
        assert oracle.digest(alg, msgs[row["name"]]).hex() == row[name], row["name"]


@pytest.mark.parametrize("alg,name", ALGS)
def test_padding_boundaries(golden, alg, name):
    for row in golden["padding"]:
        msg = bytes(i & 0xFF for i in range(row["len"]))
        assert oracle.digest(alg, msg).hex() == row[name], row["len"]


@pytest.mark.parametrize("alg,name", ALGS)
def test_differential_vs_hashlib(alg, name):
    rng = np.random.default_rng(1234 + alg)
    lens = list(range(0, 260)) + [int(x) for x in rng.integers(0, 200 * 1024, size=60)]
    for n in lens:
        msg = rng.integers(0, 256, size=n, dtype=np.uint8).tobytes()
        assert oracle.digest(alg, msg) == hashlib.new(name, msg).digest(), n


@pytest.mark.parametrize("alg,name", ALGS)
def test_update_split_points_and_midstate(alg, name):
    rng = np.random.default_rng(7)
    msg = rng.integers(0, 256, size=5000, dtype=np.uint8).tobytes()
    want = hashlib.new(name, msg).digest()
    for cut in (0, 1, 63, 64, 65, 127, 128, 4096, 4999, 5000):
        c = oracle.Ctx(alg).update(msg[:cut]).update(msg[cut:])
        assert c.final() == want, cut
    # the exported midstate after whole blocks equals the state of a fresh context fed the same blocks
    a = oracle.Ctx(alg).update(msg[:1024])
    b = oracle.Ctx(alg).update(msg[:512]).update(msg[512:1024])
    assert a.state() == b.state()


def test_generator_goldens(golden):
    g = golden["generator"]
    assert oracle.gen(0, 0, 16).tobytes().hex() == g["seed0_bytes_0_16"]
    assert oracle.gen(1, 0, 16).tobytes().hex() == g["seed1_bytes_0_16"]
    assert oracle.gen(0, 5, 16).tobytes().hex() == g["seed0_bytes_5_21"]
    K = 256 * 1024
    assert oracle.digest(oracle.SHA1, oracle.gen(0, 0, K)).hex() == g["sha1_seed0_piece0_256k"]
    assert oracle.digest(oracle.SHA1, oracle.gen(0, K, K)).hex() == g["sha1_seed0_piece1_256k"]
    assert oracle.digest(oracle.SHA1, oracle.gen(0, 4095 * K, K)).hex() == g["sha1_seed0_piece4095_256k"]
    assert oracle.digest(oracle.SHA1, oracle.gen(0, 0, 4 << 20)).hex() == g["sha1_seed0_piece0_4m"]
    assert oracle.digest(oracle.SHA1, oracle.gen(0, 0, 16384)).hex() == g["sha1_seed0_piece0_16k"]
    assert oracle.digest(oracle.SHA1, oracle.gen(0, 15, 16385)).hex() == g["sha1_seed0_unaligned_15_16385"]
    c = oracle.gen(0, 0, K).copy()
    c[12345] ^= 1
    assert oracle.digest(oracle.SHA1, c).hex() == g["sha1_seed0_piece0_256k_corrupt_12345"]
    assert oracle.digest(oracle.MD5, oracle.gen(0, 0, 1 << 20)).hex() == g["md5_seed0_1m"]
    assert oracle.digest(oracle.SHA256, oracle.gen(0, 0, 1 << 20)).hex() == g["sha256_seed0_1m"]
    for row in g["configs"]:
        L = row["piece_len"]
        d = oracle.hash_msgs(oracle.SHA1, oracle.gen(row["seed"], 0, 2 * L), L)
        assert d[0].tobytes().hex() == row["sha1_piece0"]
        assert d[1].tobytes().hex() == row["sha1_piece1"]


def test_part_goldens_64m(golden):
    g = golden["generator"]
    part = oracle.gen(0, 0, 64 << 20)
    assert oracle.digest(oracle.MD5, part).hex() == g["md5_seed0_part0_64m"]
    assert oracle.digest(oracle.SHA256, part).hex() == g["sha256_seed0_part0_64m"]


def test_s3_etag(golden):
    md5s = oracle.hash_msgs(oracle.MD5, oracle.gen(0, 0, 3 << 20), 1 << 20)
    assert oracle.s3_etag(md5s) == golden["generator"]["s3_etag_seed0_3x1m"]


def test_batched_layout_short_tail_and_mt():
    data = oracle.gen(9, 0, 10 * 16384 + 777)
    one = oracle.hash_msgs(oracle.SHA1, data, 16384, 1)
    many = oracle.hash_msgs(oracle.SHA1, data, 16384, 4)
    assert one.shape == (11, 20) and np.array_equal(one, many)
    assert one[10].tobytes() == hashlib.sha1(data[10 * 16384:].tobytes()).digest()
    assert oracle.count(0, 16384) == 0 and oracle.count(16384, 16384) == 1 and oracle.count(16385, 16384) == 2


def test_verify_bitmap_and_cfg5_rule():
    L = 16384
    data = oracle.gen(5, 0, 300 * L + 55)
    expected = oracle.hash_msgs(oracle.SHA1, data, L)
    n = expected.shape[0]
    bad = []
    dirty = data.copy()
    for i in range(n):
        plen = min(L, data.size - i * L)
        hit, off, bit = oracle.cfg5_corruption(i, plen)
        if hit or i == n - 1:
            dirty[i * L + off] ^= np.uint8(1 << bit)
            bad.append(i)
    bitmap, nbad = oracle.verify_pieces(dirty, L, expected, 2)
    assert nbad == len(bad) and nbad >= 1
    got_bad = [i for i in range(n) if not (int(bitmap[i >> 6]) >> (i & 63)) & 1]
    assert got_bad == bad
    clean_bitmap, clean_bad = oracle.verify_pieces(data, L, expected)
    assert clean_bad == 0
    # the rule itself: ~1 % of indices selected, deterministic
    hits = sum(oracle.cfg5_corruption(i, L)[0] for i in range(20000))
    assert 120 <= hits <= 280


def test_openssl_baseline_agrees_with_the_oracle():
    """oracle/gh_openssl.c (the second CPU baseline, OpenSSL EVP) against the scalar port on ragged batches."""
    import numpy as np
    import pytest
    if oracle.openssl_lib() is None:
        pytest.skip("OpenSSL headers not present: libgh_openssl.so not built")
    data = oracle.gen(17, 5, 3 * 65536 + 4321)
    for L in (1, 55, 64, 1000, 16384, 65536, data.size + 5):
        got = oracle.openssl_hash_msgs(7, data[: min(data.size, 4000 * L)], L, 3)
        for a in (oracle.SHA1, oracle.MD5, oracle.SHA256):
            assert np.array_equal(got[a], oracle.hash_msgs(a, data[: min(data.size, 4000 * L)], L, 2)), (a, L)
