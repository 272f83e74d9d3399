"""Regenerates tests/golden/vectors.json.

Independent of oracle/ and of the CUDA library on purpose: digests come from OpenSSL through
Python's hashlib, the synthetic stream from a numpy restatement of SURVEY.md Appendix C. The
reference (/root/reference) has no digest fixtures to import (its only test is
internal/process/process_test.go:13-62), so these vectors pin the oracle to the public standards
(FIPS 180-4, RFC 1321) that Go's crypto/sha1, crypto/md5 and crypto/sha256 implement.

Run:  python tests/golden/make_golden.py
"""
import hashlib
import json
import os

import numpy as np

G = np.uint64(0x9E3779B97F4A7C15)


def gen(seed, offset, length):
    """Bytes [offset, offset+length) of stream(seed); SURVEY.md Appendix C."""
    first, last = offset >> 3, (offset + length + 7) >> 3
    with np.errstate(over="ignore"):
        i = np.arange(first, last, dtype=np.uint64)
        z = np.uint64(seed) + (i + np.uint64(1)) * G
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    b = z.astype("<u8").view(np.uint8)
    s = offset - (first << 3)
    return b[s:s + length]


def hx(alg, data):
    return hashlib.new(alg, bytes(data)).hexdigest()


def main():
    v = {"source": "hashlib (OpenSSL) + numpy splitmix64; see make_golden.py"}
    kat = []
    for name, msg in (("empty", b""), ("abc", b"abc"),
                      ("448bit", b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq"),
                      ("million_a", b"a" * 1000000)):
        kat.append({"name": name, "len": len(msg), "sha1": hx("sha1", msg), "md5": hx("md5", msg),
                    "sha256": hx("sha256", msg)})
    v["kat"] = kat
    pad = []
    for n in (1, 55, 56, 57, 63, 64, 65, 119, 120, 127, 128, 16383, 16384, 16385):
        msg = bytes(i & 0xFF for i in range(n))
        pad.append({"len": n, "sha1": hx("sha1", msg), "md5": hx("md5", msg), "sha256": hx("sha256", msg)})
    v["padding"] = pad
    g = {"seed0_bytes_0_16": gen(0, 0, 16).tobytes().hex(),
         "seed1_bytes_0_16": gen(1, 0, 16).tobytes().hex(),
         "seed0_bytes_5_21": gen(0, 5, 16).tobytes().hex()}
    K = 256 * 1024
    g["sha1_seed0_piece0_256k"] = hx("sha1", gen(0, 0, K))
    g["sha1_seed0_piece1_256k"] = hx("sha1", gen(0, K, K))
    g["sha1_seed0_piece4095_256k"] = hx("sha1", gen(0, 4095 * K, K))
    g["sha1_seed0_piece0_4m"] = hx("sha1", gen(0, 0, 4 << 20))
    g["sha1_seed0_piece0_16k"] = hx("sha1", gen(0, 0, 16384))
    g["sha1_seed0_unaligned_15_16385"] = hx("sha1", gen(0, 15, 16385))
    c = bytearray(gen(0, 0, K).tobytes())
    c[12345] ^= 1
    g["sha1_seed0_piece0_256k_corrupt_12345"] = hx("sha1", c)
    g["md5_seed0_1m"] = hx("md5", gen(0, 0, 1 << 20))
    g["sha256_seed0_1m"] = hx("sha256", gen(0, 0, 1 << 20))
    g["md5_seed0_part0_64m"] = hx("md5", gen(0, 0, 64 << 20))
    g["sha256_seed0_part0_64m"] = hx("sha256", gen(0, 0, 64 << 20))
    parts = b"".join(hashlib.md5(bytes(gen(0, k << 20, 1 << 20))).digest() for k in range(3))
    g["s3_etag_seed0_3x1m"] = hashlib.md5(parts).hexdigest() + "-3"
    # first / second piece of each BASELINE.json config's stream (seeds 1..5)
    cfg = []
    for seed, L in ((1, 256 << 10), (2, 4 << 20), (4, 16 << 20), (5, 16 << 10)):
        cfg.append({"seed": seed, "piece_len": L, "sha1_piece0": hx("sha1", gen(seed, 0, L)),
                    "sha1_piece1": hx("sha1", gen(seed, L, L))})
    g["configs"] = cfg
    g["cfg3_seed3_1m"] = {"md5": hx("md5", gen(3, 0, 1 << 20)), "sha256": hx("sha256", gen(3, 0, 1 << 20))}
    v["generator"] = g
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vectors.json")
    with open(out, "w") as f:
        json.dump(v, f, indent=1)
    print("wrote", out)


if __name__ == "__main__":
    main()
