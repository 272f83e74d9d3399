"""ctypes loader for the CPU oracle (oracle/gh_oracle.c). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module. The product path (downloader-go_b200/) never does. Parity is unpinned by the reference's
own tests (it has none for this path, see gh_oracle.h); the oracle is pinned to FIPS 180-4 /
RFC 1321 known-answer vectors and to OpenSSL via hashlib in tests/test_oracle.py.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libgh_oracle.so")

SHA1, MD5, SHA256 = 1, 2, 4
DIGEST_LEN = {SHA1: 20, MD5: 16, SHA256: 32}


def build(force=False):
    src = os.path.join(_HERE, "gh_oracle.c")
    hdr = os.path.join(_HERE, "gh_oracle.h")
    if (force or not os.path.exists(_SO)
            or os.path.getmtime(_SO) < max(os.path.getmtime(src), os.path.getmtime(hdr))):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "libgh_oracle.so"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        u8p, u64 = ctypes.c_void_p, ctypes.c_uint64
        L.gho_hash.argtypes = [ctypes.c_int, u8p, ctypes.c_size_t, u8p]
        L.gho_hash.restype = None
        L.gho_splitmix64.argtypes = [u64]
        L.gho_splitmix64.restype = u64
        L.gho_gen_word.argtypes = [u64, u64]
        L.gho_gen_word.restype = u64
        L.gho_gen_fill.argtypes = [u64, u64, u64, u8p]
        L.gho_gen_fill.restype = None
        L.gho_count.argtypes = [u64, u64]
        L.gho_count.restype = u64
        L.gho_hash_msgs_mt.argtypes = [ctypes.c_int, u8p, u64, u64, u8p, ctypes.c_int]
        L.gho_hash_msgs_mt.restype = None
        L.gho_verify_pieces.argtypes = [u8p, u64, u64, u8p, u8p, ctypes.c_int]
        L.gho_verify_pieces.restype = u64
        L.gho_cfg5_corruption.argtypes = [u64, u64, ctypes.POINTER(u64), ctypes.POINTER(ctypes.c_int)]
        L.gho_cfg5_corruption.restype = ctypes.c_int
        L.gho_s3_etag.argtypes = [u8p, u64, ctypes.c_char_p]
        L.gho_s3_etag.restype = None
        L.gho_init.argtypes = [u8p, ctypes.c_int]
        L.gho_update.argtypes = [u8p, u8p, ctypes.c_size_t]
        L.gho_final.argtypes = [u8p, u8p]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data.view(np.uint8).reshape(-1))
    return np.frombuffer(bytes(data), dtype=np.uint8)


def digest(alg, data):
    a = _as_u8(data)
    out = np.empty(DIGEST_LEN[alg], dtype=np.uint8)
    lib().gho_hash(alg, _ptr(a) if a.size else None, a.size, _ptr(out))
    return out.tobytes()


class Ctx:
    """init/update/final with the midstate visible (state words, total) -- mirrors gho_ctx."""

    class _S(ctypes.Structure):
        _fields_ = [("state", ctypes.c_uint32 * 8), ("total", ctypes.c_uint64),
                    ("buf", ctypes.c_uint8 * 64), ("buflen", ctypes.c_uint32), ("alg", ctypes.c_int)]

    def __init__(self, alg):
        self.alg = alg
        self.s = Ctx._S()
        lib().gho_init(ctypes.byref(self.s), alg)

    def update(self, data):
        a = _as_u8(data)
        if a.size:
            lib().gho_update(ctypes.byref(self.s), _ptr(a), a.size)
        return self

    def state(self):
        return list(self.s.state)[: DIGEST_LEN[self.alg] // 4]

    def final(self):
        out = np.empty(DIGEST_LEN[self.alg], dtype=np.uint8)
        lib().gho_final(ctypes.byref(self.s), _ptr(out))
        return out.tobytes()


def gen(seed, offset, length):
    out = np.empty(length, dtype=np.uint8)
    if length:
        lib().gho_gen_fill(seed, offset, length, _ptr(out))
    return out


def splitmix64(x):
    return int(lib().gho_splitmix64(x & 0xFFFFFFFFFFFFFFFF))


def count(total, msg_len):
    return int(lib().gho_count(total, msg_len))


def hash_msgs(alg, data, msg_len, nthreads=1):
    a = _as_u8(data)
    n = count(a.size, msg_len)
    out = np.empty((n, DIGEST_LEN[alg]), dtype=np.uint8)
    if n:
        lib().gho_hash_msgs_mt(alg, _ptr(a), a.size, msg_len, _ptr(out), nthreads)
    return out


def verify_pieces(data, piece_len, expected, nthreads=1):
    a = _as_u8(data)
    e = _as_u8(expected)
    n = count(a.size, piece_len)
    bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
    bad = 0
    if n:
        bad = lib().gho_verify_pieces(_ptr(a), a.size, piece_len, _ptr(e), _ptr(bitmap), nthreads)
    return bitmap, int(bad)


def cfg5_corruption(piece_index, piece_len_i):
    off = ctypes.c_uint64()
    bit = ctypes.c_int()
    hit = lib().gho_cfg5_corruption(piece_index, piece_len_i, ctypes.byref(off), ctypes.byref(bit))
    return bool(hit), int(off.value), int(bit.value)


def s3_etag(part_md5s):
    a = _as_u8(part_md5s)
    buf = ctypes.create_string_buffer(64)
    lib().gho_s3_etag(_ptr(a), a.size // 16, buf)
    return buf.value.decode()


# ---- OpenSSL (SHA-NI) baseline: oracle/gh_openssl.c, optional -------------------------------------
_ossl = None


def openssl_lib():
    """libgh_openssl.so or None when it could not be built (no OpenSSL headers)."""
    global _ossl
    if _ossl is None:
        so = os.path.join(_HERE, "libgh_openssl.so")
        src = os.path.join(_HERE, "gh_openssl.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.call(["make", "-s", "-C", _HERE, "-B", "libgh_openssl.so"])
        if not os.path.exists(so):
            _ossl = False
        else:
            L = ctypes.CDLL(so)
            L.gho_openssl_hash_msgs_mt.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                                   ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
            L.gho_openssl_hash_msgs_mt.restype = ctypes.c_int
            _ossl = L
    return _ossl or None


def openssl_hash_msgs(algs, data, msg_len, nthreads=1):
    """{alg: digests} for every bit of `algs`, computed by OpenSSL EVP on `nthreads` threads."""
    L = openssl_lib()
    if L is None:
        raise RuntimeError("libgh_openssl.so is not available")
    a = _as_u8(data)
    n = count(a.size, msg_len)
    outs = {alg: np.empty((n, DIGEST_LEN[alg]), dtype=np.uint8) for alg in (SHA1, MD5, SHA256) if algs & alg}
    if n:
        L.gho_openssl_hash_msgs_mt(algs, _ptr(a), a.size, msg_len, _ptr(outs[SHA1]) if SHA1 in outs else None,
                                   _ptr(outs[MD5]) if MD5 in outs else None, _ptr(outs[SHA256]) if SHA256 in outs else None,
                                   nthreads)
    return outs
