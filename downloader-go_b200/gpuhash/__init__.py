"""gpuhash -- host-side mirror (ctypes) of the Go package `gpuhash` over libgpuhash.so.

The reference is Go, but no Go toolchain exists in this image, so the host
side that tests and benchmarks drive is this module; go/gpuhash/gpuhash.go is the same surface in
cgo for a maintainer to build. Names follow the Go package: VerifyPieces / HashParts / S3ETag.

What it replaces in the reference (which has no hashing interface of its own):
  VerifyPieces -> anacrolix/torrent Torrent.hashPiece + pieceHashed, reached from
                  internal/downloader/torrent/torrent.go:79-106 (storage seam :41)
  HashParts    -> minio-go hashMaterials (crypto/md5 + sha256-simd), reached from
                  internal/uploader/uploader.go:89

There is NO CPU fallback here: importing works anywhere (so the ABI can be inspected), but
Context() raises GpuHashError unless libgpuhash.so is built and a CUDA device is present.
"""
import ctypes
import os

import numpy as np

SHA1, MD5, SHA256 = 1, 2, 4
DIGEST_LEN = {SHA1: 20, MD5: 16, SHA256: 32}

GH_OK, GH_EINVAL, GH_ENODEVICE, GH_ENOMEM, GH_ECUDA, GH_ECTX_LOST, GH_ERANGE, GH_EIO = range(8)

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("GPUHASH_LIB") or os.path.join(_PKG, "libgpuhash.so")


class GpuHashError(RuntimeError):
    def __init__(self, status, what, detail=""):
        self.status = status
        super().__init__("%s: status %d (%s)%s" % (what, status, _strerror(status), (": " + detail) if detail else ""))


class Span(ctypes.Structure):
    _fields_ = [("host_ptr", ctypes.c_void_p), ("len", ctypes.c_uint64)]


class Config(ctypes.Structure):
    _fields_ = [("n_devices", ctypes.c_int), ("devices", ctypes.POINTER(ctypes.c_int)),
                ("slot_bytes", ctypes.c_uint64), ("slots", ctypes.c_int), ("stage_threads", ctypes.c_int)]


class File(ctypes.Structure):
    _fields_ = [("path", ctypes.c_char_p), ("len", ctypes.c_uint64)]


class BatchSlot(ctypes.Structure):
    _fields_ = [("buffer", ctypes.c_uint32), ("index", ctypes.c_uint32), ("seq", ctypes.c_uint64)]


class BatcherStats(ctypes.Structure):
    _fields_ = [("batches", ctypes.c_uint64), ("pieces", ctypes.c_uint64), ("flushed_full", ctypes.c_uint64),
                ("flushed_by_deadline", ctypes.c_uint64), ("failed_batches", ctypes.c_uint64)]


class Stats(ctypes.Structure):
    _fields_ = [("kernel_launches", ctypes.c_uint64), ("h2d_bytes", ctypes.c_uint64),
                ("d2h_bytes", ctypes.c_uint64), ("kernel_ms", ctypes.c_double), ("wall_ms", ctypes.c_double)]


_lib = None

# name -> (restype, argtypes); every symbol include/gpuhash.h declares
_VP, _U64, _INT, _SZ = ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t
ABI = {
    "gh_config_from_env": (_INT, [ctypes.POINTER(Config), ctypes.POINTER(ctypes.c_int), _INT, ctypes.POINTER(_INT)]),
    "gh_suggest_devices": (_INT, [_INT, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_double)]),
    "gh_create": (_INT, [ctypes.POINTER(_VP), ctypes.POINTER(Config)]),
    "gh_destroy": (None, [_VP]),
    "gh_strerror": (ctypes.c_char_p, [_INT]),
    "gh_last_error": (ctypes.c_char_p, [_VP]),
    "gh_device_count": (_INT, [_VP]),
    "gh_get_stats": (_INT, [_VP, ctypes.POINTER(Stats)]),
    "gh_version": (ctypes.c_char_p, []),
    "gh_last_kernel": (ctypes.c_char_p, []),
    "gh_debug_fail_nth_copy": (_INT, [_VP, ctypes.c_int64]),
    "gh_verify_pieces": (_INT, [_VP, ctypes.POINTER(Span), _SZ, _U64, _VP, _VP, _VP, ctypes.POINTER(_U64)]),
    "gh_hash_parts": (_INT, [_VP, ctypes.POINTER(Span), _SZ, _U64, _INT, _VP, _VP, _VP]),
    "gh_hex_digests": (_INT, [_VP, _U64, _INT, ctypes.c_char_p, _SZ]),
    "gh_s3_etag": (_INT, [_VP, _VP, _U64, ctypes.c_char_p, _SZ]),
    "gh_verify_files": (_INT, [_VP, ctypes.POINTER(File), _SZ, _U64, _VP, _VP, _VP, ctypes.POINTER(_U64)]),
    "gh_hash_file_parts": (_INT, [_VP, ctypes.POINTER(File), _SZ, _U64, _INT, _VP, _VP, _VP]),
    "gh_stream_open": (_INT, [_VP, _INT, _U64, _U64, _VP, ctypes.POINTER(_VP)]),
    "gh_stream_acquire": (_INT, [_VP, ctypes.POINTER(_VP), ctypes.POINTER(_U64)]),
    "gh_stream_submit": (_INT, [_VP, _U64]),
    "gh_stream_finish": (_INT, [_VP, _VP, _VP, _VP, _VP, ctypes.POINTER(_U64)]),
    "gh_stream_abort": (None, [_VP]),
    "gh_batcher_create": (_INT, [_VP, _U64, ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(_VP)]),
    "gh_batcher_verify": (_INT, [_VP, _VP, _U64, _VP, ctypes.POINTER(_INT), _VP]),
    "gh_batcher_reserve": (_INT, [_VP, _U64, ctypes.POINTER(BatchSlot), ctypes.POINTER(_VP)]),
    "gh_batcher_commit": (_INT, [_VP, ctypes.POINTER(BatchSlot), _VP, ctypes.POINTER(_INT), _VP]),
    "gh_batcher_cancel": (_INT, [_VP, ctypes.POINTER(BatchSlot)]),
    "gh_batcher_flush": (_INT, [_VP]),
    "gh_batcher_get_stats": (_INT, [_VP, ctypes.POINTER(BatcherStats)]),
    "gh_batcher_destroy": (None, [_VP]),
    "gh_hash_parts_per_file": (_INT, [_VP, ctypes.POINTER(Span), _SZ, _U64, _INT, _VP, _VP, _VP]),
    "gh_hash_file_parts_per_file": (_INT, [_VP, ctypes.POINTER(File), _SZ, _U64, _INT, _VP, _VP, _VP]),
    "gh_metainfo_parse": (_INT, [_VP, _SZ, ctypes.POINTER(_VP), ctypes.c_char_p, _SZ]),
    "gh_metainfo_free": (None, [_VP]),
    "gh_metainfo_piece_length": (_U64, [_VP]),
    "gh_metainfo_n_pieces": (_U64, [_VP]),
    "gh_metainfo_total_length": (_U64, [_VP]),
    "gh_metainfo_pieces": (_VP, [_VP]),
    "gh_metainfo_name": (ctypes.c_char_p, [_VP]),
    "gh_metainfo_n_files": (_SZ, [_VP]),
    "gh_metainfo_file": (_INT, [_VP, _SZ, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(_U64)]),
    "gh_metainfo_info_hash": (_INT, [_VP, _VP, _VP]),
    "gh_verify_torrent": (_INT, [_VP, _VP, ctypes.c_char_p, _VP, _VP, ctypes.POINTER(_U64)]),
    "gh_verify_pieces_device": (_INT, [_VP, _INT, _VP, _U64, _U64, _VP, _VP, _VP, _VP]),
    "gh_hash_parts_device": (_INT, [_VP, _INT, _VP, _U64, _U64, _INT, _VP, _VP, _VP, _VP]),
    "gh_hash_update_device": (_INT, [_VP, _INT, _INT, _VP, _U64, _U64, _U64, _U64, _U64, _U64, _VP, _VP, _VP, _VP, _VP]),
    "gh_hash_update_device_noalloc": (_INT, [_INT, _INT, _VP, _U64, _U64, _U64, _U64, _U64, _U64, _VP, _VP, _VP, _VP]),
    "gh_generate_device": (_INT, [_VP, _INT, _VP, _U64, _U64, _U64, _VP]),
    "gh_host_alloc": (_INT, [ctypes.POINTER(_VP), _SZ]),
    "gh_host_free": (_INT, [_VP]),
    "gh_host_register": (_INT, [_VP, _SZ]),
    "gh_host_unregister": (_INT, [_VP]),
    "gh_device_alloc": (_INT, [_VP, _INT, ctypes.POINTER(_VP), _SZ]),
    "gh_device_free": (_INT, [_VP, _INT, _VP]),
    "gh_memcpy_d2h": (_INT, [_VP, _INT, _VP, _VP, _SZ]),
    "gh_memcpy_h2d": (_INT, [_VP, _INT, _VP, _VP, _SZ]),
    "gh_shard_range": (None, [_U64, _INT, _INT, ctypes.POINTER(_U64), ctypes.POINTER(_U64)]),
}


def load():
    """dlopen libgpuhash.so and type every exported entry point. No device is touched."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GpuHashError(GH_ENODEVICE, "libgpuhash.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`)")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in ABI.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _strerror(status):
    try:
        return load().gh_strerror(status).decode()
    except Exception:  # library missing: still produce a readable message
        return "?"


NOALLOC_MAX_MSGS = 16384


def hash_update_device_noalloc(cuda_device, algos, d_data, stride, chunk_len, prior, n_msgs, msg_len, last_msg_len,
                               d_sha1=None, d_md5=None, d_sha256=None, stream=None):
    """gh_hash_update_device_noalloc: the resumable hasher for callers that must not allocate device
    memory (midstates sit in a fixed array inside the library, at most NOALLOC_MAX_MSGS messages).
    Needs no Context. Raises GpuHashError(GH_ERANGE) for batches beyond the fixed array."""
    vp = ctypes.c_void_p
    rc = load().gh_hash_update_device_noalloc(cuda_device, algos, vp(d_data), stride, chunk_len, prior, n_msgs, msg_len,
                                              last_msg_len, vp(d_sha1) if d_sha1 else None,
                                              vp(d_md5) if d_md5 else None, vp(d_sha256) if d_sha256 else None,
                                              vp(stream) if stream else None)
    if rc:
        raise GpuHashError(rc, "gh_hash_update_device_noalloc")


class Disabled(RuntimeError):
    """GPUHASH_DISABLE is set: the caller keeps its CPU path (the library has none of its own)."""


def config_from_env():
    """gh_config_from_env -> (devices or None, slot_bytes, slots, stage_threads, disabled)."""
    cfg, devs, dis = Config(), (ctypes.c_int * 64)(), ctypes.c_int()
    rc = load().gh_config_from_env(ctypes.byref(cfg), devs, 64, ctypes.byref(dis))
    if rc:
        raise GpuHashError(rc, "gh_config_from_env")
    return ([devs[i] for i in range(cfg.n_devices)] if cfg.n_devices else None, int(cfg.slot_bytes), int(cfg.slots),
            int(cfg.stage_threads), bool(dis.value))


def suggest_devices(n):
    """gh_suggest_devices -> (list of n CUDA ordinals, aggregate host->device GB/s measured for them)."""
    devs, gbs = (ctypes.c_int * max(n, 1))(), ctypes.c_double()
    rc = load().gh_suggest_devices(n, devs, ctypes.byref(gbs))
    if rc:
        raise GpuHashError(rc, "gh_suggest_devices(%d)" % n)
    return [devs[i] for i in range(n)], float(gbs.value)


def hex_digests(digests):
    """[n, digest_len] uint8 -> list of n lowercase hex strings (gh_hex_digests), chunk order kept."""
    a = np.ascontiguousarray(digests, dtype=np.uint8)
    n, dl = a.shape
    buf = ctypes.create_string_buffer(n * 2 * dl + 1)
    rc = load().gh_hex_digests(_ptr(a) if n else None, n, dl, buf, len(buf))
    if rc:
        raise GpuHashError(rc, "gh_hex_digests")
    txt = buf.value.decode()
    return [txt[2 * dl * i:2 * dl * (i + 1)] for i in range(n)]


def last_kernel():
    """Name of the hashing kernel variant this thread launched last (device-resident calls launch on
    the calling thread; host-buffer calls on a context with one device do too)."""
    return load().gh_last_kernel().decode()


def shard_range(n_msgs, n_shards, shard):
    b, e = ctypes.c_uint64(), ctypes.c_uint64()
    load().gh_shard_range(n_msgs, n_shards, shard, ctypes.byref(b), ctypes.byref(e))
    return int(b.value), int(e.value)


def count(total, msg_len):
    return 0 if total == 0 else (total + msg_len - 1) // msg_len


def _u8(a):
    if isinstance(a, np.ndarray):
        if a.dtype != np.uint8 or not a.flags["C_CONTIGUOUS"]:
            a = np.ascontiguousarray(a).view(np.uint8).reshape(-1)
        return a
    return np.frombuffer(bytes(a), dtype=np.uint8)


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None and a.size else None


class Metainfo:
    """Parsed BEP 3 metainfo (gh_metainfo_*): a .torrent file's bytes or a bare info dictionary."""

    def __init__(self, data):
        L = load()
        a = _u8(data)
        h = ctypes.c_void_p()
        err = ctypes.create_string_buffer(256)
        rc = L.gh_metainfo_parse(_ptr(a), a.size, ctypes.byref(h), err, 256)
        if rc:
            raise GpuHashError(rc, "gh_metainfo_parse", err.value.decode())
        self.h, self.L = h, L
        self.piece_length = int(L.gh_metainfo_piece_length(h))
        self.n_pieces = int(L.gh_metainfo_n_pieces(h))
        self.total_length = int(L.gh_metainfo_total_length(h))
        self.name = os.fsdecode(L.gh_metainfo_name(h))  # bytes from the wire: not necessarily UTF-8
        self.pieces = ctypes.string_at(L.gh_metainfo_pieces(h), 20 * self.n_pieces) if self.n_pieces else b""
        self.files = []
        for i in range(L.gh_metainfo_n_files(h)):
            p, n = ctypes.c_char_p(), ctypes.c_uint64()
            L.gh_metainfo_file(h, i, ctypes.byref(p), ctypes.byref(n))
            self.files.append((os.fsdecode(p.value), int(n.value)))

    def close(self):
        if getattr(self, "h", None):
            self.L.gh_metainfo_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


class PinnedBuffer:
    """Page-locked host memory from gh_host_alloc, viewable as a numpy uint8 array."""

    def __init__(self, nbytes):
        p = ctypes.c_void_p()
        rc = load().gh_host_alloc(ctypes.byref(p), nbytes)
        if rc:
            raise GpuHashError(rc, "gh_host_alloc(%d)" % nbytes, load().gh_last_error(None).decode())
        self.ptr, self.nbytes = p.value, nbytes
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * max(nbytes, 1)).from_address(self.ptr))[:nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            load().gh_host_free(ctypes.c_void_p(self.ptr))
            self.ptr = None


class SpanList:
    """A prebuilt gh_span array over a list of byte buffers (keeps them alive). Building it costs a few
    microseconds per buffer in Python; callers that verify the same files repeatedly build it once."""

    def __init__(self, files):
        self.arrs = [_u8(f) for f in files]
        self.n = len(self.arrs)
        self.sp = (Span * max(self.n, 1))()
        self.total = 0
        for i, a in enumerate(self.arrs):
            self.sp[i].host_ptr = a.ctypes.data if a.size else None
            self.sp[i].len = a.size
            self.total += a.size


class Stream:
    """gh_stream_*: feed a payload window by window into the pinned ring.

        st = ctx.Stream(gpuhash.SHA1, piece_len, total, expected)
        while True:
            buf = st.acquire()            # numpy uint8 view of the pinned window, or None when done
            if buf is None: break
            buf[:] = ...                  # exactly len(buf) bytes
            st.submit()
        bitmap, digests, n_bad = st.finish()
    """

    def __init__(self, ctx, algos, msg_len, total_len, expected=None):
        self.ctx, self.algos, self.n = ctx, algos, count(total_len, msg_len) if msg_len > 0 else 0
        self._exp = _u8(expected) if expected is not None else None
        h = ctypes.c_void_p()
        rc = ctx.L.gh_stream_open(ctx.h, algos, msg_len, total_len, _ptr(self._exp), ctypes.byref(h))
        ctx._check(rc, "gh_stream_open")
        self.h, self._cap = h, 0

    def acquire(self):
        buf, cap = ctypes.c_void_p(), ctypes.c_uint64()
        self.ctx._check(self.ctx.L.gh_stream_acquire(self.h, ctypes.byref(buf), ctypes.byref(cap)), "gh_stream_acquire")
        if not cap.value:
            return None
        self._cap = int(cap.value)
        return np.ctypeslib.as_array((ctypes.c_uint8 * self._cap).from_address(buf.value))

    def submit(self, n_bytes=None):
        self.ctx._check(self.ctx.L.gh_stream_submit(self.h, self._cap if n_bytes is None else n_bytes), "gh_stream_submit")

    def finish(self):
        n = self.n
        outs = {a: np.zeros((n, DIGEST_LEN[a]), dtype=np.uint8) for a in (SHA1, MD5, SHA256) if self.algos & a}
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        nbad = ctypes.c_uint64()
        rc = self.ctx.L.gh_stream_finish(self.h, _ptr(outs.get(SHA1)), _ptr(outs.get(MD5)), _ptr(outs.get(SHA256)),
                                         _ptr(bitmap), ctypes.byref(nbad))
        if rc == GH_OK:
            self.h = None
        self.ctx._check(rc, "gh_stream_finish")
        return bitmap, outs, int(nbad.value)

    def abort(self):
        if self.h:
            self.ctx.L.gh_stream_abort(self.h)
            self.h = None


class Batcher:
    """gh_batcher_*: the micro-batcher a replaced Torrent.hashPiece would call (Go: gpustorage.Batcher).
    verify() blocks until the piece's verdict is known; call it from as many threads as pieces should
    share a batch (ctypes releases the GIL for the duration). A failed batch raises GpuHashError --
    the caller's cue to hash on the CPU -- and is never reported as a mismatch."""

    def __init__(self, ctx, piece_len, max_pieces, max_delay_ms):
        self.ctx = ctx
        h = ctypes.c_void_p()
        ctx._check(ctx.L.gh_batcher_create(ctx.h, piece_len, max_pieces, max_delay_ms, ctypes.byref(h)), "gh_batcher_create")
        self.h = h

    def verify(self, data, expected20, want_digest=False):
        a = _u8(data)
        e = _u8(expected20)
        ok = ctypes.c_int(-1)
        dig = np.zeros(20, dtype=np.uint8) if want_digest else None
        rc = self.ctx.L.gh_batcher_verify(self.h, _ptr(a), a.size, _ptr(e), ctypes.byref(ok), _ptr(dig))
        if rc:
            raise GpuHashError(rc, "gh_batcher_verify", self.ctx.L.gh_last_error(self.ctx.h).decode())
        return (bool(ok.value), dig.tobytes()) if want_digest else bool(ok.value)

    def reserve(self, length):
        """-> (slot, numpy view of the pinned slot of `length` bytes) for in-place producers."""
        slot, buf = BatchSlot(), ctypes.c_void_p()
        rc = self.ctx.L.gh_batcher_reserve(self.h, length, ctypes.byref(slot), ctypes.byref(buf))
        if rc:
            raise GpuHashError(rc, "gh_batcher_reserve")
        view = np.ctypeslib.as_array((ctypes.c_uint8 * max(length, 1)).from_address(buf.value))[:length]
        return slot, view

    def commit(self, slot, expected20):
        e = _u8(expected20)
        ok = ctypes.c_int(-1)
        rc = self.ctx.L.gh_batcher_commit(self.h, ctypes.byref(slot), _ptr(e), ctypes.byref(ok), None)
        if rc:
            raise GpuHashError(rc, "gh_batcher_commit", self.ctx.L.gh_last_error(self.ctx.h).decode())
        return bool(ok.value)

    def cancel(self, slot):
        return self.ctx.L.gh_batcher_cancel(self.h, ctypes.byref(slot))

    def flush(self):
        self.ctx.L.gh_batcher_flush(self.h)

    def stats(self):
        s = BatcherStats()
        self.ctx.L.gh_batcher_get_stats(self.h, ctypes.byref(s))
        return s

    def close(self):
        if getattr(self, "h", None):
            self.ctx.L.gh_batcher_destroy(self.h)
            self.h = None


class Context:
    """One gh_ctx. `devices` is a list of CUDA ordinals (None = all visible)."""

    def __init__(self, devices=None, slot_bytes=0, slots=0, stage_threads=0):
        L = load()
        cfg = Config()
        self._dev_arr = None
        if devices is not None:
            self._dev_arr = (ctypes.c_int * len(devices))(*devices)
            cfg.n_devices = len(devices)
            cfg.devices = ctypes.cast(self._dev_arr, ctypes.POINTER(ctypes.c_int))
        cfg.slot_bytes, cfg.slots, cfg.stage_threads = slot_bytes, slots, stage_threads
        h = ctypes.c_void_p()
        rc = L.gh_create(ctypes.byref(h), ctypes.byref(cfg))
        if rc:
            raise GpuHashError(rc, "gh_create")
        self.h = h
        self.L = L

    @classmethod
    def from_env(cls):
        """The context the environment describes (GPUHASH_DEVICES / _SLOT_MIB / _SLOTS / _STAGE_THREADS);
        raises Disabled when GPUHASH_DISABLE is set -- what the Go adapters do before installing themselves."""
        devices, slot_bytes, slots, stage_threads, disabled = config_from_env()
        if disabled:
            raise Disabled("GPUHASH_DISABLE is set")
        return cls(devices=devices, slot_bytes=slot_bytes, slots=slots, stage_threads=stage_threads)

    def close(self):
        if getattr(self, "h", None):
            self.L.gh_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def _check(self, rc, what):
        if rc:
            raise GpuHashError(rc, what, self.L.gh_last_error(self.h).decode())

    @property
    def n_devices(self):
        return self.L.gh_device_count(self.h)

    def stats(self):
        s = Stats()
        self._check(self.L.gh_get_stats(self.h, ctypes.byref(s)), "gh_get_stats")
        return s

    # ---- host-buffer API (names as in the Go package) ------------------------------------
    @staticmethod
    def _spans(files):
        sl = files if isinstance(files, SpanList) else SpanList(files)
        return sl.arrs, sl.sp, sl.total

    def VerifyPieces(self, files, piece_len, expected=None, want_digests=True):
        """files: list of byte buffers (the torrent's files in info order). Returns
        (bitmap uint64[], digests uint8[n,20] or None, n_bad)."""
        arrs, sp, total = self._spans(files)
        if piece_len <= 0:
            self._check(self.L.gh_verify_pieces(self.h, sp, len(arrs), 0, None, None, None, None), "gh_verify_pieces")
        n = count(total, piece_len)
        exp = _u8(expected) if expected is not None else None
        if exp is not None and exp.size != 20 * n:
            raise GpuHashError(GH_EINVAL, "VerifyPieces: expected must hold 20*n_pieces bytes")
        dig = np.zeros((n, 20), dtype=np.uint8) if want_digests else None
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        nbad = ctypes.c_uint64()
        rc = self.L.gh_verify_pieces(self.h, sp, len(arrs), piece_len, _ptr(exp), _ptr(dig), _ptr(bitmap),
                                     ctypes.byref(nbad))
        self._check(rc, "gh_verify_pieces")
        return bitmap, dig, int(nbad.value)

    def HashParts(self, files, part_len, algos):
        """Returns {algo: uint8[n, digest_len]} for each algorithm bit set in `algos`."""
        arrs, sp, total = self._spans(files)
        n = count(total, part_len) if part_len > 0 else 0
        outs = {a: np.zeros((n, DIGEST_LEN[a]), dtype=np.uint8) for a in (SHA1, MD5, SHA256) if algos & a}
        rc = self.L.gh_hash_parts(self.h, sp, len(arrs), part_len, algos, _ptr(outs.get(SHA1)), _ptr(outs.get(MD5)),
                                  _ptr(outs.get(SHA256)))
        self._check(rc, "gh_hash_parts")
        return outs

    @staticmethod
    def _files(files):
        arr = (File * max(len(files), 1))()
        keep = []
        total = 0
        for i, (path, n) in enumerate(files):
            b = os.fsencode(path)
            keep.append(b)
            arr[i].path, arr[i].len = b, n
            total += n
        return arr, keep, total

    def VerifyFiles(self, files, piece_len, expected=None):
        """files: [(path, length)] in info order; the library reads them itself."""
        arr, keep, total = self._files(files)
        n = count(total, piece_len) if piece_len > 0 else 0
        exp = _u8(expected) if expected is not None else None
        dig = np.zeros((n, 20), dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        nbad = ctypes.c_uint64()
        rc = self.L.gh_verify_files(self.h, arr, len(files), piece_len, _ptr(exp), _ptr(dig), _ptr(bitmap),
                                    ctypes.byref(nbad))
        self._check(rc, "gh_verify_files")
        return bitmap, dig, int(nbad.value)

    def HashFileParts(self, files, part_len, algos):
        arr, keep, total = self._files(files)
        n = count(total, part_len) if part_len > 0 else 0
        outs = {a: np.zeros((n, DIGEST_LEN[a]), dtype=np.uint8) for a in (SHA1, MD5, SHA256) if algos & a}
        rc = self.L.gh_hash_file_parts(self.h, arr, len(files), part_len, algos, _ptr(outs.get(SHA1)),
                                       _ptr(outs.get(MD5)), _ptr(outs.get(SHA256)))
        self._check(rc, "gh_hash_file_parts")
        return outs

    def HashPartsPerFile(self, files, part_len, algos):
        """files: byte buffers, or (path, length) tuples for files on disk. Parts restart at every file.
        Returns ({algo: uint8[n_parts_total, digest_len]}, [n_parts of each file])."""
        on_disk = len(files) > 0 and isinstance(files[0], tuple)
        if on_disk:
            arr, keep, _ = self._files(files)
            lens = [n for _, n in files]
        else:
            keep, arr, _ = self._spans(files)
            lens = [a.size for a in keep]
        counts = [count(n, part_len) for n in lens]
        n = sum(counts)
        outs = {a: np.zeros((n, DIGEST_LEN[a]), dtype=np.uint8) for a in (SHA1, MD5, SHA256) if algos & a}
        fn = self.L.gh_hash_file_parts_per_file if on_disk else self.L.gh_hash_parts_per_file
        rc = fn(self.h, arr, len(files), part_len, algos, _ptr(outs.get(SHA1)), _ptr(outs.get(MD5)), _ptr(outs.get(SHA256)))
        self._check(rc, "gh_hash_parts_per_file")
        return outs, counts

    def InfoHash(self, metainfo):
        out = np.zeros(20, dtype=np.uint8)
        self._check(self.L.gh_metainfo_info_hash(self.h, metainfo.h, _ptr(out)), "gh_metainfo_info_hash")
        return out.tobytes()

    def VerifyTorrent(self, metainfo, base_dir):
        n = metainfo.n_pieces
        dig = np.zeros((n, 20), dtype=np.uint8)
        bitmap = np.zeros((n + 63) // 64, dtype=np.uint64)
        nbad = ctypes.c_uint64()
        rc = self.L.gh_verify_torrent(self.h, metainfo.h, os.fsencode(base_dir), _ptr(dig), _ptr(bitmap),
                                      ctypes.byref(nbad))
        self._check(rc, "gh_verify_torrent")
        return bitmap, dig, int(nbad.value)

    def Stream(self, algos, msg_len, total_len, expected=None):
        return Stream(self, algos, msg_len, total_len, expected)

    def Batcher(self, piece_len, max_pieces, max_delay_ms):
        return Batcher(self, piece_len, max_pieces, max_delay_ms)

    def S3ETag(self, part_md5s):
        a = _u8(part_md5s)
        buf = ctypes.create_string_buffer(64)
        self._check(self.L.gh_s3_etag(self.h, _ptr(a), a.size // 16, buf, 64), "gh_s3_etag")
        return buf.value.decode()

    # ---- device-resident API (raw device addresses as ints) --------------------------------
    def device_alloc(self, nbytes, dev=0):
        p = ctypes.c_void_p()
        self._check(self.L.gh_device_alloc(self.h, dev, ctypes.byref(p), nbytes), "gh_device_alloc")
        return p.value

    def device_free(self, ptr, dev=0):
        self._check(self.L.gh_device_free(self.h, dev, ctypes.c_void_p(ptr)), "gh_device_free")

    def d2h(self, ptr, nbytes, dev=0):
        out = np.empty(nbytes, dtype=np.uint8)
        if nbytes:
            self._check(self.L.gh_memcpy_d2h(self.h, dev, _ptr(out), ctypes.c_void_p(ptr), nbytes), "gh_memcpy_d2h")
        return out

    def h2d(self, ptr, data, dev=0):
        a = _u8(data)
        if a.size:
            self._check(self.L.gh_memcpy_h2d(self.h, dev, ctypes.c_void_p(ptr), _ptr(a), a.size), "gh_memcpy_h2d")

    def generate(self, d_out, seed, offset, length, dev=0, stream=None):
        self._check(self.L.gh_generate_device(self.h, dev, ctypes.c_void_p(d_out), seed, offset, length,
                                              ctypes.c_void_p(stream) if stream else None), "gh_generate_device")

    def verify_pieces_device(self, d_data, total, piece_len, d_expected=None, d_digests=None, d_ok=None, dev=0,
                             stream=None):
        vp = ctypes.c_void_p
        self._check(self.L.gh_verify_pieces_device(self.h, dev, vp(d_data), total, piece_len,
                                                   vp(d_expected) if d_expected else None,
                                                   vp(d_digests) if d_digests else None,
                                                   vp(d_ok) if d_ok else None,
                                                   vp(stream) if stream else None), "gh_verify_pieces_device")

    def hash_parts_device(self, d_data, total, part_len, algos, d_sha1=None, d_md5=None, d_sha256=None, dev=0,
                          stream=None):
        vp = ctypes.c_void_p
        self._check(self.L.gh_hash_parts_device(self.h, dev, vp(d_data), total, part_len, algos,
                                                vp(d_sha1) if d_sha1 else None, vp(d_md5) if d_md5 else None,
                                                vp(d_sha256) if d_sha256 else None,
                                                vp(stream) if stream else None), "gh_hash_parts_device")

    def hash_update_device(self, algos, d_data, stride, chunk_len, prior, n_msgs, msg_len, last_msg_len, d_state,
                           d_sha1=None, d_md5=None, d_sha256=None, dev=0, stream=None):
        vp = ctypes.c_void_p
        self._check(self.L.gh_hash_update_device(self.h, dev, algos, vp(d_data), stride, chunk_len, prior, n_msgs,
                                                 msg_len, last_msg_len, vp(d_state),
                                                 vp(d_sha1) if d_sha1 else None, vp(d_md5) if d_md5 else None,
                                                 vp(d_sha256) if d_sha256 else None,
                                                 vp(stream) if stream else None), "gh_hash_update_device")
