"""Error contract of the ring (include/gpuhash.h, "Error contract"): a failed call has drained the
device before it returns and leaves the context usable; file errors are reported as errors and not
folded into "piece failed"; and the gather paths that mix direct DMA with staged copies in one window.
All through the C ABI, digests bit-exact against the oracle."""
import ctypes
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_pinned_span_between_pageable_spans_in_one_window(gh):
    """pageable | pinned | pageable inside one window (and inside single pieces): the pinned span goes
    by direct DMA, the others through the slot's pinned mirror -- the mirror's copy must not cover the
    pinned span's bytes."""
    L = 64 * 1024
    sizes = [5 * L + 1000, 7 * L - 300, 4 * L + 77, 3 * L, 2 * L - 777]
    total = sum(sizes)
    data = oracle.gen(51, 0, total)
    pinned = [gpuhash.PinnedBuffer(sizes[1]), gpuhash.PinnedBuffer(sizes[3])]
    try:
        parts, o, k = [], 0, 0
        for i, s in enumerate(sizes):
            if i in (1, 3):
                pinned[k].array[:] = data[o:o + s]
                parts.append(pinned[k].array)
                k += 1
            else:
                parts.append(data[o:o + s].copy())
            o += s
        want = oracle.hash_msgs(oracle.SHA1, data, L, 4)
        for _ in range(2):  # second run: the mirrors now hold stale bytes of the first
            bitmap, dig, nbad = gh.VerifyPieces(parts, L, want)
            assert nbad == 0 and np.array_equal(dig, want)
        # chunk-interleaved windows (messages longer than the ring slot) with the same mix
        ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
        try:
            P = 5 * L + 64
            got = ctx.HashParts(parts, P, gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256)
            for a in (oracle.SHA1, oracle.MD5, oracle.SHA256):
                assert np.array_equal(got[a], oracle.hash_msgs(a, data, P, 4)), a
        finally:
            ctx.close()
    finally:
        for b in pinned:
            b.free()


@pytest.mark.parametrize("mode", ["whole", "interleaved", "pinned"])
def test_forced_copy_failure_leaves_context_usable(mode):
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
    store = None
    try:
        L = 16384 if mode == "whole" else 300 * 1024 + 64
        total = 40 * (1 << 20) + 12345 if mode == "whole" else 9 * L + 500
        data = oracle.gen(52, 0, total)
        src = data
        if mode == "pinned":
            store = gpuhash.PinnedBuffer(total)
            store.array[:] = data
            src = store.array
        want = oracle.hash_msgs(oracle.SHA1, data, L, 4)
        for nth in (0, 1, 2):  # every mode makes at least three payload copies per call (one per window)
            assert ctx.L.gh_debug_fail_nth_copy(ctx.h, nth) == 0
            with pytest.raises(gpuhash.GpuHashError) as e:
                ctx.VerifyPieces([src[:total // 2], src[total // 2:]], L, want)
            assert e.value.status == gpuhash.GH_ECUDA, e.value
            assert "cudaErrorUnknown" in str(e.value)
            ctx.L.gh_debug_fail_nth_copy(ctx.h, -1)
            # the same context, straight after: bit-exact, nothing stale in the ring
            bitmap, dig, nbad = ctx.VerifyPieces([src], L, want)
            assert nbad == 0 and np.array_equal(dig, want), (mode, nth)
    finally:
        if store:
            store.free()
        ctx.close()


def test_forced_failure_in_stream_submit_then_abort():
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
    try:
        L, total = 16384, 300 * 16384
        data = oracle.gen(53, 0, total)
        st = ctx.Stream(gpuhash.SHA1, L, total)
        buf = st.acquire()
        buf[:] = data[:len(buf)]
        st.submit()
        buf = st.acquire()
        buf[:] = 0
        ctx.L.gh_debug_fail_nth_copy(ctx.h, 0)
        with pytest.raises(gpuhash.GpuHashError) as e:
            st.submit()
        assert e.value.status == gpuhash.GH_ECUDA
        ctx.L.gh_debug_fail_nth_copy(ctx.h, -1)
        with pytest.raises(gpuhash.GpuHashError):  # not everything was fed: finish refuses, the stream stays abortable
            st.finish()
        st.abort()
        got = ctx.HashParts([data], L, gpuhash.SHA1)[gpuhash.SHA1]
        assert np.array_equal(got, oracle.hash_msgs(oracle.SHA1, data, L, 4))
    finally:
        ctx.close()


def test_unreadable_file_is_an_error_not_a_bad_piece(gh, tmp_path):
    L = 16384
    good = oracle.gen(54, 0, 10 * L)
    p = tmp_path / "a.bin"
    p.write_bytes(good.tobytes())
    d = tmp_path / "is_a_directory"
    d.mkdir()
    want = oracle.hash_msgs(oracle.SHA1, np.concatenate([good, np.zeros(5 * L, np.uint8)]), L, 2)
    with pytest.raises(gpuhash.GpuHashError) as e:
        gh.VerifyFiles([(str(p), 10 * L), (str(d), 5 * L)], L, want)
    assert e.value.status == gpuhash.GH_EIO and "is_a_directory" in str(e.value)
    # a file that does not exist is a hole, not an error: its pieces fail, the others pass
    bitmap, dig, nbad = gh.VerifyFiles([(str(p), 10 * L), (str(tmp_path / "missing.bin"), 5 * L)], L, want)
    assert nbad == 0  # expected digests were computed over zeros for the missing part
    want2 = want.copy()
    want2[12, 3] ^= 1
    bitmap, dig, nbad = gh.VerifyFiles([(str(p), 10 * L), (str(tmp_path / "missing.bin"), 5 * L)], L, want2)
    assert nbad == 1 and not (int(bitmap[0]) >> 12) & 1


def test_more_files_than_file_descriptors(tmp_path):
    """A torrent of many small files (BASELINE config 5 has 10 000) must not depend on RLIMIT_NOFILE:
    files are opened as windows need them, a bounded number at a time. Child process with a soft limit
    of 80 descriptors above what the process already holds, verifying 700 files, whole-message windows and chunk-interleaved ones."""
    code = textwrap.dedent("""
        import os, resource, sys
        sys.path[:0] = [%r, %r]
        import numpy as np
        import gpuhash, oracle
        base = sys.argv[1]
        sizes = [1 + (oracle.splitmix64((5 << 32) + k) %% 40000) for k in range(700)]
        data = oracle.gen(5, 0, sum(sizes))
        files, o = [], 0
        for k, s in enumerate(sizes):
            path = os.path.join(base, "f%%04d" %% k)
            with open(path, "wb") as f:
                f.write(data[o:o + s].tobytes())
            files.append((path, s))
            o += s
        hard = resource.getrlimit(resource.RLIMIT_NOFILE)[1]
        ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
        resource.setrlimit(resource.RLIMIT_NOFILE, (max(int(f) for f in os.listdir('/proc/self/fd')) + 80, hard))
        for L in (16384, 3 << 20):
            want = oracle.hash_msgs(oracle.SHA1, data, L, 4)
            bitmap, dig, nbad = ctx.VerifyFiles(files, L, want)
            assert nbad == 0 and np.array_equal(dig, want), L
        ctx.close()
        print("ok")
    """) % (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200"))
    r = subprocess.run([sys.executable, "-c", code, str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
