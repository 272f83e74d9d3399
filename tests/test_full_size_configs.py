"""BASELINE.json configs 3, 4 and 5 -- and the 64 KiB-chunk SHA-256 shape the reference's exact upload
call produces over plain HTTP -- at FULL size, through the C ABI, against the CPU oracle
(north_star: "bit-exact ... on identical synthetic inputs, including corrupted and short-tail
pieces"). Configs 1 and 2 at full size are in test_gpu_parity.py. Inputs are the SURVEY.md
Appendix C stream with the per-config seeds; every digest / the whole bitmap is compared, not a
sample, and the oracle runs on all host threads so the file stays within a couple of minutes.

  config 3: 20 GiB, 320 parts of 64 MiB, fused MD5 + SHA-256 (uploader.go:89), S3 multipart ETag
  config 4: 24 files of 2 GiB + 7919(k+1) B, 3 073 pieces of 16 MiB, 23 straddle a file boundary
            (torrent.go:41,79-106), pinned host spans one per file; windowed when host RAM is short
  config 5: 10 000 files, 16 KiB pieces, SURVEY.md 8d fault rule (1 % + the short tail + the 55/56 mod 64
            neighbours): bitmap and n_bad identical to the CPU's
  chunks  : 20 GiB as 327 680 messages of 64 KiB, SHA-256 only (streaming SigV4 over HTTP, uploader.go:32-35)
"""
import ctypes
import os

import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu

THREADS = len(os.sched_getaffinity(0))


def _torch():
    import torch
    return torch


def _mem_available():
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable:"):
            return int(line.split()[1]) * 1024
    return 0


def _gen_device(gh, torch, seed, total):
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    gh.generate(d.data_ptr(), seed, 0, total)
    return d


def _to_pinned(gh, d, offset, nbytes):
    host = gpuhash.PinnedBuffer(nbytes)
    gh.L.gh_memcpy_d2h(gh.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr() + offset), nbytes)
    return host


def test_config3_full_size_fused_md5_sha256_and_etag(gh):
    torch = _torch()
    P, total = 64 << 20, 20 << 30
    n = total // P
    d = _gen_device(gh, torch, 3, total)
    md5 = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
    sha = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    gh.hash_parts_device(d.data_ptr(), total, P, gpuhash.MD5 | gpuhash.SHA256, None, md5.data_ptr(), sha.data_ptr())
    assert gpuhash.last_kernel() == "gh_hash_split<4, 1>"  # 10 message-warps: the fused role-split kernel
    m = md5.cpu().numpy().reshape(n, 16)
    s = sha.cpu().numpy().reshape(n, 32)
    host = _to_pinned(gh, d, 0, total)
    del d
    try:
        assert np.array_equal(host.array[:4096], oracle.gen(3, 0, 4096))  # generator parity at the head ...
        assert np.array_equal(host.array[-4096:], oracle.gen(3, total - 4096, 4096))  # ... and the tail
        assert np.array_equal(m, oracle.hash_msgs(oracle.MD5, host.array, P, THREADS))
        assert np.array_equal(s, oracle.hash_msgs(oracle.SHA256, host.array, P, THREADS))
        assert gh.S3ETag(m) == oracle.s3_etag(m) and gh.S3ETag(m).endswith("-320")
        # and end to end: the same 320 parts from pinned host memory through the chunk-interleaved ring
        got = gh.HashParts([host.array], P, gpuhash.MD5 | gpuhash.SHA256)
        assert np.array_equal(got[gpuhash.MD5], m) and np.array_equal(got[gpuhash.SHA256], s)
    finally:
        host.free()


def test_sha256_64k_chunks_full_size(gh):
    """What the reference's exact call (os.File, known size, zero options) hashes client-side over
    plain HTTP: SHA-256 of every 64 KiB chunk of the body -- 327 680 independent messages per 20 GiB."""
    torch = _torch()
    C, total = 64 << 10, 20 << 30
    n = total // C
    d = _gen_device(gh, torch, 3, total)
    sha = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    gh.hash_parts_device(d.data_ptr(), total, C, gpuhash.SHA256, None, None, sha.data_ptr())
    assert gpuhash.last_kernel().startswith("gh_hash_lanes<4, 1")  # plenty of chains: one message per lane
    got = sha.cpu().numpy().reshape(n, 32)
    # the oracle over a third of the chunks (head, middle, tail windows), every chunk in them
    win = 2 << 30
    for off in (0, (total // 2 // C) * C, total - win):
        h = d[off:off + win].cpu().numpy()
        assert np.array_equal(got[off // C:(off + win) // C], oracle.hash_msgs(oracle.SHA256, h, C, THREADS)), off
    assert len({g.tobytes() for g in got[::997]}) == len(got[::997])
    del d


def test_config4_full_size_pieces_across_file_boundaries(gh):
    torch = _torch()
    L = 16 << 20
    sizes = [(2 << 30) + 7919 * (k + 1) for k in range(24)]
    total = sum(sizes)
    n = (total + L - 1) // L
    assert n == 3073 and total - (n - 1) * L == 2375700
    d = _gen_device(gh, torch, 4, total)
    # host side: the whole stream pinned when RAM allows, else two windows cut at a piece boundary
    avail = _mem_available()
    cuts = [0, total] if avail > total + (40 << 30) else [0, (n // 2) * L, total]
    digests, bitmaps = [], []
    want_all = []
    victim = 1234
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        host = _to_pinned(gh, d, lo, hi - lo)
        try:
            want = oracle.hash_msgs(oracle.SHA1, host.array, L, THREADS)
            want_all.append(want)
            if lo <= victim * L < hi:
                host.array[victim * L - lo + 99] ^= 1
            # one span per file (cut where the window cuts): pieces straddle the spans at odd offsets
            files, o = [], 0
            for s in sizes:
                a, b = max(o, lo), min(o + s, hi)
                if a < b:
                    files.append(host.array[a - lo:b - lo])
                o += s
            bitmap, dig, nbad = gh.VerifyPieces(files, L, want)
            digests.append(dig)
            bitmaps.append((bitmap, want.shape[0], nbad))
        finally:
            host.free()
    del d
    want = np.concatenate(want_all)
    dig = np.concatenate(digests)
    assert dig.shape == (n, 20)
    mask = np.ones(n, dtype=bool)
    mask[victim] = False
    assert np.array_equal(dig[mask], want[mask]) and not np.array_equal(dig[victim], want[victim])
    bad, base = [], 0
    for bitmap, cnt, nbad in bitmaps:
        bad += [base + i for i in range(cnt) if not (int(bitmap[i >> 6]) >> (i & 63)) & 1]
        base += cnt
    assert bad == [victim] and sum(b[2] for b in bitmaps) == 1
    # the 23 pieces that straddle a file boundary are among those compared: name them
    off, straddle = 0, []
    for s in sizes[:-1]:
        off += s
        if off % L:
            straddle.append(off // L)
    assert len(straddle) == 23


def test_config5_full_size_bitmap_identical_to_cpu(gh):
    torch = _torch()
    L = 16 << 10
    sizes = [1 + (oracle.splitmix64((5 << 32) + k) % 2097152) for k in range(10000)]
    total = sum(sizes)
    n = (total + L - 1) // L
    d = _gen_device(gh, torch, 5, total)
    host = _to_pinned(gh, d, 0, total)
    del d
    try:
        expected = oracle.hash_msgs(oracle.SHA1, host.array, L, THREADS)  # from clean data
        assert expected.shape[0] == n
        bad = []
        forced = {n - 1}
        lens = np.minimum(L, total - np.arange(n, dtype=np.int64) * L)
        for want_mod in (55, 56):  # one piece each whose length is 55 / 56 mod 64, if present (only the tail can be)
            hit = np.nonzero(lens % 64 == want_mod)[0]
            if hit.size:
                forced.add(int(hit[0]))
        for i in range(n):
            hit, off, bit = oracle.cfg5_corruption(i, int(lens[i]))
            if hit or i in forced:
                host.array[i * L + off] ^= np.uint8(1 << bit)
                bad.append(i)
        assert 0.005 * n < len(bad) < 0.02 * n
        files, o = [], 0
        for s in sizes:
            files.append(host.array[o:o + s])
            o += s
        bitmap, dig, nbad = gh.VerifyPieces(gpuhash.SpanList(files), L, expected)
        want_bitmap, want_bad = oracle.verify_pieces(host.array, L, expected, THREADS)
        assert nbad == want_bad == len(bad)
        assert np.array_equal(bitmap, want_bitmap)  # bit-identical to the CPU's
        got_bad = np.nonzero(np.unpackbits(bitmap.view(np.uint8), bitorder="little")[:n] == 0)[0]
        assert got_bad.tolist() == bad
        assert np.array_equal(dig, oracle.hash_msgs(oracle.SHA1, host.array, L, THREADS))
    finally:
        host.free()
