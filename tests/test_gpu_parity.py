"""GPU parity: the CUDA path, called through the C ABI (ctypes mirror in downloader-go_b200/gpuhash),
against the CPU oracle on identical seeded inputs. Bit-exact is the bar for every case (integer /
byte work; no tolerance). Layout follows what the reference's callers would present: pieces of a
concatenated multi-file stream (BEP 3; torrent.go:41-106) and parts of one file (uploader.go:89)."""
import hashlib
import os

import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu

ALGS = [(gpuhash.SHA1, "sha1"), (gpuhash.MD5, "md5"), (gpuhash.SHA256, "sha256")]


def _torch():
    import torch
    return torch


def _bits(bitmap, n):
    return [(int(bitmap[i >> 6]) >> (i & 63)) & 1 for i in range(n)]


# ------------------------------------------------------------------ known answers through the ABI
@pytest.mark.parametrize("alg,name", ALGS)
def test_known_answer_vectors(gh, golden, alg, name):
    msgs = {"abc": b"abc", "448bit": b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq",
            "million_a": b"a" * 1000000}
    for row in golden["kat"]:
        if row["name"] == "empty":
            continue  # an empty stream has no pieces; covered by test_empty_inputs
        msg = msgs[row["name"]]
        out = gh.HashParts([msg], len(msg), alg)[alg]
        assert out.shape == (1, gpuhash.DIGEST_LEN[alg])
        assert out[0].tobytes().hex() == row[name], row["name"]


@pytest.mark.parametrize("alg,name", ALGS)
def test_padding_boundaries(gh, golden, alg, name):
    for row in golden["padding"]:
        n = row["len"]
        msg = bytes(i & 0xFF for i in range(n))
        # three equal messages and a 1-byte tail: exercises the uniform-stride layout at odd strides
        out = gh.HashParts([msg * 3 + b"\x07"], n, alg)[alg]
        assert out.shape[0] == 4
        for k in range(3):
            assert out[k].tobytes().hex() == row[name], (n, k)
        assert out[3].tobytes() == hashlib.new(name, b"\x07").digest()


def test_empty_inputs(gh):
    bitmap, dig, nbad = gh.VerifyPieces([], 16384, None)
    assert bitmap.size == 0 and dig.shape == (0, 20) and nbad == 0
    bitmap, dig, nbad = gh.VerifyPieces([b"", b""], 16384, None)
    assert dig.shape == (0, 20)
    assert gh.HashParts([b""], 1 << 20, gpuhash.MD5 | gpuhash.SHA256)[gpuhash.MD5].shape == (0, 16)


def test_error_behaviour(gh):
    with pytest.raises(gpuhash.GpuHashError) as e:
        gh.VerifyPieces([b"abc"], 0, None)
    assert e.value.status == gpuhash.GH_EINVAL
    with pytest.raises(gpuhash.GpuHashError) as e:
        gh.HashParts([b"abc"], 3, 0)
    assert e.value.status == gpuhash.GH_EINVAL
    with pytest.raises(gpuhash.GpuHashError) as e:
        gh.HashParts([b"abc"], 3, 8)
    assert e.value.status == gpuhash.GH_EINVAL
    # the context stays usable after a rejected call
    assert gh.HashParts([b"abc"], 3, gpuhash.SHA1)[gpuhash.SHA1][0].tobytes() == hashlib.sha1(b"abc").digest()


# ------------------------------------------------------------------ differential vs the oracle
@pytest.mark.parametrize("algos", [gpuhash.SHA1, gpuhash.MD5, gpuhash.SHA256, gpuhash.MD5 | gpuhash.SHA256,
                                   gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256])
@pytest.mark.parametrize("msg_len,total", [(64, 64 * 100), (1000, 1000 * 77 + 13), (16384, 16384 * 130 + 55),
                                           (16400, 16400 * 40), (65536, 65536 * 33 + 56), (262144, 262144 * 9 + 1),
                                           (4096 + 16, (4096 + 16) * 64 + 63), (120, 120 * 1000 + 119)])
def test_differential_batches(gh, algos, msg_len, total):
    data = oracle.gen(77 + msg_len, 3, total)  # offset 3: host buffer content is arbitrary
    got = gh.HashParts([data], msg_len, algos)
    for alg, _ in ALGS:
        if algos & alg:
            want = oracle.hash_msgs(alg, data, msg_len, 4)
            assert np.array_equal(got[alg], want), (alg, msg_len)


def test_verify_bitmap_cfg5_rule_multi_file(gh):
    """Config 5 at reduced size: 16 KiB pieces over many small files, 1 % injected corruption plus
    the short tail and the 55/56-mod-64 neighbours; bitmap and n_bad identical to the CPU's."""
    L = 16384
    sizes = [1 + (oracle.splitmix64((5 << 32) + k) % 200000) for k in range(120)]
    total = sum(sizes)
    clean = oracle.gen(5, 0, total)
    expected = oracle.hash_msgs(oracle.SHA1, clean, L, 4)
    n = expected.shape[0]
    dirty = clean.copy()
    bad = set()
    for i in range(n):
        plen = min(L, total - i * L)
        hit, off, bit = oracle.cfg5_corruption(i, plen)
        if hit or i == n - 1 or i in (3, 64, 65):
            dirty[i * L + off] ^= np.uint8(1 << bit)
            bad.add(i)
    files, o = [], 0
    for s in sizes:
        files.append(dirty[o:o + s])
        o += s
    bitmap, dig, nbad = gh.VerifyPieces(files, L, expected)
    want_bitmap, want_bad = oracle.verify_pieces(dirty, L, expected, 4)
    assert nbad == want_bad == len(bad)
    assert np.array_equal(bitmap, want_bitmap)
    assert [i for i, b in enumerate(_bits(bitmap, n)) if not b] == sorted(bad)
    assert np.array_equal(dig, oracle.hash_msgs(oracle.SHA1, dirty, L, 4))


def test_season_pack_layout_unaligned_file_boundaries(gh):
    """Config 4 at reduced size: files of size 2^21 + 7919*(k+1), 1 MiB pieces spanning boundaries
    that fall at odd stream offsets; spans are pageable numpy buffers (staging ring path)."""
    L = 1 << 20
    sizes = [(1 << 21) + 7919 * (k + 1) for k in range(6)]
    total = sum(sizes)
    data = oracle.gen(4, 0, total)
    files, o = [], 0
    for s in sizes:
        files.append(data[o:o + s].copy())
        o += s
    expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
    bitmap, dig, nbad = gh.VerifyPieces(files, L, expected)
    assert nbad == 0 and np.array_equal(dig, expected)
    assert all(_bits(bitmap, expected.shape[0]))


def test_pinned_spans_direct_dma(gh):
    L = 256 * 1024
    total = 40 * L + 4321
    buf = gpuhash.PinnedBuffer(total)
    try:
        buf.array[:] = oracle.gen(1, 0, total)
        expected = oracle.hash_msgs(oracle.SHA1, buf.array, L, 4)
        expected[7, 0] ^= 0xFF
        bitmap, dig, nbad = gh.VerifyPieces([buf.array], L, expected)
        assert nbad == 1 and _bits(bitmap, 41)[7] == 0 and sum(_bits(bitmap, 41)) == 40
    finally:
        buf.free()


@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("algos", [gpuhash.SHA1, gpuhash.MD5 | gpuhash.SHA256, gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256])
def test_chunk_interleaved_ring(algos, pinned):
    """Messages larger than the ring window: the host ring streams one sub-chunk of every message
    per window and the kernel carries midstates between launches (config 3's regime)."""
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
    try:
        L = 300 * 1024 + 64
        sizes = [5 * L + 17, 3 * L, 2 * L + 100000]  # file boundaries inside messages
        total = sum(sizes)
        data = oracle.gen(3, 0, total)
        store = gpuhash.PinnedBuffer(total) if pinned else None
        src = data
        if pinned:
            store.array[:] = data
            src = store.array
        files, o = [], 0
        for s in sizes:
            files.append(src[o:o + s])
            o += s
        got = ctx.HashParts(files, L, algos)
        for alg, _ in ALGS:
            if algos & alg:
                assert np.array_equal(got[alg], oracle.hash_msgs(alg, data, L, 4)), alg
        if algos & gpuhash.SHA1:
            expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
            expected[2, 5] ^= 1
            bitmap, dig, nbad = ctx.VerifyPieces(files, L, expected)
            n = expected.shape[0]
            assert nbad == 1 and [i for i, b in enumerate(_bits(bitmap, n)) if not b] == [2]
        if store:
            store.free()
    finally:
        ctx.close()


def test_s3_etag(gh, golden):
    data = oracle.gen(0, 0, 3 << 20)
    md5s = gh.HashParts([data], 1 << 20, gpuhash.MD5)[gpuhash.MD5]
    assert gh.S3ETag(md5s) == golden["generator"]["s3_etag_seed0_3x1m"]


# ------------------------------------------------------------------ device-resident entry points
def test_generator_matches_oracle(gh):
    torch = _torch()
    for seed, off, n in ((0, 0, 4096), (1, 5, 100001), (5, 123456789, 1 << 20), (2, 8, 7)):
        d = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
        d.zero_()
        gh.generate(d.data_ptr() + 16, seed, off, n)
        got = d[16:16 + n].cpu().numpy()
        assert np.array_equal(got, oracle.gen(seed, off, n)), (seed, off, n)
        assert int(d[:16].sum()) == 0 and int(d[16 + n:].sum()) == 0


def test_device_verify_cfg1_full_size(gh, golden):
    """Config 1 at full size (1 GiB, 256 KiB pieces) resident in HBM, checked piece by piece."""
    torch = _torch()
    L, total = 256 << 10, 1 << 30
    n = total // L
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    gh.generate(d.data_ptr(), 1, 0, total)
    dig = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    gh.verify_pieces_device(d.data_ptr(), total, L, None, dig.data_ptr(), None)
    got = dig.cpu().numpy().reshape(n, 20)
    host = d.cpu().numpy()
    want = oracle.hash_msgs(oracle.SHA1, host, L, 8)
    assert np.array_equal(got, want)
    row = [r for r in golden["generator"]["configs"] if r["seed"] == 1][0]
    assert got[0].tobytes().hex() == row["sha1_piece0"] and got[1].tobytes().hex() == row["sha1_piece1"]
    # verification with corrupted pieces, bitmap on the device
    exp = torch.from_numpy(want.reshape(-1).copy()).cuda()
    ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
    bad = [0, 31, 32, 1000, n - 1]
    for i in bad:
        d[i * L + (i * 7919) % L] ^= 0x10
    gh.verify_pieces_device(d.data_ptr(), total, L, exp.data_ptr(), dig.data_ptr(), ok.data_ptr())
    words = ok.cpu().numpy().view(np.uint32)
    got_bad = [i for i in range(n) if not (int(words[i >> 5]) >> (i & 31)) & 1]
    assert got_bad == bad


def test_device_cfg2_full_size_properties(gh):
    """Config 2 at full size (50 GiB, 4 MiB pieces, 12 800 pieces) when HBM allows; size-independent
    checks: sampled pieces equal the oracle, a corrupted piece flips exactly its own bit, a second
    run is identical (idempotence)."""
    torch = _torch()
    L = 4 << 20
    free, _ = torch.cuda.mem_get_info()
    total = 50 << 30
    if free < total + (2 << 30):
        total = (free // 2 // L) * L
    n = total // L
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    gh.generate(d.data_ptr(), 2, 0, total)
    dig = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    gh.verify_pieces_device(d.data_ptr(), total, L, None, dig.data_ptr(), None)
    got = dig.cpu().numpy().reshape(n, 20).copy()
    for i in (0, 1, 31, 32, n // 2, n - 1):
        assert got[i].tobytes() == oracle.digest(oracle.SHA1, oracle.gen(2, i * L, L)), i
    assert len({g.tobytes() for g in got}) == n  # all distinct
    ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
    exp = dig.clone()
    victim = n - 3
    d[victim * L + 4097] ^= 1
    gh.verify_pieces_device(d.data_ptr(), total, L, exp.data_ptr(), dig.data_ptr(), ok.data_ptr())
    words = ok.cpu().numpy().view(np.uint32)
    assert [i for i in range(n) if not (int(words[i >> 5]) >> (i & 31)) & 1] == [victim]
    again = dig.cpu().numpy().reshape(n, 20)
    mask = np.ones(n, dtype=bool)
    mask[victim] = False
    assert np.array_equal(again[mask], got[mask]) and not np.array_equal(again[victim], got[victim])
    del d


def test_device_parts_cfg3_reduced_and_golden(gh, golden):
    """Config 3's shape at reduced count: 64 MiB parts, fused MD5 + SHA-256, seed 0 golden for part 0."""
    torch = _torch()
    P, nparts = 64 << 20, 6
    total = P * nparts - 12345  # short last part
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    gh.generate(d.data_ptr(), 0, 0, total)
    md5 = torch.zeros(nparts * 16, dtype=torch.uint8, device="cuda")
    sha = torch.zeros(nparts * 32, dtype=torch.uint8, device="cuda")
    gh.hash_parts_device(d.data_ptr(), total, P, gpuhash.MD5 | gpuhash.SHA256, None, md5.data_ptr(), sha.data_ptr())
    m = md5.cpu().numpy().reshape(nparts, 16)
    s = sha.cpu().numpy().reshape(nparts, 32)
    g = golden["generator"]
    assert m[0].tobytes().hex() == g["md5_seed0_part0_64m"]
    assert s[0].tobytes().hex() == g["sha256_seed0_part0_64m"]
    host = d.cpu().numpy()
    assert np.array_equal(m, oracle.hash_msgs(oracle.MD5, host, P, 6))
    assert np.array_equal(s, oracle.hash_msgs(oracle.SHA256, host, P, 6))


@pytest.mark.parametrize("algos", [gpuhash.SHA1, gpuhash.MD5, gpuhash.SHA256, gpuhash.MD5 | gpuhash.SHA256,
                                   gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256])
def test_resumable_update_device(gh, algos):
    """init/update/final with midstates on the device: three unequal windows, short last message."""
    torch = _torch()
    n, L, last = 70, 5 * 4096 + 100, 4096 + 33
    total = (n - 1) * L + last
    host = oracle.gen(11, 0, total)
    d = torch.from_numpy(host).cuda()
    state = torch.zeros(n * 17, dtype=torch.int32, device="cuda")
    outs = {a: torch.zeros(n * gpuhash.DIGEST_LEN[a], dtype=torch.uint8, device="cuda") for a, _ in ALGS}
    prior = 0
    for chunk in (4096, 8192, 2 * 4096 + 128):
        # window layout: message m's bytes for this call at base + m*chunk (packed like a ring slot)
        win = torch.zeros(n * chunk + 64, dtype=torch.uint8, device="cuda")
        for mi in range(n):
            mlen = last if mi == n - 1 else L
            lo, hi = min(prior, mlen), min(prior + chunk, mlen)
            if hi > lo:
                win[mi * chunk: mi * chunk + (hi - lo)] = d[mi * L + lo: mi * L + hi]
        gh.hash_update_device(algos, win.data_ptr(), chunk, chunk, prior, n, L, last, state.data_ptr(),
                              outs[gpuhash.SHA1].data_ptr() if algos & gpuhash.SHA1 else None,
                              outs[gpuhash.MD5].data_ptr() if algos & gpuhash.MD5 else None,
                              outs[gpuhash.SHA256].data_ptr() if algos & gpuhash.SHA256 else None)
        prior += chunk
    assert prior >= L
    for alg, _ in ALGS:
        if algos & alg:
            got = outs[alg].cpu().numpy().reshape(n, -1)
            assert np.array_equal(got, oracle.hash_msgs(alg, host, L, 4)), alg


def test_device_unaligned_base_and_stride(gh):
    """Stream offset 15, odd piece length: the byte-granular kernel (SURVEY.md Appendix C golden)."""
    torch = _torch()
    total = 16385 * 5 + 3
    d = torch.zeros(total + 64, dtype=torch.uint8, device="cuda")
    gh.generate(d.data_ptr() + 15, 0, 15, total)
    dig = torch.zeros(6 * 20, dtype=torch.uint8, device="cuda")
    gh.verify_pieces_device(d.data_ptr() + 15, total, 16385, None, dig.data_ptr(), None)
    got = dig.cpu().numpy().reshape(6, 20)
    host = oracle.gen(0, 15, total)
    assert np.array_equal(got, oracle.hash_msgs(oracle.SHA1, host, 16385))
    assert got[0].tobytes().hex() == "27a36dfb037754434ba684bcb73fc15ff2100290"


def test_n_bad_without_a_bitmap_buffer(gh):
    """Callers that only want the count may pass ok_bitmap_out = NULL."""
    import ctypes
    L = 16384
    data = oracle.gen(13, 0, 70 * L)
    expected = oracle.hash_msgs(oracle.SHA1, data, L)
    dirty = data.copy()
    dirty[3 * L] ^= 1
    dirty[69 * L + 5] ^= 1
    sl = gpuhash.SpanList([dirty])
    nbad = ctypes.c_uint64()
    rc = gh.L.gh_verify_pieces(gh.h, sl.sp, sl.n, L, ctypes.c_void_p(expected.ctypes.data), None, None, ctypes.byref(nbad))
    assert rc == 0 and nbad.value == 2


def test_stats_report_launches_and_bytes(gh):
    before = gh.stats()
    data = oracle.gen(1, 0, 1 << 20)
    gh.VerifyPieces([data], 16384, None)
    after = gh.stats()
    assert after.kernel_launches > before.kernel_launches
    assert after.h2d_bytes - before.h2d_bytes == (1 << 20)
    assert after.d2h_bytes - before.d2h_bytes == 64 * 20
    assert after.kernel_ms > 0


def test_lane_kernels_with_role_split_disabled():
    """Small batches normally take the role-split kernels (trio for SHA-1; schedule / rounds / MD5
    warps for the others); the library reads its tuning switches once per process, so the
    one-message-per-lane kernels are exercised on the same cases in a child process with both
    switched off."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, GPUHASH_SPLIT_WARPS_PER_SM="0", GPUHASH_TRIO_WARPS_PER_SM="0")
    here = os.path.abspath(__file__)
    r = subprocess.run([sys.executable, "-m", "pytest", here, "-q", "-x", "-m", "gpu", "-k",
                        "differential or chunk_interleaved or resumable or cfg5 or padding or cfg3"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_multi_device_context_shards_pieces():
    """In-process multi-GPU: one context over every visible device, pieces sharded in contiguous
    64-aligned ranges, bitmap words and digests merged on the host. Needs >= 2 GPUs."""
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("single-GPU box")
    ctx = gpuhash.Context()  # all devices
    try:
        assert ctx.n_devices == torch.cuda.device_count()
        L = 16384
        total = 1000 * L + 4097
        data = oracle.gen(8, 0, total)
        expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
        dirty = data.copy()
        bad = [0, 63, 64, 65, 500, 1000]
        for i in bad:
            dirty[i * L + 7] ^= 0x80
        bitmap, dig, nbad = ctx.VerifyPieces([dirty], L, expected)
        want_bitmap, want_bad = oracle.verify_pieces(dirty, L, expected, 4)
        assert nbad == want_bad == len(bad) and np.array_equal(bitmap, want_bitmap)
        assert np.array_equal(dig, oracle.hash_msgs(oracle.SHA1, dirty, L, 4))
        parts = ctx.HashParts([data], 1 << 20, gpuhash.MD5 | gpuhash.SHA256)
        assert np.array_equal(parts[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, data, 1 << 20, 4))
        assert np.array_equal(parts[gpuhash.SHA256], oracle.hash_msgs(oracle.SHA256, data, 1 << 20, 4))
    finally:
        ctx.close()


def test_incremental_stream_feed():
    """gh_stream_*: windows of whole pieces written in place into the pinned ring, more windows than
    slots, short last piece, one corrupted piece; and the refusals the header promises."""
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
    try:
        L = 16384
        total = 500 * L + 999  # 501 pieces, 64 per window -> 8 windows on 3 slots
        data = oracle.gen(6, 0, total)
        expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
        dirty = data.copy()
        dirty[123 * L + 5] ^= 1
        st = ctx.Stream(gpuhash.SHA1 | gpuhash.MD5, L, total, expected)
        with pytest.raises(gpuhash.GpuHashError):  # one-shot calls are refused while a stream is open
            ctx.HashParts([b"abc"], 3, gpuhash.SHA1)
        off, windows = 0, 0
        while True:
            buf = st.acquire()
            if buf is None:
                break
            assert len(buf) % L == 0 or off + len(buf) == total
            buf[:] = dirty[off:off + len(buf)]
            off += len(buf)
            st.submit()
            windows += 1
        assert off == total and windows == 8
        bitmap, outs, nbad = st.finish()
        assert nbad == 1 and [i for i, b in enumerate(_bits(bitmap, 501)) if not b] == [123]
        assert np.array_equal(outs[gpuhash.SHA1], oracle.hash_msgs(oracle.SHA1, dirty, L, 4))
        assert np.array_equal(outs[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, dirty, L, 4))
        # the context is usable again
        assert ctx.HashParts([b"abc"], 3, gpuhash.SHA1)[gpuhash.SHA1][0].tobytes() == hashlib.sha1(b"abc").digest()
        # a message longer than a slot is refused
        with pytest.raises(gpuhash.GpuHashError) as e:
            ctx.Stream(gpuhash.SHA1, (1 << 20) + 64, 10 << 20)
        assert e.value.status == gpuhash.GH_ERANGE
        # abort releases the context
        st = ctx.Stream(gpuhash.SHA256, L, total)
        st.acquire()
        st.abort()
        got = ctx.HashParts([data], L, gpuhash.SHA256)[gpuhash.SHA256]
        assert np.array_equal(got, oracle.hash_msgs(oracle.SHA256, data, L, 4))
    finally:
        ctx.close()


def test_calls_leave_the_callers_current_device_alone():
    """A context over every device touches every device, yet the calling thread's current device is
    what it was afterwards (a host sharing the CUDA runtime -- torch here -- keeps allocating where it
    did before)."""
    torch = _torch()
    last = torch.cuda.device_count() - 1
    for cur in sorted({0, last}):
        torch.cuda.set_device(cur)
        ctx = gpuhash.Context()
        try:
            assert torch.cuda.current_device() == cur
            data = oracle.gen(4, 0, 300 * 4096)
            ctx.HashParts([data], 4096, gpuhash.SHA1 | gpuhash.MD5)
            st = ctx.Stream(gpuhash.SHA1, 4096, 8192)
            st.abort()
            t = torch.zeros(16, device="cuda")
            assert t.device.index == cur and torch.cuda.current_device() == cur
        finally:
            ctx.close()
        assert torch.cuda.current_device() == cur
    torch.cuda.set_device(0)


def test_suggest_devices_returns_a_usable_set():
    """gh_suggest_devices: n distinct visible ordinals and a measured link figure; a context over them works."""
    torch = _torch()
    count = torch.cuda.device_count()
    for n in sorted({1, max(1, count // 2), count}):
        devs, gbs = gpuhash.suggest_devices(n)
        assert len(devs) == n and len(set(devs)) == n and all(0 <= d < count for d in devs)
        assert gbs > 5.0  # a PCIe Gen3 x16 link does better than this
    with pytest.raises(gpuhash.GpuHashError):
        gpuhash.suggest_devices(count + 1)
    devs, _ = gpuhash.suggest_devices(count)
    ctx = gpuhash.Context(devices=devs)
    try:
        data = oracle.gen(3, 0, 200 * 16384)
        got = ctx.HashParts([data], 16384, gpuhash.SHA1)[gpuhash.SHA1]
        assert np.array_equal(got, oracle.hash_msgs(oracle.SHA1, data, 16384, 4))
    finally:
        ctx.close()


def test_incremental_stream_feed_over_all_devices():
    """The same feed on a context over every visible device: windows go round the devices, digests
    and the bitmap are gathered from whichever device hashed a piece (one GPU: the degenerate case)."""
    ctx = gpuhash.Context(slot_bytes=1 << 20, slots=2)
    try:
        L = 16384
        total = 1000 * L + 5  # 1001 pieces, 64 per window -> 16 windows
        data = oracle.gen(16, 0, total)
        expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
        dirty = data.copy()
        bad = [0, 63, 64, 127, 128, 500, 1000]
        for i in bad:
            dirty[i * L] ^= 2
        st = ctx.Stream(gpuhash.SHA1 | gpuhash.SHA256, L, total, expected)
        off = 0
        while True:
            buf = st.acquire()
            if buf is None:
                break
            buf[:] = dirty[off:off + len(buf)]
            off += len(buf)
            st.submit()
        bitmap, outs, nbad = st.finish()
        assert nbad == len(bad) and [i for i, b in enumerate(_bits(bitmap, 1001)) if not b] == bad
        assert np.array_equal(outs[gpuhash.SHA1], oracle.hash_msgs(oracle.SHA1, dirty, L, 4))
        assert np.array_equal(outs[gpuhash.SHA256], oracle.hash_msgs(oracle.SHA256, dirty, L, 4))
    finally:
        ctx.close()


@pytest.mark.parametrize("piece_len,total", [(1, 5000), (7, 7 * 3000 + 3), (63, 63 * 40), (64, 64), (4096, 100), (4096, 4096),
                                             (16, 16 * 33), (1 << 20, (1 << 20) + 1), (48, 48 * 1025)])
def test_extreme_piece_shapes(gh, piece_len, total):
    """One-byte pieces, a stream shorter than one piece, exactly one piece, 33 pieces (two warps),
    lengths that are multiples of 16 but not of 64 (TMA rows with a sub-block tail)."""
    data = oracle.gen(31, 1, total)
    expected = oracle.hash_msgs(oracle.SHA1, data, piece_len, 4)
    bitmap, dig, nbad = gh.VerifyPieces([data], piece_len, expected)
    assert nbad == 0 and np.array_equal(dig, expected)
    n = expected.shape[0]
    assert all(_bits(bitmap, n)) and not any(_bits(bitmap, bitmap.size * 64)[n:])
    both = gh.HashParts([data], piece_len, gpuhash.MD5 | gpuhash.SHA256)
    assert np.array_equal(both[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, data, piece_len, 4))
    assert np.array_equal(both[gpuhash.SHA256], oracle.hash_msgs(oracle.SHA256, data, piece_len, 4))


def test_single_long_message_through_small_ring():
    """One 9 MiB message, ring windows of 1 MiB: a group of one chain, nine chained launches."""
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=2)
    try:
        n = 9 * (1 << 20) + 77
        data = oracle.gen(12, 0, n)
        got = ctx.HashParts([data, b"", data[:5]], n + 5, gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256)
        whole = np.concatenate([data, data[:5]])
        for alg, _ in ALGS:
            assert got[alg][0].tobytes() == oracle.digest(alg, whole)
    finally:
        ctx.close()


@pytest.mark.parametrize("slot_mib", [0, 1])
@pytest.mark.parametrize("algos", [gpuhash.MD5 | gpuhash.SHA256, gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256])
def test_parts_per_file_batch(algos, slot_mib, tmp_path):
    """All files of an upload job in one batch, parts restarting at every file (each file is its own
    S3 object, uploader.go:72-89): ragged last parts, an empty file, a file shorter than one part, a
    file that is an exact multiple; from memory and from disk; default ring and a 1 MiB ring (chunk-
    interleaved windows with a per-message length table)."""
    P = 300 * 1024 + 64
    sizes = [3 * P + 17, 0, 5, 2 * P, P - 1, 7 * P + 100000, 64]
    ctx = gpuhash.Context(devices=[0], slot_bytes=slot_mib << 20, slots=3)
    try:
        bufs = [oracle.gen(40 + i, 0, n) for i, n in enumerate(sizes)]
        want = {a: [] for a, _ in ALGS}
        for b in bufs:
            for a, _ in ALGS:
                if algos & a and b.size:
                    want[a].append(oracle.hash_msgs(a, b, P, 2))
        got, counts = ctx.HashPartsPerFile(bufs, P, algos)
        assert counts == [4, 0, 1, 2, 1, 8, 1]
        for a, _ in ALGS:
            if algos & a:
                assert np.array_equal(got[a], np.concatenate(want[a])), a
        paths = []
        for i, b in enumerate(bufs):
            p = os.path.join(str(tmp_path), "f%d.bin" % i)
            with open(p, "wb") as f:
                f.write(b.tobytes())
            paths.append((p, b.size))
        got2, counts2 = ctx.HashPartsPerFile(paths, P, algos)
        assert counts2 == counts
        for a, _ in ALGS:
            if algos & a:
                assert np.array_equal(got2[a], got[a]), a
    finally:
        ctx.close()


def test_randomised_shapes_against_oracle(gh):
    """Seeded random batches: message length, total, algorithm mask and the way the stream is cut into
    spans (including empty ones) are all drawn at random; every digest must equal the oracle's."""
    rng = np.random.default_rng(2024)
    masks = [gpuhash.SHA1, gpuhash.MD5, gpuhash.SHA256, gpuhash.MD5 | gpuhash.SHA256, 7, gpuhash.SHA1 | gpuhash.SHA256]
    for trial in range(60):
        L = int(rng.choice([1, 3, 16, 55, 56, 64, 100, 128, 1000, 4096, 16384, 65552, 100000]))
        n = int(rng.integers(1, 70))
        total = max(1, (n - 1) * L + int(rng.integers(1, L + 1)))
        if total > 6 << 20:
            continue
        algos = int(rng.choice(masks))
        data = oracle.gen(1000 + trial, int(rng.integers(0, 64)), total)
        cuts = sorted(int(x) for x in rng.integers(0, total + 1, size=int(rng.integers(0, 6))))
        spans, prev = [], 0
        for c in cuts + [total]:
            spans.append(data[prev:c])
            prev = c
        got = gh.HashParts(spans, L, algos)
        for alg, _ in ALGS:
            if algos & alg:
                assert np.array_equal(got[alg], oracle.hash_msgs(alg, data, L, 2)), (trial, L, total, algos)


def test_concurrent_callers_share_one_context(gh):
    """include/gpuhash.h promises that a gh_ctx may be used from several OS threads (calls are
    serialised inside; goroutines migrate between threads): four threads, different inputs, at once."""
    import threading
    results, errors = {}, []

    def work(i):
        try:
            L = [16384, 1000, 262144, 4096][i]
            data = oracle.gen(200 + i, 0, 40 * L + i)
            for _ in range(5):
                got = gh.HashParts([data], L, gpuhash.SHA1 | gpuhash.MD5)
            results[i] = (np.array_equal(got[gpuhash.SHA1], oracle.hash_msgs(oracle.SHA1, data, L)) and
                          np.array_equal(got[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, data, L)))
        except Exception as e:  # noqa
            errors.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors and results == {0: True, 1: True, 2: True, 3: True}
