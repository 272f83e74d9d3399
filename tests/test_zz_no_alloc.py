"""gh_hash_update_device_noalloc, the entry point for callers that must not allocate device memory
(a worker whose HBM is already fully committed to payload buffers and the staging ring, or code that
runs where cudaMalloc's device-wide synchronisation is not acceptable). Its midstates live in a fixed
array inside the library, shared by every call, so the things worth checking are: the digests are
the same as the allocating entry point's (bit for bit -- integer work, no tolerance), a second batch
right after the first is not disturbed by what the first left in the array, and an oversized batch
is refused. The file name sorts last on purpose: pytest -x has run every other test before this one.
"""
import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu

ALL = gpuhash.SHA1 | gpuhash.MD5 | gpuhash.SHA256


def _windows(torch, d, n, L, last, chunks):
    """Yield (window tensor, chunk, prior): message m's next `chunk` bytes packed at m*chunk."""
    prior = 0
    for chunk in chunks:
        win = torch.zeros(n * chunk + 64, dtype=torch.uint8, device="cuda")
        for m in range(n):
            mlen = last if m == n - 1 else L
            lo, hi = min(prior, mlen), min(prior + chunk, mlen)
            if hi > lo:
                win[m * chunk: m * chunk + (hi - lo)] = d[m * L + lo: m * L + hi]
        yield win, chunk, prior
        prior += chunk


def _run_batch(torch, ctx, seed, n, L, last, chunks, algos):
    total = (n - 1) * L + last
    host = oracle.gen(seed, 0, total)
    d = torch.from_numpy(host).cuda()
    lens = {gpuhash.SHA1: 20, gpuhash.MD5: 16, gpuhash.SHA256: 32}
    ref = {a: torch.zeros(n * k, dtype=torch.uint8, device="cuda") for a, k in lens.items()}
    new = {a: torch.zeros(n * k, dtype=torch.uint8, device="cuda") for a, k in lens.items()}
    state = torch.zeros(n * 17, dtype=torch.int32, device="cuda")

    def ptr(t, a):
        return t[a].data_ptr() if algos & a else None

    for win, chunk, prior in _windows(torch, d, n, L, last, chunks):
        # the window was filled on torch's stream; both hashers run on streams of their own
        torch.cuda.synchronize()
        ctx.hash_update_device(algos, win.data_ptr(), chunk, chunk, prior, n, L, last, state.data_ptr(),
                               ptr(ref, gpuhash.SHA1), ptr(ref, gpuhash.MD5), ptr(ref, gpuhash.SHA256))
        gpuhash.hash_update_device_noalloc(0, algos, win.data_ptr(), chunk, chunk, prior, n, L, last,
                                           ptr(new, gpuhash.SHA1), ptr(new, gpuhash.MD5), ptr(new, gpuhash.SHA256))
    torch.cuda.synchronize()
    for a, k in lens.items():
        if not algos & a:
            continue
        got = new[a].cpu().numpy().reshape(n, k)
        assert np.array_equal(got, ref[a].cpu().numpy().reshape(n, k)), "differs from gh_hash_update_device"
        assert np.array_equal(got, oracle.hash_msgs(a, host, L, 4)), "differs from the oracle"


def test_noalloc_matches_allocating_entry_point_twice_in_a_row():
    import torch
    ctx = gpuhash.Context(devices=[0])
    try:
        # first batch: 300 messages, three unequal windows, short last message
        _run_batch(torch, ctx, 21, 300, 5 * 4096 + 100, 4096 + 33, (4096, 8192, 2 * 4096 + 128), ALL)
        # second batch straight after, other data, fewer and longer messages: must not see batch one's midstates
        _run_batch(torch, ctx, 22, 70, 3 * 8192, 3 * 8192, (8192, 8192, 8192), gpuhash.SHA1)
        _run_batch(torch, ctx, 23, 33, 100000, 77, (65536, 65536), gpuhash.MD5 | gpuhash.SHA256)
    finally:
        ctx.close()


def test_noalloc_refuses_batches_beyond_the_fixed_array():
    import torch
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    with pytest.raises(gpuhash.GpuHashError) as e:
        gpuhash.hash_update_device_noalloc(0, gpuhash.SHA1, buf.data_ptr(), 64, 64, 0, gpuhash.NOALLOC_MAX_MSGS + 1,
                                           64, 64, buf.data_ptr())
    assert e.value.status == gpuhash.GH_ERANGE
    # exactly the maximum is accepted (64-byte messages packed back to back)
    n = gpuhash.NOALLOC_MAX_MSGS
    host = oracle.gen(24, 0, n * 64)
    d = torch.from_numpy(host).cuda()
    out = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    gpuhash.hash_update_device_noalloc(0, gpuhash.SHA1, d.data_ptr(), 64, 64, 0, n, 64, 64, out.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().reshape(n, 20), oracle.hash_msgs(oracle.SHA1, host, 64, 4))
