"""The SHA-1 trio kernel (gh_hash_trio: per chain one rounds-warp fed by two schedule-warps, up to three
chains per SM), which the launcher picks for staged SHA-1 batches of up to three message-warps per SM
-- BASELINE configs 1, 2 and 4 all fall there (4 096 / 12 800 / 3 073 pieces on 148 SMs). Bit-exact
against the oracle, through the C ABI: config 2's message count with block counts that exercise every
hand-off case (one block, an odd number of blocks, a sub-block tail, a short last piece), the verify
bitmap, counts around the one/two/three-chains-per-CTA thresholds, the resumable form and a per-file
length table. The general differential cases of test_gpu_parity.py run through it as well (every
SHA-1 batch there is small); its two fall-backs are exercised in child processes below."""
import os
import subprocess
import sys

import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu

N_CFG2 = 12800  # 400 message-warps: 2.7 per SM on a B200


def _torch():
    import torch
    return torch


@pytest.mark.parametrize("L,last", [(64, 64), (256, 256), (64 * 9, 64 * 9), (64 * 5 + 16, 7), (1024 + 48, 1024 + 48),
                                    (4096, 100), (64 * 13, 64 * 12 + 63)])
def test_trio_shapes_device_resident(gh, L, last):
    torch = _torch()
    n = N_CFG2
    total = (n - 1) * L + last
    host = oracle.gen(90 + L, 0, total)
    d = torch.from_numpy(host).cuda()
    dig = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    gh.verify_pieces_device(d.data_ptr(), total, L, None, dig.data_ptr(), None)
    if os.environ.get("GPUHASH_TRIO_WARPS_PER_SM") != "0":
        assert gpuhash.last_kernel().startswith("gh_hash_trio")  # the shape really is routed to the kernel under test
    want = oracle.hash_msgs(oracle.SHA1, host, L, 8)
    assert np.array_equal(dig.cpu().numpy().reshape(n, 20), want)
    # verification: corrupt pieces at warp / CTA / bitmap-word edges, bitmap and digests on the device
    bad = [0, 31, 32, 95, 96, 97, 6399, n - 2, n - 1]
    for i in bad:
        plen = last if i == n - 1 else L
        d[i * L + (i * 31) % plen] ^= 0x04
    exp = torch.from_numpy(want.reshape(-1).copy()).cuda()
    ok = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
    gh.verify_pieces_device(d.data_ptr(), total, L, exp.data_ptr(), dig.data_ptr(), ok.data_ptr())
    words = ok.cpu().numpy().view(np.uint32)
    assert [i for i in range(n) if not (int(words[i >> 5]) >> (i & 31)) & 1] == bad
    assert np.array_equal(dig.cpu().numpy().reshape(n, 20), oracle.hash_msgs(oracle.SHA1, d.cpu().numpy()[:total], L, 8))


def test_trio_message_counts_at_the_thresholds(gh):
    """One, two and three chains per CTA; counts that leave the last CTA fewer chains than the others
    or a partly filled last warp; the largest count the trio kernel takes (3 message-warps per SM) and
    the first one it leaves to the lane kernel."""
    torch = _torch()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    L = 64 * 7
    for n in (1, 33, sms * 32, sms * 32 + 1, sms * 64, sms * 64 + 1, sms * 64 + 33, sms * 64 + 65, sms * 96 - 31, sms * 96,
              sms * 96 + 1):
        host = oracle.gen(7, n, n * L)
        d = torch.from_numpy(host).cuda()
        dig = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
        gh.verify_pieces_device(d.data_ptr(), n * L, L, None, dig.data_ptr(), None)
        assert np.array_equal(dig.cpu().numpy().reshape(n, 20), oracle.hash_msgs(oracle.SHA1, host, L, 8)), n


def test_trio_resumable_windows(gh):
    torch = _torch()
    n, L, last = 10000, 64 * 11 + 20, 64 * 3 + 5
    total = (n - 1) * L + last
    host = oracle.gen(91, 0, total)
    rows = np.zeros((n, L), dtype=np.uint8)
    rows.reshape(-1)[:total] = host
    state = torch.zeros(n * 17, dtype=torch.int32, device="cuda")
    out = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    prior = 0
    for chunk in (64 * 2, 64 * 5, 64 * 4 + 64):
        win = torch.from_numpy(np.ascontiguousarray(np.pad(rows[:, prior:prior + chunk], ((0, 0), (0, max(0, prior + chunk - L)))))).cuda()
        torch.cuda.synchronize()
        gh.hash_update_device(gpuhash.SHA1, win.data_ptr(), chunk, chunk, prior, n, L, last, state.data_ptr(), out.data_ptr())
        prior += chunk
    assert prior >= L
    assert np.array_equal(out.cpu().numpy().reshape(n, 20), oracle.hash_msgs(oracle.SHA1, host, L, 8))


def test_trio_with_per_file_length_table(gh):
    """Parts restarting at every file: ragged last parts inside a batch of ~12 000 messages."""
    P = 64 * 8
    rng = np.random.default_rng(5)
    sizes = [int(x) for x in rng.integers(1, 40 * P, size=600)]
    bufs = [oracle.gen(300 + i, 0, s) for i, s in enumerate(sizes)]
    got, counts = gh.HashPartsPerFile(bufs, P, gpuhash.SHA1)
    want = np.concatenate([oracle.hash_msgs(oracle.SHA1, b, P, 2) for b in bufs])
    assert sum(counts) == want.shape[0] and 9472 < want.shape[0] <= 14208, want.shape
    assert np.array_equal(got[gpuhash.SHA1], want)


@pytest.mark.parametrize("switches", [{"GPUHASH_TRIO_WARPS_PER_SM": "0"},
                                      {"GPUHASH_TRIO_WARPS_PER_SM": "0", "GPUHASH_SPLIT_WARPS_PER_SM": "0"}])
def test_fallback_kernels_when_trio_is_off(switches):
    """With the trio kernel switched off, small SHA-1 batches go to the 1R+1S split kernel, and with
    that off as well to the one-message-per-lane kernel: this file's shapes and the general cases, in
    child processes (the switches are read once per process)."""
    env = dict(os.environ, **switches)
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_trio_kernel.py"),
                        os.path.join(here, "test_gpu_parity.py"), "-q", "-x", "-m", "gpu", "-k",
                        "trio_shapes or thresholds or trio_resumable or differential or chunk_interleaved or cfg5 or padding or extreme"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
