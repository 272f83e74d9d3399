"""gh_batcher_*: the micro-batcher that turns anacrolix's piece-at-a-time hashPiece calls (reference
torrent.go:79 -> :106) into GPU batches. Many threads each hand over one piece and block for its
verdict; batches flush on count or on deadline; pieces come in any order; the torrent's last piece
is short; a GPU failure surfaces as a status for every caller of the batch, never as a mismatch.
Verdicts and digests are compared with the oracle's, piece by piece."""
import threading
import time

import numpy as np
import pytest

import gpuhash
import oracle

pytestmark = pytest.mark.gpu


def _payload(seed, n, L, last):
    total = (n - 1) * L + last
    data = oracle.gen(seed, 0, total)
    expected = oracle.hash_msgs(oracle.SHA1, data, L, 4)
    return data, expected, total


def test_eight_threads_two_thousand_pieces_out_of_order():
    ctx = gpuhash.Context(devices=[0])
    n, L, last = 2000, 16384, 5000
    data, expected, total = _payload(61, n, L, last)
    dirty = data.copy()
    rng = np.random.default_rng(61)
    bad = set(int(i) for i in rng.choice(n, size=n // 100, replace=False)) | {n - 1}  # 1 % corrupt, and the short tail
    for i in bad:
        plen = last if i == n - 1 else L
        dirty[i * L + (i * 7) % plen] ^= 0x20
    want_digest = oracle.hash_msgs(oracle.SHA1, dirty, L, 4)
    order = rng.permutation(n)
    b = ctx.Batcher(L, 64, 20)
    verdict = {}
    digest = {}
    errors = []

    def worker(w):
        try:
            for i in order[w::8]:
                i = int(i)
                piece = dirty[i * L: min((i + 1) * L, total)]
                ok, dig = b.verify(piece, expected[i], want_digest=True)
                verdict[i] = ok
                digest[i] = dig
        except Exception as e:  # noqa
            errors.append(e)

    ts = [threading.Thread(target=worker, args=(w,)) for w in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    st = b.stats()
    b.close()
    ctx.close()
    assert not errors, errors
    assert sorted(i for i, ok in verdict.items() if not ok) == sorted(bad)
    assert len(verdict) == n and all(digest[i] == want_digest[i].tobytes() for i in range(n))
    # eight blocked callers can never fill a 64-piece batch: every batch left on the deadline
    assert st.pieces == n and st.flushed_by_deadline == st.batches > 0 and st.failed_batches == 0


def test_batches_fill_and_flush_on_count():
    ctx = gpuhash.Context(devices=[0])
    n, L = 96, 4096 + 64
    data, expected, total = _payload(62, n, L, L)
    b = ctx.Batcher(L, 32, 60000)  # the deadline is a minute away: only a full batch can release anyone
    oks = [None] * n

    def one(i):
        oks[i] = b.verify(data[i * L:(i + 1) * L], expected[i])

    t0 = time.perf_counter()
    ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    st = b.stats()
    b.close()
    ctx.close()
    assert all(oks) and dt < 30
    assert st.batches == 3 and st.flushed_full == 3 and st.flushed_by_deadline == 0 and st.pieces == n


def test_deadline_releases_a_lone_piece_and_flush_releases_it_at_once():
    ctx = gpuhash.Context(devices=[0])
    L = 262144
    data, expected, total = _payload(63, 3, L, 100)
    b = ctx.Batcher(L, 128, 300)
    t0 = time.perf_counter()
    assert b.verify(data[:L], expected[0]) is True
    waited = time.perf_counter() - t0
    assert 0.25 < waited < 5.0, waited  # the 300 ms deadline, not the 128-piece count, let it go
    # short last piece, wrong digest; released by flush() from another thread long before the deadline
    b2 = ctx.Batcher(L, 128, 60000)
    res = {}
    t = threading.Thread(target=lambda: res.update(v=b2.verify(data[2 * L:], expected[1])))
    t.start()
    time.sleep(0.3)
    b2.flush()
    t.join(timeout=30)
    assert res == {"v": False}
    b.close()
    b2.close()
    ctx.close()


def test_reserve_fill_in_place_commit_and_cancel():
    ctx = gpuhash.Context(devices=[0])
    L = 32768
    data, expected, total = _payload(64, 4, L, 77)
    b = ctx.Batcher(L, 4, 100)
    slots = []
    for i in range(4):
        plen = 77 if i == 3 else L
        slot, view = b.reserve(plen)
        view[:] = data[i * L:i * L + plen] if i != 1 else 0
        slots.append(slot)
    assert b.cancel(slots[1]) == 0  # the piece could not be read: give the slot back
    out = {}
    ts = [threading.Thread(target=lambda i=i: out.update({i: b.commit(slots[i], expected[i])})) for i in (0, 2, 3)]
    [t.start() for t in ts]
    [t.join(timeout=30) for t in ts]
    assert out == {0: True, 2: True, 3: True}
    b.close()
    ctx.close()


def test_gpu_failure_is_a_status_not_a_mismatch():
    ctx = gpuhash.Context(devices=[0])
    n, L = 16, 16384
    data, expected, total = _payload(65, n, L, L)
    b = ctx.Batcher(L, n, 60000)
    ctx.L.gh_debug_fail_nth_copy(ctx.h, 0)
    results = [None] * n

    def one(i):
        try:
            results[i] = b.verify(data[i * L:(i + 1) * L], expected[i])
        except gpuhash.GpuHashError as e:
            results[i] = e.status

    ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join(timeout=60) for t in ts]
    assert results == [gpuhash.GH_ECUDA] * n  # every caller of the failed batch: "unknown", not False
    ctx.L.gh_debug_fail_nth_copy(ctx.h, -1)
    # the batcher and the context keep working
    ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join(timeout=60) for t in ts]
    assert results == [True] * n
    st = b.stats()
    assert st.failed_batches == 1 and st.batches == 2
    b.close()
    ctx.close()


def test_batches_larger_than_64_pieces_shard_over_all_devices():
    import torch
    ctx = gpuhash.Context()  # every visible device
    n, L = 512, 65536
    data, expected, total = _payload(66, n, L, 4097)
    dirty = data.copy()
    dirty[300 * L] ^= 1
    b = ctx.Batcher(L, 256, 60000)
    res = [None] * n

    def one(i):
        res[i] = b.verify(dirty[i * L: min((i + 1) * L, total)], expected[i])

    ts = [threading.Thread(target=one, args=(i,)) for i in range(n)]
    [t.start() for t in ts]
    [t.join(timeout=120) for t in ts]
    assert [i for i, r in enumerate(res) if r is not True] == [300] and res[300] is False
    assert b.stats().batches == 2
    b.close()
    ctx.close()
    assert torch.cuda.device_count() >= 1


def test_batch_larger_than_a_ring_slot_runs_chunk_interleaved():
    """40 pieces of 8 MiB (one of them short) in one batch: 320 MiB against 256 MiB ring slots, so the
    batch takes the chunk-interleaved windows with a per-piece length table."""
    ctx = gpuhash.Context(devices=[0])
    n, L, last = 40, 8 << 20, (3 << 20) + 77
    data, expected, total = _payload(67, n, L, last)
    dirty = data.copy()
    dirty[17 * L + 12345] ^= 1
    b = ctx.Batcher(L, n, 60000)
    res = [None] * n

    def one(i):
        res[i] = b.verify(dirty[i * L: min((i + 1) * L, total)], expected[i], want_digest=True)

    order = list(range(n))[::-1]  # the short last piece arrives first
    ts = [threading.Thread(target=one, args=(i,)) for i in order]
    [t.start() for t in ts]
    [t.join(timeout=120) for t in ts]
    want = oracle.hash_msgs(oracle.SHA1, dirty, L, 4)
    assert [i for i in range(n) if not res[i][0]] == [17]
    assert all(res[i][1] == want[i].tobytes() for i in range(n))
    b.close()
    ctx.close()


def test_randomised_stress_small_batches_cancellations_and_short_deadline():
    """Sixteen threads, 2 400 pieces of random length (0 .. L bytes), an 8-piece batch limit and a 3 ms
    deadline, random pauses, one reservation in twenty given back: thousands of small batches and
    every hand-over between the filling and the in-flight buffer; each verdict against hashlib."""
    import hashlib
    import random
    ctx = gpuhash.Context(devices=[0])
    L = 40000
    b = ctx.Batcher(L, 8, 3)
    src = oracle.gen(68, 0, 4 << 20)
    errors, done = [], [0] * 16

    def worker(w):
        rnd = random.Random(1000 + w)
        try:
            for k in range(150):
                n = rnd.choice((0, 1, 55, 64, 4096, L, rnd.randrange(L + 1)))
                off = rnd.randrange(src.size - L)
                piece = src[off:off + n]
                want = hashlib.sha1(piece.tobytes()).digest()
                corrupt = rnd.random() < 0.1
                exp = bytes([want[0] ^ 1]) + want[1:] if corrupt else want
                if rnd.random() < 0.05:
                    slot, view = b.reserve(n)
                    time.sleep(rnd.random() * 0.002)
                    assert b.cancel(slot) == 0
                    continue
                if rnd.random() < 0.5:
                    slot, view = b.reserve(n)
                    view[:] = piece
                    if rnd.random() < 0.3:
                        time.sleep(rnd.random() * 0.004)  # a slow producer: the batch must wait for its commit
                    ok = b.commit(slot, exp)
                else:
                    ok, dig = b.verify(piece, exp, want_digest=True)
                    assert dig == want
                assert ok is (not corrupt), (w, k, n)
                done[w] += 1
        except Exception as e:  # noqa
            errors.append((w, repr(e)))

    ts = [threading.Thread(target=worker, args=(w,)) for w in range(16)]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    st = b.stats()
    b.close()
    ctx.close()
    assert not errors, errors[:3]
    assert sum(done) > 2000 and st.failed_batches == 0 and st.pieces >= sum(done)
