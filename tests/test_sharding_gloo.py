"""world_size-2 gloo run of the multi-GPU host logic on CPU: each rank owns the contiguous,
64-aligned piece range gh_shard_range gives it, produces that range's bitmap words (here with the
oracle standing in for the device), and the gathered words equal the single-rank bitmap. No
collective touches payload bytes -- the path shards with zero exchange steps."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gpuhash
import oracle

L = 16384
N_BYTES = 301 * L + 1234


def _dirty_stream():
    data = oracle.gen(5, 0, N_BYTES)
    expected = oracle.hash_msgs(oracle.SHA1, data, L)
    dirty = data.copy()
    n = expected.shape[0]
    for i in range(n):
        plen = min(L, N_BYTES - i * L)
        hit, off, bit = oracle.cfg5_corruption(i, plen)
        if hit or i % 37 == 5:
            dirty[i * L + off] ^= np.uint8(1 << bit)
    return dirty, expected


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dirty, expected = _dirty_stream()
    n = expected.shape[0]
    b, e = gpuhash.shard_range(n, world, rank)
    lo, hi = b * L, min(e * L, N_BYTES)
    bitmap, nbad = oracle.verify_pieces(dirty[lo:hi], L, expected[b:e])
    words = torch.zeros((n + 63) // 64, dtype=torch.int64)
    words[b // 64: b // 64 + bitmap.size] = torch.from_numpy(bitmap.view(np.int64))
    dist.all_reduce(words, op=dist.ReduceOp.BOR)  # result gather only: disjoint words per rank
    bad = torch.tensor([nbad])
    dist.all_reduce(bad)
    if rank == 0:
        q.put((words.numpy().view(np.uint64).copy(), int(bad.item())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_bitmap_equals_single_rank():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    words, bad = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    dirty, expected = _dirty_stream()
    full_bitmap, full_bad = oracle.verify_pieces(dirty, L, expected)
    assert np.array_equal(words, full_bitmap)
    assert bad == full_bad and bad > 0
