"""One-off probes of the GPU box: zero-copy hashing straight out of pinned host
memory vs staged copies, per-config kernel-only rates. Writes gpurun_out/probe_box.json."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch

import gpuhash
import oracle

out = {}
ctx = gpuhash.Context(devices=[0])
stream = torch.cuda.current_stream()
sp = stream.cuda_stream


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        fn()
        b.record(stream)
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


# ---- kernel-only rates of every BASELINE config shape (reduced totals where noted)
total = 8 << 30
d = torch.empty(total, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 1, 0, total, stream=sp)
dig = torch.zeros((total // 16384) * 32, dtype=torch.uint8, device="cuda")
dig2 = torch.zeros((total // 16384) * 32, dtype=torch.uint8, device="cuda")
rates = {}
for name, L, tot in (("cfg1_256KiB_1GiB", 256 << 10, 1 << 30), ("cfg2_4MiB_8GiB", 4 << 20, total),
                     ("cfg4_16MiB_8GiB", 16 << 20, total), ("cfg5_16KiB_8GiB", 16 << 10, total),
                     ("64KiB_8GiB", 64 << 10, total), ("1MiB_8GiB", 1 << 20, total)):
    ms = timed(lambda: ctx.verify_pieces_device(d.data_ptr(), tot, L, None, dig.data_ptr(), None, stream=sp))
    rates["sha1_" + name] = {"ms": round(ms, 3), "GBps": round(tot / ms / 1e6, 1), "msgs": tot // L}
for name, L, tot in (("cfg3_64MiB_8GiB", 64 << 20, total), ("1MiB_8GiB", 1 << 20, total), ("64KiB_8GiB", 64 << 10, total)):
    for algos, an in ((gpuhash.MD5, "md5"), (gpuhash.SHA256, "sha256"), (gpuhash.MD5 | gpuhash.SHA256, "md5+sha256")):
        ms = timed(lambda: ctx.hash_parts_device(d.data_ptr(), tot, L, algos, None, dig.data_ptr(), dig2.data_ptr(), stream=sp), reps=2)
        rates[an + "_" + name] = {"ms": round(ms, 3), "GBps": round(tot / ms / 1e6, 1), "msgs": tot // L}
out["kernel_only"] = rates
print(json.dumps(rates, indent=1), flush=True)

# ---- zero-copy: the same kernel reading pinned host memory through the TMA unit (no H2D copy)
hb = 4 << 30
host = gpuhash.PinnedBuffer(hb)
ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), hb)
zc = {}
for L in (16 << 10, 256 << 10, 4 << 20):
    n = hb // L
    try:
        ms = timed(lambda: ctx.verify_pieces_device(host.ptr, hb, L, None, dig.data_ptr(), None, stream=sp), reps=2)
        got = dig[: n * 20].cpu().numpy().reshape(n, 20)
        ctx.verify_pieces_device(d.data_ptr(), hb, L, None, dig2.data_ptr(), None, stream=sp)
        torch.cuda.synchronize()
        same = bool(np.array_equal(got, dig2[: n * 20].cpu().numpy().reshape(n, 20)))
        zc[str(L)] = {"ms": round(ms, 2), "GBps": round(hb / ms / 1e6, 2), "matches_device_copy": same}
    except Exception as e:  # noqa
        zc[str(L)] = {"error": str(e)}
out["zero_copy_pinned_host"] = zc
print(json.dumps(zc, indent=1), flush=True)

# ---- e2e through the ring for the same host buffer, a few slot sizes
e2e = {}
for slot_mib in (64, 256, 1024):
    c2 = gpuhash.Context(devices=[0], slot_bytes=slot_mib << 20, slots=3)
    for L in (16 << 10, 256 << 10, 4 << 20):
        c2.VerifyPieces([host.array], L, None, want_digests=True)
        t0 = time.perf_counter()
        c2.VerifyPieces([host.array], L, None, want_digests=True)
        dt = time.perf_counter() - t0
        e2e["slot%dMiB_L%d" % (slot_mib, L)] = {"GBps": round(hb / dt / 1e9, 2), "kernel_ms": round(c2.stats().kernel_ms, 2)}
    c2.close()
out["e2e_ring_4GiB_pinned"] = e2e
print(json.dumps(e2e, indent=1), flush=True)
# pageable source (staging pool)
pg = np.empty(1 << 30, dtype=np.uint8)
pg[:] = host.array[: 1 << 30]
for thr in (4, 16):
    c3 = gpuhash.Context(devices=[0], stage_threads=thr)
    c3.VerifyPieces([pg], 4 << 20, None)
    t0 = time.perf_counter()
    c3.VerifyPieces([pg], 4 << 20, None)
    out["e2e_pageable_1GiB_threads%d" % thr] = round((1 << 30) / (time.perf_counter() - t0) / 1e9, 2)
    c3.close()
host.free()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe_box.json"), "w"), indent=1)
print("done")
