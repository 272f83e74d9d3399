"""BASELINE config 4: season pack, 24 files x ~2 GiB, 16 MiB pieces spanning file boundaries, pieces
sharded over 1/2/4/8 GPUs by ONE process (one gh_ctx over k devices, contiguous 64-aligned piece
ranges, no collective). Strong scaling of the end-to-end call on pinned host memory; every run is
checked against expected digests (one injected fault). Writes gpurun_out/config4_multi.json."""
import ctypes, json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
import gpuhash, oracle

L = 16 << 20
sizes = [(2 << 30) + 7919 * (k + 1) for k in range(24)]
total = sum(sizes)
n = (total + L - 1) // L
ngpu = torch.cuda.device_count()
threads = len(os.sched_getaffinity(0))
host = gpuhash.PinnedBuffer(total)
lib = oracle.lib()
step = (total // threads + 63) & ~63
def fill(lo):
    hi = min(total, lo + step)
    lib.gho_gen_fill(4, lo, hi - lo, ctypes.c_void_p(host.ptr + lo))
ts = [threading.Thread(target=fill, args=(lo,)) for lo in range(0, total, step)]
[t.start() for t in ts]; [t.join() for t in ts]
files, o = [], 0
for s in sizes:
    files.append(host.array[o:o + s]); o += s
spans = gpuhash.SpanList(files)
out = {"pieces": n, "bytes": total, "files": 24, "host_threads": threads, "gpus_visible": ngpu}
expected = None
k = 1
while k <= ngpu:
    ctx = gpuhash.Context(devices=list(range(k)))
    if expected is None:
        _, dig, _ = ctx.VerifyPieces(spans, L, None)
        expected = dig.copy()
        # spot-check pieces that span file boundaries against the oracle
        off = 0
        for s in sizes[:3]:
            off += s
            i = off // L
            assert expected[i].tobytes() == oracle.digest(oracle.SHA1, host.array[i * L:(i + 1) * L]), i
        host.array[1234 * L + 99] ^= 1
    ctx.VerifyPieces(spans, L, expected)
    t0 = time.perf_counter()
    bitmap, dig, nbad = ctx.VerifyPieces(spans, L, expected)
    dt = time.perf_counter() - t0
    bad = [i for i in range(n) if not (int(bitmap[i >> 6]) >> (i & 63)) & 1]
    assert nbad == 1 and bad == [1234], (nbad, bad[:4])
    out["gpus_%d" % k] = {"e2e_GBps": round(total / dt / 1e9, 2), "ms": round(dt * 1e3, 1)}
    print(k, out["gpus_%d" % k], flush=True)
    ctx.close()
    k *= 2
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "config4_multi.json"), "w"), indent=1)
