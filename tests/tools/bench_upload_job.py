"""An upload job of 24 files x 2 GiB, 64 MiB parts, MD5 + SHA-256 (what uploader.go:72-89 does file by
file): one gh_hash_parts call per file vs ONE gh_hash_parts_per_file batch. Pinned host memory."""
import ctypes, json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import gpuhash, oracle
P = 64 << 20
nfiles = int(os.environ.get("NFILES", "24"))
sizes = [(2 << 30) + 7919 * (k + 1) for k in range(nfiles)]
total = sum(sizes)
host = gpuhash.PinnedBuffer(total)
threads = len(os.sched_getaffinity(0))
lib = oracle.lib()
step = (total // threads + 63) & ~63
def fill(lo):
    hi = min(total, lo + step)
    lib.gho_gen_fill(3, lo, hi - lo, ctypes.c_void_p(host.ptr + lo))
ts = [threading.Thread(target=fill, args=(lo,)) for lo in range(0, total, step)]
[t.start() for t in ts]; [t.join() for t in ts]
files, o = [], 0
for s in sizes:
    files.append(host.array[o:o + s]); o += s
ctx = gpuhash.Context(devices=[0])
algos = gpuhash.MD5 | gpuhash.SHA256
out = {"files": nfiles, "bytes": total, "part_len": P}
t0 = time.perf_counter()
got, counts = ctx.HashPartsPerFile(files, P, algos)
t1 = time.perf_counter()
got, counts = ctx.HashPartsPerFile(files, P, algos)
dt = time.perf_counter() - t1
out["one_batch"] = {"parts": int(sum(counts)), "GBps": round(total / dt / 1e9, 2), "kernel_ms": round(ctx.stats().kernel_ms, 1)}
# spot check: first and last (short) part of file 5
f5 = files[5]
base = sum(counts[:5])
assert got[gpuhash.MD5][base].tobytes() == oracle.digest(oracle.MD5, f5[:P])
assert got[gpuhash.SHA256][base + counts[5] - 1].tobytes() == oracle.digest(oracle.SHA256, f5[(counts[5] - 1) * P:])
t1 = time.perf_counter()
per = [ctx.HashParts([f], P, algos) for f in files[:4]]
dt4 = time.perf_counter() - t1
assert np.array_equal(per[0][gpuhash.MD5], got[gpuhash.MD5][:counts[0]])
out["file_by_file_first4"] = {"GBps": round(sum(sizes[:4]) / dt4 / 1e9, 2)}
print(json.dumps(out))
