"""Can the page cache be DMA'd from directly? mmap a cached file, cudaHostRegister windows of the
mapping (gh_host_register), verify from the registered spans (direct DMA, no staging copy), and time
register / verify / unregister per window size. One JSON line."""
import ctypes, json, mmap, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import gpuhash, oracle

total = int(os.environ.get("GIB", "4")) << 30
path = os.environ.get("PROBE_PATH", "/tmp/gh_mmap_probe.bin")
blk = oracle.gen(9, 0, 256 << 20)
with open(path, "wb") as f:
    for i in range(total // blk.size):
        f.write(blk.tobytes())
L = 4 << 20
want = oracle.hash_msgs(oracle.SHA1, blk, L, 8)
out = {"bytes": total, "fs": os.popen("df -T %s | tail -1" % path).read().split()[:2]}
lib = gpuhash.load()
ctx = gpuhash.Context(devices=[0])
fd = os.open(path, os.O_RDONLY)
for flags, name in ((mmap.MAP_SHARED, "shared"), (mmap.MAP_PRIVATE, "private")):
    mm = mmap.mmap(fd, total, flags=flags | getattr(mmap, "MAP_POPULATE", 0), prot=mmap.PROT_READ)
    arr = np.frombuffer(mm, dtype=np.uint8)
    base = arr.ctypes.data
    for win in (64 << 20, 256 << 20, 1 << 30):
        t_reg = t_unreg = 0.0
        rc_all = 0
        t0 = time.perf_counter()
        for off in range(0, total, win):
            a = time.perf_counter()
            rc = lib.gh_host_register(ctypes.c_void_p(base + off), win)
            t_reg += time.perf_counter() - a
            rc_all |= rc
            if rc:
                break
        if rc_all:
            out["%s_win%dMiB" % (name, win >> 20)] = {"register_rc": rc_all}
            break
        a = time.perf_counter()
        bm, dig, nb = ctx.VerifyPieces([arr[o:o + win] for o in range(0, total, win)], L, None)
        t_ver = time.perf_counter() - a
        ok = bool(np.array_equal(dig[: want.shape[0]], want))
        for off in range(0, total, win):
            a = time.perf_counter()
            lib.gh_host_unregister(ctypes.c_void_p(base + off))
            t_unreg += time.perf_counter() - a
        out["%s_win%dMiB" % (name, win >> 20)] = {"register_GBps": round(total / t_reg / 1e9, 2), "verify_GBps": round(total / t_ver / 1e9, 2),
                                                   "unregister_GBps": round(total / t_unreg / 1e9, 2), "digests_ok": ok}
    del arr
    mm.close()
os.close(fd)
# reference point: the library's own file path (pread into the pinned ring) on the same file
for thr in (4, 16):
    c2 = gpuhash.Context(devices=[0], stage_threads=thr)
    c2.VerifyFiles([(path, total)], L, None)
    t0 = time.perf_counter()
    c2.VerifyFiles([(path, total)], L, None)
    out["pread_ring_threads%d_GBps" % thr] = round(total / (time.perf_counter() - t0) / 1e9, 2)
    c2.close()
ctx.close()
os.remove(path)
print(json.dumps(out))
