"""gh_verify_files from the page cache: throughput vs number of staging threads."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import gpuhash, oracle
total = int(os.environ.get("GIB", "8")) << 30
path = "/tmp/gh_payload.bin"
if not os.path.exists(path) or os.path.getsize(path) != total:
    blk = oracle.gen(7, 0, 256 << 20)
    with open(path, "wb") as f:
        for i in range(total // blk.size):
            f.write(blk.tobytes())
out = {}
for thr in (4, 8, 16, 32):
    ctx = gpuhash.Context(devices=[0], stage_threads=thr)
    for L in (4 << 20, 16 << 10):
        ctx.VerifyFiles([(path, total)], L, None)
        t0 = time.perf_counter()
        bm, dig, nb = ctx.VerifyFiles([(path, total)], L, None)
        dt = time.perf_counter() - t0
        out["threads%d_L%d" % (thr, L)] = round(total / dt / 1e9, 2)
    ctx.close()
want = oracle.digest(oracle.SHA1, oracle.gen(7, 0, 16 << 10))
assert dig[0].tobytes() == want
print(json.dumps(out))
os.remove(path)
