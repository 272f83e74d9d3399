"""gh_verify_pieces on ordinary (pageable) host memory: staged through the pinned ring by the pool."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import gpuhash, oracle
total = 8 << 30
data = np.empty(total, dtype=np.uint8)
blk = oracle.gen(9, 0, 256 << 20)
for i in range(total // blk.size):
    data[i * blk.size:(i + 1) * blk.size] = blk
out = {}
for thr in (0, 4, 8):
    ctx = gpuhash.Context(devices=[0], stage_threads=thr)
    for L in (4 << 20, 16 << 10):
        ctx.VerifyPieces([data], L, None)
        t0 = time.perf_counter()
        ctx.VerifyPieces([data], L, None)
        out["threads%s_L%d" % (thr or "default", L)] = round(total / (time.perf_counter() - t0) / 1e9, 2)
    ctx.close()
print(json.dumps(out))
