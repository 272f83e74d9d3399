"""Kernel-only SHA-1 rates of the BASELINE piece shapes (payload resident in HBM, CUDA events) for one
library build / switch setting; one JSON line. The caller loops over GPUHASH_LIB / GPUHASH_* settings:
  GPUHASH_TRIO_WARPS_PER_SM=0 python tests/tools/ab_sha1_shapes.py
AB_NOCHECK=1 skips the digest check (measurement builds that leave a role idle)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import torch

import gpuhash
import oracle

SHAPES = [("cfg2_12800x4MiB", 12800, 4 << 20), ("cfg1_4096x256KiB", 4096, 256 << 10), ("cfg4_3072x16MiB", 3072, 16 << 20),
          ("8192x4MiB", 8192, 4 << 20), ("14208x1MiB", 14208, 1 << 20)]
only = set(sys.argv[1:])
ctx = gpuhash.Context(devices=[0])
stream = torch.cuda.current_stream()
sp = stream.cuda_stream
cap = max(n * L for _, n, L in SHAPES)
d = torch.empty(cap, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 2, 0, cap, stream=sp)
dig = torch.zeros(20 * max(n for _, n, _ in SHAPES), dtype=torch.uint8, device="cuda")
out = {"lib": os.environ.get("GPUHASH_LIB", "default"),
       "env": {k: v for k, v in os.environ.items() if k.startswith("GPUHASH_") and k != "GPUHASH_LIB"}}
for name, n, L in SHAPES:
    if only and name not in only:
        continue
    total = n * L

    def run():
        ctx.verify_pieces_device(d.data_ptr(), total, L, None, dig.data_ptr(), None, stream=sp)

    run()
    torch.cuda.synchronize()
    got = dig.cpu().numpy()[: n * 20].reshape(n, 20)
    for i in (() if os.environ.get("AB_NOCHECK") else (0, n // 2, n - 1)):
        assert got[i].tobytes() == oracle.digest(oracle.SHA1, oracle.gen(2, i * L, L)), (name, i)
    times = []
    for _ in range(4):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        run()
        b.record(stream)
        torch.cuda.synchronize()
        times.append(a.elapsed_time(b))
    ms = min(times)
    blocks = (L + 9 + 63) // 64
    out[name] = {"ms": round(ms, 3), "GBps": round(total / ms / 1e6, 1), "clk_per_block": round(ms * 1e-3 * 1.965e9 / blocks, 1)}
print(json.dumps(out))
