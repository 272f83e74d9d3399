"""Kernel-only rates of the BASELINE config shapes for the current GPUHASH_* tuning switches.
Usage: GPUHASH_FMA_ADDS=0|1 GPUHASH_TPS=2|4 python tests/tools/ab_kernels.py  -> one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import torch

import gpuhash

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


total = 8 << 30
d = torch.empty(total, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 1, 0, total, stream=sp)
o1 = torch.zeros((total // 16384) * 32, dtype=torch.uint8, device="cuda")
o2 = torch.zeros((total // 16384) * 32, dtype=torch.uint8, device="cuda")
out = {"env": {k: os.environ.get(k) for k in ("GPUHASH_LIB", "GPUHASH_FMA_ADDS", "GPUHASH_TPS", "GPUHASH_SPLIT_WARPS_PER_SM")}}
quick = os.environ.get("AB_QUICK") == "1"
for name, L, tot in (("sha1_16KiB", 16 << 10, total), ("sha1_256KiB_1GiB", 256 << 10, 1 << 30), ("sha1_256KiB", 256 << 10, total),
                     ("sha1_1MiB", 1 << 20, total), ("sha1_4MiB", 4 << 20, total), ("sha1_16MiB", 16 << 20, total)):
    ms = timed(lambda: ctx.verify_pieces_device(d.data_ptr(), tot, L, None, o1.data_ptr(), None, stream=sp))
    out[name] = round(tot / ms / 1e6, 1)
for name, L, tot in ((("64KiB", 64 << 10, total), ("1MiB", 1 << 20, total)) if quick else (("64KiB", 64 << 10, total), ("1MiB", 1 << 20, total), ("64MiB_2GiB", 64 << 20, 2 << 30))):
    for algos, an in ((gpuhash.MD5, "md5"), (gpuhash.SHA256, "sha256"), (gpuhash.MD5 | gpuhash.SHA256, "md5+sha256")):
        ms = timed(lambda: ctx.hash_parts_device(d.data_ptr(), tot, L, algos, None, o1.data_ptr(), o2.data_ptr(), stream=sp), reps=2)
        out[an + "_" + name] = round(tot / ms / 1e6, 1)
print(json.dumps(out))
