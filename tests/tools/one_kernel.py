"""Runs one device-resident hashing launch a few times (for ncu captures).
Usage: python tests/tools/one_kernel.py <algos mask> <msg_len> <total_bytes>"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import torch

import gpuhash

algos, L, total = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ctx = gpuhash.Context(devices=[0])
sp = torch.cuda.current_stream().cuda_stream
d = torch.empty(total, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 1, 0, total, stream=sp)
n = total // L
o = [torch.zeros(n * 32 + 64, dtype=torch.uint8, device="cuda") for _ in range(3)]
for _ in range(3):
    ctx.hash_parts_device(d.data_ptr(), total, L, algos, o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), stream=sp)
torch.cuda.synchronize()
print("ok")
