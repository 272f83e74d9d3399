"""Aggregate pinned host->device bandwidth with 1/2/4/8 GPUs copying at once: the end-to-end roofline
for multi-GPU runs. Plain cudaMemcpyAsync through torch, no hashing."""
import json, sys, time
import torch
n = torch.cuda.device_count()
size = 1 << 30
host = [torch.empty(size, dtype=torch.uint8).pin_memory() for _ in range(n)]
dev = [torch.empty(size, dtype=torch.uint8, device="cuda:%d" % i) for i in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]
out = {}
subsets = {1: [[0]], 2: [[0, 1], [0, n // 2]], 4: [list(range(4)), list(range(0, n, max(1, n // 4)))[:4]], 8: [list(range(8))]}
for k, variants in subsets.items():
    if k > n:
        continue
    for ids in variants:
        if max(ids) >= n:
            continue
        def run(reps):
            for _ in range(reps):
                for i in ids:
                    with torch.cuda.stream(streams[i]):
                        dev[i].copy_(host[i], non_blocking=True)
            for i in ids:
                streams[i].synchronize()
        run(2)
        t0 = time.perf_counter()
        run(10)
        dt = time.perf_counter() - t0
        out["gpus_" + "_".join(map(str, ids))] = round(10 * len(ids) * size / dt / 1e9, 1)
print(json.dumps(out))
