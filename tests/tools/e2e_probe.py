"""e2e vs kernel time for long-message (chunk-interleaved) shapes; GPUHASH_NO_2D=1 switches the ring
to per-row 1-D copies. Prints one JSON line."""
import ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
import gpuhash
total = 8 << 30
out = {"no2d": os.environ.get("GPUHASH_NO_2D"), "slots": os.environ.get("SLOTS"), "slot_mib": os.environ.get("SLOT_MIB")}
ctx = gpuhash.Context(devices=[0], slots=int(os.environ.get("SLOTS", "0")), slot_bytes=int(os.environ.get("SLOT_MIB", "0")) << 20)
d = torch.empty(total, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 3, 0, total)
host = gpuhash.PinnedBuffer(total)
ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), total)
for name, P, algos in (("fused_64MiB", 64 << 20, 6), ("sha1_64MiB", 64 << 20, 1), ("sha1_4MiB", 4 << 20, 1)):
    ctx.HashParts([host.array], P, algos)
    t0 = time.perf_counter()
    ctx.HashParts([host.array], P, algos)
    dt = time.perf_counter() - t0
    st = ctx.stats()
    out[name] = {"wall_ms": round(dt * 1e3, 1), "kernel_ms": round(st.kernel_ms, 1), "GBps": round(total / dt / 1e9, 2)}
print(json.dumps(out))
