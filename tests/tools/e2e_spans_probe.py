"""Config-5-like e2e: 10 000 pinned spans, 16 KiB pieces. Prints library wall / kernel time per call."""
import ctypes, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
import gpuhash, oracle
ctx = gpuhash.Context(devices=[0])
sizes = [1 + (oracle.splitmix64((5 << 32) + k) % 2097152) for k in range(10000)]
total = sum(sizes)
d = torch.empty(total, dtype=torch.uint8, device="cuda")
ctx.generate(d.data_ptr(), 5, 0, total)
host = gpuhash.PinnedBuffer(total)
ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), total)
files, o = [], 0
for s in sizes:
    files.append(host.array[o:o + s]); o += s
sl = gpuhash.SpanList(files)
one = gpuhash.SpanList([host.array])
# the same files with 4 KiB gaps between them (separate allocations in real life): cannot be merged
gap = gpuhash.PinnedBuffer(total + 4096 * len(sizes))
gfiles, o, g = [], 0, 0
for s in sizes:
    gap.array[g:g + s] = host.array[o:o + s]
    gfiles.append(gap.array[g:g + s]); o += s; g += s + 4096
gl = gpuhash.SpanList(gfiles)
print("copy streams:", os.environ.get("GPUHASH_COPY_STREAMS"))
for name, spans in (("10k_contiguous_spans", sl), ("10k_gapped_spans", gl), ("1_span", one)):
    for i in range(3):
        t0 = time.perf_counter()
        ctx.VerifyPieces(spans, 16384, None)
        dt = time.perf_counter() - t0
        st = ctx.stats()
        print(name, i, "py_wall_ms %.1f lib_wall_ms %.1f kernel_ms %.1f GB/s %.2f" % (dt * 1e3, st.wall_ms, st.kernel_ms, total / dt / 1e9), flush=True)
