"""All five BASELINE.json configs on one GPU: kernel-only (payload in HBM, CUDA events) and end to
end (pinned host spans through the ring), each checked against the CPU oracle, with the oracle port
timed on all host threads beside it. Writes gpurun_out/configs.json. Not the driver's bench (that is
bench.py, configs[1]); this is the per-config table in DESIGN.md."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch

import gpuhash
import oracle

only = set(sys.argv[1:])
ctx = gpuhash.Context(devices=[0])
stream = torch.cuda.current_stream()
sp = stream.cuda_stream
threads = len(os.sched_getaffinity(0))
out = {"host_threads": threads}


def ev_time(fn, reps=3):
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


def wall(fn, reps=2):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def to_host(d, n):
    host = gpuhash.PinnedBuffer(n)
    ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), n)
    return host


def cpu_sha1(host_arr, L, sample):
    s = min(sample, host_arr.size) // L * L
    t0 = time.perf_counter()
    oracle.hash_msgs(oracle.SHA1, host_arr[:s], L, threads)
    return round(s / (time.perf_counter() - t0) / 1e9, 2)


def openssl_rate(host_arr, L, sample, algs=("sha1",)):
    """OpenSSL EVP (SHA-NI for SHA-1 / SHA-256 where the CPU has it) on all host threads through the C
    harness oracle/gh_openssl.c: the upper bound of the two CPU figures (the oracle port is the lower)."""
    s = min(sample, host_arr.size) // L * L
    mask = sum({"sha1": 1, "md5": 2, "sha256": 4}[a] for a in algs)
    t0 = time.perf_counter()
    oracle.openssl_hash_msgs(mask, host_arr[:s], L, threads)
    return round(s / (time.perf_counter() - t0) / 1e9, 2)


def alu_frac(kernel, n_msgs, msg_len, ms):
    """Fraction of the INT32 ALU pipes' issue capacity (148 SMs x 64 lanes x 1 965 MHz) the launch used:
    blocks/s x ALU instructions per block from the shipped SASS (bench.py's roofline formula)."""
    import bench
    mix = bench.alu_instructions_per_block(kernel)
    if not mix:
        return None
    blocks = n_msgs * ((msg_len + 9 + 63) // 64)
    return round(blocks / (ms / 1e3) * mix["alu"] / (148 * 64 * 1.965e9), 3)


def sha1_config(name, seed, file_sizes, L, corrupt_cfg5=False):
    total = sum(file_sizes)
    n = (total + L - 1) // L
    d = torch.empty(total + 64, dtype=torch.uint8, device="cuda")
    ctx.generate(d.data_ptr(), seed, 0, total, stream=sp)
    exp = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    dig = torch.zeros(n * 20, dtype=torch.uint8, device="cuda")
    ok = torch.zeros((n + 31) // 32 + 1, dtype=torch.int32, device="cuda")
    ctx.verify_pieces_device(d.data_ptr(), total, L, None, exp.data_ptr(), None, stream=sp)
    torch.cuda.synchronize()
    exp_host = exp.cpu().numpy().reshape(n, 20).copy()
    res = {"pieces": n, "bytes": total, "files": len(file_sizes), "piece_len": L}
    # oracle spot checks on clean data: first, last (short) and every file-boundary piece (up to 8)
    check = {0, n - 1}
    off = 0
    for s in file_sizes[:-1][:8]:
        off += s
        check.add(off // L)
    for i in sorted(check):
        lo, hi = i * L, min((i + 1) * L, total)
        assert exp_host[i].tobytes() == oracle.digest(oracle.SHA1, oracle.gen(seed, lo, hi - lo)), (name, i)
    bad = []
    if corrupt_cfg5:
        idx, offs, bits = [], [], []
        for i in range(n):
            plen = min(L, total - i * L)
            hit, o, b = oracle.cfg5_corruption(i, plen)
            if hit or i == n - 1:
                idx.append(i)
                offs.append(i * L + o)
                bits.append(1 << b)
        bad = idx
        d[torch.tensor(offs, device="cuda")] ^= torch.tensor(bits, dtype=torch.uint8, device="cuda")
    ms = ev_time(lambda: ctx.verify_pieces_device(d.data_ptr(), total, L, exp.data_ptr(), dig.data_ptr(), ok.data_ptr(), stream=sp))
    words = ok.cpu().numpy().view(np.uint32)
    got_bad = [i for i in range(n) if not (int(words[i >> 5]) >> (i & 31)) & 1] if n < 2000000 else None
    assert got_bad == bad, (name, got_bad[:5], bad[:5])
    res["kernel_only_GBps"] = round(total / ms / 1e6, 1)
    res["kernel_ms"] = round(ms, 3)
    res["kernel"] = gpuhash.last_kernel()
    res["alu_pipe_frac"] = alu_frac(res["kernel"], n, L, ms)
    # end to end from pinned host spans (one span per file)
    host = to_host(d, total)
    files, o = [], 0
    for s in file_sizes:
        files.append(host.array[o:o + s])
        o += s
    files = gpuhash.SpanList(files)  # the C call is what is timed, not Python's per-file bookkeeping
    r = {}

    def e2e():
        r["v"] = ctx.VerifyPieces(files, L, exp_host)
    dt = wall(e2e)
    bitmap, digs, nbad = r["v"]
    assert nbad == len(bad)
    res["e2e_GBps"] = round(total / dt / 1e9, 2)
    res["e2e_kernel_ms"] = round(ctx.stats().kernel_ms, 2)
    if corrupt_cfg5:
        # deliverable of config 5: bitmap bit-identical to the CPU's, n_bad equal
        t0 = time.perf_counter()
        want_bitmap, want_bad = oracle.verify_pieces(host.array, L, exp_host, threads)
        res["cpu_port_verify_GBps_full"] = round(total / (time.perf_counter() - t0) / 1e9, 2)
        assert want_bad == nbad and np.array_equal(bitmap, want_bitmap)
        res["bitmap_identical_to_cpu"] = True
        res["n_bad"] = nbad
    res["cpu_port_GBps"] = cpu_sha1(host.array, L, 8 << 30)
    res["cpu_openssl_GBps"] = openssl_rate(host.array, L, 8 << 30)
    host.free()
    del d
    torch.cuda.empty_cache()
    out[name] = res
    print(name, json.dumps(res), flush=True)


if not only or "1" in only:
    sha1_config("config1_1GiB_256KiB", 1, [1 << 30], 256 << 10)
if not only or "4" in only:
    sha1_config("config4_24x2GiB_16MiB", 4, [(2 << 30) + 7919 * (k + 1) for k in range(24)], 16 << 20)
if not only or "5" in only:
    sizes = [1 + (oracle.splitmix64((5 << 32) + k) % 2097152) for k in range(10000)]
    sha1_config("config5_10kfiles_16KiB", 5, sizes, 16 << 10, corrupt_cfg5=True)
if not only or "3" in only:
    P, total = 64 << 20, 20 << 30
    n = total // P
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    ctx.generate(d.data_ptr(), 3, 0, total, stream=sp)
    md5 = torch.zeros(n * 16, dtype=torch.uint8, device="cuda")
    sha = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    res = {"parts": n, "bytes": total, "part_len": P}
    for algos, nm in ((gpuhash.MD5 | gpuhash.SHA256, "fused"), (gpuhash.MD5, "md5"), (gpuhash.SHA256, "sha256")):
        ms = ev_time(lambda: ctx.hash_parts_device(d.data_ptr(), total, P, algos, None, md5.data_ptr(), sha.data_ptr(), stream=sp), reps=1)
        res["kernel_only_%s_GBps" % nm] = round(total / ms / 1e6, 2)
        res["kernel_%s" % nm] = gpuhash.last_kernel()
    m = md5.cpu().numpy().reshape(n, 16)
    s = sha.cpu().numpy().reshape(n, 32)
    for i in (0, n - 1):
        part = oracle.gen(3, i * P, P)
        assert m[i].tobytes() == oracle.digest(oracle.MD5, part) and s[i].tobytes() == oracle.digest(oracle.SHA256, part)
    host = to_host(d, total)
    r = {}

    def e2e3():
        r["v"] = ctx.HashParts([host.array], P, gpuhash.MD5 | gpuhash.SHA256)
    dt = wall(e2e3, reps=1)
    assert np.array_equal(r["v"][gpuhash.MD5], m) and np.array_equal(r["v"][gpuhash.SHA256], s)
    res["e2e_fused_GBps"] = round(total / dt / 1e9, 2)
    res["e2e_kernel_ms"] = round(ctx.stats().kernel_ms, 1)
    res["s3_etag"] = ctx.S3ETag(m)
    assert res["s3_etag"] == oracle.s3_etag(m)
    smp = 4 << 30
    t0 = time.perf_counter()
    oracle.hash_msgs(oracle.MD5, host.array[:smp], P, threads)
    oracle.hash_msgs(oracle.SHA256, host.array[:smp], P, threads)
    res["cpu_port_md5_plus_sha256_GBps"] = round(smp / (time.perf_counter() - t0) / 1e9, 2)
    res["cpu_openssl_md5_plus_sha256_GBps"] = openssl_rate(host.array, P, total, ("md5", "sha256"))
    host.free()
    out["config3_20GiB_64MiB_md5_sha256"] = res
    print("config3", json.dumps(res), flush=True)
if not only or "chunks" in only:
    # what the reference's exact upload call hashes client-side over plain HTTP (uploader.go:32-35, :89):
    # SHA-256 of every 64 KiB chunk of the body (streaming SigV4) -- 327 680 independent messages per 20 GiB
    C, total = 64 << 10, 20 << 30
    n = total // C
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    ctx.generate(d.data_ptr(), 3, 0, total, stream=sp)
    sha = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    ms = ev_time(lambda: ctx.hash_parts_device(d.data_ptr(), total, C, gpuhash.SHA256, None, None, sha.data_ptr(), stream=sp))
    res = {"chunks": n, "bytes": total, "chunk_len": C, "kernel_only_GBps": round(total / ms / 1e6, 1), "kernel_ms": round(ms, 3),
           "kernel": gpuhash.last_kernel()}
    res["alu_pipe_frac"] = alu_frac(res["kernel"], n, C, ms)
    res["frac_of_sha256_alu_bound_1072GBps"] = round(res["kernel_only_GBps"] / 1072.0, 3)
    s = sha.cpu().numpy().reshape(n, 32)
    host = to_host(d, total)
    for off in (0, total - (1 << 30)):
        want = oracle.hash_msgs(oracle.SHA256, host.array[off:off + (1 << 30)], C, threads)
        assert np.array_equal(s[off // C:(off + (1 << 30)) // C], want)
    r = {}

    def e2ec():
        r["v"] = ctx.HashParts([host.array], C, gpuhash.SHA256)
    dt = wall(e2ec, reps=2)
    assert np.array_equal(r["v"][gpuhash.SHA256], s)
    res["e2e_GBps"] = round(total / dt / 1e9, 2)
    res["e2e_kernel_ms"] = round(ctx.stats().kernel_ms, 1)
    # the chunk signatures chain in order (each signs the previous signature + this chunk's hex digest):
    # the signer needs lowercase hex, chunk order -- exactly the rows of the output, hexlified
    res["first_chunk_hex"] = s[0].tobytes().hex()
    smp = 4 << 30
    t0 = time.perf_counter()
    oracle.hash_msgs(oracle.SHA256, host.array[:smp], C, threads)
    res["cpu_port_GBps"] = round(smp / (time.perf_counter() - t0) / 1e9, 2)
    res["cpu_openssl_GBps"] = openssl_rate(host.array, C, 8 << 30, ("sha256",))
    host.free()
    del d
    torch.cuda.empty_cache()
    out["sha256_64KiB_chunks_20GiB"] = res
    print("chunks", json.dumps(res), flush=True)

os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", os.environ.get("CONFIGS_OUT", "configs.json")), "w"), indent=1)
print("done")
