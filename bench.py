#!/usr/bin/env python
"""bench.py -- verified GB/s of batched SHA-1 piece verification (BASELINE.json configs[1]).

Workload (per GPU): synthetic 50 GiB torrent payload (SURVEY.md Appendix C stream, seed 2), 4 MiB
pieces, 12 800 pieces, every piece hashed with SHA-1 and compared with its expected digest on the
device. A step is one pass over the whole payload.

  value      kernel-only: payload resident in HBM, one launch per step, CUDA events on the launching
             stream (torch's current stream is handed to the C ABI).
  e2e        the same pass through gh_verify_pieces with HOST (pinned) buffers: H2D of the payload
             and D2H of digests + bitmap inside the timed region; next to it the link roofline
             measured in the same run (plain pinned cudaMemcpyAsync on every rank at once).
  roofline   the roof that binds the kernel: the INT32 ALU pipe. achieved = 64-byte blocks per
             second x ALU-pipe instructions per block (counted from the SASS of the shipped
             library for the kernel that ran), peak = SMs x 64 lanes x the SM clock sampled during
             the run. The HBM figures (algorithmic bytes / time vs MEASURED_PEAKS.json) ride along.
  cpu_baseline  the oracle's scalar C port (oracle/gh_oracle.c) on all host threads, bounded sample,
             with OpenSSL (SHA-NI) on the same threads beside it.

--impl reference times the CPU path alone (the reference's Go crypto/sha1 cannot run: no Go
toolchain in this image; the oracle port stands in, kind "port"), same config.

N>1 (torchrun, one rank per GPU): weak scaling, every rank verifies its own 50 GiB payload; pieces
are independent so there is no collective on the data path (torch.distributed only carries barriers
and the max-over-ranks of the timings). When the box shows more GPUs than ranks, the ranks are placed
on the device set with the better measured aggregate host->device bandwidth. After the weak-scaling
part rank 0 alone verifies BASELINE config 4 (24 files, 48 GiB, 16 MiB pieces spanning file
boundaries) through ONE gh_ctx over all N devices: strong scaling of the in-process sharding.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200"), os.path.join(ROOT, "tests", "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

PIECE_LEN = 4 << 20
TOTAL = 50 << 30
SEED = 2
METRIC = "verified GB/s (SHA-1 piece verification, 50 GiB / 4 MiB pieces)"
WORKLOAD = "configs[1]: synthetic 50 GiB torrent payload, 4 MiB pieces, batched SHA-1 verify"


def shared_config(total, world):
    """The part of `config` both arms print identically."""
    return {"workload": WORKLOAD, "payload_bytes_per_gpu": total, "piece_len": PIECE_LEN,
            "pieces_per_gpu": total // PIECE_LEN, "seed": SEED,
            "l2": "inputs (%.0f GiB) exceed the 126 MB L2 and every CPU cache; no flush needed" % (total / 2**30),
            "parallelism": "independent pieces, %d rank(s), no collective" % world}


def mem_available():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 0


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML)."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._t = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
            return self
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.1)

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_profile():
    """Per-launch DRAM traffic of the bench kernel from the committed ncu capture (stamped with the
    commit and kernel it was taken on), if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_inputs.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def alu_instructions_per_block(kernel):
    """ALU-pipe (LOP3/SHF/IADD3/PRMT/LEA ...) instructions one message executes per 64-byte block in
    `kernel`, read off the SASS of the shipped libgpuhash.so: the sum over the kernel's per-block
    loops (innermost loops of at least 150 ALU instructions -- one for the lane kernels, the rounds
    loop plus the schedule loop for the role-split ones)."""
    import sass_count
    table = sass_count.loop_table()
    want = kernel.replace(" ", "")
    for name, rows in table.items():
        flat = name.replace("(int)", "").replace("(bool)", "").replace(" ", "")
        if not (flat == want or flat.endswith("::" + want)):
            continue
        spans = [(int(r["from"], 16), int(r["to"], 16), r) for r in rows]
        inner = [r for lo, hi, r in spans if not any(lo <= lo2 and hi2 <= hi and (lo2, hi2) != (lo, hi) for lo2, hi2, _ in spans)]
        per_block = [r for r in inner if r["alu"] >= 150]
        if per_block:
            return {"alu": sum(r["alu"] for r in per_block), "fma": sum(r["fma"] for r in per_block),
                    "all": sum(r["instrs"] for r in per_block), "loops": len(per_block),
                    "library_sha256": sass_count.lib_sha256()[:16]}
    return None


def parallel_gen(oracle, seed, offset, ptr, n, threads):
    """Fill host memory at `ptr` with stream(seed)[offset:offset+n] using `threads` host threads."""
    step = max(1 << 20, ((n + threads - 1) // threads + 63) & ~63)

    def work(lo):
        hi = min(n, lo + step)
        oracle.lib().gho_gen_fill(seed, offset + lo, hi - lo, ctypes.c_void_p(ptr + lo))
    ts = [threading.Thread(target=work, args=(lo,)) for lo in range(0, n, step)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()


def openssl_rate(data, piece_len, threads):
    """GB/s of OpenSSL's SHA-1 (SHA-NI where the CPU has it) on `threads` threads: the C harness of
    oracle/gh_openssl.c when it could be built, else hashlib (same library, Python call overhead)."""
    import oracle
    if oracle.openssl_lib() is not None:
        t0 = time.perf_counter()
        oracle.openssl_hash_msgs(oracle.SHA1, data, piece_len, threads)
        return len(data) / (time.perf_counter() - t0) / 1e9
    import hashlib
    from concurrent.futures import ThreadPoolExecutor
    mv = memoryview(data)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda i: hashlib.sha1(mv[i:i + piece_len]).digest(), range(0, len(mv), piece_len)))
    return len(mv) / (time.perf_counter() - t0) / 1e9


def emit(fd, line):
    os.write(fd, (json.dumps(line) + "\n").encode())


def run_reference(args, out_fd):
    """The CPU arm: oracle port (no Go toolchain -> the reference's crypto/sha1 cannot run), all host
    threads, same metric and config; a step is one pass over the whole 50 GiB payload when host RAM
    and the time budget allow, else over a stated prefix of it."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    import numpy as np
    import oracle
    threads = host_threads()
    probe = np.empty(64 * PIECE_LEN, dtype=np.uint8)
    parallel_gen(oracle, SEED, 0, probe.ctypes.data, probe.size, threads)
    t0 = time.perf_counter()
    oracle.hash_msgs(oracle.SHA1, probe, PIECE_LEN, threads)
    rate = probe.size / (time.perf_counter() - t0)  # bytes/s
    total = int(args.total_gib * 2**30) // PIECE_LEN * PIECE_LEN
    budget_s = float(os.environ.get("GH_BENCH_REF_SECONDS", "150"))
    step_s = os.environ.get("GH_BENCH_REF_STEP_SECONDS")  # tests: a fixed, small amount of CPU work per step
    want = int(rate * float(step_s)) if step_s else int(rate * budget_s / max(args.steps + 1, 1))
    avail = mem_available()
    cap = max(1 << 30, avail - (8 << 30)) if avail else 4 << 30
    sample = max(16 * PIECE_LEN, min(total, cap, want) // PIECE_LEN * PIECE_LEN)
    data = np.empty(sample, dtype=np.uint8)
    parallel_gen(oracle, SEED, 0, data.ctypes.data, sample, threads)
    expected = oracle.hash_msgs(oracle.SHA1, data, PIECE_LEN, threads)
    warm = data[: min(sample, 64 * PIECE_LEN)]
    for _ in range(args.warmup):  # warm-up on a small prefix: threads and caches, not 50 GiB passes
        oracle.verify_pieces(warm, PIECE_LEN, expected[: warm.size // PIECE_LEN], threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        bitmap, bad = oracle.verify_pieces(data, PIECE_LEN, expected, threads)
        assert bad == 0
    dt = time.perf_counter() - t0
    gbs = sample * args.steps / dt / 1e9
    full = sample == total
    sample_desc = ("the whole payload per step" if full else
                   "first %.2f GiB (%d pieces) of the payload per step (%s)" % (
                       sample / 2**30, sample // PIECE_LEN,
                       "fixed CPU work per step" if step_s else "time budget %.0f s / host RAM" % budget_s))
    cpu = {"value": round(gbs, 3), "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample_desc}
    try:
        osl = data[: min(sample, 8 << 30)]
        cpu["openssl_sha_ni_gbs"] = round(openssl_rate(osl, PIECE_LEN, threads), 3)
        cpu["openssl_sample"] = "first %.2f GiB, EVP_sha1 on %d threads" % (osl.size / 2**30, threads)
    except Exception:
        pass
    config = shared_config(total, world)
    if not full:
        config["cpu_sample"] = sample_desc
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": config,
        "notes": {"cpu": "Go crypto/sha1 not runnable (no Go toolchain); scalar C port of the same algorithm, every piece "
                         "verified against its expected digest; OpenSSL/SHA-NI on the same threads in cpu_baseline"},
        "cpu_baseline": cpu,
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(out_fd, line)
    return 0


def link_probe(torch, device, reps, sync_all, max_over_ranks, world, chunk=1 << 30):
    """Aggregate host->device GB/s of plain pinned cudaMemcpyAsync, every rank copying `reps` x 1 GiB
    to `device` at the same time (barriers on both sides, the slowest rank's clock). The copies go out
    as 256 MiB pieces on four streams, the shape of the ring's own traffic, so that the figure is the
    roof of what the product can be given -- one big copy at a time gets a few per cent less of an
    uplink that several GPUs share."""
    with torch.cuda.device(device):
        src = torch.empty(chunk, dtype=torch.uint8, pin_memory=True)
        dst = torch.empty(chunk, dtype=torch.uint8, device="cuda:%d" % device)
        src.fill_(1)
        streams = [torch.cuda.Stream(device=device) for _ in range(4)]
        piece = chunk // 4

        def one_pass():
            for k, st in enumerate(streams):
                with torch.cuda.stream(st):
                    dst[k * piece:(k + 1) * piece].copy_(src[k * piece:(k + 1) * piece], non_blocking=True)

        one_pass()
        torch.cuda.synchronize(device)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(reps):
            one_pass()
        torch.cuda.synchronize(device)
        dt = max_over_ranks(time.perf_counter() - t0)
        sync_all()
        del src, dst
        torch.cuda.empty_cache()
    return world * reps * chunk / dt / 1e9


def config4_strong(gpuhash, oracle, np, devices, threads):
    """BASELINE config 4 through ONE gh_ctx over `devices`: 24 files of 2 GiB + 7919(k+1) bytes, 16 MiB
    pieces (23 of them span a file boundary), pieces sharded over the devices by the library; expected
    digests spot-checked against the oracle, one injected fault that must be the only bad piece."""
    L = 16 << 20
    sizes = [(2 << 30) + 7919 * (k + 1) for k in range(24)]
    total = sum(sizes)
    n = (total + L - 1) // L
    starts = [sum(sizes[:k]) for k in range(len(sizes))]
    # host side: one pinned buffer per file (what a downloader that pins its file buffers would hold);
    # if the driver refuses to pin that much next to the other ranks' contexts, pageable memory -- the
    # library then stages through its own pinned ring -- and the line says so
    bufs, host_mode = [], "pinned, one buffer per file"
    try:
        for sz in sizes:
            bufs.append(gpuhash.PinnedBuffer(sz))
        arrays = [b.array for b in bufs]
    except gpuhash.GpuHashError as e:
        for b in bufs:
            b.free()
        bufs = []
        arrays = [np.empty(sz, dtype=np.uint8) for sz in sizes]
        host_mode = "pageable (pinned allocation refused: %s)" % e
    try:
        for a, st in zip(arrays, starts):
            parallel_gen(oracle, 4, st, a.ctypes.data, a.size, threads)

        def stream(lo, hi):  # bytes [lo, hi) of the concatenated stream
            parts = []
            for a, st in zip(arrays, starts):
                x, y = max(lo, st), min(hi, st + a.size)
                if x < y:
                    parts.append(a[x - st:y - st])
            return parts[0] if len(parts) == 1 else np.concatenate(parts)

        spans = gpuhash.SpanList(arrays)
        ctx = gpuhash.Context(devices=list(devices))
        try:
            _, dig, _ = ctx.VerifyPieces(spans, L, None)
            expected = dig.copy()
            check = {0, n - 1}
            for st in starts[1:4]:
                check.add(st // L)  # pieces that straddle a file boundary
            for i in sorted(check):
                lo, hi = i * L, min((i + 1) * L, total)
                assert expected[i].tobytes() == oracle.digest(oracle.SHA1, stream(lo, hi)), "config 4 piece %d" % i
            victim = 1234
            pos = victim * L + 99
            k = max(j for j, st in enumerate(starts) if st <= pos)
            arrays[k][pos - starts[k]] ^= 1
            ctx.VerifyPieces(spans, L, expected)
            t0 = time.perf_counter()
            bitmap, dig, nbad = ctx.VerifyPieces(spans, L, expected)
            dt = time.perf_counter() - t0
            bad = [i for i in range(n) if not (int(bitmap[i >> 6]) >> (i & 63)) & 1]
            assert nbad == 1 and bad == [victim], (nbad, bad[:4])
            return {"e2e": round(total / dt / 1e9, 2), "unit": "GB/s", "ms": round(dt * 1e3, 1), "pieces": n, "files": len(sizes),
                    "bytes": total, "n_bad": int(nbad), "devices": list(devices), "host_memory": host_mode,
                    "how": "one gh_ctx over all ranks' devices, host spans one per file, bitmap == {only piece %d bad}" % victim}
        finally:
            ctx.close()
    finally:
        for b in bufs:
            b.free()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--total-gib", type=float, default=TOTAL / 2**30, help="debug only: shrink the payload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-config4", action="store_true", help="skip the config-4 strong-scaling leg at N>1")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    # stdout carries exactly one JSON line: everything else any library prints (NCCL's banner and
    # NCCL_DEBUG output included -- the driver's setting is left alone) goes to stderr for the whole run
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args, out_fd)

    import numpy as np
    import torch

    import gpuhash

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        print("bench.py: no CUDA device; the product path has no CPU fallback", file=sys.stderr)
        return 2
    visible = torch.cuda.device_count()
    dist = None
    if world > 1:
        import torch.distributed as dist
        # gloo carries the host-side control plane (placement probe, waiting for rank 0's config-4 leg),
        # NCCL the barriers and reductions around the timed regions (communicator created on first use,
        # on the device chosen below)
        dist.init_process_group("cpu:gloo,cuda:nccl")

    def host_barrier():
        if dist:
            dist.all_reduce(torch.zeros(1))

    def host_max(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- placement: which `world` of the `visible` devices. GPUs behind the same PCIe switch share its
    # uplink, so when there is a choice measure both the first `world` devices and an evenly strided set
    placement = {"visible": visible, "candidates": {}}
    cands = {"first": list(range(world))}
    env_dev = os.environ.get("GPUHASH_DEVICES")
    if env_dev:
        cands = {"env": [int(x) for x in env_dev.split(",")][:world]}
    elif world > 1 and visible >= 2 * world:
        cands["strided"] = [i * (visible // world) for i in range(world)]
    if len(cands) > 1:
        for name, devs in cands.items():
            placement["candidates"][name] = {"devices": devs, "h2d_gbs": round(
                link_probe(torch, devs[local_rank], 40, host_barrier, host_max, world), 1)}
        best = max(placement["candidates"], key=lambda k: placement["candidates"][k]["h2d_gbs"])
    else:
        best = next(iter(cands))
    devices = cands[best]
    placement["chosen"] = best
    placement["devices"] = devices
    dev = devices[local_rank]
    torch.cuda.set_device(dev)

    def barrier():
        if dist:
            dist.all_reduce(torch.zeros(1, device="cuda"))
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    barrier()
    total = int(args.total_gib * 2**30) // PIECE_LEN * PIECE_LEN
    n_pieces = total // PIECE_LEN
    seed = SEED + 1000 * rank  # every rank its own payload (weak scaling)

    # NUMA: allocate pinned memory from the CPUs next to this GPU when NVML can tell us which
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(dev)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass

    ctx = gpuhash.Context(devices=[dev])
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream

    # ---- payload in HBM, expected digests from a first (untimed) pass, spot-checked vs the oracle
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    ctx.generate(d.data_ptr(), seed, 0, total, stream=sp)
    expected = torch.zeros(n_pieces * 20, dtype=torch.uint8, device="cuda")
    digests = torch.zeros(n_pieces * 20, dtype=torch.uint8, device="cuda")
    ok = torch.zeros((n_pieces + 31) // 32 + 1, dtype=torch.int32, device="cuda")
    ctx.verify_pieces_device(d.data_ptr(), total, PIECE_LEN, None, expected.data_ptr(), None, stream=sp)
    kernel_name = gpuhash.last_kernel()
    torch.cuda.synchronize()
    import oracle
    exp_host = expected.cpu().numpy().reshape(n_pieces, 20)
    for i in sorted({0, 1, n_pieces // 2, n_pieces - 1}):
        want = oracle.digest(oracle.SHA1, oracle.gen(seed, i * PIECE_LEN, PIECE_LEN))
        assert exp_host[i].tobytes() == want, "device digest of piece %d differs from the oracle" % i
    # one injected fault so that the verification result is not all-ones
    victim = n_pieces // 3
    d[victim * PIECE_LEN + 12345] ^= 1

    def check_bitmap(words_u32):
        good = int(np.unpackbits(words_u32.view(np.uint8)).sum())
        assert good == n_pieces - 1, "bitmap popcount %d != %d" % (good, n_pieces - 1)
        assert not (int(words_u32[victim >> 5]) >> (victim & 31)) & 1

    def step_kernel():
        ctx.verify_pieces_device(d.data_ptr(), total, PIECE_LEN, expected.data_ptr(), digests.data_ptr(), ok.data_ptr(),
                                 stream=sp)

    for _ in range(args.warmup):
        step_kernel()
    barrier()
    launches0 = ctx.stats().kernel_launches
    sampler = ClockSampler(dev).start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record(stream)
    for a, b in evs:
        a.record(stream)
        step_kernel()
        b.record(stream)
    t_all1.record(stream)
    barrier()
    clocks = sampler.stop()
    kernel_launches = ctx.stats().kernel_launches - launches0
    total_ms = max_over_ranks(t_all0.elapsed_time(t_all1))
    launch_ms = sum(a.elapsed_time(b) for a, b in evs) / len(evs)
    check_bitmap(ok.cpu().numpy().view(np.uint32)[: (n_pieces + 31) // 32])
    value = world * total * args.steps / (total_ms / 1e3) / 1e9

    # ---- roofline of the dominant (only) kernel, from THIS run's launch time and clock
    hbm_peak, hbm_src = measured_peak()
    hbm_achieved = total / (launch_ms / 1e3) / 1e9
    prof = committed_profile()
    props = torch.cuda.get_device_properties(dev)
    sms = props.multi_processor_count
    mix = None
    try:
        mix = alu_instructions_per_block(kernel_name)
    except Exception as e:  # cuobjdump missing: keep the HBM figures, say why the ALU ones are absent
        print("bench.py: SASS instruction count unavailable: %r" % (e,), file=sys.stderr)
    blocks_per_launch = n_pieces * ((PIECE_LEN + 9 + 63) // 64)
    clock_mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965
    same_kernel = prof.get("kernel", "?").split("<")[0] == kernel_name.split("<")[0]  # DRAM traffic does not depend on the template arguments
    roofline = {"kernel": kernel_name, "algorithmic_bytes_per_launch": total, "launch_ms": round(launch_ms, 3),
                "hbm": {"achieved": round(hbm_achieved, 2), "peak": hbm_peak, "unit": "GB/s", "frac": round(hbm_achieved / hbm_peak, 4),
                        "peak_source": hbm_src},
                "traffic": prof.get("dram_bytes_per_launch") if same_kernel else None,
                "traffic_source": prof.get("source") if same_kernel else None}
    if mix:
        achieved = blocks_per_launch / (launch_ms / 1e3) * mix["alu"] / 1e12       # T thread-instructions / s
        peak = sms * 64 * clock_mhz * 1e6 / 1e12
        roofline.update({"bound": "int32_alu", "achieved": round(achieved, 3), "peak": round(peak, 3), "unit": "T thread-instr/s",
                         "frac": round(achieved / peak, 4), "alu_instr_per_block": mix["alu"], "fma_instr_per_block": mix["fma"],
                         "instr_per_block": mix["all"], "per_block_loops": mix["loops"], "library_sha256": mix["library_sha256"],
                         "peak_source": "%d SMs x 64 INT32 lanes/clk (profiles/r01_microbench_int32_h2d.json) x %d MHz sampled in this run"
                                        % (sms, clock_mhz),
                         "blocks_per_launch": blocks_per_launch})
    else:
        roofline.update({"bound": "hbm", "achieved": roofline["hbm"]["achieved"], "peak": hbm_peak, "unit": "GB/s",
                         "frac": roofline["hbm"]["frac"], "note": "ALU instruction count unavailable; kernel is INT32-ALU bound"})

    # ---- end to end: host (pinned) buffers through gh_verify_pieces
    avail = mem_available()
    budget = max(PIECE_LEN * 64, (avail - (12 << 30)) // max(world, 1)) if avail else total
    win = min(total, budget // PIECE_LEN * PIECE_LEN)
    reps = (total + win - 1) // win
    win = ((total // PIECE_LEN + reps - 1) // reps) * PIECE_LEN  # equal windows, whole pieces
    host = gpuhash.PinnedBuffer(win)
    ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), win)
    del d
    torch.cuda.empty_cache()
    spans, left = [], total
    while left > 0:
        take = min(win, left)
        spans.append(host.array[:take])
        left -= take
    if reps == 1:
        exp_e2e = exp_host
        host_desc = "whole payload in one pinned buffer"
    else:
        # the host window repeats: expected digests repeat with it (piece-aligned windows)
        per = win // PIECE_LEN
        exp_e2e = np.concatenate([exp_host[:per]] * reps)[:n_pieces]
        host_desc = "%.1f GiB pinned window presented %d times as spans (host RAM %.0f GiB available)" % (
            win / 2**30, reps, avail / 2**30)
    exp_e2e = np.ascontiguousarray(exp_e2e)
    host_has_fault = victim * PIECE_LEN + 12345 < win
    span_list = gpuhash.SpanList(spans)

    def step_e2e():
        return ctx.VerifyPieces(span_list, PIECE_LEN, exp_e2e)

    s0 = ctx.stats()
    bitmap, dig, nbad = step_e2e()
    s1 = ctx.stats()
    e2e_kernel = gpuhash.last_kernel()
    h2d_per_step = s1.h2d_bytes - s0.h2d_bytes + n_pieces * 20  # payload + expected digests
    d2h_per_step = s1.d2h_bytes - s0.d2h_bytes
    n_fault = sum(1 for r in range(reps) if host_has_fault and r * (win // PIECE_LEN) + victim < n_pieces)
    assert nbad == n_fault, "e2e n_bad %d != %d" % (nbad, n_fault)
    for _ in range(args.warmup - 1):
        step_e2e()
    barrier()
    e2e_launch0 = ctx.stats().kernel_launches
    sampler2 = ClockSampler(dev).start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        bitmap, dig, nbad = step_e2e()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    clocks_e2e = sampler2.stop()
    e2e_launches = ctx.stats().kernel_launches - e2e_launch0
    e2e_kernel_ms = ctx.stats().kernel_ms
    e2e_value = world * total * args.steps / e2e_s / 1e9
    host.free()
    ctx.close()
    del expected, digests, ok
    torch.cuda.empty_cache()

    # ---- the link roofline of this run: plain pinned cudaMemcpyAsync, all ranks at once, >= 2 s
    link_gbs = link_probe(torch, dev, 120, barrier, max_over_ranks, world)

    # ---- CPU baseline (rank 0, N=1 only): the oracle port on every host thread, bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            os.sched_setaffinity(0, range(os.cpu_count()))
        except Exception:
            pass
        threads = host_threads()
        probe = np.empty(64 * PIECE_LEN, dtype=np.uint8)
        parallel_gen(oracle, SEED, 0, probe.ctypes.data, probe.size, threads)
        t0 = time.perf_counter()
        oracle.hash_msgs(oracle.SHA1, probe, PIECE_LEN, threads)
        rate = probe.size / (time.perf_counter() - t0) / 1e9
        sample = max(64 * PIECE_LEN, min(16 << 30, int(rate * 1e9 * 10)) // PIECE_LEN * PIECE_LEN)
        data = np.empty(sample, dtype=np.uint8)
        parallel_gen(oracle, SEED, 0, data.ctypes.data, sample, threads)
        t0 = time.perf_counter()
        oracle.hash_msgs(oracle.SHA1, data, PIECE_LEN, threads)
        dt = time.perf_counter() - t0
        cpu = {"value": round(sample / dt / 1e9, 3), "unit": "GB/s", "cores": threads, "kind": "port",
               "sample": "first %.2f GiB of the seed-%d stream, %.1f s" % (sample / 2**30, SEED, dt)}
        try:  # OpenSSL (SHA-NI) on the same sample: an upper bound for any CPU SHA-1
            cpu["openssl_sha_ni_gbs"] = round(openssl_rate(data, PIECE_LEN, threads), 3)
        except Exception:
            pass
        del data

    # ---- N>1: config 4 through one context over every rank's device (rank 0; the others have freed theirs)
    c4 = None
    host_barrier()
    if world > 1 and not args.no_config4:
        if rank == 0:
            try:
                os.sched_setaffinity(0, range(os.cpu_count()))
            except Exception:
                pass
            try:
                c4 = config4_strong(gpuhash, oracle, np, devices, host_threads())
            except (MemoryError, gpuhash.GpuHashError, AssertionError) as e:  # reported, never fatal for the weak-scaling line
                c4 = {"failed": repr(e)[:300]}
        host_barrier()

    if rank == 0:
        config = shared_config(total, world)
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": config,
            "notes": {"timing": "value: CUDA events on the launching stream; e2e: host clock around the synchronous C-ABI calls",
                      "e2e_host_buffer": host_desc, "e2e_kernel": e2e_kernel, "placement": placement},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": {"value": round(e2e_value, 2), "unit": "GB/s", "h2d_bytes_per_step": int(h2d_per_step),
                    "d2h_bytes_per_step": int(d2h_per_step), "ms_per_step": round(e2e_s / args.steps * 1e3, 2),
                    "kernel_ms_last_step": round(e2e_kernel_ms, 2), "gpu_launches": int(e2e_launches),
                    "link_roofline_gbs": round(link_gbs, 2), "frac_of_link": round(e2e_value / link_gbs, 4),
                    "link_how": "pinned cudaMemcpyAsync, 120 x 1 GiB per rank as 256 MiB pieces on 4 streams, all %d rank(s) at once, same devices" % world,
                    "clocks": clocks_e2e},
            "gpu_launches": int(kernel_launches),
            "clocks": clocks,
        }
        if c4 is not None:
            line["config4_strong"] = c4
        emit(out_fd, line)
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
