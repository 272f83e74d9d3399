#!/usr/bin/env python
"""bench.py -- verified GB/s of batched SHA-1 piece verification (BASELINE.json configs[1]).

Workload (N=1): synthetic 50 GiB torrent payload (SURVEY.md Appendix C stream, seed 2), 4 MiB pieces,
12 800 pieces, every piece hashed with SHA-1 and compared with its expected digest on the device.
A step is one pass over the whole payload.

  value      kernel-only: payload resident in HBM, one launch per step, CUDA events on the launching
             stream (torch's current stream is handed to the C ABI).
  e2e        the same pass through gh_verify_pieces with HOST (pinned) buffers: H2D of the payload
             and D2H of digests + bitmap inside the timed region.
  roofline   HBM: algorithmic bytes (1 byte read per payload byte) / kernel time vs the measured copy
             bandwidth in MEASURED_PEAKS.json. The kernel is integer-pipe bound, not HBM bound; the
             int_pipe block says how far from that bound it is.
  cpu_baseline  the oracle's scalar C port (oracle/gh_oracle.c) on all host threads, bounded sample.

--impl reference times the CPU path alone (the reference's Go crypto/sha1 cannot run: no Go toolchain
in this image; the oracle port stands in, kind "port").

N>1 (torchrun, one rank per GPU): weak scaling, every rank verifies its own 50 GiB payload; pieces are
independent so there is no collective on the data path (torch.distributed only carries the barrier
and the max-over-ranks of the timings).
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "downloader-go_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

PIECE_LEN = 4 << 20
TOTAL = 50 << 30
SEED = 2
METRIC = "verified GB/s (SHA-1 piece verification, 50 GiB / 4 MiB pieces)"


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
    """Per-launch DRAM traffic and integer-pipe figures from the committed ncu summary, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_inputs.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def parallel_gen(oracle, seed, offset, out, threads):
    """Fill `out` (numpy uint8) with stream(seed)[offset:offset+len] using `threads` host threads."""
    n = out.size
    step = max(1 << 20, ((n + threads - 1) // threads + 63) & ~63)
    def work(lo):
        hi = min(n, lo + step)
        oracle.lib().gho_gen_fill(seed, offset + lo, hi - lo, ctypes.c_void_p(out.ctypes.data + lo))
    ts = [threading.Thread(target=work, args=(lo,)) for lo in range(0, n, step)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()


def cpu_rate(oracle, data, piece_len, threads, seconds_target=None):
    """GB/s of the oracle port hashing `data` once with `threads` threads."""
    t0 = time.perf_counter()
    oracle.hash_msgs(oracle.SHA1, data, piece_len, threads)
    dt = time.perf_counter() - t0
    return data.size / dt / 1e9, dt


def run_reference(args):
    """The CPU arm: oracle port (no Go toolchain -> the reference's crypto/sha1 cannot run), all host
    threads, same metric/config, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import numpy as np
    import oracle
    threads = host_threads()
    probe = np.empty(64 * PIECE_LEN, dtype=np.uint8)
    parallel_gen(oracle, SEED, 0, probe, threads)
    rate, _ = cpu_rate(oracle, probe, PIECE_LEN, threads)
    want = int(rate * 1e9 * float(os.environ.get("GH_BENCH_REF_STEP_SECONDS", "2.0")))  # CPU work per step
    avail = mem_available()
    cap = max(1 << 30, min(16 << 30, avail // 4 if avail else 4 << 30))
    sample = max(16 * PIECE_LEN, min(cap, want) // PIECE_LEN * PIECE_LEN)
    data = np.empty(sample, dtype=np.uint8)
    parallel_gen(oracle, SEED, 0, data, threads)
    expected = oracle.hash_msgs(oracle.SHA1, data, PIECE_LEN, threads)
    for _ in range(args.warmup):
        oracle.verify_pieces(data, PIECE_LEN, expected, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        bitmap, bad = oracle.verify_pieces(data, PIECE_LEN, expected, threads)
        assert bad == 0
    dt = time.perf_counter() - t0
    gbs = sample * args.steps / dt / 1e9
    sample_desc = "first %.2f GiB (%d pieces) of the seed-%d stream per step" % (sample / 2**30, sample // PIECE_LEN, SEED)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(gbs, 3), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "configs[1]: 50 GiB torrent payload, 4 MiB pieces, SHA-1 verify", "piece_len": PIECE_LEN,
                   "sample": sample_desc,
                   "note": "Go crypto/sha1 not runnable (no Go toolchain); scalar C port of the same algorithm, verified"},
        "cpu_baseline": {"value": round(gbs, 3), "unit": "GB/s", "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": round(gbs, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--total-gib", type=float, default=TOTAL / 2**30, help="debug only: shrink the payload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch

    import gpuhash

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        print("bench.py: no CUDA device; the product path has no CPU fallback", file=sys.stderr)
        return 2
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL announces its version on stdout; stdout is reserved for the one JSON line
        os.environ["NCCL_DEBUG"] = "WARN"
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    total = int(args.total_gib * 2**30) // PIECE_LEN * PIECE_LEN
    n_pieces = total // PIECE_LEN
    seed = SEED + 1000 * rank  # every rank its own payload (weak scaling)

    # NUMA: allocate pinned memory from the CPUs next to this GPU when NVML can tell us which
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass

    ctx = gpuhash.Context(devices=[local_rank])
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream

    # ---- payload in HBM, expected digests from a first (untimed) pass, spot-checked vs the oracle
    d = torch.empty(total, dtype=torch.uint8, device="cuda")
    ctx.generate(d.data_ptr(), seed, 0, total, stream=sp)
    expected = torch.zeros(n_pieces * 20, dtype=torch.uint8, device="cuda")
    digests = torch.zeros(n_pieces * 20, dtype=torch.uint8, device="cuda")
    ok = torch.zeros((n_pieces + 31) // 32 + 1, dtype=torch.int32, device="cuda")
    ctx.verify_pieces_device(d.data_ptr(), total, PIECE_LEN, None, expected.data_ptr(), None, stream=sp)
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
    sampler = ClockSampler(local_rank).start()
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
    peak, peak_src = measured_peak()
    achieved = total / (launch_ms / 1e3) / 1e9
    prof = committed_profile()
    roofline = {"bound": "hbm", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": prof.get("dram_bytes_per_launch"),
                "peak_source": peak_src, "kernel": "gh_hash_lanes<SHA1, staged>",
                "algorithmic_bytes_per_launch": total,
                "note": "integer-pipe bound kernel (SHA-1 ~10 ALU ops/byte); see int_pipe"}
    if prof.get("int_pipe"):
        roofline["int_pipe"] = prof["int_pipe"]

    # ---- end to end: host (pinned) buffers through gh_verify_pieces
    avail = mem_available()
    budget = max(PIECE_LEN * 64, (avail - (12 << 30)) // max(world, 1)) if avail else total
    win = min(total, budget // PIECE_LEN * PIECE_LEN)
    reps = (total + win - 1) // win
    win = ((total // PIECE_LEN + reps - 1) // reps) * PIECE_LEN  # equal windows, whole pieces
    host = gpuhash.PinnedBuffer(win)
    ctx.L.gh_memcpy_d2h(ctx.h, 0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(d.data_ptr()), win)
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

    def step_e2e():
        return ctx.VerifyPieces(spans, PIECE_LEN, exp_e2e)

    s0 = ctx.stats()
    bitmap, dig, nbad = step_e2e()
    s1 = ctx.stats()
    h2d_per_step = s1.h2d_bytes - s0.h2d_bytes + n_pieces * 20  # payload + expected digests
    d2h_per_step = s1.d2h_bytes - s0.d2h_bytes
    n_fault = sum(1 for r in range(reps) if host_has_fault and r * (win // PIECE_LEN) + victim < n_pieces)
    assert nbad == n_fault, "e2e n_bad %d != %d" % (nbad, n_fault)
    for _ in range(args.warmup - 1):
        step_e2e()
    barrier()
    e2e_launch0 = ctx.stats().kernel_launches
    sampler2 = ClockSampler(local_rank).start()
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

    # ---- CPU baseline (rank 0, N=1 only): the oracle port on every host thread, bounded sample
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        try:
            os.sched_setaffinity(0, range(os.cpu_count()))
            threads = host_threads()
        except Exception:
            pass
        probe = np.empty(64 * PIECE_LEN, dtype=np.uint8)
        parallel_gen(oracle, SEED, 0, probe, threads)
        rate, _ = cpu_rate(oracle, probe, PIECE_LEN, threads)
        sample = max(64 * PIECE_LEN, min(8 << 30, int(rate * 1e9 * 10)) // PIECE_LEN * PIECE_LEN)
        data = np.empty(sample, dtype=np.uint8)
        parallel_gen(oracle, SEED, 0, data, threads)
        rate, dt = cpu_rate(oracle, data, PIECE_LEN, threads)
        cpu = {"value": round(rate, 3), "unit": "GB/s", "cores": threads, "kind": "port",
               "sample": "first %.2f GiB of the seed-%d stream, %.1f s" % (sample / 2**30, SEED, dt)}
        try:  # OpenSSL (SHA-NI) on the same sample, informational: an upper bound for any CPU SHA-1
            import hashlib
            from concurrent.futures import ThreadPoolExecutor
            mv = memoryview(data)
            t0 = time.perf_counter()
            with ThreadPoolExecutor(threads) as ex:
                list(ex.map(lambda i: hashlib.sha1(mv[i:i + PIECE_LEN]).digest(), range(0, sample, PIECE_LEN)))
            cpu["openssl_sha_ni_gbs"] = round(sample / (time.perf_counter() - t0) / 1e9, 3)
        except Exception:
            pass

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "configs[1]: synthetic 50 GiB torrent payload, 4 MiB pieces, batched SHA-1 verify",
                       "payload_bytes_per_gpu": total, "piece_len": PIECE_LEN, "pieces_per_gpu": n_pieces,
                       "seed": SEED, "l2": "inputs (%.0f GiB) exceed the 126 MB L2; no flush needed" % (total / 2**30),
                       "timing": "value: CUDA events on the launching stream; e2e: host clock around the synchronous C-ABI calls",
                       "e2e_host_buffer": host_desc, "parallelism": "independent pieces, %d rank(s), no collective" % world},
            "roofline": roofline,
            "cpu_baseline": cpu,
            "e2e": {"value": round(e2e_value, 2), "unit": "GB/s", "h2d_bytes_per_step": int(h2d_per_step),
                    "d2h_bytes_per_step": int(d2h_per_step), "ms_per_step": round(e2e_s / args.steps * 1e3, 2),
                    "kernel_ms_last_step": round(e2e_kernel_ms, 2), "gpu_launches": int(e2e_launches),
                    "clocks": clocks_e2e},
            "gpu_launches": int(kernel_launches),
            "clocks": clocks,
        }
        print(json.dumps(line))
    ctx.close()
    if dist:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
