// gh_api.cu -- the C ABI of libgpuhash.so (include/gpuhash.h): contexts, the pinned staging ring,
// host-span gather, multi-GPU sharding and the device-resident entry points.
//
// Host-side design: per device a ring of `slots` device buffers, each
// with its own COPY stream, and one COMPUTE stream for every hashing kernel. A window's copies wait
// for the kernel that last read the slot (event `consumed`), its kernel waits for the copies (event
// `h2d_done`); so the H2D of windows b+1.. runs on the copy engine while the kernel of window b runs
// on the SMs, and kernels are naturally ordered (they may share midstates). Spans that already live in pinned memory are DMA'd straight from
// the caller's buffer; pageable spans are first copied into the slot's pinned mirror by a small host
// thread pool. Pieces / parts are independent, so devices get contiguous 64-aligned message ranges and
// there is no collective of any kind.
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gpuhash.h"
#include "gh_internal.h"

void gh_stage_copy(uint8_t* dst, const uint8_t* src, size_t n);  // gh_copy.cpp: non-temporal stores into the pinned mirror

namespace {

// The library makes the device it needs current on every entry (goroutines migrate between OS
// threads), but it gives the calling thread its own current device back on the way out: a caller
// that shares the CUDA runtime with the library (PyTorch in the tests and bench.py, any C++ host)
// must not find its allocations landing on another GPU after a gh_* call.
struct DeviceRestore {
    int prev = -1;
    DeviceRestore() {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            prev = -1;
            cudaGetLastError();
        }
    }
    ~DeviceRestore() {
        if (prev >= 0 && cudaSetDevice(prev) != cudaSuccess) cudaGetLastError();
    }
};

constexpr uint64_t kDefaultSlotBytes = 256ull << 20;
constexpr int kDefaultSlots = 4;
constexpr int kDefaultStageThreads = 4;
constexpr uint64_t kMinChunk = 64ull << 10;  // smallest per-message chunk of the chunk-interleaved mode

// ------------------------------------------------------------------------------ host thread pool
class Pool {
   public:
    explicit Pool(int n) {
        for (int i = 0; i < n; i++) threads_.emplace_back([this] { run(); });
    }
    ~Pool() {
        {
            std::lock_guard<std::mutex> g(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    // memcpy split over the pool; returns when every byte is copied.
    void copy(uint8_t* dst, const uint8_t* src, uint64_t n) {
        const uint64_t grain = 2ull << 20;
        if (threads_.empty() || n <= grain) {
            gh_stage_copy(dst, src, n);
            return;
        }
        std::atomic<int> pending{0};
        std::mutex dmu;
        std::condition_variable dcv;
        const uint64_t parts = std::min<uint64_t>((n + grain - 1) / grain, threads_.size() * 4);
        const uint64_t each = ((n + parts - 1) / parts + 63) & ~63ull;
        pending = static_cast<int>((n + each - 1) / each);
        for (uint64_t off = 0; off < n; off += each) {
            const uint64_t len = std::min(each, n - off);
            std::lock_guard<std::mutex> g(mu_);
            q_.push_back([=, &pending, &dmu, &dcv] {
                gh_stage_copy(dst + off, src + off, len);
                if (pending.fetch_sub(1) == 1) {
                    std::lock_guard<std::mutex> g2(dmu);
                    dcv.notify_one();
                }
            });
            cv_.notify_one();
        }
        std::unique_lock<std::mutex> lk(dmu);
        dcv.wait(lk, [&] { return pending.load() == 0; });
    }

    // fn(i) for i in [0, n), spread over the pool; returns when all are done.
    void parallel_for(uint64_t n, const std::function<void(uint64_t)>& fn) {
        if (n == 0) return;
        if (threads_.empty() || n == 1) {
            for (uint64_t i = 0; i < n; i++) fn(i);
            return;
        }
        std::atomic<uint64_t> next{0};
        std::atomic<int> pending{0};
        std::mutex dmu;
        std::condition_variable dcv;
        const int workers = static_cast<int>(std::min<uint64_t>(n, threads_.size()));
        pending = workers;
        for (int w = 0; w < workers; w++) {
            std::lock_guard<std::mutex> g(mu_);
            q_.push_back([&] {
                for (;;) {
                    const uint64_t i = next.fetch_add(1);
                    if (i >= n) break;
                    fn(i);
                }
                if (pending.fetch_sub(1) == 1) {
                    std::lock_guard<std::mutex> g2(dmu);
                    dcv.notify_one();
                }
            });
            cv_.notify_one();
        }
        std::unique_lock<std::mutex> lk(dmu);
        dcv.wait(lk, [&] { return pending.load() == 0; });
    }
    // pread of [off, off+n) into dst, in 4 MiB pieces over the pool. Bytes the file does not have
    // (fd < 0: file missing; or a short file) read as zeros, so the pieces that touch them simply fail
    // verification. A read ERROR is something else: the first errno is returned (0 = fine).
    static int pread_full(int fd, uint64_t off, uint8_t* p, uint64_t len) {
        while (len) {
            const ssize_t r = fd >= 0 ? pread(fd, p, len, static_cast<off_t>(off)) : 0;
            if (r < 0) {
                if (errno == EINTR) continue;
                return errno;
            }
            if (r == 0) {
                memset(p, 0, len);
                break;
            }
            p += r; off += r; len -= r;
        }
        return 0;
    }
    int read(int fd, uint64_t off, uint8_t* dst, uint64_t n) {
        const uint64_t grain = 4ull << 20;
        std::atomic<int> err{0};
        parallel_for((n + grain - 1) / grain, [&](uint64_t i) {
            const uint64_t o = i * grain, len = std::min(grain, n - o);
            const int e = pread_full(fd, off + o, dst + o, len);
            if (e) err.store(e);
        });
        return err.load();
    }
    int read2d(int fd, uint64_t off, uint64_t spitch, uint8_t* dst, uint64_t dpitch, uint64_t width, uint64_t rows) {
        std::atomic<int> err{0};
        parallel_for(rows, [&](uint64_t r) {
            const int e = pread_full(fd, off + r * spitch, dst + r * dpitch, width);
            if (e) err.store(e);
        });
        return err.load();
    }

    // strided copy: rows of `width` bytes, pitches in bytes; rows are spread over the pool.
    void copy2d(uint8_t* dst, uint64_t dpitch, const uint8_t* src, uint64_t spitch, uint64_t width, uint64_t rows) {
        if (rows == 0 || width == 0) return;
        const uint64_t nt = threads_.empty() ? 1 : threads_.size();
        const uint64_t per = std::max<uint64_t>(1, (rows + nt * 4 - 1) / (nt * 4));
        if (threads_.empty() || rows * width <= (2ull << 20)) {
            for (uint64_t r = 0; r < rows; r++) gh_stage_copy(dst + r * dpitch, src + r * spitch, width);
            return;
        }
        std::atomic<int> pending{static_cast<int>((rows + per - 1) / per)};
        std::mutex dmu;
        std::condition_variable dcv;
        for (uint64_t r0 = 0; r0 < rows; r0 += per) {
            const uint64_t r1 = std::min(rows, r0 + per);
            std::lock_guard<std::mutex> g(mu_);
            q_.push_back([=, &pending, &dmu, &dcv] {
                for (uint64_t r = r0; r < r1; r++) gh_stage_copy(dst + r * dpitch, src + r * spitch, width);
                if (pending.fetch_sub(1) == 1) {
                    std::lock_guard<std::mutex> g2(dmu);
                    dcv.notify_one();
                }
            });
            cv_.notify_one();
        }
        std::unique_lock<std::mutex> lk(dmu);
        dcv.wait(lk, [&] { return pending.load() == 0; });
    }

   private:
    void run() {
        for (;;) {
            std::function<void()> f;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (stop_ && q_.empty()) return;
                f = std::move(q_.front());
                q_.erase(q_.begin());
            }
            f();
        }
    }
    std::vector<std::thread> threads_;
    std::vector<std::function<void()>> q_;
    std::mutex mu_;
    std::condition_variable cv_;
    bool stop_ = false;
};

struct Slot {
    uint8_t* d_buf = nullptr;
    uint8_t* h_pin = nullptr;        // pinned mirror, allocated on first pageable span
    cudaStream_t stream = nullptr;   // copy stream of this slot
    cudaEvent_t h2d_done = nullptr;  // the window's bytes are in d_buf (and h_pin may be refilled)
    cudaEvent_t consumed = nullptr;  // the kernel reading d_buf has finished
    bool h2d_pending = false;
    bool in_use = false;             // a kernel was queued on d_buf since the call started
};

struct Device {
    int ordinal = -1;
    int sm_count = 0;
    cudaStream_t stream = nullptr;  // stream of the device-resident entry points and of result copies
    std::vector<Slot> slots;
    // grow-only device scratch reused across calls
    uint8_t* d_out = nullptr;   uint64_t d_out_cap = 0;   // digests (+ expected)
    uint32_t* d_ok = nullptr;   uint64_t d_ok_cap = 0;
    uint32_t* d_state = nullptr; uint64_t d_state_cap = 0;
    uint64_t* d_lens = nullptr;  uint64_t d_lens_cap = 0;  // per-message lengths (table batches)
    uint8_t* h_out = nullptr;   uint64_t h_out_cap = 0;   // pinned mirror of results
    std::vector<cudaEvent_t> ev_pool;                     // start/stop pairs for kernel timing
    size_t ev_used = 0;
};

}  // namespace

struct gh_ctx {
    std::mutex mu;
    std::vector<Device> devs;
    uint64_t slot_bytes = kDefaultSlotBytes;
    int n_slots = kDefaultSlots;
    Pool* pool = nullptr;
    std::string last_error;
    gh_stats stats{};
    std::atomic<uint64_t> launches{0}, h2d{0}, d2h{0};
    std::atomic<bool> lost{false};
    bool stream_open = false;  // a gh_stream owns the ring until it is finished or aborted
    std::mutex err_mu;         // last_error is written by one host thread per device
    std::atomic<int64_t> fail_copy_in{-1};  // gh_debug_fail_nth_copy: >= 0 = payload copies left before a forced failure
};

namespace {

int fail(gh_ctx* c, int code, const std::string& msg) {
    if (c) {
        std::lock_guard<std::mutex> g(c->err_mu);
        c->last_error = msg;
    }
    return code;
}

int cuda_fail(gh_ctx* c, cudaError_t e, const char* what) {
    std::string msg = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    cudaGetLastError();  // the error is reported through the status; do not leave it for the caller's next CUDA call
    bool lost = false;
    if (c) {
        if (e == cudaErrorIllegalAddress || e == cudaErrorLaunchFailure || e == cudaErrorHardwareStackError ||
            e == cudaErrorIllegalInstruction || e == cudaErrorMisalignedAddress)
            c->lost = true;
        lost = c->lost;
        std::lock_guard<std::mutex> g(c->err_mu);
        c->last_error = msg;
    }
    if (e == cudaErrorMemoryAllocation) return GH_ENOMEM;
    return lost ? GH_ECTX_LOST : GH_ECUDA;
}

// Payload copies go through here so that tests can make the n-th one fail (gh_debug_fail_nth_copy).
cudaError_t payload_copy_gate(gh_ctx* c) {
    if (c->fail_copy_in.load(std::memory_order_relaxed) < 0) return cudaSuccess;
    return c->fail_copy_in.fetch_sub(1) == 0 ? cudaErrorUnknown : cudaSuccess;
}

#define GH_CUDA(ctx, call)                                           \
    do {                                                             \
        cudaError_t _e = (call);                                     \
        if (_e != cudaSuccess) return cuda_fail((ctx), _e, #call);   \
    } while (0)

uint64_t count_msgs(uint64_t total, uint64_t msg_len) {
    if (total == 0 || msg_len == 0) return 0;
    return (total + msg_len - 1) / msg_len;
}

int digest_len(int alg) { return alg == GH_SHA1 ? 20 : alg == GH_MD5 ? 16 : 32; }

template <class T>
int grow(gh_ctx* c, T** p, uint64_t* cap, uint64_t need_bytes, bool pinned_host = false) {
    if (need_bytes <= *cap && *p) return GH_OK;
    if (*p) {
        if (pinned_host) cudaFreeHost(*p);
        else cudaFree(*p);
        *p = nullptr;
        *cap = 0;
    }
    uint64_t want = std::max<uint64_t>(need_bytes, 4096);
    void* q = nullptr;
    cudaError_t e = pinned_host ? cudaMallocHost(&q, want) : cudaMalloc(&q, want);
    if (e != cudaSuccess) return cuda_fail(c, e, pinned_host ? "cudaMallocHost" : "cudaMalloc");
    *p = static_cast<T*>(q);
    *cap = want;
    return GH_OK;
}

// Splits `algos` into the kernel variants that exist: SHA-1 alone, and MD5/SHA-256 alone or fused.
void split_algos(int algos, int out[2], int* n) {
    *n = 0;
    if (algos & GH_SHA1) out[(*n)++] = GH_SHA1;
    if (algos & (GH_MD5 | GH_SHA256)) out[(*n)++] = algos & (GH_MD5 | GH_SHA256);
}

struct EvPair { cudaEvent_t a, b; };

int next_events(gh_ctx* c, Device& d, EvPair* out) {
    if (d.ev_used + 2 > d.ev_pool.size()) {
        for (int i = 0; i < 2; i++) {
            cudaEvent_t e;
            GH_CUDA(c, cudaEventCreate(&e));
            d.ev_pool.push_back(e);
        }
    }
    out->a = d.ev_pool[d.ev_used++];
    out->b = d.ev_pool[d.ev_used++];
    return GH_OK;
}

// Launch all kernel variants for one window on `stream`, bracketed by a timing event pair.
int launch_window(gh_ctx* c, Device& d, int algos, GhJob job, cudaStream_t stream, bool timed) {
    int parts[2], np;
    split_algos(algos, parts, &np);
    EvPair ev{};
    if (timed) {
        int rc = next_events(c, d, &ev);
        if (rc) return rc;
        GH_CUDA(c, cudaEventRecord(ev.a, stream));
    }
    for (int i = 0; i < np; i++) {
        GhJob j = job;
        if (parts[i] != GH_SHA1) { j.expected = nullptr; j.ok_words = nullptr; }
        GH_CUDA(c, gh_launch_hash(parts[i], j, d.sm_count, stream));
        c->launches++;
    }
    if (timed) GH_CUDA(c, cudaEventRecord(ev.b, stream));
    return GH_OK;
}

// --------------------------------------------------------------------------- span gather
// The logical stream as a list of runs: each run is host memory (ptr) or a byte range of a file (fd).
struct SpanIndex {
    size_t n = 0;
    std::vector<const uint8_t*> ptr;  // null for file-backed runs
    std::vector<const char*> path;    // file-backed runs: the caller's path (valid for the duration of the call)
    std::vector<char> is_file;
    std::vector<uint64_t> start;      // stream offset of each run, plus the total at [n]
    std::vector<char> pinned;
    uint64_t total() const { return start[n]; }
};

// Files are opened when a window first needs them and at most kMaxOpen stay open per device thread:
// a torrent of 10 000 files (BASELINE config 5) must not run into RLIMIT_NOFILE. A file that does not
// exist is a hole (fd -1: its bytes read as zeros and the pieces over it fail verification); any
// other failure to open -- EMFILE, EACCES, EIO ... -- is an error of the call, reported with errno.
class FdCache {
   public:
    static constexpr size_t kMaxOpen = 64;
    ~FdCache() {
        for (auto& e : open_)
            if (e.fd >= 0) close(e.fd);
    }
    // 0 = ok (*fd may be -1 for a missing file), else errno
    int get(const SpanIndex& ix, size_t i, int* fd) {
        for (auto& e : open_)
            if (e.idx == i) {
                e.stamp = ++clock_;
                *fd = e.fd;
                return 0;
            }
        int f = -1;
        if (ix.path[i]) {
            do f = open(ix.path[i], O_RDONLY | O_CLOEXEC);
            while (f < 0 && errno == EINTR);
            if (f < 0 && errno != ENOENT && errno != ENOTDIR) return errno;
        }
        if (open_.size() >= kMaxOpen) {
            size_t victim = 0;
            for (size_t k = 1; k < open_.size(); k++)
                if (open_[k].stamp < open_[victim].stamp) victim = k;
            if (open_[victim].fd >= 0) close(open_[victim].fd);
            open_[victim] = open_.back();
            open_.pop_back();
        }
        open_.push_back({i, f, ++clock_});
        *fd = f;
        return 0;
    }

   private:
    struct Entry { size_t idx; int fd; uint64_t stamp; };
    std::vector<Entry> open_;
    uint64_t clock_ = 0;
};

// Byte ranges of a slot's pinned mirror that were filled by the host and still need their H2D copy.
// Kept as separate runs: a pinned span DMA'd straight into the slot may sit between two staged ones,
// and one copy over the hull would overwrite it with whatever the mirror held there.
struct StagedRuns {
    struct Run { uint64_t off, width, rows, pitch; };  // rows of `width` bytes, `pitch` apart (rows == 1: plain range)
    std::vector<Run> runs;
    void add(uint64_t lo, uint64_t hi) {
        if (lo >= hi) return;
        if (!runs.empty() && runs.back().rows == 1 && runs.back().off + runs.back().width == lo) runs.back().width = hi - runs.back().off;
        else runs.push_back({lo, hi - lo, 1, 0});
    }
    void add_rows(uint64_t off, uint64_t pitch, uint64_t width, uint64_t rows) {
        if (!width || !rows) return;
        if (width == pitch || rows == 1) add(off, off + (rows - 1) * pitch + width);
        else runs.push_back({off, width, rows, pitch});
    }
};

int io_fail(gh_ctx* c, int err, const char* path) {
    return fail(c, GH_EIO, std::string("reading ") + (path ? path : "?") + ": " + strerror(err));
}

// [lo, hi) of the pinned allocation that contains p, or false. cudaPointerGetAttributes costs a few
// microseconds; with thousands of spans cut from a few pinned buffers (10 000 small files) asking once
// per allocation instead of once per span is worth ~40 ms per call.
bool pinned_range(const void* p, uintptr_t* lo, uintptr_t* hi) {
    cudaPointerAttributes a{};
    const bool pinned = cudaPointerGetAttributes(&a, p) == cudaSuccess && a.type == cudaMemoryTypeHost;
    cudaGetLastError();  // older drivers flag unregistered memory as an error; clear it
    if (!pinned) return false;
    *lo = reinterpret_cast<uintptr_t>(p);
    *hi = *lo + 1;
    typedef int (*GetAttrFn)(void*, int, unsigned long long);
    static GetAttrFn get_attr = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<GetAttrFn>(f);
    }();
    if (get_attr) {
        unsigned long long start = 0;
        size_t size = 0;
        const unsigned long long addr = reinterpret_cast<uintptr_t>(p);
        // CU_POINTER_ATTRIBUTE_RANGE_START_ADDR = 11, CU_POINTER_ATTRIBUTE_RANGE_SIZE = 12
        if (get_attr(&start, 11, addr) == 0 && get_attr(&size, 12, addr) == 0 && start && size && start <= addr &&
            addr < start + size) {
            *lo = static_cast<uintptr_t>(start);
            *hi = *lo + size;
        }
    }
    return true;
}

void build_index(const gh_span* spans, size_t n, SpanIndex* ix, bool merge = true) {
    ix->ptr.clear();
    ix->start.clear();
    ix->pinned.clear();
    uint64_t off = 0;
    uintptr_t known_lo = 0, known_hi = 0;  // the pinned allocation the previous span was found in
    for (size_t i = 0; i < n; i++) {
        if (!spans[i].len) continue;
        const uint8_t* p = static_cast<const uint8_t*>(spans[i].host_ptr);
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        bool pinned = false;
        if (a >= known_lo && a + spans[i].len <= known_hi) {
            pinned = true;
        } else if (pinned_range(p, &known_lo, &known_hi)) {
            // trust the range only if the whole span lies in it; a span reaching past a registered
            // region is staged like pageable memory
            pinned = a + spans[i].len <= known_hi || known_hi == known_lo + 1;
            if (known_hi == known_lo + 1) known_lo = known_hi = 0;
        } else {
            known_lo = known_hi = 0;
        }
        // Files read back to back into one arena are one run of bytes: merge spans that continue the
        // previous one in memory, so 10 000 small files cost a handful of DMAs per window, not 10 000.
        if (merge && !ix->ptr.empty() && ix->pinned.back() == static_cast<char>(pinned) &&
            ix->ptr.back() + (off - ix->start.back()) == p) {
            off += spans[i].len;
            continue;
        }
        ix->ptr.push_back(p);
        ix->start.push_back(off);
        ix->pinned.push_back(pinned);
        off += spans[i].len;
    }
    ix->n = ix->ptr.size();
    ix->start.push_back(off);
    ix->path.assign(ix->n, nullptr);
    ix->is_file.assign(ix->n, 0);
}

void build_index_files(const gh_file* files, size_t n, SpanIndex* ix) {
    ix->n = n;
    ix->ptr.assign(n, nullptr);
    ix->path.assign(n, nullptr);
    ix->is_file.assign(n, 1);
    ix->start.resize(n + 1);
    ix->pinned.assign(n, 0);
    uint64_t off = 0;
    for (size_t i = 0; i < n; i++) {
        ix->start[i] = off;
        off += files[i].len;
        if (files[i].len) ix->path[i] = files[i].path;
    }
    ix->start[n] = off;
}

// Copies stream bytes [a, b) into slot memory at d_dst (device) / h_dst (pinned mirror, same offset).
// Pinned spans are DMA'd directly; pageable and file-backed ones are staged in the mirror, and
// `staged` collects the mirror ranges that need their H2D afterwards (flush_staged).
int gather(gh_ctx* c, const SpanIndex& ix, FdCache& fds, uint64_t a, uint64_t b, Slot& s, uint64_t dst_off,
           StagedRuns* staged) {
    if (a >= b) return GH_OK;
    size_t i = std::upper_bound(ix.start.begin(), ix.start.begin() + ix.n, a) - ix.start.begin() - 1;
    uint64_t x = a;
    while (x < b && i < ix.n) {
        const uint64_t s_lo = ix.start[i], s_hi = ix.start[i + 1];
        if (s_hi <= x) { i++; continue; }
        const uint64_t hi = std::min(b, s_hi);
        const uint64_t len = hi - x;
        const uint64_t doff = dst_off + (x - a);
        GH_CUDA(c, payload_copy_gate(c));
        if (ix.pinned[i]) {
            GH_CUDA(c, cudaMemcpyAsync(s.d_buf + doff, ix.ptr[i] + (x - s_lo), len, cudaMemcpyHostToDevice, s.stream));
        } else {
            if (!s.h_pin) GH_CUDA(c, cudaMallocHost(reinterpret_cast<void**>(&s.h_pin), c->slot_bytes + 256));
            if (s.h2d_pending) {  // the previous window's DMA out of this mirror must have drained
                GH_CUDA(c, cudaEventSynchronize(s.h2d_done));
                s.h2d_pending = false;
            }
            if (ix.is_file[i]) {
                int fd = -1, err = fds.get(ix, i, &fd);
                if (!err) err = c->pool->read(fd, x - s_lo, s.h_pin + doff, len);
                if (err) return io_fail(c, err, ix.path[i]);
            } else {
                c->pool->copy(s.h_pin + doff, ix.ptr[i] + (x - s_lo), len);
            }
            staged->add(doff, doff + len);
        }
        c->h2d += len;
        x = hi;
        i++;
    }
    return GH_OK;
}

int flush_staged(gh_ctx* c, Slot& s, const StagedRuns& staged) {
    for (const auto& r : staged.runs) {
        if (r.rows == 1)
            GH_CUDA(c, cudaMemcpyAsync(s.d_buf + r.off, s.h_pin + r.off, r.width, cudaMemcpyHostToDevice, s.stream));
        else
            GH_CUDA(c, cudaMemcpy2DAsync(s.d_buf + r.off, r.pitch, s.h_pin + r.off, r.pitch, r.width, r.rows,
                                         cudaMemcpyHostToDevice, s.stream));
        s.h2d_pending = true;
    }
    return GH_OK;
}

// Before a window's copies: the slot's previous kernel must have drained it.
int slot_begin(gh_ctx* c, Slot& s) {
    if (s.in_use) GH_CUDA(c, cudaStreamWaitEvent(s.stream, s.consumed, 0));
    return GH_OK;
}

// After a window's copies: queue its kernel(s) on the compute stream behind them.
int slot_launch(gh_ctx* c, Device& d, Slot& s, int algos, const GhJob& j);

// Where each message of a batch lives in the logical stream. Uniform: message m = stream[m*L, ...),
// only the last one short (pieces of a torrent, parts of one file). Table: explicit offset/length per
// message (parts of MANY files in one batch: every file starts a new part, so every file's last part
// may be short); on the device every message still owns a row of L bytes.
struct MsgTable {
    uint64_t L = 0, total = 0, n = 0;
    std::vector<uint64_t> off_, len_;  // empty in the uniform case
    bool uniform() const { return off_.empty(); }
    uint64_t off(uint64_t m) const { return uniform() ? m * L : off_[m]; }
    uint64_t len(uint64_t m) const { return uniform() ? std::min(L, total - m * L) : len_[m]; }
};

// Chunk-interleaved window: for every message m of the group [g0, g0+g) bring stream bytes
// [(g0+m)*L + prior, +min(chunk, len_m - prior)) to slot offset m*chunk. Consecutive rows that lie
// inside one span move as one strided copy (DMA for pinned spans, pooled memcpy for pageable ones);
// a row that straddles a span (file) boundary falls back to the byte-exact 1-D gather.
int gather_rows(gh_ctx* c, const SpanIndex& ix, FdCache& fds, const MsgTable& T, Slot& s, uint64_t g0, uint64_t g,
                uint64_t prior, uint64_t chunk) {
    StagedRuns staged;
    size_t si = 0;
    uint64_t m = 0;
    const uint64_t L = T.L;
    while (m < g) {
        const uint64_t mlen = T.len(g0 + m);
        if (prior >= mlen) { m++; continue; }
        const uint64_t w = std::min(chunk, mlen - prior);
        const uint64_t lo = T.off(g0 + m) + prior;
        while (si + 1 < ix.n && ix.start[si + 1] <= lo) si++;
        while (si > 0 && ix.start[si] > lo) si--;
        const bool inside = si < ix.n && ix.start[si] <= lo && lo + w <= ix.start[si + 1];
        if (!inside) {
            int rc = gather(c, ix, fds, lo, lo + w, s, m * chunk, &staged);
            if (rc) return rc;
            m++;
            continue;
        }
        // extend over following rows of the same width, one message length apart, inside span si
        uint64_t rows = 1;
        while (m + rows < g) {
            const uint64_t m2 = g0 + m + rows;
            const uint64_t mlen2 = T.len(m2);
            if (prior >= mlen2 || std::min(chunk, mlen2 - prior) != w) break;
            if (T.off(m2) != T.off(m2 - 1) + L) break;
            if (T.off(m2) + prior + w > ix.start[si + 1]) break;
            rows++;
        }
        GH_CUDA(c, payload_copy_gate(c));
        if (ix.pinned[si]) {
            const uint8_t* src = ix.ptr[si] + (lo - ix.start[si]);
            static const bool no2d = getenv("GPUHASH_NO_2D") != nullptr;
            if (no2d) {
                for (uint64_t r = 0; r < rows; r++)
                    GH_CUDA(c, cudaMemcpyAsync(s.d_buf + (m + r) * chunk, src + r * L, w, cudaMemcpyHostToDevice, s.stream));
            } else {
                GH_CUDA(c, cudaMemcpy2DAsync(s.d_buf + m * chunk, chunk, src, L, w, rows, cudaMemcpyHostToDevice, s.stream));
            }
        } else {
            if (!s.h_pin) GH_CUDA(c, cudaMallocHost(reinterpret_cast<void**>(&s.h_pin), c->slot_bytes + 256));
            if (s.h2d_pending) {
                GH_CUDA(c, cudaEventSynchronize(s.h2d_done));
                s.h2d_pending = false;
            }
            if (ix.is_file[si]) {
                int fd = -1, err = fds.get(ix, si, &fd);
                if (!err) err = c->pool->read2d(fd, lo - ix.start[si], L, s.h_pin + m * chunk, chunk, w, rows);
                if (err) return io_fail(c, err, ix.path[si]);
            } else {
                c->pool->copy2d(s.h_pin + m * chunk, chunk, ix.ptr[si] + (lo - ix.start[si]), L, w, rows);
            }
            staged.add_rows(m * chunk, chunk, w, rows);
        }
        c->h2d += w * rows;
        m += rows;
    }
    return flush_staged(c, s, staged);
}

int slot_launch(gh_ctx* c, Device& d, Slot& s, int algos, const GhJob& j) {
    GH_CUDA(c, cudaEventRecord(s.h2d_done, s.stream));
    GH_CUDA(c, cudaStreamWaitEvent(d.stream, s.h2d_done, 0));
    int rc = launch_window(c, d, algos, j, d.stream, true);
    if (rc) return rc;
    GH_CUDA(c, cudaEventRecord(s.consumed, d.stream));
    s.in_use = true;
    return GH_OK;
}

// --------------------------------------------------------------------------- one device's share
struct BatchArgs {
    const SpanIndex* ix;
    const MsgTable* T;
    uint64_t msg_len;
    int algos;
    const uint8_t* expected;  // host, 20*n_total or null
    uint8_t* out[3];          // host: sha1, md5, sha256 (whole-call arrays) or null
    uint64_t* bitmap;         // host, whole-call, or null
    uint64_t n_total;
};

// Leaves the device idle and the ring in its initial state: every queued copy and kernel of this
// context has finished, no slot is marked busy. Called on every error exit, so that (a) a one-shot
// call never returns while a DMA may still read the caller's buffers -- the promise the header makes
// for all exits, not only the successful one -- and (b) the next call cannot overwrite a slot that a
// stale kernel is still reading.
void drain_device(Device& d) {
    if (cudaSetDevice(d.ordinal) != cudaSuccess) {
        cudaGetLastError();
        return;
    }
    for (auto& s : d.slots) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        s.in_use = false;
        s.h2d_pending = false;
    }
    if (d.stream) cudaStreamSynchronize(d.stream);
    cudaGetLastError();
}

int run_device_body(gh_ctx* c, Device& d, const BatchArgs& A, uint64_t m0, uint64_t m1, double* kernel_ms);

int run_device(gh_ctx* c, Device& d, const BatchArgs& A, uint64_t m0, uint64_t m1, double* kernel_ms) {
    *kernel_ms = 0;
    if (m0 >= m1) return GH_OK;
    const int rc = run_device_body(c, d, A, m0, m1, kernel_ms);
    if (rc != GH_OK) drain_device(d);
    return rc;
}

int run_device_body(gh_ctx* c, Device& d, const BatchArgs& A, uint64_t m0, uint64_t m1, double* kernel_ms) {
    GH_CUDA(c, cudaSetDevice(d.ordinal));
    FdCache fds;
    const MsgTable& T = *A.T;
    const uint64_t n = m1 - m0, L = A.msg_len;
    const bool table = !T.uniform();

    // result layout on the device: [sha1 20n][md5 16n][sha256 32n][expected 20n], each 256-aligned
    auto al = [](uint64_t v) { return (v + 255) & ~255ull; };
    const uint64_t o_sha1 = 0, o_md5 = al(20 * n), o_sha256 = o_md5 + al(16 * n), o_exp = o_sha256 + al(32 * n);
    const uint64_t out_bytes = o_exp + al(20 * n);
    int rc;
    if ((rc = grow(c, &d.d_out, &d.d_out_cap, out_bytes))) return rc;
    if ((rc = grow(c, &d.h_out, &d.h_out_cap, out_bytes, true))) return rc;
    const uint64_t ok_words = (n + 31) / 32;
    if ((rc = grow(c, &d.d_ok, &d.d_ok_cap, (ok_words + 1) * 4))) return rc;
    const bool want_ok = A.expected != nullptr && (A.algos & GH_SHA1);
    if (want_ok) {
        memcpy(d.h_out + o_exp, A.expected + 20 * m0, 20 * n);
        GH_CUDA(c, cudaMemcpyAsync(d.d_out + o_exp, d.h_out + o_exp, 20 * n, cudaMemcpyHostToDevice, d.stream));
        GH_CUDA(c, cudaMemsetAsync(d.d_ok, 0, (ok_words + 1) * 4, d.stream));
    }
    if (table) {  // per-message lengths for the kernels (staged through the pinned result mirror)
        if ((rc = grow(c, &d.d_lens, &d.d_lens_cap, n * 8))) return rc;
        GH_CUDA(c, cudaMemcpyAsync(d.d_lens, T.len_.data() + m0, n * 8, cudaMemcpyHostToDevice, d.stream));
    }
    for (auto& sl : d.slots) sl.in_use = false;
    const size_t first_timing_event = d.ev_used;

    GhJob job{};
    job.msg_len = L;
    job.sha1 = (A.algos & GH_SHA1) ? d.d_out + o_sha1 : nullptr;
    job.md5 = (A.algos & GH_MD5) ? d.d_out + o_md5 : nullptr;
    job.sha256 = (A.algos & GH_SHA256) ? d.d_out + o_sha256 : nullptr;

    const uint64_t S = c->slot_bytes;
    const uint64_t bytes_here = table ? n * L : T.off(m1 - 1) + T.len(m1 - 1) - T.off(m0);
    if (bytes_here <= S || L <= kMinChunk) {
        // whole messages per window (everything fits one window, or messages are small enough that a
        // window still carries thousands of independent chains)
        const uint64_t per = std::max<uint64_t>(1, S / L);
        uint64_t b = 0;
        for (uint64_t q = m0; q < m1; q += per, b++) {
            Slot& s = d.slots[b % d.slots.size()];
            const uint64_t k = std::min(per, m1 - q);
            StagedRuns staged;
            if ((rc = slot_begin(c, s))) return rc;
            // runs of messages that follow each other in the stream at full length land back to back
            // in the slot as well: one gather per run (the whole window in the uniform case)
            for (uint64_t r0 = q; r0 < q + k;) {
                uint64_t r1 = r0 + 1;
                while (r1 < q + k && T.len(r1 - 1) == L && T.off(r1) == T.off(r1 - 1) + L) r1++;
                if ((rc = gather(c, *A.ix, fds, T.off(r0), T.off(r1 - 1) + T.len(r1 - 1), s, (r0 - q) * L, &staged)))
                    return rc;
                r0 = r1;
            }
            if ((rc = flush_staged(c, s, staged))) return rc;
            GhJob j = job;
            j.data = s.d_buf;
            j.stride = L;
            j.chunk_len = L;
            j.prior = 0;
            j.n_msgs = k;
            j.last_msg_len = T.len(q + k - 1);
            const uint64_t rel = q - m0;
            if (table) j.lens = d.d_lens + rel;
            if (j.sha1) j.sha1 += 20 * rel;
            if (j.md5) j.md5 += 16 * rel;
            if (j.sha256) j.sha256 += 32 * rel;
            if (want_ok) {
                j.expected = d.d_out + o_exp + 20 * rel;
                j.ok_words = d.d_ok;
                j.ok_base = rel;
            }
            if ((rc = slot_launch(c, d, s, A.algos, j))) return rc;
        }
    } else {
        // chunk-interleaved: every window carries the next `chunk` bytes of every message of a group,
        // so that all chains advance together; midstates stay in HBM. Rows that
        // lie inside one pinned span travel as a single strided (2-D) DMA.
        if ((rc = grow(c, &d.d_state, &d.d_state_cap, n * GH_STATE_WORDS * 4))) return rc;
        // groups of equal size (a ragged last group of a few hundred chains would run chain-starved):
        // as many groups as the per-message minimum chunk requires, messages spread evenly, whole warps
        const uint64_t gmax = std::max<uint64_t>(1, S / kMinChunk);
        const uint64_t n_groups = (n + gmax - 1) / gmax;
        const uint64_t gsz = std::min<uint64_t>(gmax, (((n + n_groups - 1) / n_groups) + 31) & ~31ull);
        uint64_t b = 0;
        for (uint64_t g0 = m0; g0 < m1; g0 += gsz) {
            const uint64_t g = std::min(gsz, m1 - g0);
            const uint64_t chunk = std::min<uint64_t>((S / g) & ~63ull, (L + 63) & ~63ull);
            const uint64_t rel = g0 - m0;
            const uint64_t g_last_len = T.len(g0 + g - 1);
            for (uint64_t prior = 0; prior < L; prior += chunk, b++) {
                Slot& s = d.slots[b % d.slots.size()];
                if ((rc = slot_begin(c, s))) return rc;
                if ((rc = gather_rows(c, *A.ix, fds, T, s, g0, g, prior, chunk))) return rc;
                GhJob j = job;
                j.data = s.d_buf;
                j.stride = chunk;
                j.chunk_len = chunk;
                j.prior = prior;
                j.n_msgs = g;
                j.last_msg_len = g_last_len;
                j.state = d.d_state;  // group-local [word][g]; kernels run in order on one stream, so groups may share it
                if (table) j.lens = d.d_lens + rel;
                if (j.sha1) j.sha1 += 20 * rel;
                if (j.md5) j.md5 += 16 * rel;
                if (j.sha256) j.sha256 += 32 * rel;
                if (want_ok) {
                    j.expected = d.d_out + o_exp + 20 * rel;
                    j.ok_words = d.d_ok;
                    j.ok_base = rel;
                }
                if ((rc = slot_launch(c, d, s, A.algos, j))) return rc;
            }
        }
    }
    const size_t last_timing_event = d.ev_used;

    // results back: every kernel ran on d.stream, so the D2H copies simply follow them there
    struct { int alg; uint64_t off; int idx; } outs[3] = {{GH_SHA1, o_sha1, 0}, {GH_MD5, o_md5, 1}, {GH_SHA256, o_sha256, 2}};
    for (auto& o : outs) {
        if (!(A.algos & o.alg) || !A.out[o.idx]) continue;
        const uint64_t bytes = static_cast<uint64_t>(digest_len(o.alg)) * n;
        GH_CUDA(c, cudaMemcpyAsync(d.h_out + o.off, d.d_out + o.off, bytes, cudaMemcpyDeviceToHost, d.stream));
        c->d2h += bytes;
    }
    uint32_t* h_ok = reinterpret_cast<uint32_t*>(d.h_out + o_exp);  // expected is no longer needed on the host
    if (want_ok && A.bitmap) {
        GH_CUDA(c, cudaMemcpyAsync(h_ok, d.d_ok, ok_words * 4, cudaMemcpyDeviceToHost, d.stream));
        c->d2h += ok_words * 4;
    }
    GH_CUDA(c, cudaStreamSynchronize(d.stream));
    for (auto& s : d.slots) s.h2d_pending = false;
    for (auto& o : outs) {
        if (!(A.algos & o.alg) || !A.out[o.idx]) continue;
        const uint64_t dl = digest_len(o.alg);
        memcpy(A.out[o.idx] + dl * m0, d.h_out + o.off, dl * n);
    }
    if (want_ok && A.bitmap) {  // m0 is 64-aligned: this device owns whole 64-bit words
        uint32_t* dst = reinterpret_cast<uint32_t*>(A.bitmap) + m0 / 32;
        memcpy(dst, h_ok, ok_words * 4);
        if (ok_words & 1) dst[ok_words] = 0;
    }
    for (size_t e = first_timing_event; e + 1 < last_timing_event; e += 2) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, d.ev_pool[e], d.ev_pool[e + 1]) == cudaSuccess) *kernel_ms += ms;
    }
    return GH_OK;
}

int run_index(gh_ctx* c, SpanIndex& ix, uint64_t msg_len, int algos, const uint8_t* expected, uint8_t* sha1_out,
              uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad, bool per_run = false);
int run_table(gh_ctx* c, SpanIndex& ix, const MsgTable& T, int algos, const uint8_t* expected, uint8_t* sha1_out,
              uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad);

int run_batch(gh_ctx* c, const gh_span* spans, size_t n_spans, uint64_t msg_len, int algos, const uint8_t* expected,
              uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad,
              bool per_file = false) {
    DeviceRestore restore_device;
    if (!c) return GH_EINVAL;
    std::lock_guard<std::mutex> g(c->mu);
    if (n_spans && !spans) return fail(c, GH_EINVAL, "spans is NULL");
    for (size_t i = 0; i < n_spans; i++)
        if (spans[i].len && !spans[i].host_ptr) return fail(c, GH_EINVAL, "span with NULL pointer and non-zero length");
    SpanIndex ix;
    build_index(spans, n_spans, &ix, !per_file);
    return run_index(c, ix, msg_len, algos, expected, sha1_out, md5_out, sha256_out, bitmap, n_bad, per_file);
}

int run_files(gh_ctx* c, const gh_file* files, size_t n_files, uint64_t msg_len, int algos, const uint8_t* expected,
              uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad,
              bool per_file = false) {
    DeviceRestore restore_device;
    if (!c) return GH_EINVAL;
    std::lock_guard<std::mutex> g(c->mu);
    if (n_files && !files) return fail(c, GH_EINVAL, "files is NULL");
    for (size_t i = 0; i < n_files; i++)
        if (files[i].len && !files[i].path) return fail(c, GH_EINVAL, "file with NULL path and non-zero length");
    SpanIndex ix;
    build_index_files(files, n_files, &ix);
    return run_index(c, ix, msg_len, algos, expected, sha1_out, md5_out, sha256_out, bitmap, n_bad, per_file);
}

// caller holds c->mu. per_run: messages restart at every run (file) of the index instead of running
// through the concatenated stream.
int run_index(gh_ctx* c, SpanIndex& ix, uint64_t msg_len, int algos, const uint8_t* expected, uint8_t* sha1_out,
              uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad, bool per_run) {
    if (c->lost) return fail(c, GH_ECTX_LOST, "context lost after a device fault");
    if (c->stream_open) return fail(c, GH_EINVAL, "a gh_stream is open on this context; finish or abort it first");
    if (msg_len == 0) return fail(c, GH_EINVAL, "message length must be > 0");
    if (!(algos & (GH_SHA1 | GH_MD5 | GH_SHA256)) || (algos & ~(GH_SHA1 | GH_MD5 | GH_SHA256)))
        return fail(c, GH_EINVAL, "algos must be a mask of GH_SHA1|GH_MD5|GH_SHA256");
    MsgTable T;
    T.L = msg_len;
    T.total = ix.total();
    if (per_run) {
        for (size_t i = 0; i < ix.n; i++) {
            const uint64_t lo = ix.start[i], hi = ix.start[i + 1];
            for (uint64_t o = lo; o < hi; o += msg_len) {
                T.off_.push_back(o);
                T.len_.push_back(std::min(msg_len, hi - o));
            }
        }
        T.n = T.off_.size();
    } else {
        T.n = count_msgs(T.total, msg_len);
    }
    return run_table(c, ix, T, algos, expected, sha1_out, md5_out, sha256_out, bitmap, n_bad);
}

// caller holds c->mu and has checked lost / stream_open / algos. The batch proper: shard the table's
// messages over the context's devices, run, merge.
int run_table(gh_ctx* c, SpanIndex& ix, const MsgTable& T, int algos, const uint8_t* expected, uint8_t* sha1_out,
              uint8_t* md5_out, uint8_t* sha256_out, uint64_t* bitmap, uint64_t* n_bad) {
    const auto t0 = std::chrono::steady_clock::now();
    const uint64_t msg_len = T.L;
    const uint64_t n = T.n;
    if (n_bad) *n_bad = 0;
    if (n == 0) {
        c->stats.kernel_ms = 0;
        c->stats.wall_ms = 0;
        return GH_OK;
    }
    // n_bad is counted from the bitmap: keep a private one when the caller wants only the count
    std::vector<uint64_t> own_bitmap;
    if (!bitmap && n_bad && expected && (algos & GH_SHA1)) {
        own_bitmap.assign((n + 63) / 64 + 1, 0);
        bitmap = own_bitmap.data();
    }
    BatchArgs A{&ix, &T, msg_len, algos, expected, {sha1_out, md5_out, sha256_out}, bitmap, n};
    const int nd = static_cast<int>(c->devs.size());
    std::vector<int> rcs(nd, GH_OK);
    std::vector<double> kms(nd, 0.0);
    std::vector<std::string> errs(nd);
    for (auto& d : c->devs) d.ev_used = 0;
    if (nd == 1) {
        rcs[0] = run_device(c, c->devs[0], A, 0, n, &kms[0]);
    } else {
        std::vector<std::thread> th;
        for (int i = 0; i < nd; i++) {
            uint64_t b, e;
            gh_shard_range(n, nd, i, &b, &e);
            th.emplace_back([&, i, b, e] { rcs[i] = run_device(c, c->devs[i], A, b, e, &kms[i]); });
        }
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < nd; i++)
        if (rcs[i]) return rcs[i];
    if (bitmap && expected && (algos & GH_SHA1)) {
        uint64_t good = 0;
        const uint64_t words = (n + 63) / 64;
        for (uint64_t w = 0; w < words; w++) good += __builtin_popcountll(bitmap[w]);
        if (n_bad) *n_bad = n - good;
    } else if (bitmap) {
        memset(bitmap, 0, ((n + 63) / 64) * 8);
    }
    c->stats.kernel_ms = *std::max_element(kms.begin(), kms.end());
    c->stats.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return GH_OK;
}

Device* get_dev(gh_ctx* c, int idx) {
    if (!c || idx < 0 || idx >= static_cast<int>(c->devs.size())) return nullptr;
    return &c->devs[idx];
}

}  // namespace

// ================================================================================ C ABI
extern "C" {

const char* gh_version(void) { return "gpuhash 0.1 (sm_100a)"; }

const char* gh_strerror(int status) {
    switch (status) {
        case GH_OK: return "ok";
        case GH_EINVAL: return "invalid argument";
        case GH_ENODEVICE: return "no usable CUDA device";
        case GH_ENOMEM: return "out of memory";
        case GH_ECUDA: return "CUDA error";
        case GH_ECTX_LOST: return "context lost after a device fault";
        case GH_ERANGE: return "size beyond limit";
        case GH_EIO: return "file could not be opened or read";
        default: return "unknown status";
    }
}

// context-free entry points (gh_host_alloc ...) leave their failure text here, per thread
static thread_local std::string g_tls_error;
static int tls_cuda_fail(cudaError_t e, const char* what) {
    g_tls_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? GH_ENOMEM : GH_ECUDA;
}

const char* gh_last_error(gh_ctx* ctx) {
    if (!ctx) return g_tls_error.c_str();
    static thread_local std::string copy;  // valid until this thread's next gh_last_error call
    std::lock_guard<std::mutex> g(ctx->err_mu);
    copy = ctx->last_error;
    return copy.c_str();
}

int gh_debug_fail_nth_copy(gh_ctx* ctx, int64_t n) {
    if (!ctx) return GH_EINVAL;
    ctx->fail_copy_in.store(n);
    return GH_OK;
}

void gh_shard_range(uint64_t n_msgs, int n_shards, int shard, uint64_t* begin, uint64_t* end) {
    if (n_shards < 1) n_shards = 1;
    const uint64_t words = (n_msgs + 63) / 64;  // shard in units of 64 messages = one bitmap word
    const uint64_t base = words / n_shards, extra = words % n_shards;
    const uint64_t s = static_cast<uint64_t>(shard);
    const uint64_t w0 = s * base + std::min<uint64_t>(s, extra);
    const uint64_t w1 = w0 + base + (s < extra ? 1 : 0);
    *begin = std::min(n_msgs, w0 * 64);
    *end = std::min(n_msgs, w1 * 64);
}

int gh_create(gh_ctx** out, const gh_config* cfg) {
    DeviceRestore restore_device;
    if (!out) return GH_EINVAL;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return GH_ENODEVICE;  // no CPU fallback on purpose
    gh_ctx* c = new gh_ctx();
    int nd = cfg && cfg->n_devices > 0 ? cfg->n_devices : count;
    if (cfg && cfg->slot_bytes) c->slot_bytes = std::max<uint64_t>((cfg->slot_bytes + 255) & ~255ull, 1ull << 20);
    if (cfg && cfg->slots) c->n_slots = std::max(2, cfg->slots);
    // default: one staging thread per host core, between 4 and 16 (pread/memcpy into the pinned ring
    // scales to ~44 GB/s at 16 threads on the test box, 26 GB/s at 4)
    int st = static_cast<int>(std::thread::hardware_concurrency());
    st = std::max(kDefaultStageThreads, std::min(16, st));
    if (cfg && cfg->stage_threads > 0) st = cfg->stage_threads;
    for (int i = 0; i < nd; i++) {
        const int ord = (cfg && cfg->devices && cfg->n_devices > 0) ? cfg->devices[i] : i;
        if (ord < 0 || ord >= count) {
            gh_destroy(c);
            return GH_ENODEVICE;
        }
        Device d;
        d.ordinal = ord;
        c->devs.push_back(d);
    }
    for (auto& d : c->devs) {
        bool ok = cudaSetDevice(d.ordinal) == cudaSuccess;
        ok = ok && cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, d.ordinal) == cudaSuccess;
        ok = ok && gh_kernels_init_device() == cudaSuccess;
        ok = ok && cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) == cudaSuccess;
        d.slots.resize(c->n_slots);
        for (auto& s : d.slots) {
            ok = ok && cudaMalloc(reinterpret_cast<void**>(&s.d_buf), c->slot_bytes + 256) == cudaSuccess;
            ok = ok && cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) == cudaSuccess;
            ok = ok && cudaEventCreateWithFlags(&s.h2d_done, cudaEventDisableTiming) == cudaSuccess;
            ok = ok && cudaEventCreateWithFlags(&s.consumed, cudaEventDisableTiming) == cudaSuccess;
        }
        if (!ok) {
            cudaError_t le = cudaGetLastError();
            gh_destroy(c);
            return le == cudaErrorMemoryAllocation ? GH_ENOMEM : GH_ECUDA;
        }
    }
    c->pool = new Pool(st);
    *out = c;
    return GH_OK;
}

void gh_destroy(gh_ctx* c) {
    DeviceRestore restore_device;
    if (!c) return;
    for (auto& d : c->devs) {
        if (cudaSetDevice(d.ordinal) != cudaSuccess) continue;
        cudaDeviceSynchronize();
        for (auto& s : d.slots) {
            if (s.d_buf) cudaFree(s.d_buf);
            if (s.h_pin) cudaFreeHost(s.h_pin);
            if (s.stream) cudaStreamDestroy(s.stream);
            if (s.h2d_done) cudaEventDestroy(s.h2d_done);
            if (s.consumed) cudaEventDestroy(s.consumed);
        }
        if (d.d_out) cudaFree(d.d_out);
        if (d.d_ok) cudaFree(d.d_ok);
        if (d.d_state) cudaFree(d.d_state);
        if (d.d_lens) cudaFree(d.d_lens);
        if (d.h_out) cudaFreeHost(d.h_out);
        for (auto e : d.ev_pool) cudaEventDestroy(e);
        if (d.stream) cudaStreamDestroy(d.stream);
    }
    delete c->pool;
    delete c;
}

int gh_device_count(gh_ctx* ctx) { return ctx ? static_cast<int>(ctx->devs.size()) : 0; }

int gh_get_stats(gh_ctx* ctx, gh_stats* out) {
    if (!ctx || !out) return GH_EINVAL;
    *out = ctx->stats;
    out->kernel_launches = ctx->launches.load();
    out->h2d_bytes = ctx->h2d.load();
    out->d2h_bytes = ctx->d2h.load();
    return GH_OK;
}

int gh_verify_pieces(gh_ctx* ctx, const gh_span* spans, size_t n_spans, uint64_t piece_len, const uint8_t* expected,
                     uint8_t* digests_out, uint64_t* ok_bitmap_out, uint64_t* n_bad_out) {
    return run_batch(ctx, spans, n_spans, piece_len, GH_SHA1, expected, digests_out, nullptr, nullptr, ok_bitmap_out,
                     n_bad_out);
}

int gh_hash_parts(gh_ctx* ctx, const gh_span* spans, size_t n_spans, uint64_t part_len, int algos, uint8_t* sha1_out,
                  uint8_t* md5_out, uint8_t* sha256_out) {
    return run_batch(ctx, spans, n_spans, part_len, algos, nullptr, sha1_out, md5_out, sha256_out, nullptr, nullptr);
}

int gh_verify_files(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t piece_len, const uint8_t* expected,
                    uint8_t* digests_out, uint64_t* ok_bitmap_out, uint64_t* n_bad_out) {
    return run_files(ctx, files, n_files, piece_len, GH_SHA1, expected, digests_out, nullptr, nullptr, ok_bitmap_out,
                     n_bad_out);
}

int gh_hash_file_parts(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t part_len, int algos,
                       uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out) {
    return run_files(ctx, files, n_files, part_len, algos, nullptr, sha1_out, md5_out, sha256_out, nullptr, nullptr);
}

int gh_hash_parts_per_file(gh_ctx* ctx, const gh_span* files, size_t n_files, uint64_t part_len, int algos,
                           uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out) {
    return run_batch(ctx, files, n_files, part_len, algos, nullptr, sha1_out, md5_out, sha256_out, nullptr, nullptr, true);
}

int gh_hash_file_parts_per_file(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t part_len, int algos,
                                uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out) {
    return run_files(ctx, files, n_files, part_len, algos, nullptr, sha1_out, md5_out, sha256_out, nullptr, nullptr, true);
}

int gh_s3_etag(gh_ctx* ctx, const uint8_t* part_md5s, uint64_t n_parts, char* out, size_t out_len) {
    if (!ctx || !out || out_len < 56 || (n_parts && !part_md5s)) return GH_EINVAL;
    uint8_t d[16];
    gh_span sp{part_md5s, n_parts * 16};
    int rc;
    if (n_parts == 0) {
        // md5 of the empty string: hash one zero-length message through the device path
        return fail(ctx, GH_EINVAL, "an S3 multipart upload has at least one part");
    }
    rc = gh_hash_parts(ctx, &sp, 1, n_parts * 16, GH_MD5, nullptr, d, nullptr);
    if (rc) return rc;
    for (int i = 0; i < 16; i++) snprintf(out + 2 * i, 3, "%02x", d[i]);
    snprintf(out + 32, out_len - 32, "-%llu", static_cast<unsigned long long>(n_parts));
    return GH_OK;
}

// ---- incremental feed ----------------------------------------------------------------------
struct gh_stream {
    gh_ctx* c = nullptr;
    int algos = 0;
    uint64_t L = 0, total = 0, n = 0, per = 0;
    uint64_t next_msg = 0, window = 0, fed = 0;
    bool want_ok = false, acquired = false;
    uint64_t o_sha1 = 0, o_md5 = 0, o_sha256 = 0, o_exp = 0, ok_words = 0;
    uint64_t cap = 0;  // bytes the acquired window must receive
    struct Win { int dev; uint64_t first, count; };
    std::vector<Win> wins;  // which device hashed which messages (windows go round the context's devices)
};

static void stream_release(gh_stream* st) {
    if (!st) return;
    {
        std::lock_guard<std::mutex> g(st->c->mu);
        st->c->stream_open = false;
    }
    delete st;
}

// window w of a stream lives on device w % n_devices, in that device's ring slot (w / n_devices) % slots
static Device& stream_dev(gh_stream* st, uint64_t w) { return st->c->devs[w % st->c->devs.size()]; }
static Slot& stream_slot(gh_stream* st, uint64_t w) {
    Device& d = stream_dev(st, w);
    return d.slots[(w / st->c->devs.size()) % d.slots.size()];
}
static void stream_drain_all(gh_ctx* c) {
    for (auto& d : c->devs) drain_device(d);
}

int gh_stream_open(gh_ctx* ctx, int algos, uint64_t msg_len, uint64_t total_len, const uint8_t* expected,
                   gh_stream** out) {
    DeviceRestore restore_device;
    if (!ctx || !out) return GH_EINVAL;
    *out = nullptr;
    std::lock_guard<std::mutex> g(ctx->mu);
    if (ctx->lost) return GH_ECTX_LOST;
    if (ctx->stream_open) return fail(ctx, GH_EINVAL, "a gh_stream is already open on this context");
    if (msg_len == 0) return fail(ctx, GH_EINVAL, "message length must be > 0");
    if (!(algos & 7) || (algos & ~7)) return fail(ctx, GH_EINVAL, "bad algos mask");
    if (msg_len > ctx->slot_bytes)
        return fail(ctx, GH_ERANGE, "message longer than a ring slot: use gh_hash_parts / gh_hash_file_parts, which interleave");
    gh_stream* st = new gh_stream();
    st->c = ctx;
    st->algos = algos;
    st->L = msg_len;
    st->total = total_len;
    st->n = count_msgs(total_len, msg_len);
    st->per = std::max<uint64_t>(1, ctx->slot_bytes / msg_len);
    st->want_ok = expected != nullptr && (algos & GH_SHA1);
    auto al = [](uint64_t v) { return (v + 255) & ~255ull; };
    const uint64_t n = std::max<uint64_t>(st->n, 1);
    st->o_sha1 = 0;
    st->o_md5 = al(20 * n);
    st->o_sha256 = st->o_md5 + al(16 * n);
    st->o_exp = st->o_sha256 + al(32 * n);
    st->ok_words = (n + 31) / 32;
    const uint64_t out_bytes = st->o_exp + al(20 * n);
    // every device gets result arrays for the whole stream (digests are 20-68 bytes per message) and the
    // expected digests; at finish each window's slice is read back from the device that hashed it
    for (auto& d : ctx->devs) {
        int rc = GH_OK;
        cudaError_t e = cudaSetDevice(d.ordinal);
        if (e != cudaSuccess) rc = cuda_fail(ctx, e, "cudaSetDevice");
        if (!rc) rc = grow(ctx, &d.d_out, &d.d_out_cap, out_bytes);
        if (!rc) rc = grow(ctx, &d.h_out, &d.h_out_cap, out_bytes, true);
        if (!rc) rc = grow(ctx, &d.d_ok, &d.d_ok_cap, (st->ok_words + 1) * 4);
        if (!rc && st->want_ok) {
            memcpy(d.h_out + st->o_exp, expected, 20 * st->n);
            e = cudaMemcpyAsync(d.d_out + st->o_exp, d.h_out + st->o_exp, 20 * st->n, cudaMemcpyHostToDevice, d.stream);
            if (e == cudaSuccess) e = cudaMemsetAsync(d.d_ok, 0, (st->ok_words + 1) * 4, d.stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream);  // h_out is reused for results later
            if (e != cudaSuccess) rc = cuda_fail(ctx, e, "gh_stream_open");
        }
        if (rc) {
            delete st;
            return rc;
        }
        for (auto& sl : d.slots) sl.in_use = false;
        d.ev_used = 0;
    }
    ctx->stream_open = true;
    *out = st;
    return GH_OK;
}

int gh_stream_acquire(gh_stream* st, void** buf, uint64_t* cap) {
    DeviceRestore restore_device;
    if (!st || !buf || !cap) return GH_EINVAL;
    gh_ctx* c = st->c;
    std::lock_guard<std::mutex> g(c->mu);
    if (st->acquired) return fail(c, GH_EINVAL, "gh_stream_acquire: the previous window was not submitted");
    if (st->next_msg >= st->n) {
        *buf = nullptr;
        *cap = 0;
        return GH_OK;  // everything has been fed
    }
    Device& d = stream_dev(st, st->window);
    GH_CUDA(c, cudaSetDevice(d.ordinal));
    Slot& s = stream_slot(st, st->window);
    if (!s.h_pin) GH_CUDA(c, cudaHostAlloc(reinterpret_cast<void**>(&s.h_pin), c->slot_bytes + 256, cudaHostAllocPortable));
    if (s.h2d_pending) {  // the DMA that last read this pinned buffer must have drained
        GH_CUDA(c, cudaEventSynchronize(s.h2d_done));
        s.h2d_pending = false;
    }
    const uint64_t k = std::min(st->per, st->n - st->next_msg);
    const uint64_t lo = st->next_msg * st->L, hi = std::min((st->next_msg + k) * st->L, st->total);
    st->cap = hi - lo;
    st->acquired = true;
    *buf = s.h_pin;
    *cap = st->cap;
    return GH_OK;
}

int gh_stream_submit(gh_stream* st, uint64_t n_bytes) {
    DeviceRestore restore_device;
    if (!st) return GH_EINVAL;
    gh_ctx* c = st->c;
    std::lock_guard<std::mutex> g(c->mu);
    if (!st->acquired) return fail(c, GH_EINVAL, "gh_stream_submit without gh_stream_acquire");
    if (n_bytes != st->cap) return fail(c, GH_EINVAL, "gh_stream_submit: a window must be filled completely (cap bytes)");
    Device& d = stream_dev(st, st->window);
    GH_CUDA(c, cudaSetDevice(d.ordinal));
    Slot& s = stream_slot(st, st->window);
    int rc = slot_begin(c, s);
    StagedRuns staged;
    staged.add(0, n_bytes);
    if (!rc) {
        const cudaError_t ge = payload_copy_gate(c);
        if (ge != cudaSuccess) rc = cuda_fail(c, ge, "gh_stream_submit: copy");
    }
    if (!rc) rc = flush_staged(c, s, staged);
    if (rc) {  // nothing of this stream may still be running when the caller sees the error and aborts
        stream_drain_all(c);
        return rc;
    }
    c->h2d += n_bytes;
    const uint64_t k = std::min(st->per, st->n - st->next_msg);
    GhJob j{};
    j.data = s.d_buf;
    j.stride = st->L;
    j.chunk_len = st->L;
    j.n_msgs = k;
    j.msg_len = st->L;
    j.last_msg_len = (st->next_msg + k == st->n) ? st->total - (st->n - 1) * st->L : st->L;
    const uint64_t rel = st->next_msg;
    if (st->algos & GH_SHA1) j.sha1 = d.d_out + st->o_sha1 + 20 * rel;
    if (st->algos & GH_MD5) j.md5 = d.d_out + st->o_md5 + 16 * rel;
    if (st->algos & GH_SHA256) j.sha256 = d.d_out + st->o_sha256 + 32 * rel;
    if (st->want_ok) {
        j.expected = d.d_out + st->o_exp + 20 * rel;
        j.ok_words = d.d_ok;
        j.ok_base = rel;
    }
    if ((rc = slot_launch(c, d, s, st->algos, j))) {
        stream_drain_all(c);
        return rc;
    }
    st->wins.push_back({static_cast<int>(st->window % c->devs.size()), rel, k});
    st->next_msg += k;
    st->fed += n_bytes;
    st->window++;
    st->acquired = false;
    return GH_OK;
}

int gh_stream_finish(gh_stream* st, uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out, uint64_t* ok_bitmap_out,
                     uint64_t* n_bad_out) {
    DeviceRestore restore_device;
    if (!st) return GH_EINVAL;
    gh_ctx* c = st->c;
    int rc = GH_OK;
    {
        std::lock_guard<std::mutex> g(c->mu);
        if (st->next_msg != st->n || st->acquired) {
            rc = fail(c, GH_EINVAL, "gh_stream_finish before every window was submitted (use gh_stream_abort to give up)");
        } else {
            const uint64_t n = st->n;
            struct { int alg; uint64_t off; uint8_t* dst; int dl; } outs[3] = {
                {GH_SHA1, st->o_sha1, sha1_out, 20}, {GH_MD5, st->o_md5, md5_out, 16}, {GH_SHA256, st->o_sha256, sha256_out, 32}};
            cudaError_t e = cudaSuccess;
            // queue every window's slices (and every device's bitmap words) for read-back, then wait per device
            for (const auto& w : st->wins) {
                Device& d = c->devs[w.dev];
                if (e == cudaSuccess) e = cudaSetDevice(d.ordinal);
                for (auto& o : outs)
                    if ((st->algos & o.alg) && o.dst && e == cudaSuccess) {
                        e = cudaMemcpyAsync(d.h_out + o.off + o.dl * w.first, d.d_out + o.off + o.dl * w.first, o.dl * w.count,
                                            cudaMemcpyDeviceToHost, d.stream);
                        c->d2h += o.dl * w.count;
                    }
            }
            for (auto& d : c->devs) {
                if (e == cudaSuccess) e = cudaSetDevice(d.ordinal);
                uint32_t* h_ok = reinterpret_cast<uint32_t*>(d.h_out + st->o_exp);
                if (st->want_ok && n && e == cudaSuccess)
                    e = cudaMemcpyAsync(h_ok, d.d_ok, st->ok_words * 4, cudaMemcpyDeviceToHost, d.stream);
                if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream);
                for (auto& sl : d.slots) sl.h2d_pending = false;
            }
            if (e != cudaSuccess) {
                rc = cuda_fail(c, e, "gh_stream_finish");
                stream_drain_all(c);
            } else {
                for (const auto& w : st->wins)
                    for (auto& o : outs)
                        if ((st->algos & o.alg) && o.dst)
                            memcpy(o.dst + o.dl * w.first, c->devs[w.dev].h_out + o.off + o.dl * w.first, o.dl * w.count);
                if (n_bad_out) *n_bad_out = 0;
                if (ok_bitmap_out && n) {
                    memset(ok_bitmap_out, 0, ((n + 63) / 64) * 8);
                    if (st->want_ok) {
                        // a device's bitmap holds 1s only for pieces it hashed itself: OR them together
                        uint32_t* dst = reinterpret_cast<uint32_t*>(ok_bitmap_out);
                        for (auto& d : c->devs) {
                            const uint32_t* h_ok = reinterpret_cast<const uint32_t*>(d.h_out + st->o_exp);
                            for (uint64_t k = 0; k < st->ok_words; k++) dst[k] |= h_ok[k];
                        }
                        uint64_t good = 0;
                        for (uint64_t w = 0; w < (n + 63) / 64; w++) good += __builtin_popcountll(ok_bitmap_out[w]);
                        if (n_bad_out) *n_bad_out = n - good;
                    }
                }
            }
        }
    }
    if (rc == GH_OK || rc == GH_ECUDA || rc == GH_ECTX_LOST) stream_release(st);
    return rc;
}

void gh_stream_abort(gh_stream* st) {
    DeviceRestore restore_device;
    if (!st) return;
    gh_ctx* c = st->c;
    {
        std::lock_guard<std::mutex> g(c->mu);
        stream_drain_all(c);
    }
    stream_release(st);
}

// ---- micro-batcher for live pieces ------------------------------------------------------------
// anacrolix/torrent hashes a piece the moment its last chunk arrives, about two at a time, from
// its own goroutines (reference torrent.go:79 t.DownloadAll() ... :106 client.WaitAll()). A GPU wants
// hundreds of pieces per launch. The batcher is the adapter between the two: any number of threads
// hand it ONE piece each and block; it collects them in a pinned buffer and runs the whole lot as one
// batch when `max_pieces` are waiting or the oldest has waited `max_delay_ms`; every caller wakes with
// the verdict of ITS piece. Pieces arrive in any order and with any index; only their bytes, length
// and expected digest matter. A failure of the batch (CUDA error, lost context) reaches every caller
// of that batch as a STATUS, never as "hash mismatch": the caller falls back to hashing on the CPU,
// it must not re-download good data.
struct gh_batcher {
    gh_ctx* c = nullptr;
    uint64_t piece_len = 0;
    uint32_t max_pieces = 0;
    uint32_t max_delay_ms = 0;

    struct Batch {
        uint8_t* buf = nullptr;            // pinned, max_pieces * piece_len
        std::vector<uint64_t> len;         // per slot
        std::vector<uint8_t> expected;     // 20 per slot
        std::vector<uint8_t> digest;       // 20 per slot (result)
        std::vector<uint64_t> bitmap;      // result
        uint32_t reserved = 0;             // slots handed out
        uint32_t committed = 0;            // slots whose bytes and expected digest are in place (or that were cancelled)
        uint32_t readers = 0;              // callers that still have to pick up their verdict
        uint64_t seq = 0;                  // which use of this buffer (verdicts are picked up by (buffer, seq))
        bool closed = false;               // no more reservations: queued for or in flight on the GPU
        bool done = false;                 // results are in
        int status = GH_OK;                // of the whole batch
        std::chrono::steady_clock::time_point first;
    };
    Batch b[2];
    int filling = 0;
    std::mutex mu;
    std::condition_variable cv_callers;  // space available / verdict ready
    std::condition_variable cv_flusher;  // a batch became runnable, or its deadline needs re-arming
    std::thread flusher;
    bool stop = false;
    uint32_t active_calls = 0;
    gh_batcher_stats st{};
};

namespace {

// One batch through the ring: messages k = 0..n-1 at buf + k*piece_len with their own lengths.
int batcher_run(gh_batcher* B, gh_batcher::Batch& bt) {
    gh_ctx* c = B->c;
    const uint32_t n = bt.reserved;
    std::lock_guard<std::mutex> g(c->mu);
    if (c->lost) return fail(c, GH_ECTX_LOST, "context lost after a device fault");
    if (c->stream_open) return fail(c, GH_EINVAL, "a gh_stream is open on this context; finish or abort it first");
    SpanIndex ix;
    gh_span sp{bt.buf, static_cast<uint64_t>(n) * B->piece_len};
    build_index(&sp, 1, &ix);
    MsgTable T;
    T.L = B->piece_len;
    T.total = sp.len;
    T.n = n;
    for (uint32_t k = 0; k < n; k++) {
        T.off_.push_back(static_cast<uint64_t>(k) * B->piece_len);
        T.len_.push_back(bt.len[k]);
    }
    std::fill(bt.bitmap.begin(), bt.bitmap.end(), 0);
    return run_table(c, ix, T, GH_SHA1, bt.expected.data(), bt.digest.data(), nullptr, nullptr, bt.bitmap.data(), nullptr);
}

void batcher_loop(gh_batcher* B) {
    std::unique_lock<std::mutex> lk(B->mu);
    for (;;) {
        gh_batcher::Batch& bt = B->b[B->filling];
        // runnable: something is waiting, every reserved slot has been filled, and the batch is full,
        // overdue, or was closed by gh_batcher_flush / destroy
        auto runnable = [&] {
            if (bt.done || bt.reserved == 0 || bt.committed < bt.reserved) return false;  // done: waiting for its readers
            if (bt.closed || bt.reserved >= B->max_pieces || B->stop) return true;
            return std::chrono::steady_clock::now() >= bt.first + std::chrono::milliseconds(B->max_delay_ms);
        };
        while (!runnable()) {
            if (B->stop && bt.reserved == 0) return;
            if (!bt.done && bt.reserved > 0 && bt.committed == bt.reserved && !bt.closed)
                B->cv_flusher.wait_until(lk, bt.first + std::chrono::milliseconds(B->max_delay_ms));
            else
                B->cv_flusher.wait(lk);
        }
        const bool by_deadline = !bt.closed && bt.reserved < B->max_pieces;
        bt.closed = true;
        // new pieces go to the other buffer while this one is on the GPU (once its verdicts are picked up)
        B->filling ^= 1;
        B->cv_callers.notify_all();
        lk.unlock();
        const int rc = batcher_run(B, bt);
        lk.lock();
        bt.status = rc;
        bt.done = true;
        B->st.batches++;
        B->st.pieces += bt.reserved;
        if (by_deadline) B->st.flushed_by_deadline++;
        else B->st.flushed_full++;
        if (rc != GH_OK) B->st.failed_batches++;
        if (bt.readers == 0) {  // only cancelled slots: nobody will pick anything up
            bt.reserved = bt.committed = 0;
            bt.closed = bt.done = false;
            bt.seq++;
        }
        B->cv_callers.notify_all();
    }
}

}  // namespace

int gh_batcher_create(gh_ctx* ctx, uint64_t piece_len, uint32_t max_pieces, uint32_t max_delay_ms, gh_batcher** out) {
    if (!ctx || !out) return GH_EINVAL;
    *out = nullptr;
    if (piece_len == 0 || max_pieces == 0) return fail(ctx, GH_EINVAL, "gh_batcher_create: piece_len and max_pieces must be > 0");
    const uint64_t bytes = piece_len * max_pieces;
    if (bytes / max_pieces != piece_len || bytes > (4ull << 30))
        return fail(ctx, GH_ERANGE, "gh_batcher_create: max_pieces * piece_len is limited to 4 GiB per batch buffer");
    gh_batcher* B = new gh_batcher();
    B->c = ctx;
    B->piece_len = piece_len;
    B->max_pieces = max_pieces;
    B->max_delay_ms = max_delay_ms;
    for (auto& bt : B->b) {
        if (cudaHostAlloc(reinterpret_cast<void**>(&bt.buf), bytes, cudaHostAllocPortable) != cudaSuccess) {
            cudaGetLastError();
            for (auto& x : B->b)
                if (x.buf) cudaFreeHost(x.buf);
            delete B;
            return fail(ctx, GH_ENOMEM, "gh_batcher_create: pinned batch buffers");
        }
        bt.len.assign(max_pieces, 0);
        bt.expected.assign(20ull * max_pieces, 0);
        bt.digest.assign(20ull * max_pieces, 0);
        bt.bitmap.assign((max_pieces + 63) / 64 + 1, 0);
    }
    B->flusher = std::thread(batcher_loop, B);
    *out = B;
    return GH_OK;
}

int gh_batcher_reserve(gh_batcher* B, uint64_t len, gh_batch_slot* slot, void** buf) {
    if (!B || !slot || !buf) return GH_EINVAL;
    if (len > B->piece_len) return GH_ERANGE;
    std::unique_lock<std::mutex> lk(B->mu);
    if (B->stop) return GH_EINVAL;
    B->active_calls++;
    for (;;) {
        gh_batcher::Batch& bt = B->b[B->filling];
        if (!bt.closed && bt.reserved < B->max_pieces) {
            const uint32_t k = bt.reserved++;
            if (k == 0) {
                bt.first = std::chrono::steady_clock::now();
                bt.status = GH_OK;
            }
            bt.len[k] = len;
            slot->buffer = B->filling;
            slot->index = k;
            slot->seq = bt.seq;
            *buf = bt.buf + static_cast<uint64_t>(k) * B->piece_len;
            B->active_calls--;
            return GH_OK;
        }
        if (B->stop) {
            B->active_calls--;
            B->cv_callers.notify_all();
            return GH_EINVAL;
        }
        B->cv_callers.wait(lk);  // both buffers busy: the GPU batch or its readers are still at work
    }
}

static int batcher_finish_slot(gh_batcher* B, const gh_batch_slot* slot, const uint8_t* expected20, bool cancel, int* ok_out,
                               uint8_t* digest_out) {
    if (!B || !slot || slot->buffer > 1) return GH_EINVAL;
    std::unique_lock<std::mutex> lk(B->mu);
    gh_batcher::Batch& bt = B->b[slot->buffer];
    if (slot->seq != bt.seq || slot->index >= bt.reserved || bt.done) return GH_EINVAL;
    if (cancel) {
        bt.len[slot->index] = 0;  // hashed as an empty message, verdict ignored
        memset(&bt.expected[20ull * slot->index], 0, 20);
    } else {
        if (!expected20) return GH_EINVAL;
        memcpy(&bt.expected[20ull * slot->index], expected20, 20);
        bt.readers++;
    }
    bt.committed++;
    B->cv_flusher.notify_one();
    if (cancel) return GH_OK;
    B->active_calls++;
    const uint64_t seq = bt.seq;
    B->cv_callers.wait(lk, [&] { return bt.done && bt.seq == seq; });
    const int rc = bt.status;
    if (rc == GH_OK) {
        if (ok_out) *ok_out = static_cast<int>((bt.bitmap[slot->index >> 6] >> (slot->index & 63)) & 1);
        if (digest_out) memcpy(digest_out, &bt.digest[20ull * slot->index], 20);
    }
    if (--bt.readers == 0) {  // last verdict picked up: the buffer can fill again
        bt.reserved = bt.committed = 0;
        bt.closed = bt.done = false;
        bt.seq++;
    }
    B->active_calls--;
    B->cv_callers.notify_all();
    return rc;
}

int gh_batcher_commit(gh_batcher* B, const gh_batch_slot* slot, const uint8_t* expected20, int* ok_out, uint8_t* digest_out) {
    return batcher_finish_slot(B, slot, expected20, false, ok_out, digest_out);
}

int gh_batcher_cancel(gh_batcher* B, const gh_batch_slot* slot) {
    return batcher_finish_slot(B, slot, nullptr, true, nullptr, nullptr);
}

int gh_batcher_verify(gh_batcher* B, const void* data, uint64_t len, const uint8_t* expected20, int* ok_out,
                      uint8_t* digest_out) {
    if (!B || (len && !data) || !expected20) return GH_EINVAL;
    gh_batch_slot slot;
    void* buf = nullptr;
    int rc = gh_batcher_reserve(B, len, &slot, &buf);
    if (rc) return rc;
    memcpy(buf, data, len);  // on the caller's thread: many callers copy in parallel
    return gh_batcher_commit(B, &slot, expected20, ok_out, digest_out);
}

int gh_batcher_flush(gh_batcher* B) {
    if (!B) return GH_EINVAL;
    std::lock_guard<std::mutex> g(B->mu);
    gh_batcher::Batch& bt = B->b[B->filling];
    if (bt.reserved > 0) bt.closed = true;  // runs as soon as every reserved slot is committed
    B->cv_flusher.notify_one();
    return GH_OK;
}

int gh_batcher_get_stats(gh_batcher* B, gh_batcher_stats* out) {
    if (!B || !out) return GH_EINVAL;
    std::lock_guard<std::mutex> g(B->mu);
    *out = B->st;
    return GH_OK;
}

void gh_batcher_destroy(gh_batcher* B) {
    if (!B) return;
    {
        std::unique_lock<std::mutex> lk(B->mu);
        B->stop = true;  // no new reservations; what is waiting is flushed so that no caller hangs
        B->cv_flusher.notify_all();
        B->cv_callers.notify_all();
        B->cv_callers.wait(lk, [&] {
            for (auto& bt : B->b)
                if (bt.readers || bt.reserved != bt.committed || (bt.reserved && !bt.done)) return false;
            return B->active_calls == 0;
        });
    }
    B->cv_flusher.notify_all();
    if (B->flusher.joinable()) B->flusher.join();
    for (auto& bt : B->b)
        if (bt.buf) cudaFreeHost(bt.buf);
    delete B;
}

// ---- device-resident -----------------------------------------------------------------------
static int device_call(gh_ctx* ctx, int dev_index, int algos, const GhJob& job_in, void* cuda_stream) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d) return fail(ctx, GH_EINVAL, "bad device index");
    std::lock_guard<std::mutex> g(ctx->mu);
    if (ctx->lost) return GH_ECTX_LOST;
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : d->stream;
    if (job_in.ok_words && job_in.prior == 0)
        GH_CUDA(ctx, cudaMemsetAsync(job_in.ok_words, 0, ((job_in.n_msgs + 31) / 32) * 4, st));
    d->ev_used = 0;
    const auto t0 = std::chrono::steady_clock::now();
    int rc = launch_window(ctx, *d, algos, job_in, st, cuda_stream == nullptr);
    if (rc) return rc;
    if (!cuda_stream) {
        GH_CUDA(ctx, cudaStreamSynchronize(st));
        float ms = 0;
        if (d->ev_used >= 2) cudaEventElapsedTime(&ms, d->ev_pool[0], d->ev_pool[1]);
        ctx->stats.kernel_ms = ms;
        ctx->stats.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return GH_OK;
}

int gh_verify_pieces_device(gh_ctx* ctx, int dev_index, const void* d_data, uint64_t total, uint64_t piece_len,
                            const uint8_t* d_expected, uint8_t* d_digests, uint32_t* d_ok_words, void* cuda_stream) {
    if (!ctx) return GH_EINVAL;
    if (piece_len == 0) return fail(ctx, GH_EINVAL, "piece length must be > 0");
    const uint64_t n = count_msgs(total, piece_len);
    if (n == 0) return GH_OK;
    if (!d_data) return fail(ctx, GH_EINVAL, "d_data is NULL");
    GhJob j{};
    j.data = static_cast<const uint8_t*>(d_data);
    j.stride = piece_len;
    j.chunk_len = piece_len;
    j.n_msgs = n;
    j.msg_len = piece_len;
    j.last_msg_len = total - (n - 1) * piece_len;
    j.sha1 = d_digests;
    j.expected = d_expected;
    j.ok_words = d_expected ? d_ok_words : nullptr;
    return device_call(ctx, dev_index, GH_SHA1, j, cuda_stream);
}

int gh_hash_parts_device(gh_ctx* ctx, int dev_index, const void* d_data, uint64_t total, uint64_t part_len, int algos,
                         uint8_t* d_sha1, uint8_t* d_md5, uint8_t* d_sha256, void* cuda_stream) {
    if (!ctx) return GH_EINVAL;
    if (part_len == 0) return fail(ctx, GH_EINVAL, "part length must be > 0");
    if (!(algos & 7) || (algos & ~7)) return fail(ctx, GH_EINVAL, "bad algos mask");
    const uint64_t n = count_msgs(total, part_len);
    if (n == 0) return GH_OK;
    if (!d_data) return fail(ctx, GH_EINVAL, "d_data is NULL");
    GhJob j{};
    j.data = static_cast<const uint8_t*>(d_data);
    j.stride = part_len;
    j.chunk_len = part_len;
    j.n_msgs = n;
    j.msg_len = part_len;
    j.last_msg_len = total - (n - 1) * part_len;
    j.sha1 = d_sha1;
    j.md5 = d_md5;
    j.sha256 = d_sha256;
    return device_call(ctx, dev_index, algos, j, cuda_stream);
}

int gh_hash_update_device(gh_ctx* ctx, int dev_index, int algos, const void* d_data, uint64_t stride,
                          uint64_t chunk_len, uint64_t prior_bytes, uint64_t n_msgs, uint64_t msg_len,
                          uint64_t last_msg_len, uint32_t* d_state, uint8_t* d_sha1, uint8_t* d_md5, uint8_t* d_sha256,
                          void* cuda_stream) {
    if (!ctx) return GH_EINVAL;
    if (!(algos & 7) || (algos & ~7)) return fail(ctx, GH_EINVAL, "bad algos mask");
    if (n_msgs == 0) return GH_OK;
    if (!d_data || !d_state) return fail(ctx, GH_EINVAL, "d_data / d_state is NULL");
    if (prior_bytes & 63) return fail(ctx, GH_EINVAL, "prior_bytes must be a multiple of 64");
    if ((chunk_len & 63) && prior_bytes + chunk_len < msg_len)
        return fail(ctx, GH_EINVAL, "chunk_len must be a multiple of 64 unless every message ends in this call");
    if (last_msg_len > msg_len) return fail(ctx, GH_EINVAL, "last_msg_len > msg_len");
    GhJob j{};
    j.data = static_cast<const uint8_t*>(d_data);
    j.stride = stride;
    j.chunk_len = chunk_len;
    j.prior = prior_bytes;
    j.n_msgs = n_msgs;
    j.msg_len = msg_len;
    j.last_msg_len = last_msg_len;
    j.state = d_state;
    j.sha1 = d_sha1;
    j.md5 = d_md5;
    j.sha256 = d_sha256;
    return device_call(ctx, dev_index, algos, j, cuda_stream);
}

int gh_generate_device(gh_ctx* ctx, int dev_index, void* d_out, uint64_t seed, uint64_t offset, uint64_t len,
                       void* cuda_stream) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d) return fail(ctx, GH_EINVAL, "bad device index");
    if (len && !d_out) return fail(ctx, GH_EINVAL, "d_out is NULL");
    std::lock_guard<std::mutex> g(ctx->mu);
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : d->stream;
    GH_CUDA(ctx, gh_launch_generate(static_cast<uint8_t*>(d_out), seed, offset, len, st));
    if (!cuda_stream) GH_CUDA(ctx, cudaStreamSynchronize(st));
    return GH_OK;
}

// ---- device placement ------------------------------------------------------------------------
// End to end the path is bound by the host->device link, and on multi-GPU hosts several GPUs often
// share one PCIe uplink (the 8xB200 box of DESIGN.md section 6: GPUs 0-3 together get 115 GB/s, GPUs
// 0,2,4,6 get 213). Which n of the visible devices to use is therefore measured, not guessed: plain
// pinned cudaMemcpyAsync on all devices of a candidate set at once, for the first n devices and for
// an evenly strided set.
namespace {
double probe_h2d(const std::vector<int>& devs, double seconds) {
    const size_t chunk = 128ull << 20;
    const int n = static_cast<int>(devs.size());
    std::vector<double> bytes(n, 0.0);
    std::atomic<int> ready{0}, failed{0};
    std::atomic<bool> go{false};
    std::vector<std::thread> th;
    std::chrono::steady_clock::time_point t0, t1;
    for (int i = 0; i < n; i++)
        th.emplace_back([&, i] {
            void *h = nullptr, *d = nullptr;
            cudaStream_t st = nullptr;
            bool ok = cudaSetDevice(devs[i]) == cudaSuccess && cudaHostAlloc(&h, chunk, cudaHostAllocPortable) == cudaSuccess &&
                      cudaMalloc(&d, chunk) == cudaSuccess && cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess;
            if (ok) {
                memset(h, 1, chunk);
                ok = cudaMemcpyAsync(d, h, chunk, cudaMemcpyHostToDevice, st) == cudaSuccess && cudaStreamSynchronize(st) == cudaSuccess;
            }
            if (!ok) failed++;
            ready++;
            while (!go.load()) std::this_thread::yield();
            const auto end = std::chrono::steady_clock::now() + std::chrono::duration<double>(seconds);
            while (ok && std::chrono::steady_clock::now() < end) {
                for (int k = 0; k < 4 && ok; k++) ok = cudaMemcpyAsync(d, h, chunk, cudaMemcpyHostToDevice, st) == cudaSuccess;
                ok = ok && cudaStreamSynchronize(st) == cudaSuccess;
                if (ok) bytes[i] += 4.0 * chunk;
            }
            if (st) cudaStreamDestroy(st);
            if (d) cudaFree(d);
            if (h) cudaFreeHost(h);
            cudaGetLastError();
        });
    while (ready.load() < n) std::this_thread::yield();
    t0 = std::chrono::steady_clock::now();
    go = true;
    for (auto& t : th) t.join();
    t1 = std::chrono::steady_clock::now();
    if (failed.load()) return -1.0;
    double total = 0;
    for (double b : bytes) total += b;
    return total / std::chrono::duration<double>(t1 - t0).count() / 1e9;
}
}  // namespace

int gh_suggest_devices(int n, int* devices_out, double* h2d_gbs_out) {
    if (!devices_out || n <= 0) return GH_EINVAL;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return GH_ENODEVICE;
    }
    if (n > count) return GH_ENODEVICE;
    std::vector<std::vector<int>> cands(1);
    for (int i = 0; i < n; i++) cands[0].push_back(i);
    if (count >= 2 * n) {
        cands.emplace_back();
        for (int i = 0; i < n; i++) cands[1].push_back(i * (count / n));
    }
    size_t best = 0;
    double best_gbs = 0;
    if (cands.size() > 1 || h2d_gbs_out) {
        for (size_t c = 0; c < cands.size(); c++) {
            const double g = probe_h2d(cands[c], 0.4);
            if (g < 0) return GH_ECUDA;
            if (g > best_gbs * 1.03) {  // a later candidate must be clearly better to displace the plain first-n choice
                best_gbs = g;
                best = c;
            }
        }
    }
    for (int i = 0; i < n; i++) devices_out[i] = cands[best][i];
    if (h2d_gbs_out) *h2d_gbs_out = best_gbs;
    return GH_OK;
}

// ---- memory helpers ------------------------------------------------------------------------
int gh_host_alloc(void** out, size_t bytes) {
    if (!out) return GH_EINVAL;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) return tls_cuda_fail(e, "cudaHostAlloc");
    return GH_OK;
}
int gh_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? GH_OK : GH_ECUDA; }
int gh_host_register(void* p, size_t bytes) {
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) return tls_cuda_fail(e, "cudaHostRegister");
    return GH_OK;
}
int gh_host_unregister(void* p) { return cudaHostUnregister(p) == cudaSuccess ? GH_OK : GH_ECUDA; }

int gh_device_alloc(gh_ctx* ctx, int dev_index, void** out, size_t bytes) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d || !out) return GH_EINVAL;
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    GH_CUDA(ctx, cudaMalloc(out, bytes ? bytes : 1));
    return GH_OK;
}
int gh_device_free(gh_ctx* ctx, int dev_index, void* p) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d) return GH_EINVAL;
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    GH_CUDA(ctx, cudaFree(p));
    return GH_OK;
}
int gh_memcpy_d2h(gh_ctx* ctx, int dev_index, void* dst_host, const void* src_dev, size_t bytes) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d) return GH_EINVAL;
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    GH_CUDA(ctx, cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
    return GH_OK;
}
int gh_memcpy_h2d(gh_ctx* ctx, int dev_index, void* dst_dev, const void* src_host, size_t bytes) {
    DeviceRestore restore_device;
    Device* d = get_dev(ctx, dev_index);
    if (!d) return GH_EINVAL;
    GH_CUDA(ctx, cudaSetDevice(d->ordinal));
    GH_CUDA(ctx, cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
    return GH_OK;
}

}  // extern "C"
