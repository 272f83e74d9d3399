// gh_copy.cpp -- host copy into the pinned staging ring with non-temporal stores.
//
// Pageable spans (a Go []byte is one: runtime.Pinner pins for the garbage collector, not for the DMA
// engine) reach the device through the slot's pinned mirror, so every such byte is copied once on
// the host. A plain memcpy costs three memory transfers per byte -- read the source, read the
// destination line for ownership, write it back -- and the destination is never read by a CPU again
// (the DMA engine reads it from DRAM). Streaming stores skip the read-for-ownership and keep the
// mirror out of the caches: two transfers per byte. GPUHASH_NT_COPY=0 falls back to memcpy.
#include <immintrin.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

namespace {

__attribute__((target("avx2"))) void copy_nt_avx2(uint8_t* dst, const uint8_t* src, size_t n) {
    // head: up to the next 32-byte boundary of dst
    size_t head = (32 - (reinterpret_cast<uintptr_t>(dst) & 31)) & 31;
    if (head > n) head = n;
    memcpy(dst, src, head);
    dst += head; src += head; n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 64));
        const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i*>(dst + i + 96), d);
    }
    _mm_sfence();  // the DMA that follows must see the streamed lines
    memcpy(dst + i, src + i, n - i);
}

bool use_nt() {
    static const bool v = [] {
        const char* e = getenv("GPUHASH_NT_COPY");
        if (e && *e == '0') return false;
        return __builtin_cpu_supports("avx2") != 0;
    }();
    return v;
}

}  // namespace

// Copy for staging into pinned memory that only the DMA engine will read.
void gh_stage_copy(uint8_t* dst, const uint8_t* src, size_t n) {
    if (n >= 4096 && use_nt()) copy_nt_avx2(dst, src, n);
    else memcpy(dst, src, n);
}
