// gh_kernels.cu -- sm_100a kernels of libgpuhash: the lane-per-message hasher and the stream generator.
//
// Mapping (north_star): one independent message (torrent piece / S3 part) per
// lane, 32 messages per warp, warps fully independent of one another (no __syncthreads anywhere).
// HBM is read exactly once. The batch is described to the TMA unit as a 2-D tensor
// [message][bytes of that message in this launch] (row pitch = message stride); one elected lane per
// warp issues cp.async.bulk.tensor.2d box loads of 32 rows x 128 bytes -- the next 128 bytes of each
// of the warp's 32 messages -- into a 128B-swizzled shared-memory tile, completion on a per-warp
// mbarrier, double-buffered so that stage s+1 is in flight while stage s is being compressed. Rows
// past the end of the batch are zero-filled by the TMA unit, so there is no ragged-edge code in the
// load path. Each lane then reads its own row with 128-bit LDS (the hardware swizzle makes the
// 8-lane phases conflict-free), byte-swaps with PRMT and runs the compression with the chaining state
// and the 16-word schedule ring in registers. Messages longer than one launch's window carry their
// midstate in HBM between launches (GhJob::state), which is what lets the host ring stream 64 MiB
// parts chunk-interleaved. Digest compare + pass/fail bitmap are fused into the same kernel
// (ballot -> one atomicOr per warp).
//
// What it replaces: anacrolix/torrent v1.15.2 Torrent.hashPiece / pieceHashed (crypto/sha1), reached
// from reference internal/downloader/torrent/torrent.go:79-106, and minio-go v6.0.55 hashMaterials
// (crypto/md5 + sha256-simd), reached from reference internal/uploader/uploader.go:89.
#include "gh_internal.h"

#include <cuda.h>
#include <stdlib.h>

#include <algorithm>
#include <string>
#include <string.h>

#include "../../include/gpuhash.h"
#include "gh_hash.cuh"

namespace {

using namespace ghd;

constexpr int kTileRowBytes = 128;                       // inner box extent = the swizzle span
constexpr int kTileBytes = 32 * kTileRowBytes;           // one TMA box: 32 messages x 128 B
constexpr int kStages = 2;
constexpr int kMaxWarpsPerBlock = 4;
// TPS = tiles per stage: 2 (256 B = 4 blocks of every message per stage, 16 KiB of shared memory per
// warp, 13 warps/SM) when there are many warps, 4 (8 blocks per stage, half the barrier traffic) when
// warps are few and shared memory is not what limits residency.
__host__ __device__ constexpr int warp_smem(int tps) { return kStages * tps * kTileBytes; }  // tiles only; mbarriers follow all tiles

// ---------------------------------------------------------------------------- mbarrier / TMA PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "GH_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GH_DONE;\n"
        "bra GH_WAIT;\n"
        "GH_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// TMA box load: 32 rows x 32 uint32 starting at element x of row y of the batch tensor.
__device__ __forceinline__ void tma_load_box(uint32_t dst, const CUtensorMap* tmap, uint32_t x, uint32_t y,
                                             uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
            "r"(dst),
        "l"(tmap), "r"(x), "r"(y), "r"(bar)
        : "memory");
}

// ---------------------------------------------------------------------------- per-lane hash state
template <int ALGOS, bool FMA>
struct Lane {
    uint32_t s1[5];
    uint32_t m5[4];
    uint32_t s2[8];
    uint32_t one;  // the kernel parameter 1, multiplier of the IMAD-based adds (gh_hash.cuh addf)

    __device__ __forceinline__ void init() {
        if constexpr (ALGOS & GH_SHA1) sha1_init(s1);
        if constexpr (ALGOS & GH_MD5) md5_init(m5);
        if constexpr (ALGOS & GH_SHA256) sha256_init(s2);
    }
    // midstates: [word][message] so that a warp's loads and stores are coalesced
    __device__ __forceinline__ void load(const uint32_t* st, uint64_t n, uint64_t m) {
        if constexpr (ALGOS & GH_SHA1) {
#pragma unroll
            for (int i = 0; i < 5; i++) s1[i] = st[(12 + i) * n + m];
        }
        if constexpr (ALGOS & GH_MD5) {
#pragma unroll
            for (int i = 0; i < 4; i++) m5[i] = st[i * n + m];
        }
        if constexpr (ALGOS & GH_SHA256) {
#pragma unroll
            for (int i = 0; i < 8; i++) s2[i] = st[(4 + i) * n + m];
        }
    }
    __device__ __forceinline__ void store(uint32_t* st, uint64_t n, uint64_t m) const {
        if constexpr (ALGOS & GH_SHA1) {
#pragma unroll
            for (int i = 0; i < 5; i++) st[(12 + i) * n + m] = s1[i];
        }
        if constexpr (ALGOS & GH_MD5) {
#pragma unroll
            for (int i = 0; i < 4; i++) st[i * n + m] = m5[i];
        }
        if constexpr (ALGOS & GH_SHA256) {
#pragma unroll
            for (int i = 0; i < 8; i++) st[(4 + i) * n + m] = s2[i];
        }
    }
    // r: one 64-byte block as sixteen words in memory (little-endian load) order
    __device__ __forceinline__ void absorb(const uint32_t (&r)[16]) {
        if constexpr (ALGOS & GH_MD5) md5_compress(m5, r);
        if constexpr (ALGOS & GH_SHA256) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; i++) w[i] = bswap(r[i]);
            sha256_compress<FMA>(s2, w, one);
        }
        if constexpr (ALGOS & GH_SHA1) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; i++) w[i] = bswap(r[i]);
            sha1_compress<FMA>(s1, w, one);
        }
    }
    // the block that carries the 64-bit message bit length in bytes 56..63 (words 14, 15 of r ignored)
    __device__ __forceinline__ void absorb_with_length(const uint32_t (&r)[16], uint64_t bits) {
        const uint32_t lo = static_cast<uint32_t>(bits), hi = static_cast<uint32_t>(bits >> 32);
        if constexpr (ALGOS & GH_MD5) {
            uint32_t x[16];
#pragma unroll
            for (int i = 0; i < 14; i++) x[i] = r[i];
            x[14] = lo; x[15] = hi;  // RFC 1321 3.2: low-order word first
            md5_compress(m5, x);
        }
        if constexpr (ALGOS & GH_SHA256) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 14; i++) w[i] = bswap(r[i]);
            w[14] = hi; w[15] = lo;  // FIPS 180-4 5.1.1: big-endian 64-bit length
            sha256_compress<FMA>(s2, w, one);
        }
        if constexpr (ALGOS & GH_SHA1) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 14; i++) w[i] = bswap(r[i]);
            w[14] = hi; w[15] = lo;
            sha1_compress<FMA>(s1, w, one);
        }
    }
};

// Padding (FIPS 180-4 5.1.1 / RFC 1321 3.1): tail = the rem (< 64) bytes after the last full block.
template <int ALGOS, bool FMA>
__device__ __noinline__ void finalize(Lane<ALGOS, FMA>& L, const uint8_t* tail, uint32_t rem, uint64_t total_len) {
    uint32_t r[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const uint32_t k = 4 * i + b;
            if (k < rem) v |= static_cast<uint32_t>(tail[k]) << (8 * b);
            else if (k == rem) v |= 0x80u << (8 * b);
        }
        r[i] = v;
    }
    const uint64_t bits = total_len * 8;
    if (rem >= 56) {  // no room for the length: it goes into a second, otherwise empty block
        L.absorb(r);
#pragma unroll
        for (int i = 0; i < 16; i++) r[i] = 0;
    }
    L.absorb_with_length(r, bits);
}

template <int ALGOS, bool FMA>
__device__ __forceinline__ bool emit(const Lane<ALGOS, FMA>& L, const GhJob& job, uint64_t m) {
    bool ok = false;
    if constexpr (ALGOS & GH_SHA1) {
        uint32_t d[5];
#pragma unroll
        for (int i = 0; i < 5; i++) d[i] = bswap(L.s1[i]);
        if (job.sha1) {
            uint32_t* o = reinterpret_cast<uint32_t*>(job.sha1 + 20 * m);
#pragma unroll
            for (int i = 0; i < 5; i++) o[i] = d[i];
        }
        if (job.expected) {
            const uint32_t* e = reinterpret_cast<const uint32_t*>(job.expected + 20 * m);
            uint32_t diff = 0;
#pragma unroll
            for (int i = 0; i < 5; i++) diff |= d[i] ^ e[i];
            ok = diff == 0;
        }
    }
    if constexpr (ALGOS & GH_MD5) {
        if (job.md5) {
            uint4* o = reinterpret_cast<uint4*>(job.md5 + 16 * m);
            *o = make_uint4(L.m5[0], L.m5[1], L.m5[2], L.m5[3]);
        }
    }
    if constexpr (ALGOS & GH_SHA256) {
        if (job.sha256) {
            uint4* o = reinterpret_cast<uint4*>(job.sha256 + 32 * m);
            o[0] = make_uint4(bswap(L.s2[0]), bswap(L.s2[1]), bswap(L.s2[2]), bswap(L.s2[3]));
            o[1] = make_uint4(bswap(L.s2[4]), bswap(L.s2[5]), bswap(L.s2[6]), bswap(L.s2[7]));
        }
    }
    return ok;
}

// 64 bytes at p (any alignment) -> sixteen memory-order words. The byte-granular fallback path.
__device__ __forceinline__ void load_block_direct(const uint8_t* p, uint32_t (&r)[16]) {
    if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
        const uint32_t* q = reinterpret_cast<const uint32_t*>(p);
#pragma unroll
        for (int i = 0; i < 16; i++) r[i] = __ldg(q + i);
    } else {
#pragma unroll
        for (int i = 0; i < 16; i++)
            r[i] = static_cast<uint32_t>(__ldg(p + 4 * i)) | (static_cast<uint32_t>(__ldg(p + 4 * i + 1)) << 8) |
                   (static_cast<uint32_t>(__ldg(p + 4 * i + 2)) << 16) |
                   (static_cast<uint32_t>(__ldg(p + 4 * i + 3)) << 24);
    }
}

// One 64-byte block of this lane's row out of a 128B-swizzled tile: 16-byte chunk c of row r sits at
// chunk (c ^ (r & 7)). Shared-space addresses are 32-bit, so no generic-address arithmetic per load.
__device__ __forceinline__ void lds_block(uint32_t row_addr, uint32_t c0, uint32_t sw, uint32_t (&q)[16]) {
#pragma unroll
    for (uint32_t c = 0; c < 4; c++)
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(q[4 * c]), "=r"(q[4 * c + 1]), "=r"(q[4 * c + 2]), "=r"(q[4 * c + 3])
                     : "r"(row_addr + (((c0 + c) ^ sw) << 4))
                     : "memory");
}

// ---------------------------------------------------------------------------- the hasher
template <int ALGOS, bool STAGED, bool FMA, int TPS>
__global__ void __launch_bounds__(32 * kMaxWarpsPerBlock)
gh_hash_lanes(const __grid_constant__ CUtensorMap tmap, const GhJob job) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t wpb = blockDim.x >> 5;
    const uint64_t m_warp = job.first_msg + (static_cast<uint64_t>(blockIdx.x) * wpb + warp) * 32u;
    const uint64_t m = m_warp + lane;
    const uint64_t m_end = job.first_msg + job.count;

    const bool active = m < m_end;
    const uint64_t total_m = !active ? 0 : job.lens ? job.lens[m] : (m == job.n_msgs - 1 ? job.last_msg_len : job.msg_len);
    const bool done_before = !active || (job.prior > 0 && job.prior >= total_m);
    const uint64_t remain = done_before ? 0 : total_m - job.prior;
    const bool ends_now = !done_before && remain <= job.chunk_len;
    const uint64_t avail = remain < job.chunk_len ? remain : job.chunk_len;
    const uint32_t nfull = static_cast<uint32_t>(avail >> 6);
    const uint32_t rem = ends_now ? static_cast<uint32_t>(avail & 63u) : 0u;
    const uint8_t* src = job.data + m * job.stride;

    Lane<ALGOS, FMA> L;
    L.one = job.one;
    if (!done_before) {
        if (job.prior == 0) L.init();
        else L.load(job.state, job.n_msgs, m);
    }

    if constexpr (STAGED) {
        constexpr uint32_t kChunk = TPS * kTileRowBytes;   // bytes of every message per stage
        constexpr uint32_t kBPS = kChunk / 64;             // blocks per stage
        constexpr uint32_t kWarpSmem = warp_smem(TPS);
        const uint32_t tiles_u32 = smem_u32(smem + warp * kWarpSmem);
        const uint32_t bar0 = smem_u32(smem + wpb * kWarpSmem) + warp * (8 * kStages);
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < kStages; s++) mbar_init(bar0 + 8 * s, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();

        const uint32_t n_stages = (__reduce_max_sync(0xffffffffu, nfull) + kBPS - 1) / kBPS;  // warp-uniform
        const uint32_t row0 = static_cast<uint32_t>(m_warp);  // tensor row of lane 0's message

        // one lane arms the stage barrier with the bytes of the boxes it is about to request
        auto issue = [&](uint32_t s) {
            const uint32_t b = s % kStages;
            const uint32_t bar = bar0 + 8 * b;
            uint32_t ntile = 0;
#pragma unroll
            for (uint32_t t = 0; t < TPS; t++)
                if ((s * kChunk + t * kTileRowBytes) / 4 < job.tma_inner) ntile++;
            mbar_arrive_expect_tx(bar, ntile * kTileBytes);
#pragma unroll
            for (uint32_t t = 0; t < TPS; t++) {
                const uint32_t x = (s * kChunk + t * kTileRowBytes) / 4;
                if (x < job.tma_inner) tma_load_box(tiles_u32 + (b * TPS + t) * kTileBytes, &tmap, x, row0, bar);
            }
        };
        if (lane == 0)
            for (uint32_t s = 0; s < kStages && s < n_stages; s++) issue(s);

        const uint32_t row_base = tiles_u32 + lane * kTileRowBytes;
        const uint32_t sw = lane & 7u;
        for (uint32_t s = 0; s < n_stages; s++) {
            const uint32_t b = s % kStages;
            mbar_wait(bar0 + 8 * b, (s / kStages) & 1u);
            const uint32_t first = s * kBPS;
            const uint32_t nb = nfull > first ? min(kBPS, nfull - first) : 0u;
#pragma unroll 1
            for (uint32_t j = 0; j < nb; j++) {
                uint32_t r[16];
                lds_block(row_base + (b * TPS + (j >> 1)) * kTileBytes, (j & 1u) * 4u, sw, r);
                L.absorb(r);
            }
            // every lane has consumed its row of buffer b: hand it back to the TMA unit for stage s+kStages
            __syncwarp();
            if (lane == 0 && s + kStages < n_stages) issue(s + kStages);
        }
    } else {
#pragma unroll 1
        for (uint32_t j = 0; j < nfull; j++) {
            uint32_t r[16];
            load_block_direct(src + static_cast<uint64_t>(j) * 64, r);
            L.absorb(r);
        }
    }

    bool ok = false;
    if (!done_before) {
        if (ends_now) {
            finalize<ALGOS, FMA>(L, src + static_cast<uint64_t>(nfull) * 64, rem, total_m);
            ok = emit<ALGOS, FMA>(L, job, m);
        } else {
            L.store(job.state, job.n_msgs, m);
        }
    }
    __syncwarp();
    if (job.ok_words) {
        const uint32_t okmask = __ballot_sync(0xffffffffu, ok);
        if (lane == 0 && okmask) {  // the warp's 32 messages may straddle two bitmap words
            const uint64_t bit0 = job.ok_base + m;
            const uint32_t sh = static_cast<uint32_t>(bit0 & 31u);
            atomicOr(job.ok_words + (bit0 >> 5), okmask << sh);
            if (sh && (okmask >> (32 - sh))) atomicOr(job.ok_words + (bit0 >> 5) + 1, okmask >> (32 - sh));
        }
    }
}

// ---------------------------------------------------------------------------- role-split hasher
// For batches with too few messages to give every sub-partition a warp (BASELINE configs 1-4 are all
// like that: 4 096 / 12 800 / 320 / 3 073 messages on 592 sub-partitions), a lone warp runs at the
// rate one ALU pipe issues it -- ~1 140 cycles per SHA-1 block, ~2 700 per SHA-256 block -- and idle
// lanes or idle sub-partitions do not help (measured: profiles/). What does help is taking the part
// of a block that does not depend on the chaining state off the chain's warp: the message schedule
// (byte swap, W[16..] expansion, + K[t]) is ~40 % of SHA-1's and ~45 % of SHA-256's instructions.
// So here 32 messages are served by a CTA of two (three) warps on different sub-partitions:
//   warp S  "schedule": drives the TMA ring, reads the raw block, expands W, adds K and stores the
//           round inputs WK[t] into a shared-memory buffer (one STS.128 per 4 rounds, conflict-free);
//   warp R  "rounds":   owns the chaining state, reads WK[t] and runs only the rounds;
//   warp M  (fused MD5+SHA-256 only): MD5 over the same TMA tiles.
// Hand-off is by mbarrier pairs (full/empty) on two WK buffers and on the TMA stages.
#ifndef GH_HANDOFF
#define GH_HANDOFF 4
#endif
#ifndef GH_SPLIT_R_PAIRS
#define GH_SPLIT_R_PAIRS 1
#endif
#ifndef GH_R_IADD3
#define GH_R_IADD3 1  // SHA-1 rounds warps: f + e + WK as one IADD3 (see sha1_rounds); 0 = the two-IMAD form
#endif
#ifndef GH_TRIO_R_LOW
#define GH_TRIO_R_LOW 0  // 1 = rounds warps take the low warp ids of their sub-partition (A/B of the issue arbiter's preference)
#endif
constexpr int kHandoff = GH_HANDOFF;  // 64-byte blocks per S -> R hand-off (one barrier round trip each)
constexpr int kSplitStages = 3;
constexpr int kSplitTPS = 2;
constexpr uint32_t kSplitChunk = kSplitTPS * kTileRowBytes;
constexpr uint32_t kSplitBPS = kSplitChunk / 64;
template <int ALGO>
struct SplitCfg {
    static constexpr int kWords = ALGO == GH_SHA1 ? 80 : 64;       // round inputs per block
    static constexpr int kWkBytes = kWords * 32 * 4;                // one WK buffer: [word/4][lane][4]
};
template <int ALGO>
__host__ __device__ constexpr int split_smem() {
    return kSplitStages * kSplitTPS * kTileBytes + 2 * kHandoff * SplitCfg<ALGO>::kWkBytes + 128;
}

__device__ __forceinline__ uint32_t sha1_k(int t) {
    return t < 20 ? 0x5a827999u : t < 40 ? 0x6ed9eba1u : t < 60 ? 0x8f1bbcdcu : 0xca62c1d6u;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void mbar_arrive1(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// SHA-1 rounds over precomputed WK[t] = W[t] + K[t] (FIPS 180-4 6.1.2 step 3 without the schedule)
// `one` (the kernel parameter 1) makes e + WK an IMAD that ptxas can neither fold nor re-associate:
// that sum is ready rounds ahead, and keeping it a separate operand leaves exactly one dependent
// instruction (LEA.HI = rotl5(a) + t) between consecutive values of a.
__device__ __forceinline__ void sha1_rounds(uint32_t (&s)[5], uint32_t wk_addr, uint32_t one) {
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4];
    uint4 v[3];
    v[0] = lds128(wk_addr);
    v[1] = lds128(wk_addr + 512);
    v[2] = lds128(wk_addr + 1024);
#pragma unroll
    for (int ch = 0; ch < 20; ch++) {
        const uint4 cur = v[ch % 3];
        if (ch + 3 < 20) v[ch % 3] = lds128(wk_addr + (ch + 3) * 512);
        const uint32_t wk[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int t = 4 * ch + i;
            uint32_t f;
            if (t < 20) f = f_ch(b, c, d);
            else if (t < 40) f = f_xor3(b, c, d);
            else if (t < 60) f = f_maj(b, c, d);
            else f = f_xor3(b, c, d);
#if GH_R_IADD3
            // f + e + WK as ONE three-input add on the ALU pipe: a fourth ALU instruction per round, but no
            // cross-pipe hop on the path a[t] -> f -> sum -> a[t+2] (static schedule 682 vs 749 clocks per block)
            const uint32_t pre = e + wk[i];
            const uint32_t tmp = rotl(a, 5) + (f + pre);
#else
            const uint32_t pre = addf<true>(e, wk[i], one);
            const uint32_t tmp = rotl(a, 5) + (f + pre);
#endif
            e = d; d = c; c = rotl(b, 30); b = a; a = tmp;
        }
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e;
}

// NB consecutive blocks whose round inputs lie back to back (block j at wk_addr + j * 80 * 128): one
// rolling prefetch of three LDS.128 runs across the block boundaries, so only the first block of the
// batch pays the load latency at its start.
template <int NB>
__device__ __forceinline__ void sha1_rounds_n(uint32_t (&s)[5], uint32_t wk_addr) {
    uint4 v[3];
    v[0] = lds128(wk_addr);
    v[1] = lds128(wk_addr + 512);
    v[2] = lds128(wk_addr + 1024);
#pragma unroll
    for (int j = 0; j < NB; j++) {
        uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4];
#pragma unroll
        for (int ch = 0; ch < 20; ch++) {
            const int g = 20 * j + ch;
            const uint4 cur = v[g % 3];
            if (g + 3 < 20 * NB) v[g % 3] = lds128(wk_addr + (g + 3) * 512);
            const uint32_t wk[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int t = 4 * ch + i;
                uint32_t f;
                if (t < 20) f = f_ch(b, c, d);
                else if (t < 40) f = f_xor3(b, c, d);
                else if (t < 60) f = f_maj(b, c, d);
                else f = f_xor3(b, c, d);
                const uint32_t pre = e + wk[i];
                const uint32_t tmp = rotl(a, 5) + (f + pre);
                e = d; d = c; c = rotl(b, 30); b = a; a = tmp;
            }
        }
        s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e;
    }
}

__device__ __forceinline__ void sha256_rounds(uint32_t (&s)[8], uint32_t wk_addr, uint32_t one) {
    uint32_t x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = s[i];
    uint4 v[3];
    v[0] = lds128(wk_addr);
    v[1] = lds128(wk_addr + 512);
    v[2] = lds128(wk_addr + 1024);
#pragma unroll
    for (int ch = 0; ch < 16; ch++) {
        const uint4 cur = v[ch % 3];
        if (ch + 3 < 16) v[ch % 3] = lds128(wk_addr + (ch + 3) * 512);
        const uint32_t wk[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int t = 4 * ch + i;
            uint32_t& a = x[(0 - t) & 7]; uint32_t& b = x[(1 - t) & 7]; uint32_t& c = x[(2 - t) & 7];
            uint32_t& d = x[(3 - t) & 7]; uint32_t& e = x[(4 - t) & 7]; uint32_t& f = x[(5 - t) & 7];
            uint32_t& g = x[(6 - t) & 7]; uint32_t& h = x[(7 - t) & 7];
            // h + WK is ready rounds ahead (IMAD, not re-associable); Sigma1(e), the deepest term, is added last
            const uint32_t hk = addf<true>(h, wk[i], one);
            const uint32_t xc = hk + f_ch(e, f, g);
            const uint32_t s1e = S256_S1(e);
            const uint32_t t2 = S256_S0(a) + f_maj(a, b, c);
            d = (d + xc) + s1e;
            h = (xc + t2) + s1e;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) s[i] += x[i];
}

// schedule warp: w = the block as big-endian-decoded words; writes WK[0..kWords) for its lane
template <int ALGO>
__device__ __forceinline__ void expand_and_store(uint32_t (&w)[16], uint32_t wk_addr) {
    uint32_t o[4];
    if constexpr (ALGO == GH_SHA1) {
#pragma unroll
        for (int t = 0; t < 80; t++) {
            uint32_t wt;
            if (t < 16) wt = w[t];
            else { wt = rotl(f_xor3(w[(t - 3) & 15], w[(t - 8) & 15], w[(t - 14) & 15]) ^ w[t & 15], 1); w[t & 15] = wt; }
            o[t & 3] = wt + sha1_k(t);
            if ((t & 3) == 3) sts128(wk_addr + (t >> 2) * 512, o[0], o[1], o[2], o[3]);
        }
    } else {
#pragma unroll
        for (int t = 0; t < 64; t++) {
            uint32_t wt;
            if (t < 16) wt = w[t];
            else {
                wt = S256_s1(w[(t - 2) & 15]) + w[(t - 7) & 15] + S256_s0(w[(t - 15) & 15]) + w[t & 15];
                w[t & 15] = wt;
            }
            o[t & 3] = wt + K256[t];
            if ((t & 3) == 3) sts128(wk_addr + (t >> 2) * 512, o[0], o[1], o[2], o[3]);
        }
    }
}

template <int ALGO, bool WITH_MD5>
__global__ void __launch_bounds__(WITH_MD5 ? 96 : 64)
gh_hash_split(const __grid_constant__ CUtensorMap tmap, const GhJob job) {
    extern __shared__ __align__(1024) uint8_t smem[];
    using Cfg = SplitCfg<ALGO>;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t role = threadIdx.x >> 5;  // 0 = rounds, 1 = schedule, 2 = MD5
    const uint64_t m_warp = job.first_msg + static_cast<uint64_t>(blockIdx.x) * 32u;
    const uint64_t m = m_warp + lane;
    const bool active = m < job.first_msg + job.count;

    // Every message this grid covers has the same number of bytes present in this launch (the launcher
    // excludes a last message with fewer), so trip counts are uniform across the CTA; whether a
    // message ENDS in this launch is still a per-lane matter (the last one may be shorter overall).
    const uint64_t common_remain = job.prior >= job.msg_len ? 0 : job.msg_len - job.prior;
    const uint32_t n_blocks = static_cast<uint32_t>((common_remain < job.chunk_len ? common_remain : job.chunk_len) >> 6);
    const uint32_t n_stages = (n_blocks + kSplitBPS - 1) / kSplitBPS;
    const uint64_t total_m = !active ? 0 : job.lens ? job.lens[m] : (m == job.n_msgs - 1 ? job.last_msg_len : job.msg_len);
    const bool done_before = !active || (job.prior > 0 && job.prior >= total_m);
    const uint64_t remain = done_before ? 0 : total_m - job.prior;
    const bool ends_now = !done_before && remain <= job.chunk_len;
    const uint64_t avail = remain < job.chunk_len ? remain : job.chunk_len;
    const uint32_t my_blocks = static_cast<uint32_t>(avail >> 6);  // == n_blocks unless a length table says less
    const uint32_t rem = ends_now ? static_cast<uint32_t>(avail & 63u) : 0u;
    const uint8_t* src = job.data + m * job.stride;

    const uint32_t tiles_u32 = smem_u32(smem);
    const uint32_t wk_u32 = tiles_u32 + kSplitStages * kSplitTPS * kTileBytes;
    const uint32_t bars = wk_u32 + 2 * kHandoff * Cfg::kWkBytes;
    const uint32_t tma_full = bars, tma_empty = bars + 8 * kSplitStages;
    const uint32_t w_full = bars + 16 * kSplitStages, w_empty = w_full + 16;
    constexpr uint32_t kTileReaders = WITH_MD5 ? 2 : 1;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < kSplitStages; s++) {
            mbar_init(tma_full + 8 * s, 1);
            mbar_init(tma_empty + 8 * s, kTileReaders);
        }
#pragma unroll
        for (int b = 0; b < 2; b++) {
            mbar_init(w_full + 8 * b, 1);
            mbar_init(w_empty + 8 * b, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();  // the only CTA-wide barrier: after this the warps meet through mbarriers only

    const uint32_t row_base = tiles_u32 + lane * kTileRowBytes;
    const uint32_t sw = lane & 7u;

    if (role == 1) {
        // ------------------------------------------------------------------ schedule warp
        const uint32_t row0 = static_cast<uint32_t>(m_warp);
        auto issue = [&](uint32_t s) {
            const uint32_t b = s % kSplitStages;
            uint32_t ntile = 0;
#pragma unroll
            for (uint32_t t = 0; t < kSplitTPS; t++)
                if ((s * kSplitChunk + t * kTileRowBytes) / 4 < job.tma_inner) ntile++;
            mbar_arrive_expect_tx(tma_full + 8 * b, ntile * kTileBytes);
#pragma unroll
            for (uint32_t t = 0; t < kSplitTPS; t++) {
                const uint32_t x = (s * kSplitChunk + t * kTileRowBytes) / 4;
                if (x < job.tma_inner)
                    tma_load_box(tiles_u32 + (b * kSplitTPS + t) * kTileBytes, &tmap, x, row0, tma_full + 8 * b);
            }
        };
        if (lane == 0)
            for (uint32_t s = 0; s < kSplitStages && s < n_stages; s++) issue(s);
        uint32_t jb = 0;
        for (uint32_t s = 0; s < n_stages; s++) {
            const uint32_t b = s % kSplitStages;
            mbar_wait(tma_full + 8 * b, (s / kSplitStages) & 1u);
            const uint32_t nb = min(kSplitBPS, n_blocks - s * kSplitBPS);
#pragma unroll 1
            for (uint32_t j = 0; j < nb; j++, jb++) {
                uint32_t r[16], w[16];
                lds_block(row_base + (b * kSplitTPS + (j >> 1)) * kTileBytes, (j & 1u) * 4u, sw, r);
#pragma unroll
                for (int i = 0; i < 16; i++) w[i] = bswap(r[i]);
                const uint32_t pair = jb / kHandoff, half = jb % kHandoff;
                const uint32_t wb = pair & 1u, use = pair >> 1;
                if (half == 0 && use) mbar_wait(w_empty + 8 * wb, (use - 1) & 1u);  // rounds warp has drained this buffer
                expand_and_store<ALGO>(w, wk_u32 + (wb * kHandoff + half) * Cfg::kWkBytes + lane * 16);
                if (half == kHandoff - 1 || jb == n_blocks - 1) {
                    __syncwarp();  // orders every lane's stores before lane 0's release-arrive
                    if (lane == 0) mbar_arrive1(w_full + 8 * wb);
                }
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive1(tma_empty + 8 * b);
                if (s + kSplitStages < n_stages) {
                    mbar_wait(tma_empty + 8 * b, (s / kSplitStages) & 1u);  // the MD5 warp is done with it too
                    issue(s + kSplitStages);
                }
            }
        }
    } else if (role == 0) {
        // ------------------------------------------------------------------ rounds warp
        Lane<ALGO, false> L;
        L.one = 1;
        if (job.prior == 0 || done_before) L.init();
        else L.load(job.state, job.n_msgs, m);
#pragma unroll 1
        for (uint32_t jb = 0; jb < n_blocks; jb += kHandoff) {
            const uint32_t pair = jb / kHandoff, wb = pair & 1u, use = pair >> 1;
            mbar_wait(w_full + 8 * wb, use & 1u);
#pragma unroll 1
            for (uint32_t half = 0; half < kHandoff && jb + half < n_blocks; half++) {
                const uint32_t wk_addr = wk_u32 + (wb * kHandoff + half) * Cfg::kWkBytes + lane * 16;
                if constexpr (ALGO == GH_SHA1 && GH_SPLIT_R_PAIRS) {
                    // two blocks as one unrolled body, WK prefetch rolling across the boundary (the rounds
                    // warp has its sub-partition to itself here; see gh_hash_trio<1, true>)
                    if (job.lens == nullptr && jb + half + 2 <= n_blocks && half + 2 <= kHandoff) {
                        sha1_rounds_n<2>(L.s1, wk_addr);
                        half++;
                        continue;
                    }
                }
                if (jb + half < my_blocks) {
                    if constexpr (ALGO == GH_SHA1) sha1_rounds(L.s1, wk_addr, job.one);
                    else sha256_rounds(L.s2, wk_addr, job.one);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive1(w_empty + 8 * wb);
        }
        bool ok = false;
        if (active && !done_before) {
            if (ends_now) {
                finalize<ALGO, false>(L, src + static_cast<uint64_t>(my_blocks) * 64, rem, total_m);
                ok = emit<ALGO, false>(L, job, m);
            } else {
                L.store(job.state, job.n_msgs, m);
            }
        }
        __syncwarp();
        if (job.ok_words) {
            const uint32_t okmask = __ballot_sync(0xffffffffu, ok);
            if (lane == 0 && okmask) {
                const uint64_t bit0 = job.ok_base + m;
                const uint32_t sh = static_cast<uint32_t>(bit0 & 31u);
                atomicOr(job.ok_words + (bit0 >> 5), okmask << sh);
                if (sh && (okmask >> (32 - sh))) atomicOr(job.ok_words + (bit0 >> 5) + 1, okmask >> (32 - sh));
            }
        }
    } else if (WITH_MD5) {
        // ------------------------------------------------------------------ MD5 warp
        Lane<GH_MD5, false> L;
        L.one = 1;
        if (job.prior == 0 || done_before) L.init();
        else L.load(job.state, job.n_msgs, m);
        for (uint32_t s = 0; s < n_stages; s++) {
            const uint32_t b = s % kSplitStages;
            mbar_wait(tma_full + 8 * b, (s / kSplitStages) & 1u);
            const uint32_t nb = min(kSplitBPS, n_blocks - s * kSplitBPS);
#pragma unroll 1
            for (uint32_t j = 0; j < nb; j++) {
                if (s * kSplitBPS + j >= my_blocks) break;
                uint32_t r[16];
                lds_block(row_base + (b * kSplitTPS + (j >> 1)) * kTileBytes, (j & 1u) * 4u, sw, r);
                L.absorb(r);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive1(tma_empty + 8 * b);
        }
        if (active && !done_before) {
            if (ends_now) {
                finalize<GH_MD5, false>(L, src + static_cast<uint64_t>(my_blocks) * 64, rem, total_m);
                emit<GH_MD5, false>(L, job, m);
            } else {
                L.store(job.state, job.n_msgs, m);
            }
        }
    }
}

// ---------------------------------------------------------------------------- trio hasher (SHA-1)
// What the two kernels above leave on the table (measured on B200: profiles/r02_ab_sha1_kernels.jsonl,
// profiles/r02_ab_team_kernel_rejected.jsonl): a warp that is alone on its sub-partition issues at
// 0.45-0.7 IPC whatever it runs -- its own dependent instructions are 4-5 clocks apart and nothing
// fills the gaps -- so neither the lane kernel (one chain = one warp = 1 185 clk per 64-byte block at
// config 2) nor a rounds-warp fed by ONE schedule warp (which, alone on its sub-partition, needs ~490
// clk per block and is the bottleneck as soon as it serves more than one chain) gets near the ALU
// pipes' capacity.
//
// The schedule of block b does not depend on block b-1, so one chain can feed on SEVERAL schedule
// warps. Here a chain (32 messages) is served by three warps:
//   R    rounds: owns the chaining state, reads W+K from shared memory, runs only the 80 rounds;
//   S_a  schedule (TMA tiles -> byte swap -> W[0..79] + K -> shared memory) of RATIO of every RATIO+1
//        hand-off batches;
//   S_b  schedule of the remaining one.
// A CTA carries one chain (three warps on three sub-partitions) or three: nine working warps on the
// four sub-partitions of one SM. With warp w on sub-partition w % 4 the roles are laid out so that
// sub-partitions 0-2 each hold one R and its chain's S_b and sub-partition 3 holds the three S_a
// warps; RATIO = 2 evens the ALU issue load out (R 650 + S_b 142 clocks per block-step against
// 3 x 284). Every (chain, buffer) has exactly one producer warp and one consumer warp, so the
// full/empty mbarrier pairs never have two waiters a phase apart; every TMA stage belongs to one
// S warp, its only reader, so the tile ring needs no "empty" barrier.
#ifndef GH_TRIO_RATIO3
#define GH_TRIO_RATIO3 2  // S_a : S_b share in the three-chains-per-CTA layout (2:1 measured +5 % over 1:1 at config 2)
#endif
#ifndef GH_TRIO_DEBUG
#define GH_TRIO_DEBUG 0  // measurement builds only: 1 = rounds warps idle, 2 = schedule warps idle (digests are then wrong)
#endif
constexpr int kTrioMaxChains = 3;
constexpr int kTrioH = 2;          // blocks per hand-off batch = one TMA tile row of 128 bytes per message
constexpr uint32_t kTrioWkBlock = 80 * 32 * 4;                       // [t/4][lane][4]
// RATIO = how many hand-off batches S_a takes for each one S_b takes. S_b shares its sub-partition with
// R in the three-chain layout, S_a's sit together on the fourth: 2:1 evens the ALU load out (R 650 +
// S_b 142 issue clocks per block-step against 3 x 284 on the S_a sub-partition; 1:1 is 863 against 639).
// Batch q goes to WK buffer q % (RATIO+1); S_b owns the last buffer, S_a the others; TMA stage q % stages
// with stages a multiple of RATIO+1, so every buffer and every stage has exactly one producer warp.
template <int RATIO>
struct TrioCfg {
    static constexpr uint32_t kBufs = RATIO + 1;
    static constexpr uint32_t kStages = RATIO == 1 ? 4 : RATIO + 1;
    static constexpr uint32_t kTilesPerChain = kStages * kTileBytes;
    static constexpr uint32_t kWkPerChain = kBufs * kTrioH * kTrioWkBlock;
    static constexpr uint32_t kBarsPerChain = 8 * (kStages + 2 * kBufs);
    __host__ __device__ static constexpr uint32_t smem(int chains) { return chains * (kTilesPerChain + kWkPerChain + kBarsPerChain) + 64; }
};
__host__ __device__ constexpr int trio_warps(int chains) { return chains == 1 ? 3 : chains == 2 ? 8 : 12; }

// role of warp w in a CTA of `chains` chains: 0 = rounds, 1 = schedule (even batches), 2 = schedule (odd), 3 = none
__device__ __forceinline__ void trio_role(uint32_t w, uint32_t chains, uint32_t* role, uint32_t* chain) {
    *role = 3;
    *chain = 0;
    if (chains == 1) {  // three warps, three sub-partitions
        *role = w;
    } else if (chains == 2) {  // R0 R1 on 0,1 with S_b of the same chain below them; S_a on 2,3
        if (w == 4 || w == 5) { *role = 0; *chain = w - 4; }
        else if (w == 0 || w == 1) { *role = 2; *chain = w; }
        else if (w == 2 || w == 3) { *role = 1; *chain = w - 2; }
        else if (w == 6 || w == 7) { *role = 3; }
    } else {
#if GH_TRIO_R_LOW
        if (w <= 2) { *role = 0; *chain = w; }
        else if (w >= 4 && w <= 6) { *role = 2; *chain = w - 4; }
#else
        if (w >= 4 && w <= 6) { *role = 0; *chain = w - 4; }
        else if (w <= 2) { *role = 2; *chain = w; }
#endif
        else if ((w & 3u) == 3u) { *role = 1; *chain = w >> 2; }
    }
}

template <int RATIO, bool RBATCH>
__global__ void __launch_bounds__(32 * 12)
gh_hash_trio(const __grid_constant__ CUtensorMap tmap, const GhJob job) {
    extern __shared__ __align__(1024) uint8_t smem[];
    using Cfg = TrioCfg<RATIO>;
    constexpr uint32_t kBufs = Cfg::kBufs, kTrioStages = Cfg::kStages;
    constexpr uint32_t kTrioTilesPerChain = Cfg::kTilesPerChain, kTrioWkPerChain = Cfg::kWkPerChain, kTrioBarsPerChain = Cfg::kBarsPerChain;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t chains = job.chains_per_cta;
    uint32_t role, ci;
    trio_role(threadIdx.x >> 5, chains, &role, &ci);
    const uint64_t mw0 = static_cast<uint64_t>(blockIdx.x) * chains;  // first message-warp (chain) of this CTA
    const uint64_t n_mw_total = (job.count + 31) / 32;
    const uint32_t n_mw = static_cast<uint32_t>(n_mw_total - mw0 < chains ? n_mw_total - mw0 : chains);

    // As in gh_hash_split: every message of the grid has the same number of bytes in this launch.
    const uint64_t common_remain = job.prior >= job.msg_len ? 0 : job.msg_len - job.prior;
    const uint32_t n_blocks = static_cast<uint32_t>((common_remain < job.chunk_len ? common_remain : job.chunk_len) >> 6);
    const uint32_t n_batches = (n_blocks + kTrioH - 1) / kTrioH;

    // shared memory: [tiles of chain 0..][WK of chain 0..][barriers of chain 0..]
    const uint32_t base = smem_u32(smem);
    const uint32_t tiles = base + ci * kTrioTilesPerChain;
    const uint32_t wk = base + chains * kTrioTilesPerChain + ci * kTrioWkPerChain;
    const uint32_t bars = base + chains * (kTrioTilesPerChain + kTrioWkPerChain) + ci * kTrioBarsPerChain;
    const uint32_t tile_full = bars, w_full = bars + 8 * kTrioStages, w_empty = w_full + 8 * kBufs;
    if (threadIdx.x == 0) {
        const uint32_t b0 = base + chains * (kTrioTilesPerChain + kTrioWkPerChain);
        for (uint32_t k = 0; k < chains * (kTrioStages + 2 * kBufs); k++) mbar_init(b0 + 8 * k, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();  // the only CTA-wide barrier
    if (role == 3 || ci >= n_mw) return;

    if (role != 0) {
        // ------------------------------------------------------------------ schedule warp
        // S_b (role 2) takes the batches q = kBufs-1 (mod kBufs), S_a (role 1) all others; a batch's WK buffer
        // is q % kBufs and its TMA stage q % kTrioStages, both owned by the warp that owns the batch.
        const bool is_b = role == 2;
        auto mine = [&](uint32_t q) { return ((q % kBufs) == kBufs - 1) == is_b; };
        const uint32_t sw = lane & 7u;
        const uint32_t row0 = static_cast<uint32_t>(job.first_msg + (mw0 + ci) * 32u);
        const uint32_t row_base = tiles + lane * kTileRowBytes;
        auto issue = [&](uint32_t q) {  // batch q = bytes [128 q, 128 q + 128) of every message
            const uint32_t st = q % kTrioStages;
            mbar_arrive_expect_tx(tile_full + 8 * st, kTileBytes);
            tma_load_box(tiles + st * kTileBytes, &tmap, q * (kTileRowBytes / 4), row0, tile_full + 8 * st);
        };
        if (lane == 0)
            for (uint32_t q = 0; q < kTrioStages && q < n_batches; q++)
                if (mine(q)) issue(q);
#pragma unroll 1
        for (uint32_t q = 0; q < n_batches; q++) {
            if (!mine(q)) continue;
            const uint32_t st = q % kTrioStages, buf = q % kBufs, use = q / kBufs;
            mbar_wait(tile_full + 8 * st, (q / kTrioStages) & 1u);
            if (use) mbar_wait(w_empty + 8 * buf, (use - 1) & 1u);  // R has drained this buffer
            const uint32_t nb = min(static_cast<uint32_t>(kTrioH), n_blocks - q * kTrioH);
            const uint32_t wk0 = wk + buf * kTrioH * kTrioWkBlock + lane * 16;
#pragma unroll 1
            for (uint32_t j = 0; j < nb; j++) {
#if GH_TRIO_DEBUG != 2
                uint32_t r[16], w[16];
                lds_block(row_base + st * kTileBytes, j * 4u, sw, r);
#pragma unroll
                for (int k = 0; k < 16; k++) w[k] = bswap(r[k]);
                expand_and_store<GH_SHA1>(w, wk0 + j * kTrioWkBlock);
#endif
            }
            __syncwarp();  // every lane's stores and tile reads precede lane 0's release-arrive / re-issue
            if (lane == 0) {
                mbar_arrive1(w_full + 8 * buf);
                if (q + kTrioStages < n_batches) issue(q + kTrioStages);
            }
        }
        return;
    }

    // ---------------------------------------------------------------------- rounds warp
    const uint64_t m = job.first_msg + (mw0 + ci) * 32u + lane;
    const bool active = m < job.first_msg + job.count;
    const uint64_t total_m = !active ? 0 : job.lens ? job.lens[m] : (m == job.n_msgs - 1 ? job.last_msg_len : job.msg_len);
    const bool done_before = !active || (job.prior > 0 && job.prior >= total_m);
    const uint64_t remain = done_before ? 0 : total_m - job.prior;
    const bool ends_now = !done_before && remain <= job.chunk_len;
    const uint64_t avail = remain < job.chunk_len ? remain : job.chunk_len;
    const uint32_t my_blocks = static_cast<uint32_t>(avail >> 6);
    const uint32_t rem = ends_now ? static_cast<uint32_t>(avail & 63u) : 0u;
    const uint8_t* src = job.data + m * job.stride;

    Lane<GH_SHA1, false> L;
    L.one = 1;
    if (job.prior == 0 || done_before) L.init();
    else L.load(job.state, job.n_msgs, m);
#pragma unroll 1
    for (uint32_t q = 0; q < n_batches; q++) {
        const uint32_t p = q % kBufs;
        mbar_wait(w_full + 8 * p, (q / kBufs) & 1u);
        const uint32_t wk0 = wk + p * kTrioH * kTrioWkBlock + lane * 16;
        // RBATCH (one chain per CTA): a whole hand-off batch as one unrolled body. Without a length table
        // every row of the grid has all n_blocks blocks (the launcher keeps a shorter last message out
        // of the tensor); idle lanes compute on zeros and never store.
        if (RBATCH && job.lens == nullptr && (q + 1) * kTrioH <= n_blocks) {
            sha1_rounds_n<kTrioH>(L.s1, wk0);
        } else {
#pragma unroll 1
            for (uint32_t j = 0; j < kTrioH && q * kTrioH + j < n_blocks; j++)
#if GH_TRIO_DEBUG != 1
                if (q * kTrioH + j < my_blocks) sha1_rounds(L.s1, wk0 + j * kTrioWkBlock, job.one);
#endif
        }
        __syncwarp();
        if (lane == 0) mbar_arrive1(w_empty + 8 * p);
    }
    bool ok = false;
    if (active && !done_before) {
        if (ends_now) {
            finalize<GH_SHA1, false>(L, src + static_cast<uint64_t>(my_blocks) * 64, rem, total_m);
            ok = emit<GH_SHA1, false>(L, job, m);
        } else {
            L.store(job.state, job.n_msgs, m);
        }
    }
    __syncwarp();
    if (job.ok_words) {
        const uint32_t okmask = __ballot_sync(0xffffffffu, ok);
        if (lane == 0 && okmask) {
            const uint64_t bit0 = job.ok_base + m;
            const uint32_t sh = static_cast<uint32_t>(bit0 & 31u);
            atomicOr(job.ok_words + (bit0 >> 5), okmask << sh);
            if (sh && (okmask >> (32 - sh))) atomicOr(job.ok_words + (bit0 >> 5) + 1, okmask >> (32 - sh));
        }
    }
}

// ---------------------------------------------------------------------------- synthetic stream
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// stream byte k = byte (k & 7) of mix64(seed + (k/8 + 1) * golden), little-endian.
__global__ void gh_generate(uint8_t* out, uint64_t seed, uint64_t offset, uint64_t len) {
    const uint64_t first_word = offset >> 3;
    const uint64_t n_words = ((offset + len + 7) >> 3) - first_word;
    const uint64_t step = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    const bool aligned = ((reinterpret_cast<uintptr_t>(out) - (offset & 7)) & 7) == 0;
    for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_words; i += step) {
        const uint64_t wi = first_word + i;
        const uint64_t w = mix64(seed + (wi + 1) * 0x9E3779B97F4A7C15ull);
        const uint64_t k0 = wi << 3;  // stream offset of this word's byte 0
        if (aligned && k0 >= offset && k0 + 8 <= offset + len) {
            *reinterpret_cast<uint64_t*>(out + (k0 - offset)) = w;
        } else {
#pragma unroll
            for (int b = 0; b < 8; b++) {
                const uint64_t k = k0 + b;
                if (k >= offset && k < offset + len) out[k - offset] = static_cast<uint8_t>(w >> (8 * b));
            }
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// The batch as a 2-D uint32 tensor: row m = the 64-byte-block part of message m's bytes in this launch.
bool make_tmap(const GhJob& job, uint64_t rows, uint64_t inner_words, CUtensorMap* out) {
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return false;
    const cuuint64_t dims[2] = {inner_words, rows};
    const cuuint64_t strides[1] = {job.stride};
    const cuuint32_t box[2] = {kTileRowBytes / 4, 32};
    const cuuint32_t estr[2] = {1, 1};
    return enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint8_t*>(job.data), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Tuning switches (read once): GPUHASH_FMA_ADDS=0 keeps every add on the ALU pipe, GPUHASH_TPS=2|4
// forces the stage depth. Defaults are what measured best on B200 (profiles/).
int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}
// -1 = per-algorithm default (measured on B200, profiles/r01_ab_*.jsonl): SHA-1 moves its off-chain
// adds to the FMA pipe (+10..13 % both for a lone warp and saturated), SHA-256 moves all of them
// (+12..15 % saturated; lone warps use the role-split kernels instead), fused MD5+SHA-256 only when
// warps are plentiful.
int fma_adds_override() {
    static const int v = env_int("GPUHASH_FMA_ADDS", -1);
    return v;
}
// Role-split kernels are used while message-warps <= this many per SM (0 disables). Measured on B200:
// at 64..128 message-warps the split is +27..48 %; at 400 (2.7 per SM) its 800 warps oversubscribe the
// 592 sub-partitions and it loses 7 %, so the default is 2 per SM = one warp per sub-partition.
int split_max_warps_per_sm() {
    static const int v = env_int("GPUHASH_SPLIT_WARPS_PER_SM", 2);
    return v;
}
// The trio kernel (per chain one rounds-warp fed by two schedule-warps, up to three chains per SM)
// serves staged SHA-1 batches of at most this many message-warps per SM (0 disables it; the lane
// and 1R+1S split kernels then take over as before).
int trio_max_warps_per_sm() {
    static const int v = env_int("GPUHASH_TRIO_WARPS_PER_SM", kTrioMaxChains);
    return v;
}
// Share of the schedule work between a chain's two schedule warps (S_a : S_b). GPUHASH_TRIO_RATIO=1|2
// forces it; default per layout (measured, profiles/r02_ab_sha1_kernels.jsonl).
int trio_ratio(int chains) {
    static const int v = env_int("GPUHASH_TRIO_RATIO", 0);
    if (v == 1 || v == 2) return v;
    return chains == 3 ? GH_TRIO_RATIO3 : 1;
}
int forced_tps() {
    static const int v = env_int("GPUHASH_TPS", 0);
    return v;
}

// Name of the hashing kernel this host thread launched last (gh_last_kernel): tests assert on the
// path taken, bench.py looks the kernel's instruction mix up in the shipped SASS under this name.
thread_local const char* g_last_kernel = "";
template <int ALGOS, bool STAGED, bool FMA, int TPS>
const char* lanes_name() {
    static const std::string n = "gh_hash_lanes<" + std::to_string(ALGOS) + ", " + (STAGED ? "1" : "0") + ", " + (FMA ? "1" : "0") +
                                 ", " + std::to_string(TPS) + ">";
    return n.c_str();
}

template <int ALGOS, bool FMA, int TPS>
cudaError_t launch_staged(const GhJob& job, const CUtensorMap& tmap, unsigned blocks, int wpb, cudaStream_t stream) {
    const size_t smem = static_cast<size_t>(wpb) * (warp_smem(TPS) + 8 * kStages);
    gh_hash_lanes<ALGOS, true, FMA, TPS><<<dim3(blocks), dim3(32 * wpb), smem, stream>>>(tmap, job);
    g_last_kernel = lanes_name<ALGOS, true, FMA, TPS>();
    return cudaGetLastError();
}

template <int ALGO, bool WITH_MD5>
cudaError_t launch_split(const GhJob& job_in, uint64_t count, const CUtensorMap& tmap, uint32_t inner_words,
                         cudaStream_t stream) {
    GhJob job = job_in;
    job.first_msg = 0;
    job.count = count;
    job.tma_inner = inner_words;
    job.one = 1;
    const uint64_t nblocks = (count + 31) / 32;
    if (nblocks > 0x7fffffffull) return cudaErrorInvalidValue;
    gh_hash_split<ALGO, WITH_MD5><<<dim3(static_cast<unsigned>(nblocks)), dim3(WITH_MD5 ? 96 : 64), split_smem<ALGO>(),
                                    stream>>>(tmap, job);
    g_last_kernel = ALGO == GH_SHA1 ? "gh_hash_split<1, 0>" : WITH_MD5 ? "gh_hash_split<4, 1>" : "gh_hash_split<4, 0>";
    return cudaGetLastError();
}

cudaError_t launch_trio(const GhJob& job_in, uint64_t count, int chains, const CUtensorMap& tmap, uint32_t inner_words,
                        cudaStream_t stream) {
    GhJob job = job_in;
    job.first_msg = 0;
    job.count = count;
    job.tma_inner = inner_words;
    job.one = 1;
    job.chains_per_cta = static_cast<uint32_t>(chains);
    const uint64_t n_mw = (count + 31) / 32;
    const uint64_t nblocks = (n_mw + chains - 1) / chains;
    if (nblocks > 0x7fffffffull) return cudaErrorInvalidValue;
    // one chain per CTA: rounds warp alone on its sub-partition, whole batches unrolled (+3.6 %); three
    // chains: the doubled loop body costs more in instruction fetch than it saves (-4.5 %), 2:1 schedule split
    const dim3 grid(static_cast<unsigned>(nblocks)), block(32 * trio_warps(chains));
    const int ratio = trio_ratio(chains);
    if (chains == 1 && ratio == 1) {
        gh_hash_trio<1, true><<<grid, block, TrioCfg<1>::smem(chains), stream>>>(tmap, job);
        g_last_kernel = "gh_hash_trio<1, 1>";
    } else if (ratio == 2) {
        gh_hash_trio<2, false><<<grid, block, TrioCfg<2>::smem(chains), stream>>>(tmap, job);
        g_last_kernel = "gh_hash_trio<2, 0>";
    } else {
        gh_hash_trio<1, false><<<grid, block, TrioCfg<1>::smem(chains), stream>>>(tmap, job);
        g_last_kernel = "gh_hash_trio<1, 0>";
    }
    return cudaGetLastError();
}

template <int ALGOS>
cudaError_t launch_range(const GhJob& job_in, uint64_t first, uint64_t count, bool staged, const CUtensorMap& tmap,
                         uint32_t inner_words, int wpb, int tps, cudaStream_t stream) {
    if (count == 0) return cudaSuccess;
    GhJob job = job_in;
    job.first_msg = first;
    job.count = count;
    job.tma_inner = inner_words;
    job.one = 1;
    const uint64_t warps = (count + 31) / 32;
    const uint64_t nblocks = (warps + wpb - 1) / wpb;
    if (nblocks > 0x7fffffffull) return cudaErrorInvalidValue;
    const unsigned blocks = static_cast<unsigned>(nblocks);
    bool fma = ALGOS == GH_SHA1 || ALGOS == GH_SHA256 || (ALGOS == (GH_MD5 | GH_SHA256) && wpb > 1);
    if (ALGOS == GH_MD5) fma = false;  // MD5 has no FMA variant
    if (fma_adds_override() >= 0 && ALGOS != GH_MD5) fma = fma_adds_override() != 0;
    if (staged) {
        if (fma) return tps == 4 ? launch_staged<ALGOS, true, 4>(job, tmap, blocks, wpb, stream)
                                 : launch_staged<ALGOS, true, 2>(job, tmap, blocks, wpb, stream);
        return tps == 4 ? launch_staged<ALGOS, false, 4>(job, tmap, blocks, wpb, stream)
                        : launch_staged<ALGOS, false, 2>(job, tmap, blocks, wpb, stream);
    }
    if (fma) gh_hash_lanes<ALGOS, false, true, 2><<<dim3(blocks), dim3(32 * wpb), 0, stream>>>(tmap, job);
    else gh_hash_lanes<ALGOS, false, false, 2><<<dim3(blocks), dim3(32 * wpb), 0, stream>>>(tmap, job);
    if (count > 1) g_last_kernel = fma ? lanes_name<ALGOS, false, true, 2>() : lanes_name<ALGOS, false, false, 2>();  // not the 1-message tail launch
    return cudaGetLastError();
}

template <int ALGOS>
cudaError_t launch_algo(const GhJob& job, int sm_count, cudaStream_t stream) {
    if (job.n_msgs == 0) return cudaSuccess;
    // bytes of an ordinary (not last) message present in this launch, and of the last one
    auto present = [&](uint64_t len) {
        if (job.prior >= len) return uint64_t(0);
        return len - job.prior < job.chunk_len ? len - job.prior : job.chunk_len;
    };
    const uint64_t common = present(job.msg_len), last = job.lens ? common : present(job.last_msg_len);
    const uint64_t inner_words = (common >> 6) * 16;
    const bool aligned = ((reinterpret_cast<uintptr_t>(job.data) | job.stride) & 15u) == 0 && job.stride < (1ull << 40);
    // The TMA rows all have the common extent; a shorter last message (BEP 3 tail piece) would make
    // its row run past the end of the caller's buffer, so it is hashed by the byte-granular kernel.
    const bool last_short = last < common;
    const uint64_t tma_rows = last_short ? job.n_msgs - 1 : job.n_msgs;
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    bool staged = aligned && inner_words > 0 && inner_words <= 0xffffffffull && tma_rows > 0 &&
                  tma_rows <= 0xffffffffull && make_tmap(job, tma_rows, inner_words, &tmap);
    // few messages: one warp per block so that warps spread over all SMs / sub-partitions, deep stages;
    // many messages: 4 warps per block (one per sub-partition), shallow stages so 13 warps fit an SM
    const uint64_t warps = (job.n_msgs + 31) / 32;
    const uint64_t sms = static_cast<uint64_t>(sm_count > 0 ? sm_count : 148);
    const int wpb = warps > sms * 8 ? kMaxWarpsPerBlock : 1;
    int tps = warps > sms * 6 ? 2 : 4;
    if (forced_tps() == 2 || forced_tps() == 4) tps = forced_tps();
    cudaError_t e;
    if constexpr (ALGOS == GH_SHA1) {
        const uint64_t tw = static_cast<uint64_t>(trio_max_warps_per_sm());
        // One chain per SM: rounds warp and both schedule warps each have a sub-partition to themselves.
        // Three per SM (more than two message-warps per SM, where the 1R+1S split below no longer has a
        // sub-partition for every warp): rounds and schedule warps share sub-partitions, still ahead of the
        // lane kernel. In between (two per SM) the 1R+1S split is faster -- two CTAs of two warps give every
        // warp its own sub-partition (measured: profiles/r02_ab_sha1_kernels.jsonl).
        const int chains = static_cast<int>(std::min<uint64_t>(kTrioMaxChains, (warps + sms - 1) / sms));
        if (staged && tw && warps <= sms * std::min<uint64_t>(tw, kTrioMaxChains) &&
            (chains != 2 || split_max_warps_per_sm() < 2)) {
            e = launch_trio(job, tma_rows, chains, tmap, static_cast<uint32_t>(inner_words), stream);
            if (e != cudaSuccess) return e;
            if (last_short) e = launch_range<ALGOS>(job, job.n_msgs - 1, 1, false, tmap, 0, 1, tps, stream);
            return e;
        }
    }
    if (staged && ALGOS != GH_MD5 && warps <= sms * static_cast<uint64_t>(split_max_warps_per_sm())) {
        // too few chains to fill the machine: spend the idle sub-partitions on the message schedule
        if constexpr (ALGOS == GH_SHA1) e = launch_split<GH_SHA1, false>(job, tma_rows, tmap, static_cast<uint32_t>(inner_words), stream);
        else if constexpr (ALGOS == GH_SHA256) e = launch_split<GH_SHA256, false>(job, tma_rows, tmap, static_cast<uint32_t>(inner_words), stream);
        else if constexpr (ALGOS == (GH_MD5 | GH_SHA256)) e = launch_split<GH_SHA256, true>(job, tma_rows, tmap, static_cast<uint32_t>(inner_words), stream);
        else e = cudaErrorInvalidValue;
        if (e != cudaSuccess) return e;
        if (last_short) e = launch_range<ALGOS>(job, job.n_msgs - 1, 1, false, tmap, 0, 1, tps, stream);
        return e;
    }
    if (staged) {
        e = launch_range<ALGOS>(job, 0, tma_rows, true, tmap, static_cast<uint32_t>(inner_words), wpb, tps, stream);
        if (e != cudaSuccess) return e;
        if (last_short) e = launch_range<ALGOS>(job, job.n_msgs - 1, 1, false, tmap, 0, 1, tps, stream);
        return e;
    }
    return launch_range<ALGOS>(job, 0, job.n_msgs, false, tmap, 0, wpb, tps, stream);
}

template <int ALGOS, bool FMA, int TPS>
cudaError_t init_one() {
    cudaError_t e = cudaFuncSetAttribute(gh_hash_lanes<ALGOS, true, FMA, TPS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kMaxWarpsPerBlock * (warp_smem(TPS) + 8 * kStages));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(gh_hash_lanes<ALGOS, true, FMA, TPS>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int ALGO, bool WITH_MD5>
cudaError_t init_split() {
    cudaError_t e = cudaFuncSetAttribute(gh_hash_split<ALGO, WITH_MD5>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         split_smem<ALGO>());
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(gh_hash_split<ALGO, WITH_MD5>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int RATIO, bool RBATCH>
cudaError_t init_trio() {
    cudaError_t e = cudaFuncSetAttribute(gh_hash_trio<RATIO, RBATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         TrioCfg<RATIO>::smem(kTrioMaxChains));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(gh_hash_trio<RATIO, RBATCH>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

template <int ALGOS>
cudaError_t init_algo() {
    cudaError_t e;
    if ((e = init_one<ALGOS, true, 2>()) != cudaSuccess) return e;
    if ((e = init_one<ALGOS, true, 4>()) != cudaSuccess) return e;
    if ((e = init_one<ALGOS, false, 2>()) != cudaSuccess) return e;
    return init_one<ALGOS, false, 4>();
}

}  // namespace

cudaError_t gh_kernels_init_device() {
    cudaError_t e;
    if ((e = init_algo<GH_SHA1>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_MD5>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_SHA256>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_MD5 | GH_SHA256>()) != cudaSuccess) return e;
    if ((e = init_split<GH_SHA1, false>()) != cudaSuccess) return e;
    if ((e = init_split<GH_SHA256, false>()) != cudaSuccess) return e;
    if ((e = init_split<GH_SHA256, true>()) != cudaSuccess) return e;
    if ((e = init_trio<1, true>()) != cudaSuccess) return e;
    if ((e = init_trio<1, false>()) != cudaSuccess) return e;
    if ((e = init_trio<2, false>()) != cudaSuccess) return e;
    return encode_tiled() ? cudaSuccess : cudaErrorNotSupported;
}

cudaError_t gh_launch_hash(int algos, const GhJob& job, int sm_count, cudaStream_t stream) {
    switch (algos) {
        case GH_SHA1: return launch_algo<GH_SHA1>(job, sm_count, stream);
        case GH_MD5: return launch_algo<GH_MD5>(job, sm_count, stream);
        case GH_SHA256: return launch_algo<GH_SHA256>(job, sm_count, stream);
        case GH_MD5 | GH_SHA256: return launch_algo<GH_MD5 | GH_SHA256>(job, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}

extern "C" const char* gh_last_kernel(void) { return g_last_kernel; }

cudaError_t gh_launch_generate(uint8_t* d_out, uint64_t seed, uint64_t offset, uint64_t len, cudaStream_t stream) {
    if (len == 0) return cudaSuccess;
    const uint64_t n_words = ((offset + len + 7) >> 3) - (offset >> 3);
    uint64_t blocks = (n_words + 255) / 256;
    if (blocks > 148ull * 64) blocks = 148ull * 64;
    gh_generate<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(d_out, seed, offset, len);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------- fixed-scratch entry point
// gh_hash_update_device_noalloc (include/gpuhash.h): the resumable hasher with its midstates in a
// statically allocated device array instead of a caller- or library-allocated buffer. The array is
// part of the library image: every GPU that loads libgpuhash.so holds these
// GH_NOALLOC_MAX_MSGS * GH_STATE_WORDS * 4 bytes (1 088 KiB) until the process exits.
#include <mutex>

static __device__ __align__(256) uint32_t g_noalloc_state[GH_NOALLOC_MAX_MSGS * GH_STATE_WORDS];

namespace {
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
uint32_t* g_noalloc_ptr[kNoallocMaxDevices];  // device address of g_noalloc_state per CUDA ordinal; null = not set up
int g_noalloc_sms[kNoallocMaxDevices];
}  // namespace

extern "C" int gh_hash_update_device_noalloc(int cuda_device, int algos, const void* d_data, uint64_t stride,
                                             uint64_t chunk_len, uint64_t prior_bytes, uint64_t n_msgs,
                                             uint64_t msg_len, uint64_t last_msg_len, uint8_t* d_sha1,
                                             uint8_t* d_md5, uint8_t* d_sha256, void* cuda_stream) {
    if (cuda_device < 0 || cuda_device >= kNoallocMaxDevices) return GH_ENODEVICE;
    if (!(algos & (GH_SHA1 | GH_MD5 | GH_SHA256)) || (algos & ~(GH_SHA1 | GH_MD5 | GH_SHA256))) return GH_EINVAL;
    if (n_msgs == 0) return GH_OK;
    if (n_msgs > GH_NOALLOC_MAX_MSGS) return GH_ERANGE;  // the static array holds no more midstates than that
    if (!d_data) return GH_EINVAL;
    if (prior_bytes & 63) return GH_EINVAL;
    if ((chunk_len & 63) && prior_bytes + chunk_len < msg_len) return GH_EINVAL;
    if (last_msg_len > msg_len) return GH_EINVAL;
    if (cudaSetDevice(cuda_device) != cudaSuccess) {
        cudaGetLastError();
        return GH_ENODEVICE;
    }
    uint32_t* state = nullptr;
    int sms = 0;
    {
        // Once per device, with that device current: the array's name is only a host-side handle, the
        // kernel needs the address the symbol has on THIS device.
        std::lock_guard<std::mutex> g(g_noalloc_mu);
        if (!g_noalloc_ptr[cuda_device]) {
            if (gh_kernels_init_device() != cudaSuccess) {
                cudaGetLastError();
                return GH_ECUDA;
            }
            void* p = nullptr;
            if (cudaGetSymbolAddress(&p, g_noalloc_state) != cudaSuccess || !p) {
                cudaGetLastError();
                return GH_ECUDA;
            }
            int n = 0;
            if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, cuda_device) != cudaSuccess) {
                cudaGetLastError();
                return GH_ECUDA;
            }
            g_noalloc_sms[cuda_device] = n;
            g_noalloc_ptr[cuda_device] = static_cast<uint32_t*>(p);
        }
        state = g_noalloc_ptr[cuda_device];
        sms = g_noalloc_sms[cuda_device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);  // null = the device's default stream
    // On a failed launch: clear the runtime's error so the next CUDA call of this thread does not
    // inherit it, and -- when the call owns the synchronisation (no caller stream) -- wait for what
    // was already queued, so nothing of this call is still touching the caller's buffers on return.
    auto failed = [&]() {
        cudaGetLastError();
        if (!cuda_stream) {
            cudaStreamSynchronize(st);
            cudaGetLastError();
        }
        return GH_ECUDA;
    };
    if (prior_bytes == 0) {
        // A new batch starts: drop whatever the previous batch left in the part of the array this one uses.
        if (cudaMemsetAsync(state, 0, n_msgs * GH_STATE_WORDS * sizeof(uint32_t), st) != cudaSuccess) return failed();
    }
    GhJob j{};
    j.data = static_cast<const uint8_t*>(d_data);
    j.stride = stride;
    j.chunk_len = chunk_len;
    j.prior = prior_bytes;
    j.n_msgs = n_msgs;
    j.msg_len = msg_len;
    j.last_msg_len = last_msg_len;
    j.state = state;
    j.sha1 = d_sha1;
    j.md5 = d_md5;
    j.sha256 = d_sha256;
    if (algos & GH_SHA1)
        if (gh_launch_hash(GH_SHA1, j, sms, st) != cudaSuccess) return failed();
    if (algos & (GH_MD5 | GH_SHA256))
        if (gh_launch_hash(algos & (GH_MD5 | GH_SHA256), j, sms, st) != cudaSuccess) return failed();
    if (!cuda_stream && cudaStreamSynchronize(st) != cudaSuccess) {
        cudaGetLastError();
        return GH_ECUDA;
    }
    return GH_OK;
}
