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
#include <string.h>

#include "../../include/gpuhash.h"
#include "gh_hash.cuh"

namespace {

using namespace ghd;

constexpr int kTileRowBytes = 128;                       // inner box extent = the swizzle span
constexpr int kTileBytes = 32 * kTileRowBytes;           // one TMA box: 32 messages x 128 B
constexpr int kTilesPerStage = 2;                        // 256 B (4 blocks) of every message per stage
constexpr int kChunk = kTilesPerStage * kTileRowBytes;
constexpr int kStages = 2;
constexpr int kWarpTiles = kStages * kTilesPerStage;
constexpr int kWarpSmem = kWarpTiles * kTileBytes;       // tiles only; mbarriers live after all tiles
constexpr int kMaxWarpsPerBlock = 4;

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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
template <int ALGOS>
struct Lane {
    uint32_t s1[5];
    uint32_t m5[4];
    uint32_t s2[8];

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
            sha256_compress(s2, w);
        }
        if constexpr (ALGOS & GH_SHA1) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; i++) w[i] = bswap(r[i]);
            sha1_compress(s1, w);
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
            sha256_compress(s2, w);
        }
        if constexpr (ALGOS & GH_SHA1) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 14; i++) w[i] = bswap(r[i]);
            w[14] = hi; w[15] = lo;
            sha1_compress(s1, w);
        }
    }
};

// Padding (FIPS 180-4 5.1.1 / RFC 1321 3.1): tail = the rem (< 64) bytes after the last full block.
template <int ALGOS>
__device__ __noinline__ void finalize(Lane<ALGOS>& L, const uint8_t* tail, uint32_t rem, uint64_t total_len) {
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

template <int ALGOS>
__device__ __forceinline__ bool emit(const Lane<ALGOS>& L, const GhJob& job, uint64_t m) {
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

// ---------------------------------------------------------------------------- the hasher
template <int ALGOS, bool STAGED>
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
    const uint64_t total_m = active ? (m == job.n_msgs - 1 ? job.last_msg_len : job.msg_len) : 0;
    const bool done_before = !active || (job.prior > 0 && job.prior >= total_m);
    const uint64_t remain = done_before ? 0 : total_m - job.prior;
    const bool ends_now = !done_before && remain <= job.chunk_len;
    const uint64_t avail = remain < job.chunk_len ? remain : job.chunk_len;
    const uint32_t nfull = static_cast<uint32_t>(avail >> 6);
    const uint32_t rem = ends_now ? static_cast<uint32_t>(avail & 63u) : 0u;
    const uint8_t* src = job.data + m * job.stride;

    Lane<ALGOS> L;
    if (!done_before) {
        if (job.prior == 0) L.init();
        else L.load(job.state, job.n_msgs, m);
    }

    if constexpr (STAGED) {
        uint8_t* tiles = smem + warp * kWarpSmem;
        const uint32_t tiles_u32 = smem_u32(tiles);
        const uint32_t bar0 = smem_u32(smem + wpb * kWarpSmem) + warp * (8 * kStages);
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < kStages; s++) mbar_init(bar0 + 8 * s, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();

        constexpr uint32_t kBlocksPerStage = kChunk / 64;
        const uint32_t my_stages = (nfull + kBlocksPerStage - 1) / kBlocksPerStage;
        const uint32_t n_stages = __reduce_max_sync(0xffffffffu, my_stages);
        const uint32_t row0 = static_cast<uint32_t>(m_warp);  // tensor row of lane 0's message

        // one lane arms the stage barrier with the bytes of the boxes it is about to request
        auto issue = [&](uint32_t s) {
            const uint32_t b = s % kStages;
            const uint32_t bar = bar0 + 8 * b;
            uint32_t ntile = 0;
#pragma unroll
            for (uint32_t t = 0; t < kTilesPerStage; t++)
                if ((s * kChunk + t * kTileRowBytes) / 4 < job.tma_inner) ntile++;
            mbar_arrive_expect_tx(bar, ntile * kTileBytes);
#pragma unroll
            for (uint32_t t = 0; t < kTilesPerStage; t++) {
                const uint32_t x = (s * kChunk + t * kTileRowBytes) / 4;
                if (x < job.tma_inner)
                    tma_load_box(tiles_u32 + (b * kTilesPerStage + t) * kTileBytes, &tmap, x, row0, bar);
            }
        };

        if (lane == 0)
            for (uint32_t s = 0; s < kStages && s < n_stages; s++) issue(s);

        // 128B swizzle: 16-byte chunk c of row r sits at chunk (c ^ (r & 7)) of that row
        const uint32_t row_off = lane * kTileRowBytes;
        const uint32_t sw = lane & 7u;
        for (uint32_t s = 0; s < n_stages; s++) {
            const uint32_t b = s % kStages;
            mbar_wait(bar0 + 8 * b, (s / kStages) & 1u);
            const uint32_t first = s * kBlocksPerStage;
            const uint32_t nb = nfull > first ? min(kBlocksPerStage, nfull - first) : 0u;
#pragma unroll 1
            for (uint32_t j = 0; j < nb; j++) {
                const uint8_t* tile = tiles + (b * kTilesPerStage + (j >> 1)) * kTileBytes + row_off;
                const uint32_t c0 = (j & 1u) * 4u;
                const uint4 q0 = *reinterpret_cast<const uint4*>(tile + (((c0 + 0) ^ sw) << 4));
                const uint4 q1 = *reinterpret_cast<const uint4*>(tile + (((c0 + 1) ^ sw) << 4));
                const uint4 q2 = *reinterpret_cast<const uint4*>(tile + (((c0 + 2) ^ sw) << 4));
                const uint4 q3 = *reinterpret_cast<const uint4*>(tile + (((c0 + 3) ^ sw) << 4));
                const uint32_t r[16] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w,
                                        q2.x, q2.y, q2.z, q2.w, q3.x, q3.y, q3.z, q3.w};
                L.absorb(r);
            }
            // every lane has consumed its row of buffer b (all LDS results were used above): the
            // buffer can go back to the TMA unit for stage s + kStages
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
            finalize<ALGOS>(L, src + static_cast<uint64_t>(nfull) * 64, rem, total_m);
            ok = emit<ALGOS>(L, job, m);
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

template <int ALGOS>
cudaError_t launch_range(const GhJob& job_in, uint64_t first, uint64_t count, bool staged, const CUtensorMap& tmap,
                         uint32_t inner_words, int wpb, cudaStream_t stream) {
    if (count == 0) return cudaSuccess;
    GhJob job = job_in;
    job.first_msg = first;
    job.count = count;
    job.tma_inner = inner_words;
    const uint64_t warps = (count + 31) / 32;
    const uint64_t blocks = (warps + wpb - 1) / wpb;
    if (blocks > 0x7fffffffull) return cudaErrorInvalidValue;
    const dim3 grid(static_cast<unsigned>(blocks)), block(32 * wpb);
    if (staged) {
        const size_t smem = static_cast<size_t>(wpb) * (kWarpSmem + 8 * kStages);
        gh_hash_lanes<ALGOS, true><<<grid, block, smem, stream>>>(tmap, job);
    } else {
        gh_hash_lanes<ALGOS, false><<<grid, block, 0, stream>>>(tmap, job);
    }
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
    const uint64_t common = present(job.msg_len), last = present(job.last_msg_len);
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
    // few messages: one warp per block so that warps spread over all SMs / sub-partitions;
    // many messages: 4 warps per block (one per sub-partition) to stay under the 32 blocks/SM limit
    const uint64_t warps = (job.n_msgs + 31) / 32;
    const int wpb = warps > static_cast<uint64_t>(sm_count > 0 ? sm_count : 148) * 8 ? kMaxWarpsPerBlock : 1;
    cudaError_t e;
    if (staged) {
        e = launch_range<ALGOS>(job, 0, tma_rows, true, tmap, static_cast<uint32_t>(inner_words), wpb, stream);
        if (e != cudaSuccess) return e;
        if (last_short) e = launch_range<ALGOS>(job, job.n_msgs - 1, 1, false, tmap, 0, 1, stream);
        return e;
    }
    return launch_range<ALGOS>(job, 0, job.n_msgs, false, tmap, 0, wpb, stream);
}

template <int ALGOS>
cudaError_t init_algo() {
    cudaError_t e = cudaFuncSetAttribute(gh_hash_lanes<ALGOS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kMaxWarpsPerBlock * (kWarpSmem + 8 * kStages));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(gh_hash_lanes<ALGOS, true>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                cudaSharedmemCarveoutMaxShared);
}

}  // namespace

cudaError_t gh_kernels_init_device() {
    cudaError_t e;
    if ((e = init_algo<GH_SHA1>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_MD5>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_SHA256>()) != cudaSuccess) return e;
    if ((e = init_algo<GH_MD5 | GH_SHA256>()) != cudaSuccess) return e;
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
            if (gh_kernels_init_device() != cudaSuccess) return GH_ECUDA;
            void* p = nullptr;
            if (cudaGetSymbolAddress(&p, g_noalloc_state) != cudaSuccess || !p) {
                cudaGetLastError();
                return GH_ECUDA;
            }
            int n = 0;
            if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, cuda_device) != cudaSuccess) return GH_ECUDA;
            g_noalloc_sms[cuda_device] = n;
            g_noalloc_ptr[cuda_device] = static_cast<uint32_t*>(p);
        }
        state = g_noalloc_ptr[cuda_device];
        sms = g_noalloc_sms[cuda_device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);  // null = the device's default stream
    if (prior_bytes == 0) {
        // A new batch starts: drop whatever the previous batch left in the part of the array this one uses.
        if (cudaMemsetAsync(state, 0, n_msgs * GH_STATE_WORDS * sizeof(uint32_t), st) != cudaSuccess) return GH_ECUDA;
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
        if (gh_launch_hash(GH_SHA1, j, sms, st) != cudaSuccess) return GH_ECUDA;
    if (algos & (GH_MD5 | GH_SHA256))
        if (gh_launch_hash(algos & (GH_MD5 | GH_SHA256), j, sms, st) != cudaSuccess) return GH_ECUDA;
    if (!cuda_stream && cudaStreamSynchronize(st) != cudaSuccess) return GH_ECUDA;
    return GH_OK;
}
