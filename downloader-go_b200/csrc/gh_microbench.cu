// gh_microbench.cu -- the two rooflines SURVEY.md section 7 step 2 asks for and MEASURED_PEAKS.json
// lacks: (i) issue rates of the integer instructions the hash kernels are made of (per SM per clock,
// at several warp counts), (ii) host->device link bandwidth for contiguous and strided (2-D) pinned
// copies. Standalone binary; prints one JSON object. Not part of libgpuhash.so.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "gh_hash.cuh"

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e_ = (x);                                                         \
        if (e_ != cudaSuccess) {                                                      \
            fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                  \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

enum { OP_IADD3, OP_LOP3, OP_SHF, OP_PRMT, OP_IMAD, OP_LOP3_IMAD, OP_ALU_MIX, OP_IMADHI, OP_LEAHI, OP_ROT_FMA, OP_SHA1, OP_MD5, OP_SHA256, OP_SHA1_FMA, OP_SHA256_FMA, OP_COUNT };
static const char* kNames[OP_COUNT] = {"iadd3", "lop3", "shf", "prmt", "imad", "lop3+imad", "shf+lop3+iadd3", "imad.hi", "rotl+add (lea.hi)", "rotr via imad+imad.hi",
                                       "sha1_block", "md5_block", "sha256_block", "sha1_block_fma_adds", "sha256_block_fma_adds"};

template <int OP>
__device__ __forceinline__ void body(uint32_t (&x)[8], uint32_t a, uint32_t b) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
        if (OP == OP_IADD3) asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(x[i]) : "r"(a), "r"(b));
        if (OP == OP_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (OP == OP_SHF) asm volatile("shf.l.wrap.b32 %0, %0, %0, 5;" : "+r"(x[i]));
        if (OP == OP_PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x0123;" : "+r"(x[i]) : "r"(a));
        if (OP == OP_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (OP == OP_LOP3_IMAD) {
            if (i & 1) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
            else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
        }
        if (OP == OP_IMADHI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (OP == OP_LEAHI) asm volatile("{.reg .u32 t; shf.l.wrap.b32 t, %0, %0, 5; add.u32 %0, t, %1;}" : "+r"(x[i]) : "r"(a));
        if (OP == OP_ROT_FMA) asm volatile("{.reg .u32 t; mul.lo.u32 t, %0, %1; mad.hi.u32 %0, %0, %1, t;}" : "+r"(x[i]) : "r"(a));
        if (OP == OP_ALU_MIX) {
            if (i % 3 == 0) asm volatile("shf.l.wrap.b32 %0, %0, %0, 5;" : "+r"(x[i]));
            else if (i % 3 == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
            else asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(x[i]) : "r"(a), "r"(b));
        }
    }
}

// ops per thread per iteration: 8 chains x 8 unrolled
template <int OP>
__global__ void k_ops(uint32_t* out, uint32_t a, uint32_t b, int iters, long long* cycles) {
    uint32_t x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 8 + i + a;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) body<OP>(x, a, b);
    }
    const long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// whole compression functions on register-resident data: cycles per 64-byte block per warp
template <int OP>
__global__ void k_hash(uint32_t* out, uint32_t a, int iters, long long* cycles, uint32_t one, uint32_t lane_mask) {
    if (!((lane_mask >> (threadIdx.x & 31)) & 1)) return;  // partially filled warps: does the pipe skip idle halves?
    uint32_t w[16];
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = threadIdx.x * 16 + i + a;
    uint32_t s1[5], m5[4], s2[8];
    ghd::sha1_init(s1);
    ghd::md5_init(m5);
    ghd::sha256_init(s2);
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        uint32_t v[16];
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = w[i] ^ it;
        if (OP == OP_SHA1) ghd::sha1_compress(s1, v);
        if (OP == OP_MD5) ghd::md5_compress(m5, v);
        if (OP == OP_SHA256) ghd::sha256_compress(s2, v);
        if (OP == OP_SHA1_FMA) ghd::sha1_compress<true>(s1, v, one);
        if (OP == OP_SHA256_FMA) ghd::sha256_compress<true>(s2, v, one);
    }
    const long long t1 = clock64();
    uint32_t s = 0;
    for (int i = 0; i < 5; i++) s ^= s1[i];
    for (int i = 0; i < 4; i++) s ^= m5[i];
    for (int i = 0; i < 8; i++) s ^= s2[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static uint32_t g_lane_mask = 0xffffffffu;
template <int OP>
static double run_ops(int sms, int threads, uint32_t* d_out, long long* d_cyc, int iters) {
    if (OP >= OP_SHA1) k_hash<OP><<<sms, threads>>>(d_out, 3, iters, d_cyc, 1, g_lane_mask);
    else k_ops<OP><<<sms, threads>>>(d_out, 0x9e3779b9u, 0x7f4a7c15u, iters, d_cyc);
    CK(cudaDeviceSynchronize());
    std::vector<long long> cyc(sms);
    CK(cudaMemcpy(cyc.data(), d_cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost));
    long long mx = 0;
    for (auto c : cyc) mx = c > mx ? c : mx;
    return static_cast<double>(mx);
}

template <int OP>
static void report(int sms, uint32_t* d_out, long long* d_cyc, bool last) {
    printf("  \"%s\": {", kNames[OP]);
    const int tcounts[5] = {32, 128, 256, 512, 1024};
    for (int ti = 0; ti < 5; ti++) {
        const int threads = tcounts[ti];
        const int iters = OP >= OP_SHA1 ? 200 : 400;
        run_ops<OP>(sms, threads, d_out, d_cyc, 10);
        const double cyc = run_ops<OP>(sms, threads, d_out, d_cyc, iters);
        if (OP >= OP_SHA1) {
            // cycles per 64-byte block per warp, and blocks per clock per SM
            printf("\"%d\": {\"cycles_per_block\": %.1f, \"blocks_per_clk_per_sm\": %.5f}%s", threads, cyc / iters,
                   (threads / 32.0) * iters / cyc, ti < 4 ? ", " : "");
        } else {
            const double lane_ops = static_cast<double>(threads) * iters * 64.0 * (OP == OP_IADD3 ? 1 : 1);
            printf("\"%d\": %.2f%s", threads, lane_ops / cyc, ti < 4 ? ", " : "");
        }
    }
    printf("}%s\n", last ? "" : ",");
}

int main(int argc, char** argv) {
    int dev = argc > 1 ? atoi(argv[1]) : 0;
    CK(cudaSetDevice(dev));
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, dev));
    const int sms = p.multiProcessorCount;
    uint32_t* d_out;
    long long* d_cyc;
    CK(cudaMalloc(&d_out, sizeof(uint32_t) * sms * 1024));
    CK(cudaMalloc(&d_cyc, sizeof(long long) * sms));
    printf("{\n \"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d,\n", p.name, sms, p.clockRate);
    printf(" \"int32_lane_ops_per_clk_per_sm_by_threads_per_sm\": {\n");
    printf("  \"_note\": \"iadd3 = two dependent 2-input adds fused by ptxas or not: check SASS; every figure is thread-instructions/clk/SM\",\n");
    report<OP_IADD3>(sms, d_out, d_cyc, false);
    report<OP_LOP3>(sms, d_out, d_cyc, false);
    report<OP_SHF>(sms, d_out, d_cyc, false);
    report<OP_PRMT>(sms, d_out, d_cyc, false);
    report<OP_IMAD>(sms, d_out, d_cyc, false);
    report<OP_LOP3_IMAD>(sms, d_out, d_cyc, false);
    report<OP_ALU_MIX>(sms, d_out, d_cyc, false);
    report<OP_IMADHI>(sms, d_out, d_cyc, false);
    report<OP_LEAHI>(sms, d_out, d_cyc, false);
    report<OP_ROT_FMA>(sms, d_out, d_cyc, true);
    printf(" },\n \"compress_by_threads_per_sm\": {\n");
    report<OP_SHA1>(sms, d_out, d_cyc, false);
    report<OP_MD5>(sms, d_out, d_cyc, false);
    report<OP_SHA256>(sms, d_out, d_cyc, false);
    report<OP_SHA1_FMA>(sms, d_out, d_cyc, false);
    report<OP_SHA256_FMA>(sms, d_out, d_cyc, true);
    printf(" },\n \"compress_16_of_32_lanes_active\": {\n");
    g_lane_mask = 0x0000ffffu;
    report<OP_SHA1>(sms, d_out, d_cyc, false);
    report<OP_SHA256>(sms, d_out, d_cyc, true);
    printf(" },\n \"compress_8_of_32_lanes_active\": {\n");
    g_lane_mask = 0x000000ffu;
    report<OP_SHA1>(sms, d_out, d_cyc, true);
    g_lane_mask = 0xffffffffu;
    printf(" },\n");

    // ---- host -> device link
    const size_t bytes = 1ull << 30;
    uint8_t *h, *d;
    CK(cudaHostAlloc(reinterpret_cast<void**>(&h), bytes, cudaHostAllocPortable));
    memset(h, 1, bytes);
    CK(cudaMalloc(reinterpret_cast<void**>(&d), bytes));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    auto timed = [&](auto fn) {
        fn();
        CK(cudaDeviceSynchronize());
        float best = 1e30f;
        for (int r = 0; r < 3; r++) {
            CK(cudaEventRecord(e0));
            fn();
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            best = ms < best ? ms : best;
        }
        return best;
    };
    const float ms1d = timed([&] { CK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice)); });
    const float msd2h = timed([&] { CK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost)); });
    printf(" \"h2d_gbs_1d_1GiB\": %.2f, \"d2h_gbs_1d_1GiB\": %.2f,\n", bytes / ms1d / 1e6, bytes / msd2h / 1e6);
    // strided: rows of `w` bytes at pitch 4 MiB (what a chunk-interleaved ring window of 4 MiB pieces reads)
    const size_t pitch = 4ull << 20;
    const size_t widths[4] = {4096, 16384, 65536, 262144};
    printf(" \"h2d_gbs_2d_pitch4MiB_by_row_bytes\": {");
    for (int i = 0; i < 4; i++) {
        const size_t w = widths[i], rows = bytes / pitch;
        const float ms = timed([&] { CK(cudaMemcpy2DAsync(d, w, h, pitch, w, rows, cudaMemcpyHostToDevice)); });
        printf("\"%zu\": %.2f%s", w, w * rows / ms / 1e6, i < 3 ? ", " : "");
    }
    printf("}\n}\n");
    return 0;
}
