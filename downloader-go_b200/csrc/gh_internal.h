// gh_internal.h -- launch-side interface between the C ABI (gh_api.cu) and the kernels (gh_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef GH_STATE_WORDS
#define GH_STATE_WORDS 17
#endif
/* rows per message: MD5 0..3, SHA-256 4..11, SHA-1 12..16 (disjoint: any mix of algorithms may share a batch) */

// One launch of the lane-per-message hasher. Message m's bytes for this launch start at
// data + m*stride; `prior` bytes of every message were absorbed by earlier launches.
struct GhJob {
    const uint8_t* data;
    uint64_t stride;
    uint64_t chunk_len;     // bytes of each message present in this launch (the tail message may have fewer)
    uint64_t prior;         // multiple of 64
    uint64_t n_msgs;
    uint64_t msg_len;       // full length of every message but the last
    uint64_t last_msg_len;  // full length of message n_msgs-1
    const uint64_t* lens;   // optional: full length of every message (<= msg_len); then msg_len/last_msg_len only
                            // bound the loop, and every row must be addressable for its chunk (ring slots are)
    uint32_t* state;        // [GH_STATE_WORDS][n_msgs] midstates (needed only when a message spans launches)
    uint8_t* sha1;          // digests out (20 / 16 / 32 bytes per message), may be null
    uint8_t* md5;
    uint8_t* sha256;
    const uint8_t* expected;  // 20-byte SHA-1 per message or null
    uint32_t* ok_words;       // atomicOr'ed: bit ((ok_base+m) & 31) of word ((ok_base+m) >> 5); caller zeroes first
    uint64_t ok_base;         // bitmap index of message 0 of the batch
    // filled in by gh_launch_hash: the slice [first_msg, first_msg+count) this grid covers and the
    // row extent (uint32 words) of the TMA tensor
    uint64_t first_msg;
    uint64_t count;
    uint32_t tma_inner;
    uint32_t chains_per_cta;  // trio kernel only: chains (message-warps) per CTA, 1..3
    uint32_t one;  // always 1; a kernel parameter so that ptxas keeps x*one+y as an IMAD (gh_hash.cuh addf)
};

// Launches the hasher for `algos` (GH_SHA1, GH_MD5, GH_SHA256 or GH_MD5|GH_SHA256) on `stream`.
// Picks the staged (TMA bulk copy -> shared memory) kernel when data/stride are 16-byte aligned,
// else the byte-granular kernel. Returns the cudaError of the launch.
cudaError_t gh_launch_hash(int algos, const GhJob& job, int sm_count, cudaStream_t stream);

cudaError_t gh_launch_generate(uint8_t* d_out, uint64_t seed, uint64_t offset, uint64_t len,
                               cudaStream_t stream);

// One-time per-device attribute set-up (max dynamic shared memory for the staged kernels).
cudaError_t gh_kernels_init_device();
