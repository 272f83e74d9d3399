/*
 * gh_oracle.h -- CPU restatement of the batched content-hashing path. TEST INFRASTRUCTURE ONLY.
 *
 * This file and gh_oracle.c are the checker for the CUDA library in downloader-go_b200/csrc.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it. The product library (libgpuhash.so) never links, loads or calls anything in oracle/.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference holds no hashing code, no digest fixture and
 * no known-answer test (its only test, internal/process/process_test.go:13-62, filters zero-byte
 * files). The arithmetic the reference executes lives in un-vendored modules:
 *   github.com/anacrolix/torrent v1.15.2   (go.mod:6)   -> Go crypto/sha1   (piece verification)
 *   github.com/minio/minio-go/v6 v6.0.55   (go.mod:11)  -> Go crypto/md5, minio/sha256-simd v0.1.1
 * reached from internal/downloader/torrent/torrent.go:41,79,106 and internal/uploader/uploader.go:89.
 * Those modules implement the public standards, which is what is restated here:
 *   SHA-1   FIPS 180-4 section 6.1      SHA-256  FIPS 180-4 section 6.2      MD5  RFC 1321
 *   piece layout  BEP 3 (stream = files concatenated; piece i = stream[i*L, min((i+1)L, total)))
 *   S3 multipart ETag = hex(md5(concat(binary part md5s))) + "-" + N
 * The oracle is pinned to the standards' known-answer vectors and differentially to OpenSSL
 * (Python hashlib) in tests/test_oracle.py; see tests/golden/.
 */
#ifndef GH_ORACLE_H
#define GH_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { GHO_SHA1 = 1, GHO_MD5 = 2, GHO_SHA256 = 4 };

/* Streaming context with an exported midstate (state[] + total byte count + buffered tail). */
typedef struct {
    uint32_t state[8];   /* 5 words used by SHA-1, 4 by MD5, 8 by SHA-256 */
    uint64_t total;      /* bytes absorbed so far */
    uint8_t  buf[64];    /* pending partial block */
    uint32_t buflen;
    int      alg;
} gho_ctx;

int    gho_digest_len(int alg);                       /* 20 / 16 / 32, or -1 */
void   gho_init  (gho_ctx* c, int alg);
void   gho_update(gho_ctx* c, const void* data, size_t len);
void   gho_final (gho_ctx* c, uint8_t* out);          /* writes gho_digest_len(alg) bytes */
void   gho_hash  (int alg, const void* data, size_t len, uint8_t* out);

/* Synthetic stream of SURVEY.md Appendix C: byte k of stream(seed). */
uint64_t gho_splitmix64(uint64_t x);                  /* the finaliser alone (z -> mix(z)) */
uint64_t gho_gen_word(uint64_t seed, uint64_t i);     /* word i of the stream */
void     gho_gen_fill(uint64_t seed, uint64_t offset, uint64_t len, uint8_t* out);

/* Batched forms. message i = data[i*msg_len, min((i+1)*msg_len, total)). */
uint64_t gho_count(uint64_t total, uint64_t msg_len); /* ceil(total/msg_len); 0 for total==0 */
void gho_hash_msgs(int alg, const uint8_t* data, uint64_t total, uint64_t msg_len, uint8_t* out);
/* Same, spread over nthreads pthreads by an atomic counter (used only for the timed CPU baseline). */
void gho_hash_msgs_mt(int alg, const uint8_t* data, uint64_t total, uint64_t msg_len, uint8_t* out,
                      int nthreads);
/* Compare 20-byte SHA-1 of each piece with expected; bit i of bitmap = piece i passed. */
uint64_t gho_verify_pieces(const uint8_t* data, uint64_t total, uint64_t piece_len,
                           const uint8_t* expected, uint64_t* bitmap, int nthreads);

/* SURVEY.md section 8(d) config 5 fault injection: returns 1 and the (byte, bit) to flip when piece i
 * of length len is selected by the 1 % rule h = splitmix64(0xBAD<<32 | i), h % 100 == 0. */
int gho_cfg5_corruption(uint64_t piece_index, uint64_t piece_len_i, uint64_t* byte_off, int* bit);

/* S3 multipart ETag from n binary part MD5s; out needs 32 + 1 + 20 + 1 bytes. */
void gho_s3_etag(const uint8_t* part_md5s, uint64_t n, char* out);

#ifdef __cplusplus
}
#endif
#endif
