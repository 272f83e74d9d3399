/* gh_openssl.c -- TEST / BENCH INFRASTRUCTURE ONLY: the second CPU baseline.
 *
 * OpenSSL's EVP digests (SHA-NI for SHA-1 / SHA-256 where the CPU has it, assembly MD5) over the same
 * batched layout the oracle port hashes, one worker thread per host thread, messages handed out by
 * an atomic counter. It is the UPPER bound of the two CPU figures reported beside the GPU numbers
 * (SURVEY.md section 8d): the reference's Go 1.14 toolchain has no SHA-NI path, so Go's crypto/sha1
 * sits between this and the scalar port (oracle/gh_oracle.c, the lower bound). Also an independent
 * implementation to cross-check the oracle against. Never linked into libgpuhash.so. */
#include <openssl/evp.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdlib.h>

typedef struct {
    int algs; /* bit 0 SHA-1, bit 1 MD5, bit 2 SHA-256 (the library's GH_* mask) */
    const uint8_t* data;
    uint64_t total, msg_len, n;
    uint8_t *sha1, *md5, *sha256;
    atomic_ullong next;
} job_t;

static void one(EVP_MD_CTX* c, const EVP_MD* md, const uint8_t* p, uint64_t len, uint8_t* out) {
    unsigned int dl;
    EVP_DigestInit_ex(c, md, NULL);
    EVP_DigestUpdate(c, p, len);
    EVP_DigestFinal_ex(c, out, &dl);
}

static void* worker(void* p) {
    job_t* j = (job_t*)p;
    EVP_MD_CTX* c = EVP_MD_CTX_new();
    for (;;) {
        const uint64_t i = atomic_fetch_add(&j->next, 1);
        if (i >= j->n) break;
        const uint64_t off = i * j->msg_len;
        const uint64_t len = off + j->msg_len <= j->total ? j->msg_len : j->total - off;
        if (j->algs & 1) one(c, EVP_sha1(), j->data + off, len, j->sha1 + 20 * i);
        if (j->algs & 2) one(c, EVP_md5(), j->data + off, len, j->md5 + 16 * i);
        if (j->algs & 4) one(c, EVP_sha256(), j->data + off, len, j->sha256 + 32 * i);
    }
    EVP_MD_CTX_free(c);
    return NULL;
}

/* message k = data[k*msg_len, min((k+1)*msg_len, total)); outputs of unselected algorithms may be NULL */
int gho_openssl_hash_msgs_mt(int algs, const uint8_t* data, uint64_t total, uint64_t msg_len, uint8_t* sha1, uint8_t* md5,
                             uint8_t* sha256, int nthreads) {
    if (!msg_len || !total) return 0;
    job_t j = {algs, data, total, msg_len, (total + msg_len - 1) / msg_len, sha1, md5, sha256, 0};
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, worker, &j);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    return 0;
}
