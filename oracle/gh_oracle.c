/*
 * gh_oracle.c -- scalar C restatement of SHA-1 / MD5 / SHA-256 and the piece/part batching around
 * them. TEST INFRASTRUCTURE ONLY (see gh_oracle.h for who may load this and for the "parity
 * unpinned by the reference" statement).
 *
 * Follows, function by function:
 *   sha1_block    FIPS 180-4 6.1.2  -- what anacrolix/torrent v1.15.2 Torrent.hashPiece runs via
 *                                      crypto/sha1 between reference internal/downloader/torrent/torrent.go:79
 *                                      (DownloadAll) and :106 (WaitAll)
 *   md5_block     RFC 1321 3.4      -- minio-go v6.0.55 hashMaterials -> crypto/md5, under
 *                                      reference internal/uploader/uploader.go:89
 *   sha256_block  FIPS 180-4 6.2.2  -- minio-go v6.0.55 hashMaterials -> sha256-simd, same call site
 *   padding       FIPS 180-4 5.1.1 / RFC 1321 3.1-3.2
 */
#include "gh_oracle.h"

#include <pthread.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline uint32_t rol(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static inline uint32_t ror(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static inline uint32_t be32(const uint8_t* p) {
    return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
static inline uint32_t le32(const uint8_t* p) {
    return ((uint32_t)p[3] << 24) | ((uint32_t)p[2] << 16) | ((uint32_t)p[1] << 8) | p[0];
}
static inline void put_be32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v;
}
static inline void put_le32(uint8_t* p, uint32_t v) {
    p[3] = (uint8_t)(v >> 24); p[2] = (uint8_t)(v >> 16); p[1] = (uint8_t)(v >> 8); p[0] = (uint8_t)v;
}

/* ---------------------------------------------------------------- SHA-1, FIPS 180-4 6.1.2 */
/* 16-word schedule ring: W[t] = ROTL1(W[t-3] ^ W[t-8] ^ W[t-14] ^ W[t-16]) (6.1.3 alternate method) */
#define SHA1_W(t) (w[(t) & 15] = rol(w[((t) - 3) & 15] ^ w[((t) - 8) & 15] ^ w[((t) - 14) & 15] ^ w[(t) & 15], 1))
#define SHA1_RND(f, k, wt)                         \
    do {                                           \
        uint32_t tmp = rol(a, 5) + (f) + e + (k) + (wt); \
        e = d; d = c; c = rol(b, 30); b = a; a = tmp;    \
    } while (0)

static void sha1_block(uint32_t st[5], const uint8_t* p) {
    uint32_t w[16];
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4];
    int t;
    for (t = 0; t < 16; t++) { w[t] = be32(p + 4 * t); SHA1_RND((b & c) ^ (~b & d), 0x5a827999u, w[t]); } /* Ch */
    for (; t < 20; t++) SHA1_RND((b & c) ^ (~b & d), 0x5a827999u, SHA1_W(t));                         /* Ch     */
    for (; t < 40; t++) SHA1_RND(b ^ c ^ d, 0x6ed9eba1u, SHA1_W(t));                                  /* Parity */
    for (; t < 60; t++) SHA1_RND((b & c) ^ (b & d) ^ (c & d), 0x8f1bbcdcu, SHA1_W(t));                /* Maj    */
    for (; t < 80; t++) SHA1_RND(b ^ c ^ d, 0xca62c1d6u, SHA1_W(t));                                  /* Parity */
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e;
}

/* ---------------------------------------------------------------- MD5, RFC 1321 3.4 */
static const uint8_t md5_s[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                                  5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                                  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                                  6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
/* T[i] = floor(2^32 * |sin(i+1)|), RFC 1321 3.4 */
static const uint32_t md5_t[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};

#define MD5_RND(f, g, i)                                        \
    do {                                                        \
        uint32_t nb = b + rol(a + (f) + md5_t[i] + x[g], md5_s[i]); \
        a = d; d = c; c = b; b = nb;                            \
    } while (0)

static void md5_block(uint32_t st[4], const uint8_t* p) {
    uint32_t x[16];
    for (int i = 0; i < 16; i++) x[i] = le32(p + 4 * i);
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    int i;
    for (i = 0; i < 16; i++) MD5_RND((b & c) | (~b & d), i, i);                 /* F, RFC 1321 round 1 */
    for (; i < 32; i++) MD5_RND((b & d) | (c & ~d), (5 * i + 1) & 15, i);       /* G, round 2 */
    for (; i < 48; i++) MD5_RND(b ^ c ^ d, (3 * i + 5) & 15, i);                /* H, round 3 */
    for (; i < 64; i++) MD5_RND(c ^ (b | ~d), (7 * i) & 15, i);                 /* I, round 4 */
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

/* ---------------------------------------------------------------- SHA-256, FIPS 180-4 6.2.2 */
static const uint32_t sha256_k[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

static void sha256_block(uint32_t st[8], const uint8_t* p) {
    uint32_t w[64];
    for (int t = 0; t < 16; t++) w[t] = be32(p + 4 * t);
    for (int t = 16; t < 64; t++) {
        uint32_t s0 = ror(w[t - 15], 7) ^ ror(w[t - 15], 18) ^ (w[t - 15] >> 3);
        uint32_t s1 = ror(w[t - 2], 17) ^ ror(w[t - 2], 19) ^ (w[t - 2] >> 10);
        w[t] = s1 + w[t - 7] + s0 + w[t - 16];
    }
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int t = 0; t < 64; t++) {
        uint32_t S1 = ror(e, 6) ^ ror(e, 11) ^ ror(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + sha256_k[t] + w[t];
        uint32_t S0 = ror(a, 2) ^ ror(a, 13) ^ ror(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

/* ---------------------------------------------------------------- init / update / final */
int gho_digest_len(int alg) {
    switch (alg) {
        case GHO_SHA1: return 20;
        case GHO_MD5: return 16;
        case GHO_SHA256: return 32;
        default: return -1;
    }
}

void gho_init(gho_ctx* c, int alg) {
    memset(c, 0, sizeof(*c));
    c->alg = alg;
    if (alg == GHO_SHA1) {              /* FIPS 180-4 5.3.1 */
        c->state[0] = 0x67452301u; c->state[1] = 0xefcdab89u; c->state[2] = 0x98badcfeu;
        c->state[3] = 0x10325476u; c->state[4] = 0xc3d2e1f0u;
    } else if (alg == GHO_MD5) {        /* RFC 1321 3.3 */
        c->state[0] = 0x67452301u; c->state[1] = 0xefcdab89u; c->state[2] = 0x98badcfeu;
        c->state[3] = 0x10325476u;
    } else {                            /* FIPS 180-4 5.3.3 */
        static const uint32_t iv[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au,
                                       0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
        memcpy(c->state, iv, sizeof(iv));
    }
}

static void block(gho_ctx* c, const uint8_t* p) {
    if (c->alg == GHO_SHA1) sha1_block(c->state, p);
    else if (c->alg == GHO_MD5) md5_block(c->state, p);
    else sha256_block(c->state, p);
}

void gho_update(gho_ctx* c, const void* data, size_t len) {
    const uint8_t* p = (const uint8_t*)data;
    c->total += len;
    if (c->buflen) {
        size_t take = 64 - c->buflen;
        if (take > len) take = len;
        memcpy(c->buf + c->buflen, p, take);
        c->buflen += (uint32_t)take; p += take; len -= take;
        if (c->buflen < 64) return;
        block(c, c->buf);
        c->buflen = 0;
    }
    while (len >= 64) { block(c, p); p += 64; len -= 64; }
    if (len) { memcpy(c->buf, p, len); c->buflen = (uint32_t)len; }
}

void gho_final(gho_ctx* c, uint8_t* out) {
    uint8_t pad[72];
    uint64_t bits = c->total * 8;
    uint32_t padlen = (c->buflen < 56) ? (56 - c->buflen) : (120 - c->buflen);
    memset(pad, 0, sizeof(pad));
    pad[0] = 0x80;
    uint8_t lenb[8];
    if (c->alg == GHO_MD5) { put_le32(lenb, (uint32_t)bits); put_le32(lenb + 4, (uint32_t)(bits >> 32)); }
    else                   { put_be32(lenb, (uint32_t)(bits >> 32)); put_be32(lenb + 4, (uint32_t)bits); }
    uint64_t keep = c->total;
    gho_update(c, pad, padlen);
    gho_update(c, lenb, 8);
    c->total = keep;
    int nw = gho_digest_len(c->alg) / 4;
    for (int i = 0; i < nw; i++) {
        if (c->alg == GHO_MD5) put_le32(out + 4 * i, c->state[i]);
        else put_be32(out + 4 * i, c->state[i]);
    }
}

void gho_hash(int alg, const void* data, size_t len, uint8_t* out) {
    gho_ctx c;
    gho_init(&c, alg);
    gho_update(&c, data, len);
    gho_final(&c, out);
}

/* ---------------------------------------------------------------- synthetic stream */
static inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
uint64_t gho_splitmix64(uint64_t x) { return mix64(x + 0x9E3779B97F4A7C15ull); }
uint64_t gho_gen_word(uint64_t seed, uint64_t i) { return mix64(seed + (i + 1) * 0x9E3779B97F4A7C15ull); }

void gho_gen_fill(uint64_t seed, uint64_t offset, uint64_t len, uint8_t* out) {
    uint64_t k = offset, end = offset + len;
    while (k < end && (k & 7)) { *out++ = (uint8_t)(gho_gen_word(seed, k >> 3) >> (8 * (k & 7))); k++; }
    while (k + 8 <= end) {
        uint64_t w = gho_gen_word(seed, k >> 3);
        for (int b = 0; b < 8; b++) out[b] = (uint8_t)(w >> (8 * b));
        out += 8; k += 8;
    }
    while (k < end) { *out++ = (uint8_t)(gho_gen_word(seed, k >> 3) >> (8 * (k & 7))); k++; }
}

/* ---------------------------------------------------------------- batched forms */
uint64_t gho_count(uint64_t total, uint64_t msg_len) {
    if (total == 0 || msg_len == 0) return 0;
    return (total + msg_len - 1) / msg_len;
}

void gho_hash_msgs(int alg, const uint8_t* data, uint64_t total, uint64_t msg_len, uint8_t* out) {
    uint64_t n = gho_count(total, msg_len);
    int dl = gho_digest_len(alg);
    for (uint64_t i = 0; i < n; i++) {
        uint64_t off = i * msg_len;
        uint64_t len = (off + msg_len <= total) ? msg_len : (total - off);
        gho_hash(alg, data + off, len, out + i * dl);
    }
}

typedef struct {
    int alg; const uint8_t* data; uint64_t total, msg_len, n; uint8_t* out;
    atomic_ullong* next;
} mt_job;

static void* mt_worker(void* arg) {
    mt_job* j = (mt_job*)arg;
    int dl = gho_digest_len(j->alg);
    for (;;) {
        uint64_t i = atomic_fetch_add(j->next, 1);
        if (i >= j->n) break;
        uint64_t off = i * j->msg_len;
        uint64_t len = (off + j->msg_len <= j->total) ? j->msg_len : (j->total - off);
        gho_hash(j->alg, j->data + off, len, j->out + i * dl);
    }
    return NULL;
}

void gho_hash_msgs_mt(int alg, const uint8_t* data, uint64_t total, uint64_t msg_len, uint8_t* out,
                      int nthreads) {
    if (nthreads <= 1) { gho_hash_msgs(alg, data, total, msg_len, out); return; }
    atomic_ullong next = 0;
    mt_job job = {alg, data, total, msg_len, gho_count(total, msg_len), out, &next};
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)nthreads);
    int started = 0;
    for (int t = 0; t < nthreads; t++)
        if (pthread_create(&th[started], NULL, mt_worker, &job) == 0) started++;
    if (started == 0) mt_worker(&job);
    for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
    free(th);
}

uint64_t gho_verify_pieces(const uint8_t* data, uint64_t total, uint64_t piece_len,
                           const uint8_t* expected, uint64_t* bitmap, int nthreads) {
    uint64_t n = gho_count(total, piece_len);
    uint8_t* dig = (uint8_t*)malloc(n ? n * 20 : 1);
    gho_hash_msgs_mt(GHO_SHA1, data, total, piece_len, dig, nthreads);
    uint64_t bad = 0;
    memset(bitmap, 0, ((n + 63) / 64) * 8);
    for (uint64_t i = 0; i < n; i++) {
        if (memcmp(dig + 20 * i, expected + 20 * i, 20) == 0) bitmap[i >> 6] |= 1ull << (i & 63);
        else bad++;
    }
    free(dig);
    return bad;
}

int gho_cfg5_corruption(uint64_t piece_index, uint64_t piece_len_i, uint64_t* byte_off, int* bit) {
    if (piece_len_i == 0) return 0;
    uint64_t h = gho_splitmix64((0xBADull << 32) + piece_index);
    *byte_off = h % piece_len_i;
    *bit = (int)(h % 8);
    return (h % 100) == 0;
}

void gho_s3_etag(const uint8_t* part_md5s, uint64_t n, char* out) {
    uint8_t d[16];
    gho_hash(GHO_MD5, part_md5s, (size_t)(n * 16), d);
    for (int i = 0; i < 16; i++) sprintf(out + 2 * i, "%02x", d[i]);
    sprintf(out + 32, "-%llu", (unsigned long long)n);
}
