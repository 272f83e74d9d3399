// gh_hash.cuh -- register-resident compression functions for sm_100a.
//
// One message per lane: the chaining state and the 16-word schedule ring live in registers; every
// Boolean round function is a single LOP3, every rotate a single SHF (funnel shift), byte swaps are
// PRMT. Written from the public standards that Go's crypto/sha1, crypto/md5 and crypto/sha256 (and
// minio/sha256-simd) implement -- FIPS 180-4 6.1.2 / 6.2.2 and RFC 1321 3.4 -- which is the arithmetic
// reached from reference internal/downloader/torrent/torrent.go:79-106 and
// internal/uploader/uploader.go:89.
#pragma once
#include <stdint.h>

namespace ghd {

__device__ __forceinline__ uint32_t rotl(uint32_t x, int n) { return __funnelshift_l(x, x, n); }
__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bswap(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

// lop3 truth-table immediates are evaluated with a = 0xF0, b = 0xCC, c = 0xAA.
template <int IMM>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(r) : "r"(a), "r"(b), "r"(c), "n"(IMM));
    return r;
}
__device__ __forceinline__ uint32_t f_ch(uint32_t x, uint32_t y, uint32_t z) { return lop3<0xCA>(x, y, z); }   // (x&y)|(~x&z)
__device__ __forceinline__ uint32_t f_maj(uint32_t x, uint32_t y, uint32_t z) { return lop3<0xE8>(x, y, z); }  // majority
__device__ __forceinline__ uint32_t f_xor3(uint32_t x, uint32_t y, uint32_t z) { return lop3<0x96>(x, y, z); } // x^y^z
__device__ __forceinline__ uint32_t f_md5i(uint32_t x, uint32_t y, uint32_t z) { return lop3<0x39>(x, y, z); } // y^(x|~z)

// Two-input add on the FMA pipe. The integer ALU pipe (LOP3/SHF/IADD3/PRMT, 64 lanes/clk/SM) is what
// bounds these kernels while the FMA pipe (IMAD, another 64 lanes/clk/SM) idles; x*one+y with `one` a
// kernel parameter that ptxas cannot fold is an IMAD, so the additions of a round move over there and
// the ALU pipe keeps only the Boolean functions and rotates. FMA_ADDS=false gives the plain form.
template <bool FMA_ADDS>
__device__ __forceinline__ uint32_t addf(uint32_t x, uint32_t y, uint32_t one) {
    if constexpr (FMA_ADDS) {
        uint32_t r;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(one), "r"(y));
        return r;
    } else {
        return x + y;
    }
}

// ------------------------------------------------------------------------------------ SHA-1
__device__ __forceinline__ void sha1_init(uint32_t (&s)[5]) {
    s[0] = 0x67452301u; s[1] = 0xefcdab89u; s[2] = 0x98badcfeu; s[3] = 0x10325476u; s[4] = 0xc3d2e1f0u;
}

// w: the block as sixteen big-endian-decoded words; used as the schedule ring and clobbered.
template <bool FMA_ADDS = false>
__device__ __forceinline__ void sha1_compress(uint32_t (&s)[5], uint32_t (&w)[16], uint32_t one = 1) {
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4];
#pragma unroll
    for (int t = 0; t < 80; t++) {
        uint32_t wt;
        if (t < 16) {
            wt = w[t];
        } else {
            wt = rotl(f_xor3(w[(t - 3) & 15], w[(t - 8) & 15], w[(t - 14) & 15]) ^ w[t & 15], 1);
            w[t & 15] = wt;
        }
        uint32_t f, k;
        if (t < 20)      { f = f_ch(b, c, d);   k = 0x5a827999u; }
        else if (t < 40) { f = f_xor3(b, c, d); k = 0x6ed9eba1u; }
        else if (t < 60) { f = f_maj(b, c, d);  k = 0x8f1bbcdcu; }
        else             { f = f_xor3(b, c, d); k = 0xca62c1d6u; }
        uint32_t tmp;
        if constexpr (FMA_ADDS) {
            // e + k is known four rounds ahead and wt one schedule step ahead: (e + k) + wt as two
            // FMA-pipe adds whose only late input, wt, is a single IMAD away from the result. The add
            // of f stays a plain add (ptxas may not re-associate across the IMAD), and the last add
            // fuses with rotl5 into LEA.HI, so the chain through a is unchanged while the ALU pipe
            // sheds one instruction per round.
            const uint32_t pre = addf<true>(wt, e + k, one);
            tmp = rotl(a, 5) + (f + pre);
        } else {
            tmp = rotl(a, 5) + f + (e + k + wt);
        }
        e = d; d = c; c = rotl(b, 30); b = a; a = tmp;
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e;
}

// ------------------------------------------------------------------------------------ MD5
__device__ __forceinline__ void md5_init(uint32_t (&s)[4]) {
    s[0] = 0x67452301u; s[1] = 0xefcdab89u; s[2] = 0x98badcfeu; s[3] = 0x10325476u;
}

#define GH_MD5_F(x, y, z) f_ch(x, y, z)
#define GH_MD5_G(x, y, z) f_ch(z, x, y)
#define GH_MD5_H(x, y, z) f_xor3(x, y, z)
#define GH_MD5_I(x, y, z) f_md5i(x, y, z)
#define GH_MD5_STEP(F, a, b, c, d, xk, s, ti) \
    a = rotl(a + F(b, c, d) + ((xk) + (ti)), s) + b;

// x: the block as sixteen little-endian words (read only).
__device__ __forceinline__ void md5_compress(uint32_t (&st)[4], const uint32_t (&x)[16]) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    GH_MD5_STEP(GH_MD5_F, a, b, c, d, x[0], 7, 0xd76aa478u)  GH_MD5_STEP(GH_MD5_F, d, a, b, c, x[1], 12, 0xe8c7b756u)
    GH_MD5_STEP(GH_MD5_F, c, d, a, b, x[2], 17, 0x242070dbu) GH_MD5_STEP(GH_MD5_F, b, c, d, a, x[3], 22, 0xc1bdceeeu)
    GH_MD5_STEP(GH_MD5_F, a, b, c, d, x[4], 7, 0xf57c0fafu)  GH_MD5_STEP(GH_MD5_F, d, a, b, c, x[5], 12, 0x4787c62au)
    GH_MD5_STEP(GH_MD5_F, c, d, a, b, x[6], 17, 0xa8304613u) GH_MD5_STEP(GH_MD5_F, b, c, d, a, x[7], 22, 0xfd469501u)
    GH_MD5_STEP(GH_MD5_F, a, b, c, d, x[8], 7, 0x698098d8u)  GH_MD5_STEP(GH_MD5_F, d, a, b, c, x[9], 12, 0x8b44f7afu)
    GH_MD5_STEP(GH_MD5_F, c, d, a, b, x[10], 17, 0xffff5bb1u) GH_MD5_STEP(GH_MD5_F, b, c, d, a, x[11], 22, 0x895cd7beu)
    GH_MD5_STEP(GH_MD5_F, a, b, c, d, x[12], 7, 0x6b901122u) GH_MD5_STEP(GH_MD5_F, d, a, b, c, x[13], 12, 0xfd987193u)
    GH_MD5_STEP(GH_MD5_F, c, d, a, b, x[14], 17, 0xa679438eu) GH_MD5_STEP(GH_MD5_F, b, c, d, a, x[15], 22, 0x49b40821u)

    GH_MD5_STEP(GH_MD5_G, a, b, c, d, x[1], 5, 0xf61e2562u)  GH_MD5_STEP(GH_MD5_G, d, a, b, c, x[6], 9, 0xc040b340u)
    GH_MD5_STEP(GH_MD5_G, c, d, a, b, x[11], 14, 0x265e5a51u) GH_MD5_STEP(GH_MD5_G, b, c, d, a, x[0], 20, 0xe9b6c7aau)
    GH_MD5_STEP(GH_MD5_G, a, b, c, d, x[5], 5, 0xd62f105du)  GH_MD5_STEP(GH_MD5_G, d, a, b, c, x[10], 9, 0x02441453u)
    GH_MD5_STEP(GH_MD5_G, c, d, a, b, x[15], 14, 0xd8a1e681u) GH_MD5_STEP(GH_MD5_G, b, c, d, a, x[4], 20, 0xe7d3fbc8u)
    GH_MD5_STEP(GH_MD5_G, a, b, c, d, x[9], 5, 0x21e1cde6u)  GH_MD5_STEP(GH_MD5_G, d, a, b, c, x[14], 9, 0xc33707d6u)
    GH_MD5_STEP(GH_MD5_G, c, d, a, b, x[3], 14, 0xf4d50d87u) GH_MD5_STEP(GH_MD5_G, b, c, d, a, x[8], 20, 0x455a14edu)
    GH_MD5_STEP(GH_MD5_G, a, b, c, d, x[13], 5, 0xa9e3e905u) GH_MD5_STEP(GH_MD5_G, d, a, b, c, x[2], 9, 0xfcefa3f8u)
    GH_MD5_STEP(GH_MD5_G, c, d, a, b, x[7], 14, 0x676f02d9u) GH_MD5_STEP(GH_MD5_G, b, c, d, a, x[12], 20, 0x8d2a4c8au)

    GH_MD5_STEP(GH_MD5_H, a, b, c, d, x[5], 4, 0xfffa3942u)  GH_MD5_STEP(GH_MD5_H, d, a, b, c, x[8], 11, 0x8771f681u)
    GH_MD5_STEP(GH_MD5_H, c, d, a, b, x[11], 16, 0x6d9d6122u) GH_MD5_STEP(GH_MD5_H, b, c, d, a, x[14], 23, 0xfde5380cu)
    GH_MD5_STEP(GH_MD5_H, a, b, c, d, x[1], 4, 0xa4beea44u)  GH_MD5_STEP(GH_MD5_H, d, a, b, c, x[4], 11, 0x4bdecfa9u)
    GH_MD5_STEP(GH_MD5_H, c, d, a, b, x[7], 16, 0xf6bb4b60u) GH_MD5_STEP(GH_MD5_H, b, c, d, a, x[10], 23, 0xbebfbc70u)
    GH_MD5_STEP(GH_MD5_H, a, b, c, d, x[13], 4, 0x289b7ec6u) GH_MD5_STEP(GH_MD5_H, d, a, b, c, x[0], 11, 0xeaa127fau)
    GH_MD5_STEP(GH_MD5_H, c, d, a, b, x[3], 16, 0xd4ef3085u) GH_MD5_STEP(GH_MD5_H, b, c, d, a, x[6], 23, 0x04881d05u)
    GH_MD5_STEP(GH_MD5_H, a, b, c, d, x[9], 4, 0xd9d4d039u)  GH_MD5_STEP(GH_MD5_H, d, a, b, c, x[12], 11, 0xe6db99e5u)
    GH_MD5_STEP(GH_MD5_H, c, d, a, b, x[15], 16, 0x1fa27cf8u) GH_MD5_STEP(GH_MD5_H, b, c, d, a, x[2], 23, 0xc4ac5665u)

    GH_MD5_STEP(GH_MD5_I, a, b, c, d, x[0], 6, 0xf4292244u)  GH_MD5_STEP(GH_MD5_I, d, a, b, c, x[7], 10, 0x432aff97u)
    GH_MD5_STEP(GH_MD5_I, c, d, a, b, x[14], 15, 0xab9423a7u) GH_MD5_STEP(GH_MD5_I, b, c, d, a, x[5], 21, 0xfc93a039u)
    GH_MD5_STEP(GH_MD5_I, a, b, c, d, x[12], 6, 0x655b59c3u) GH_MD5_STEP(GH_MD5_I, d, a, b, c, x[3], 10, 0x8f0ccc92u)
    GH_MD5_STEP(GH_MD5_I, c, d, a, b, x[10], 15, 0xffeff47du) GH_MD5_STEP(GH_MD5_I, b, c, d, a, x[1], 21, 0x85845dd1u)
    GH_MD5_STEP(GH_MD5_I, a, b, c, d, x[8], 6, 0x6fa87e4fu)  GH_MD5_STEP(GH_MD5_I, d, a, b, c, x[15], 10, 0xfe2ce6e0u)
    GH_MD5_STEP(GH_MD5_I, c, d, a, b, x[6], 15, 0xa3014314u) GH_MD5_STEP(GH_MD5_I, b, c, d, a, x[13], 21, 0x4e0811a1u)
    GH_MD5_STEP(GH_MD5_I, a, b, c, d, x[4], 6, 0xf7537e82u)  GH_MD5_STEP(GH_MD5_I, d, a, b, c, x[11], 10, 0xbd3af235u)
    GH_MD5_STEP(GH_MD5_I, c, d, a, b, x[2], 15, 0x2ad7d2bbu) GH_MD5_STEP(GH_MD5_I, b, c, d, a, x[9], 21, 0xeb86d391u)
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}

// ------------------------------------------------------------------------------------ SHA-256
__device__ __forceinline__ void sha256_init(uint32_t (&s)[8]) {
    s[0] = 0x6a09e667u; s[1] = 0xbb67ae85u; s[2] = 0x3c6ef372u; s[3] = 0xa54ff53au;
    s[4] = 0x510e527fu; s[5] = 0x9b05688cu; s[6] = 0x1f83d9abu; s[7] = 0x5be0cd19u;
}

__constant__ uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

__device__ __forceinline__ uint32_t S256_S0(uint32_t x) { return f_xor3(rotr(x, 2), rotr(x, 13), rotr(x, 22)); }
__device__ __forceinline__ uint32_t S256_S1(uint32_t x) { return f_xor3(rotr(x, 6), rotr(x, 11), rotr(x, 25)); }
__device__ __forceinline__ uint32_t S256_s0(uint32_t x) { return f_xor3(rotr(x, 7), rotr(x, 18), x >> 3); }
__device__ __forceinline__ uint32_t S256_s1(uint32_t x) { return f_xor3(rotr(x, 17), rotr(x, 19), x >> 10); }

// w: the block as sixteen big-endian-decoded words; used as the schedule ring and clobbered.
template <bool FMA_ADDS = false>
__device__ __forceinline__ void sha256_compress(uint32_t (&s)[8], uint32_t (&w)[16], uint32_t one = 1) {
    uint32_t v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = s[i];
#pragma unroll
    for (int t = 0; t < 64; t++) {
        uint32_t wt;
        if (t < 16) {
            wt = w[t];
        } else {
            if constexpr (FMA_ADDS)
                wt = addf<true>(addf<true>(S256_s1(w[(t - 2) & 15]), w[(t - 7) & 15], one),
                                addf<true>(S256_s0(w[(t - 15) & 15]), w[t & 15], one), one);
            else
                wt = S256_s1(w[(t - 2) & 15]) + w[(t - 7) & 15] + S256_s0(w[(t - 15) & 15]) + w[t & 15];
            w[t & 15] = wt;
        }
        // working variables a..h are v[(0 - t) & 7] .. v[(7 - t) & 7]: rotate by renaming, not by moves
        uint32_t& a = v[(0 - t) & 7]; uint32_t& b = v[(1 - t) & 7]; uint32_t& c = v[(2 - t) & 7];
        uint32_t& d = v[(3 - t) & 7]; uint32_t& e = v[(4 - t) & 7]; uint32_t& f = v[(5 - t) & 7];
        uint32_t& g = v[(6 - t) & 7]; uint32_t& h = v[(7 - t) & 7];
        if constexpr (FMA_ADDS) {
            // h + K + wt is ready long before e settles; Sigma1/Ch/Sigma0/Maj stay on the ALU pipe
            const uint32_t hk = addf<true>(addf<true>(wt, K256[t], one), h, one);
            const uint32_t t1 = addf<true>(addf<true>(hk, f_ch(e, f, g), one), S256_S1(e), one);
            const uint32_t t2 = addf<true>(S256_S0(a), f_maj(a, b, c), one);
            d = addf<true>(d, t1, one);
            h = addf<true>(t1, t2, one);
        } else {
            uint32_t t1 = h + S256_S1(e) + f_ch(e, f, g) + (K256[t] + wt);
            uint32_t t2 = S256_S0(a) + f_maj(a, b, c);
            d += t1;
            h = t1 + t2;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) s[i] += v[i];
}

}  // namespace ghd
