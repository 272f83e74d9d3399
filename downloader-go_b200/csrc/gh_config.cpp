// gh_config.cpp -- configuration from the environment, the way the reference configures itself
// (RABBITMQ_ENDPOINT, S3_ENDPOINT ... read in reference cmd/downloader/downloader.go:54-58 and
// internal/uploader/uploader.go:25). Host-only: needs no device, so it is unit-tested on the CPU.
#include <errno.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/gpuhash.h"

namespace {

// whole string must be a non-negative decimal number <= max
bool parse_u64(const char* s, unsigned long long max, unsigned long long* out) {
    if (!s || !*s) return false;
    for (const char* p = s; *p; p++)
        if (*p < '0' || *p > '9') return false;
    errno = 0;
    const unsigned long long v = strtoull(s, nullptr, 10);
    if (errno || v > max) return false;
    *out = v;
    return true;
}

bool truthy(const char* s) { return s && *s && strcmp(s, "0") != 0 && strcmp(s, "false") != 0 && strcmp(s, "no") != 0; }

}  // namespace

extern "C" int gh_config_from_env(gh_config* cfg, int* devices_buf, int devices_cap, int* disabled) {
    if (!cfg) return GH_EINVAL;
    memset(cfg, 0, sizeof(*cfg));
    if (disabled) *disabled = truthy(getenv("GPUHASH_DISABLE")) ? 1 : 0;
    unsigned long long v;
    if (const char* s = getenv("GPUHASH_DEVICES")) {
        if (*s) {
            if (!devices_buf || devices_cap <= 0) return GH_EINVAL;
            int n = 0;
            const char* p = s;
            while (*p) {
                char tok[24];
                size_t k = 0;
                while (*p && *p != ',' && k + 1 < sizeof(tok)) tok[k++] = *p++;
                tok[k] = 0;
                if ((*p && *p != ',') || !parse_u64(tok, 1023, &v)) return GH_EINVAL;  // "0,x", "0,,1", over-long token
                if (n >= devices_cap) return GH_ERANGE;
                for (int i = 0; i < n; i++)
                    if (devices_buf[i] == static_cast<int>(v)) return GH_EINVAL;  // the same device twice
                devices_buf[n++] = static_cast<int>(v);
                if (*p == ',') {
                    p++;
                    if (!*p) return GH_EINVAL;  // trailing comma
                }
            }
            cfg->n_devices = n;
            cfg->devices = devices_buf;
        }
    }
    if (const char* s = getenv("GPUHASH_SLOT_MIB")) {
        if (*s) {
            if (!parse_u64(s, 1ull << 20, &v) || v == 0) return GH_EINVAL;
            cfg->slot_bytes = v << 20;
        }
    }
    if (const char* s = getenv("GPUHASH_SLOTS")) {
        if (*s) {
            if (!parse_u64(s, 64, &v) || v < 2) return GH_EINVAL;
            cfg->slots = static_cast<int>(v);
        }
    }
    if (const char* s = getenv("GPUHASH_STAGE_THREADS")) {
        if (*s) {
            if (!parse_u64(s, 256, &v) || v == 0) return GH_EINVAL;
            cfg->stage_threads = static_cast<int>(v);
        }
    }
    return GH_OK;
}

// Lowercase hex of n digests of digest_len bytes each, in order, 2*digest_len characters per digest, no
// separators, NUL after the last: the form SigV4 wants (X-Amz-Content-Sha256, and the per-chunk
// string-to-sign of the streaming signature, which chains chunk k's signature on hex(sha256(chunk k))).
extern "C" int gh_hex_digests(const uint8_t* digests, uint64_t n, int digest_len, char* out, size_t out_len) {
    if (!out || digest_len <= 0 || (n && !digests)) return GH_EINVAL;
    const uint64_t need = n * 2ull * static_cast<uint64_t>(digest_len) + 1;
    if (out_len < need) return GH_ERANGE;
    static const char kHex[] = "0123456789abcdef";
    const uint64_t bytes = n * static_cast<uint64_t>(digest_len);
    for (uint64_t i = 0; i < bytes; i++) {
        out[2 * i] = kHex[digests[i] >> 4];
        out[2 * i + 1] = kHex[digests[i] & 15];
    }
    out[2 * bytes] = 0;
    return GH_OK;
}
