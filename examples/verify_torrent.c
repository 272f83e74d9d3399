/* verify_torrent.c -- the C ABI used from plain C: parse a .torrent, verify the payload under a
 * directory on the GPU(s), print the failed pieces.
 *
 *   gcc -std=c99 -Iinclude examples/verify_torrent.c -Ldownloader-go_b200 -lgpuhash \
 *       -Wl,-rpath,$PWD/downloader-go_b200 -o verify_torrent
 *   ./verify_torrent some.torrent /path/to/baseDir
 *
 * This is what the storage wrapper of INTEGRATION.md section 2 does through cgo at
 * internal/downloader/torrent/torrent.go:41. Exit status: 0 all pieces good, 1 some bad, 2 error. */
#include <stdio.h>
#include <stdlib.h>

#include "gpuhash.h"

static unsigned char* slurp(const char* path, size_t* len) {
    FILE* f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char* buf = (unsigned char*)malloc(n > 0 ? (size_t)n : 1);
    if (buf && fread(buf, 1, (size_t)n, f) != (size_t)n) {
        free(buf);
        buf = NULL;
    }
    fclose(f);
    *len = (size_t)n;
    return buf;
}

int main(int argc, char** argv) {
    if (argc != 3) {
        fprintf(stderr, "usage: %s file.torrent base_dir\n", argv[0]);
        return 2;
    }
    size_t len = 0;
    unsigned char* raw = slurp(argv[1], &len);
    if (!raw) {
        fprintf(stderr, "cannot read %s\n", argv[1]);
        return 2;
    }
    gh_metainfo* mi = NULL;
    char err[256];
    if (gh_metainfo_parse(raw, len, &mi, err, sizeof err) != GH_OK) {
        fprintf(stderr, "bad metainfo: %s\n", err);
        return 2;
    }
    gh_ctx* ctx = NULL;
    int rc = gh_create(&ctx, NULL);
    if (rc != GH_OK) {
        fprintf(stderr, "gh_create: %s\n", gh_strerror(rc));
        return 2;
    }
    uint64_t n = gh_metainfo_n_pieces(mi), bad = 0;
    uint64_t* bitmap = (uint64_t*)calloc((size_t)((n + 63) / 64) + 1, 8);
    unsigned char ih[20];
    gh_metainfo_info_hash(ctx, mi, ih);
    printf("name %s, %llu pieces of %llu bytes, %zu file(s), info-hash ", gh_metainfo_name(mi),
           (unsigned long long)n, (unsigned long long)gh_metainfo_piece_length(mi), gh_metainfo_n_files(mi));
    for (int i = 0; i < 20; i++) printf("%02x", ih[i]);
    printf("\n");
    rc = gh_verify_torrent(ctx, mi, argv[2], NULL, bitmap, &bad);
    if (rc != GH_OK) {
        fprintf(stderr, "gh_verify_torrent: %s (%s)\n", gh_strerror(rc), gh_last_error(ctx));
        return 2;
    }
    gh_stats st;
    gh_get_stats(ctx, &st);
    printf("%llu bad piece(s); %.1f ms wall, %.1f ms in kernels\n", (unsigned long long)bad, st.wall_ms, st.kernel_ms);
    for (uint64_t i = 0; i < n && bad; i++)
        if (!((bitmap[i >> 6] >> (i & 63)) & 1)) printf("  piece %llu failed\n", (unsigned long long)i);
    free(bitmap);
    gh_destroy(ctx);
    gh_metainfo_free(mi);
    free(raw);
    return bad ? 1 : 0;
}
