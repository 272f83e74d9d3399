// gh_metainfo.cpp -- BEP 3 metainfo (.torrent / info dictionary) reader: the data format on the input
// side of piece verification. Host-only; no CUDA in this file.
//
// In the reference the info dictionary arrives over the wire after client.AddMagnet
// (internal/downloader/torrent/torrent.go:58) and is awaited at :69 (<-t.GotInfo()); anacrolix's
// metainfo package then supplies piece length, the concatenated 20-byte piece hashes and the file
// list that its file storage (torrent.go:41) maps pieces onto. This is that step for the GPU path:
// bencode -> (piece length, pieces, files in info order) -> gh_verify_files.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/gpuhash.h"

struct gh_metainfo {
    std::string name;
    uint64_t piece_length = 0;
    std::string pieces;  // 20 * n bytes
    struct File { std::string path; uint64_t len; };
    std::vector<File> files;  // relative paths below base_dir, in info order
    bool multi_file = false;
    uint64_t total = 0;
    size_t info_begin = 0, info_end = 0;  // byte range of the bencoded info dictionary in the input
    std::string raw_info;
    std::string error;
};

namespace {

struct Parser {
    const uint8_t* p;
    size_t n, i = 0;
    std::string err;
    int depth = 0;

    bool fail(const char* m) {
        if (err.empty()) err = std::string(m) + " at byte " + std::to_string(i);
        return false;
    }
    bool integer(int64_t* out) {  // i<digits>e
        if (i >= n || p[i] != 'i') return fail("expected integer");
        i++;
        bool neg = false;
        if (i < n && p[i] == '-') { neg = true; i++; }
        if (i >= n || p[i] < '0' || p[i] > '9') return fail("bad integer");
        uint64_t v = 0;
        size_t digits = 0;
        while (i < n && p[i] >= '0' && p[i] <= '9') {
            if (v > (UINT64_MAX - 9) / 10) return fail("integer overflow");
            v = v * 10 + (p[i] - '0');
            i++;
            digits++;
        }
        if (i >= n || p[i] != 'e') return fail("unterminated integer");
        i++;
        if (v > static_cast<uint64_t>(INT64_MAX)) return fail("integer overflow");
        (void)digits;
        *out = neg ? -static_cast<int64_t>(v) : static_cast<int64_t>(v);
        return true;
    }
    bool bytes(const uint8_t** s, size_t* len) {  // <len>:<bytes>
        if (i >= n || p[i] < '0' || p[i] > '9') return fail("expected string");
        uint64_t v = 0;
        while (i < n && p[i] >= '0' && p[i] <= '9') {
            if (v > (1ull << 40)) return fail("string too long");
            v = v * 10 + (p[i] - '0');
            i++;
        }
        if (i >= n || p[i] != ':') return fail("bad string header");
        i++;
        if (v > n - i) return fail("string runs past the end");
        *s = p + i;
        *len = static_cast<size_t>(v);
        i += static_cast<size_t>(v);
        return true;
    }
    bool skip() {  // any value
        if (i >= n) return fail("unexpected end");
        if (++depth > 64) return fail("nesting too deep");
        bool ok = true;
        if (p[i] == 'i') {
            int64_t v;
            ok = integer(&v);
        } else if (p[i] == 'l' || p[i] == 'd') {
            const bool dict = p[i] == 'd';
            i++;
            while (ok && i < n && p[i] != 'e') {
                if (dict) {
                    const uint8_t* k = nullptr;
                    size_t kl = 0;
                    ok = bytes(&k, &kl);
                    if (!ok) break;
                }
                ok = skip();
            }
            if (ok && i >= n) ok = fail("unterminated container");
            if (ok) i++;
        } else {
            const uint8_t* s = nullptr;
            size_t l = 0;
            ok = bytes(&s, &l);
        }
        depth--;
        return ok;
    }
};

bool key_is(const uint8_t* k, size_t kl, const char* name) { return kl == strlen(name) && memcmp(k, name, kl) == 0; }

bool safe_component(const std::string& c) {
    if (c.empty() || c == "." || c == "..") return false;
    for (char ch : c)
        if (ch == '/' || ch == '\\' || ch == '\0') return false;
    return true;
}

bool parse_files(Parser& ps, gh_metainfo* m) {
    if (ps.i >= ps.n || ps.p[ps.i] != 'l') return ps.fail("files is not a list");
    ps.i++;
    while (ps.i < ps.n && ps.p[ps.i] != 'e') {
        if (ps.p[ps.i] != 'd') return ps.fail("file entry is not a dictionary");
        ps.i++;
        int64_t len = -1;
        std::string path;
        bool have_path = false;
        while (ps.i < ps.n && ps.p[ps.i] != 'e') {
            const uint8_t* k = nullptr;
            size_t kl = 0;
            if (!ps.bytes(&k, &kl)) return false;
            if (key_is(k, kl, "length")) {
                if (!ps.integer(&len)) return false;
            } else if (key_is(k, kl, "path")) {
                if (ps.i >= ps.n || ps.p[ps.i] != 'l') return ps.fail("path is not a list");
                ps.i++;
                while (ps.i < ps.n && ps.p[ps.i] != 'e') {
                    const uint8_t* s = nullptr;
                    size_t sl = 0;
                    if (!ps.bytes(&s, &sl)) return false;
                    std::string comp(reinterpret_cast<const char*>(s), sl);
                    if (!safe_component(comp)) return ps.fail("unsafe path component");
                    if (!path.empty()) path += '/';
                    path += comp;
                    have_path = true;
                }
                if (ps.i >= ps.n) return ps.fail("unterminated path");
                ps.i++;
            } else if (!ps.skip()) {
                return false;
            }
        }
        if (ps.i >= ps.n) return ps.fail("unterminated file entry");
        ps.i++;
        if (len < 0 || !have_path) return ps.fail("file entry without length or path");
        m->files.push_back({path, static_cast<uint64_t>(len)});
    }
    if (ps.i >= ps.n) return ps.fail("unterminated files list");
    ps.i++;
    return true;
}

bool parse_info(Parser& ps, gh_metainfo* m) {
    if (ps.i >= ps.n || ps.p[ps.i] != 'd') return ps.fail("info is not a dictionary");
    m->info_begin = ps.i;
    ps.i++;
    int64_t piece_length = -1, length = -1;
    bool have_pieces = false, have_files = false, have_name = false;
    while (ps.i < ps.n && ps.p[ps.i] != 'e') {
        const uint8_t* k = nullptr;
        size_t kl = 0;
        if (!ps.bytes(&k, &kl)) return false;
        if (key_is(k, kl, "piece length")) {
            if (!ps.integer(&piece_length)) return false;
        } else if (key_is(k, kl, "length")) {
            if (!ps.integer(&length)) return false;
        } else if (key_is(k, kl, "pieces")) {
            const uint8_t* s = nullptr;
            size_t sl = 0;
            if (!ps.bytes(&s, &sl)) return false;
            m->pieces.assign(reinterpret_cast<const char*>(s), sl);
            have_pieces = true;
        } else if (key_is(k, kl, "name")) {
            const uint8_t* s = nullptr;
            size_t sl = 0;
            if (!ps.bytes(&s, &sl)) return false;
            m->name.assign(reinterpret_cast<const char*>(s), sl);
            have_name = true;
        } else if (key_is(k, kl, "files")) {
            if (!parse_files(ps, m)) return false;
            have_files = true;
        } else if (!ps.skip()) {
            return false;
        }
    }
    if (ps.i >= ps.n) return ps.fail("unterminated info dictionary");
    ps.i++;
    m->info_end = ps.i;
    if (piece_length <= 0) return ps.fail("missing or non-positive piece length");
    if (!have_pieces || m->pieces.size() % 20) return ps.fail("pieces missing or not a multiple of 20 bytes");
    if (!have_name || !safe_component(m->name)) return ps.fail("missing or unsafe name");
    if (have_files == (length >= 0)) return ps.fail("exactly one of length / files is required");
    m->piece_length = static_cast<uint64_t>(piece_length);
    m->multi_file = have_files;
    if (have_files) {
        for (auto& f : m->files) f.path = m->name + "/" + f.path;  // anacrolix file storage: baseDir/name/path...
    } else {
        m->files.push_back({m->name, static_cast<uint64_t>(length)});
    }
    m->total = 0;
    for (auto& f : m->files) {
        if (f.len > (1ull << 50) || m->total + f.len < m->total) return ps.fail("file sizes overflow");
        m->total += f.len;
    }
    const uint64_t want = m->total == 0 ? 0 : (m->total + m->piece_length - 1) / m->piece_length;
    if (want != m->pieces.size() / 20) return ps.fail("piece count does not match the total length");
    return true;
}

}  // namespace

extern "C" {

int gh_metainfo_parse(const uint8_t* data, size_t len, gh_metainfo** out, char* err, size_t err_len) {
    if (!out) return GH_EINVAL;
    *out = nullptr;
    if (err && err_len) err[0] = 0;
    if (!data || !len) return GH_EINVAL;
    Parser ps{data, len};
    gh_metainfo* m = new gh_metainfo();
    bool ok = false, found = false;
    if (data[0] == 'd') {
        // a .torrent file: top-level dictionary with an "info" key; a bare info dictionary (what a
        // magnet link's metadata exchange delivers) has "piece length" at the top level instead
        ps.i = 1;
        ok = true;
        while (ok && ps.i < ps.n && ps.p[ps.i] != 'e') {
            const uint8_t* k = nullptr;
            size_t kl = 0;
            ok = ps.bytes(&k, &kl);
            if (!ok) break;
            if (key_is(k, kl, "info")) {
                ok = parse_info(ps, m);
                found = ok;
            } else if (key_is(k, kl, "piece length") || key_is(k, kl, "pieces")) {
                ps = Parser{data, len};  // bare info dictionary: start over at byte 0
                *m = gh_metainfo();
                ok = parse_info(ps, m);
                found = ok;
                break;
            } else {
                ok = ps.skip();
            }
        }
        if (ok && !found) ok = ps.fail("no info dictionary");
    } else {
        ps.fail("not a bencoded dictionary");
    }
    if (!ok) {
        if (err && err_len) snprintf(err, err_len, "%s", ps.err.c_str());
        delete m;
        return GH_EINVAL;
    }
    m->raw_info.assign(reinterpret_cast<const char*>(data + m->info_begin), m->info_end - m->info_begin);
    *out = m;
    return GH_OK;
}

void gh_metainfo_free(gh_metainfo* m) { delete m; }
uint64_t gh_metainfo_piece_length(const gh_metainfo* m) { return m ? m->piece_length : 0; }
uint64_t gh_metainfo_n_pieces(const gh_metainfo* m) { return m ? m->pieces.size() / 20 : 0; }
uint64_t gh_metainfo_total_length(const gh_metainfo* m) { return m ? m->total : 0; }
const uint8_t* gh_metainfo_pieces(const gh_metainfo* m) { return m ? reinterpret_cast<const uint8_t*>(m->pieces.data()) : nullptr; }
const char* gh_metainfo_name(const gh_metainfo* m) { return m ? m->name.c_str() : ""; }
size_t gh_metainfo_n_files(const gh_metainfo* m) { return m ? m->files.size() : 0; }
int gh_metainfo_file(const gh_metainfo* m, size_t i, const char** rel_path, uint64_t* len) {
    if (!m || i >= m->files.size()) return GH_EINVAL;
    if (rel_path) *rel_path = m->files[i].path.c_str();
    if (len) *len = m->files[i].len;
    return GH_OK;
}

int gh_metainfo_info_hash(gh_ctx* ctx, const gh_metainfo* m, uint8_t out[20]) {
    if (!ctx || !m || !out) return GH_EINVAL;
    gh_span sp{m->raw_info.data(), m->raw_info.size()};
    return gh_hash_parts(ctx, &sp, 1, m->raw_info.size(), GH_SHA1, out, nullptr, nullptr);
}

int gh_verify_torrent(gh_ctx* ctx, const gh_metainfo* m, const char* base_dir, uint8_t* digests_out,
                      uint64_t* ok_bitmap_out, uint64_t* n_bad_out) {
    if (!ctx || !m || !base_dir) return GH_EINVAL;
    std::vector<std::string> paths(m->files.size());
    std::vector<gh_file> files(m->files.size());
    for (size_t i = 0; i < m->files.size(); i++) {
        paths[i] = std::string(base_dir) + "/" + m->files[i].path;
        files[i].path = paths[i].c_str();
        files[i].len = m->files[i].len;
    }
    return gh_verify_files(ctx, files.data(), files.size(), m->piece_length,
                           reinterpret_cast<const uint8_t*>(m->pieces.data()), digests_out, ok_bitmap_out, n_bad_out);
}

}  // extern "C"
