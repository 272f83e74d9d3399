"""BEP 3 metainfo reader (gh_metainfo_*): host-only parsing checked on CPU against a tiny bencode
writer; verification of a torrent laid out on disk checked on the GPU against the oracle."""
import hashlib
import os

import numpy as np
import pytest

import gpuhash
import oracle


def benc(x):
    if isinstance(x, int):
        return b"i%de" % x
    if isinstance(x, bytes):
        return b"%d:%s" % (len(x), x)
    if isinstance(x, str):
        return benc(x.encode())
    if isinstance(x, list):
        return b"l" + b"".join(benc(v) for v in x) + b"e"
    if isinstance(x, dict):
        items = sorted((k.encode() if isinstance(k, str) else k, v) for k, v in x.items())
        return b"d" + b"".join(benc(k) + benc(v) for k, v in items) + b"e"
    raise TypeError(x)


def make_info(name, piece_len, files, data):
    """files: [(path components, length)] or None for a single-file torrent of len(data)."""
    pieces = b"".join(hashlib.sha1(data[i:i + piece_len]).digest() for i in range(0, len(data), piece_len))
    info = {"name": name, "piece length": piece_len, "pieces": pieces}
    if files is None:
        info["length"] = len(data)
    else:
        info["files"] = [{"length": n, "path": list(p)} for p, n in files]
    return info


def test_parse_multi_file_torrent():
    data = oracle.gen(7, 0, 100000).tobytes()
    files = [(("Season 1", "e01.mkv"), 40000), (("Season 1", "e02.mkv"), 35000), (("extras", "a", "notes.txt"), 25000)]
    info = make_info("pack", 16384, files, data)
    tor = benc({"announce": "http://tracker.invalid/announce", "creation date": 1700000000, "info": info,
                "url-list": ["http://a.invalid/", "http://b.invalid/"], "zz": {"nested": [1, 2, {"x": b"y"}]}})
    m = gpuhash.Metainfo(tor)
    assert (m.name, m.piece_length, m.n_pieces, m.total_length) == ("pack", 16384, 7, 100000)
    assert m.pieces == info["pieces"]
    assert m.files == [("pack/Season 1/e01.mkv", 40000), ("pack/Season 1/e02.mkv", 35000),
                       ("pack/extras/a/notes.txt", 25000)]
    # a bare info dictionary (what the magnet metadata exchange delivers) parses to the same thing
    m2 = gpuhash.Metainfo(benc(info))
    assert (m2.files, m2.pieces, m2.piece_length) == (m.files, m.pieces, m.piece_length)


def test_parse_single_file_and_empty():
    data = oracle.gen(8, 0, 50000).tobytes()
    m = gpuhash.Metainfo(benc({"info": make_info("movie.mp4", 32768, None, data)}))
    assert m.files == [("movie.mp4", 50000)] and m.n_pieces == 2
    e = gpuhash.Metainfo(benc({"info": make_info("empty.bin", 16384, None, b"")}))
    assert e.n_pieces == 0 and e.total_length == 0


@pytest.mark.parametrize("mutate,why", [
    (lambda i: i.update({"name": ".."}), "name"),
    (lambda i: i["files"].__setitem__(0, {"length": 10, "path": ["..", "etc", "passwd"]}), "path component"),
    (lambda i: i["files"].__setitem__(0, {"length": 10, "path": ["a/b"]}), "path component"),
    (lambda i: i.update({"pieces": i["pieces"][:-1]}), "multiple of 20"),
    (lambda i: i.update({"piece length": 0}), "piece length"),
    (lambda i: i.update({"length": 5}), "exactly one"),
    (lambda i: i.update({"pieces": i["pieces"] + bytes(20)}), "piece count"),
    (lambda i: i.pop("name"), "name"),
])
def test_rejects_malformed_or_hostile_metainfo(mutate, why):
    data = oracle.gen(9, 0, 60000).tobytes()
    info = make_info("ok", 16384, [(("a.bin",), 30000), (("b.bin",), 30000)], data)
    mutate(info)
    with pytest.raises(gpuhash.GpuHashError) as e:
        gpuhash.Metainfo(benc({"info": info}))
    assert e.value.status == gpuhash.GH_EINVAL and why in str(e.value)


@pytest.mark.parametrize("raw", [b"", b"i5e", b"d4:infoi1ee", b"d4:info", b"d4:infod4:name5:xe", b"l" * 100 + b"e" * 100,
                                 b"d3:fooli99999999999999999999eee", b"d4:infod6:pieces99999999:abce"])
def test_truncated_and_garbage_input(raw):
    with pytest.raises(gpuhash.GpuHashError):
        gpuhash.Metainfo(raw)


@pytest.mark.gpu
def test_verify_torrent_on_disk(gh, tmp_path):
    """A downloaded multi-file torrent under baseDir (anacrolix file-storage layout, torrent.go:41):
    clean -> all pieces pass; then one flipped byte, one truncated file and one missing file ->
    exactly the pieces the oracle says fail, with the hole reading as zeros."""
    L = 32768
    sizes = [150000 + 7919 * (k + 1) for k in range(5)]
    total = sum(sizes)
    data = oracle.gen(4, 0, total)
    files = [(("cd%d" % k, "track.bin"), s) for k, s in enumerate(sizes)]
    info = make_info("album", L, files, data.tobytes())
    m = gpuhash.Metainfo(benc({"info": info}))
    base = str(tmp_path)
    off = 0
    for (comps, n) in files:
        p = os.path.join(base, "album", *comps)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, "wb") as f:
            f.write(data[off:off + n].tobytes())
        off += n
    assert gh.InfoHash(m) == hashlib.sha1(benc(info)).digest()
    bitmap, dig, nbad = gh.VerifyTorrent(m, base)
    assert nbad == 0 and dig.tobytes() == m.pieces
    # damage: flip a byte in file 0, truncate file 2 by 1000 bytes, delete file 4
    dirty = data.copy()
    with open(os.path.join(base, "album", "cd0", "track.bin"), "r+b") as f:
        f.seek(12345)
        f.write(bytes([data[12345] ^ 0x20]))
    dirty[12345] ^= 0x20
    s2 = sum(sizes[:2])
    os.truncate(os.path.join(base, "album", "cd2", "track.bin"), sizes[2] - 1000)
    dirty[s2 + sizes[2] - 1000: s2 + sizes[2]] = 0
    os.remove(os.path.join(base, "album", "cd4", "track.bin"))
    dirty[sum(sizes[:4]):] = 0
    bitmap, dig, nbad = gh.VerifyTorrent(m, base)
    want_bitmap, want_bad = oracle.verify_pieces(dirty, L, np.frombuffer(m.pieces, dtype=np.uint8))
    assert nbad == want_bad and nbad >= 3 and np.array_equal(bitmap, want_bitmap)
    assert np.array_equal(dig, oracle.hash_msgs(oracle.SHA1, dirty, L))


@pytest.mark.gpu
def test_hash_file_parts_matches_buffers(gh, tmp_path):
    """Upload side: parts of a file on disk (uploader.go:73-89 hands minio an *os.File)."""
    n = 5 * (1 << 20) + 4321
    data = oracle.gen(3, 0, n)
    p = os.path.join(str(tmp_path), "upload.bin")
    with open(p, "wb") as f:
        f.write(data.tobytes())
    got = gh.HashFileParts([(p, n)], 1 << 20, gpuhash.MD5 | gpuhash.SHA256)
    assert np.array_equal(got[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, data, 1 << 20))
    assert np.array_equal(got[gpuhash.SHA256], oracle.hash_msgs(oracle.SHA256, data, 1 << 20))
    # chunk-interleaved ring over a file: window smaller than a part
    ctx = gpuhash.Context(devices=[0], slot_bytes=1 << 20, slots=3)
    try:
        got = ctx.HashFileParts([(p, n)], 3 << 19, gpuhash.SHA1 | gpuhash.MD5)
        assert np.array_equal(got[gpuhash.SHA1], oracle.hash_msgs(oracle.SHA1, data, 3 << 19))
        assert np.array_equal(got[gpuhash.MD5], oracle.hash_msgs(oracle.MD5, data, 3 << 19))
    finally:
        ctx.close()


@pytest.mark.gpu
def test_c_example_verifies_a_torrent_on_disk(tmp_path):
    """examples/verify_torrent.c, compiled with gcc -std=c99, against a payload with one bad piece."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "verify_torrent")
    libdir = os.path.dirname(gpuhash.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-I" + os.path.join(root, "include"), os.path.join(root, "examples", "verify_torrent.c"),
                           "-L" + libdir, "-lgpuhash", "-Wl,-rpath," + libdir, "-o", exe])
    data = oracle.gen(2, 0, 300000)
    info = make_info("clip.mkv", 16384, None, data.tobytes())
    tor = tmp_path / "clip.torrent"
    tor.write_bytes(benc({"announce": "http://t.invalid/a", "info": info}))
    (tmp_path / "clip.mkv").write_bytes(data.tobytes())
    r = subprocess.run([exe, str(tor), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "0 bad piece(s)" in r.stdout and hashlib.sha1(benc(info)).hexdigest() in r.stdout, r.stdout + r.stderr
    bad = bytearray(data.tobytes())
    bad[5 * 16384 + 9] ^= 4
    (tmp_path / "clip.mkv").write_bytes(bytes(bad))
    r = subprocess.run([exe, str(tor), str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1 and "piece 5 failed" in r.stdout and "1 bad piece(s)" in r.stdout, r.stdout + r.stderr


def test_parser_survives_mutated_torrents():
    """Fuzz: single-byte mutations, truncations and random splices of a valid .torrent either parse
    or are rejected with GH_EINVAL -- never crash, never hang (the parser faces bytes from peers)."""
    rng = np.random.default_rng(99)
    data = oracle.gen(5, 0, 70000).tobytes()
    files = [(("a", "b.bin"), 30000), (("c.bin",), 40000)]
    base = bytearray(benc({"announce": "x", "info": make_info("t", 16384, files, data), "z": [1, [2, {"k": b"v"}]]}))
    ok = bad = 0
    for trial in range(3000):
        m = bytearray(base)
        kind = trial % 4
        if kind == 0:
            m[int(rng.integers(0, len(m)))] = int(rng.integers(0, 256))
        elif kind == 1:
            m = m[: int(rng.integers(0, len(m)))]
        elif kind == 2:
            i, j = sorted(int(x) for x in rng.integers(0, len(m), size=2))
            m = m[:i] + m[j:]
        else:
            i = int(rng.integers(0, len(m)))
            m = m[:i] + bytes(rng.integers(0, 256, size=int(rng.integers(1, 12)), dtype=np.uint8)) + m[i:]
        try:
            mi = gpuhash.Metainfo(bytes(m))
            assert mi.piece_length > 0 and len(mi.pieces) == 20 * mi.n_pieces
            for path, n in mi.files:
                assert ".." not in path.split("/") and not path.startswith("/") and n >= 0
            mi.close()
            ok += 1
        except gpuhash.GpuHashError as e:
            assert e.status == gpuhash.GH_EINVAL
            bad += 1
    assert ok > 0 and bad > 0
