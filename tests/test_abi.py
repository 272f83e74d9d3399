"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/gpuhash.h declares,
and its host-only logic (sharding, error strings, refusing to run without a device) behaves."""
import ctypes
import os
import re

import pytest

import gpuhash

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpuhash.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gh_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = ctypes.CDLL(gpuhash.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), name
    # and the Python mirror binds exactly the declared set
    assert sorted(gpuhash.ABI) == names


def test_strerror_and_version():
    L = gpuhash.load()
    assert L.gh_strerror(0) == b"ok"
    assert b"argument" in L.gh_strerror(gpuhash.GH_EINVAL)
    assert b"device" in L.gh_strerror(gpuhash.GH_ENODEVICE)
    assert b"sm_100a" in L.gh_version()


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="a device is present")
def test_create_without_device_fails_loudly():
    with pytest.raises(gpuhash.GpuHashError) as e:
        gpuhash.Context()
    assert e.value.status == gpuhash.GH_ENODEVICE


@pytest.mark.parametrize("n,shards", [(0, 1), (1, 1), (63, 2), (64, 2), (65, 2), (4096, 8), (12800, 8), (3073, 4),
                                      (640001, 8), (100, 8)])
def test_shard_ranges_partition_on_word_boundaries(n, shards):
    prev = 0
    for s in range(shards):
        b, e = gpuhash.shard_range(n, shards, s)
        assert b == prev and b <= e <= n
        assert b % 64 == 0 or b == n
        prev = e
    assert prev == n
    sizes = [gpuhash.shard_range(n, shards, s) for s in range(shards)]
    words = [(e - b + 63) // 64 for b, e in sizes]
    assert max(words) - min(words) <= 1


def test_header_is_plain_c_and_example_links(tmp_path):
    """include/gpuhash.h is the FFI contract: it must compile as C99 (cgo and any C caller see it
    that way), and the plain-C example must link against the shared library with nothing else."""
    import subprocess
    exe = str(tmp_path / "verify_torrent")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "examples", "verify_torrent.c"), "-L" + os.path.dirname(gpuhash.LIB_PATH), "-lgpuhash",
           "-Wl,-rpath," + os.path.dirname(gpuhash.LIB_PATH), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr
