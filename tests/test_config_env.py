"""gh_config_from_env: the library reads its configuration from the environment like the reference
does for everything else (cmd/downloader/downloader.go:54-58, uploader.go:25). Host-only logic."""
import os
import subprocess
import sys

import pytest

import gpuhash

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ("GPUHASH_DEVICES", "GPUHASH_SLOT_MIB", "GPUHASH_SLOTS", "GPUHASH_STAGE_THREADS", "GPUHASH_DISABLE")


@pytest.fixture
def env(monkeypatch):
    for k in KEYS:
        monkeypatch.delenv(k, raising=False)
    return monkeypatch


def test_defaults_when_nothing_is_set(env):
    assert gpuhash.config_from_env() == (None, 0, 0, 0, False)
    env.setenv("GPUHASH_DEVICES", "")
    env.setenv("GPUHASH_SLOTS", "")
    assert gpuhash.config_from_env() == (None, 0, 0, 0, False)


def test_every_variable(env):
    env.setenv("GPUHASH_DEVICES", "0,2,4,6")
    env.setenv("GPUHASH_SLOT_MIB", "128")
    env.setenv("GPUHASH_SLOTS", "6")
    env.setenv("GPUHASH_STAGE_THREADS", "12")
    assert gpuhash.config_from_env() == ([0, 2, 4, 6], 128 << 20, 6, 12, False)
    env.setenv("GPUHASH_DEVICES", "3")
    assert gpuhash.config_from_env()[0] == [3]


@pytest.mark.parametrize("key,value", [("GPUHASH_DEVICES", "0,x"), ("GPUHASH_DEVICES", "0,,1"), ("GPUHASH_DEVICES", "0,1,"),
                                       ("GPUHASH_DEVICES", "-1"), ("GPUHASH_DEVICES", "1,1"), ("GPUHASH_DEVICES", " 0"),
                                       ("GPUHASH_SLOT_MIB", "0"), ("GPUHASH_SLOT_MIB", "64MiB"), ("GPUHASH_SLOTS", "1"),
                                       ("GPUHASH_SLOTS", "999"), ("GPUHASH_STAGE_THREADS", "0"),
                                       ("GPUHASH_STAGE_THREADS", "four")])
def test_malformed_values_are_refused(env, key, value):
    env.setenv(key, value)
    with pytest.raises(gpuhash.GpuHashError) as e:
        gpuhash.config_from_env()
    assert e.value.status == gpuhash.GH_EINVAL


def test_too_many_devices_is_a_range_error(env):
    env.setenv("GPUHASH_DEVICES", ",".join(str(i) for i in range(65)))
    with pytest.raises(gpuhash.GpuHashError) as e:
        gpuhash.config_from_env()
    assert e.value.status == gpuhash.GH_ERANGE


@pytest.mark.parametrize("value,disabled", [("1", True), ("true", True), ("yes", True), ("0", False), ("false", False),
                                            ("no", False), ("", False)])
def test_disable_switch(env, value, disabled):
    env.setenv("GPUHASH_DISABLE", value)
    assert gpuhash.config_from_env()[4] is disabled
    if disabled:
        with pytest.raises(gpuhash.Disabled):
            gpuhash.Context.from_env()


def test_c_caller_sees_the_same(env, tmp_path):
    """The same entry point from plain C (what the cgo package calls)."""
    src = tmp_path / "cfg.c"
    src.write_text('#include <stdio.h>\n#include "gpuhash.h"\nint main(void){gh_config c;int d[8],dis;'
                   'int rc=gh_config_from_env(&c,d,8,&dis);printf("%d %d %d %llu %d %d %d\\n",rc,c.n_devices,'
                   'c.n_devices?d[1]:-1,(unsigned long long)c.slot_bytes,c.slots,c.stage_threads,dis);return 0;}\n')
    exe = str(tmp_path / "cfg")
    lib = os.path.dirname(gpuhash.LIB_PATH)
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src),
                           "-L" + lib, "-lgpuhash", "-Wl,-rpath," + lib, "-o", exe])
    out = subprocess.run([exe], env=dict(os.environ, GPUHASH_DEVICES="1,5", GPUHASH_SLOT_MIB="32", GPUHASH_SLOTS="3",
                                         GPUHASH_DISABLE="1"), capture_output=True, text=True).stdout.split()
    assert out == ["0", "2", "5", str(32 << 20), "3", "0", "1"]


def test_hex_digests_for_the_signer():
    """gh_hex_digests: lowercase hex, chunk order, what X-Amz-Content-Sha256 / the chunk string-to-sign take."""
    import hashlib
    import ctypes
    import numpy as np
    chunks = [bytes([i]) * (100 + i) for i in range(5)]
    dig = np.array([list(hashlib.sha256(c).digest()) for c in chunks], dtype=np.uint8)
    assert gpuhash.hex_digests(dig) == [hashlib.sha256(c).hexdigest() for c in chunks]
    assert gpuhash.hex_digests(np.zeros((0, 32), dtype=np.uint8)) == []
    buf = ctypes.create_string_buffer(10)
    assert gpuhash.load().gh_hex_digests(ctypes.c_void_p(dig.ctypes.data), 5, 32, buf, 10) == gpuhash.GH_ERANGE
