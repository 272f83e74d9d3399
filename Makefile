# Top-level targets. Everything builds in-tree; nothing is installed.
#   make lib        libgpuhash.so for sm_100a (nvcc cross-compiles without a GPU)
#   make oracle     the CPU oracle (test infrastructure)
#   make example    the plain-C example against the shared library (C99, no CUDA headers needed)
#   make check      CPU test suite (no GPU needed)
#   make gpucheck   GPU parity suite (needs a B200)
PKG := downloader-go_b200
PY ?= python

all: lib oracle example

lib:
	$(MAKE) -s -C $(PKG)/csrc ../libgpuhash.so

oracle:
	$(MAKE) -s -C oracle

example: lib
	gcc -std=c99 -Wall -Wextra -pedantic -Iinclude examples/verify_torrent.c -L$(PKG) -lgpuhash -Wl,-rpath,'$$ORIGIN/../$(PKG)' -o examples/verify_torrent
	@examples/verify_torrent 2>&1 | grep -q usage && echo "examples/verify_torrent built (prints its usage without arguments)"

check: lib oracle
	$(PY) -m pytest tests -x -q -m "not gpu"

gpucheck: lib oracle
	$(PY) -m pytest tests -x -q -m gpu

clean:
	$(MAKE) -s -C $(PKG)/csrc clean
	$(MAKE) -s -C oracle clean
	rm -f examples/verify_torrent

.PHONY: all lib oracle example check gpucheck clean
