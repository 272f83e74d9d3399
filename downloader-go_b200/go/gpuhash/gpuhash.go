// Package gpuhash binds libgpuhash.so (include/gpuhash.h) for the downloader worker.
//
// SOURCE ONLY: there is no Go toolchain in the build image or on the GPU box (SURVEY.md section 0.4),
// so this file has never been compiled. The behaviour it exposes is exercised through the same C ABI
// by the Python mirror in downloader-go_b200/gpuhash and tests/. Build needs CGO_ENABLED=1, glibc and
// the CUDA runtime (the reference ships CGO_ENABLED=0 on alpine: Makefile:12, Dockerfile:14-16).
//
// VerifyPieces replaces, batch-wise, anacrolix/torrent's Torrent.hashPiece/pieceHashed (crypto/sha1)
// that run under internal/downloader/torrent/torrent.go:79-106; HashParts replaces minio-go's
// hashMaterials (crypto/md5 + sha256-simd) under internal/uploader/uploader.go:89.
package gpuhash

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../.. -lgpuhash -Wl,-rpath,${SRCDIR}/../..
#include <stdlib.h>
#include "gpuhash.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"runtime"
	"unsafe"
)

// Algo is a bitmask of digests to compute.
type Algo int

const (
	SHA1   Algo = C.GH_SHA1
	MD5    Algo = C.GH_MD5
	SHA256 Algo = C.GH_SHA256
)

// FileSpan is one file's bytes (or any run of stream bytes) in memory. The logical stream is the
// concatenation of the spans in order, i.e. BEP 3's file order.
type FileSpan []byte

// PartSum holds the digests of one upload part (nil for unselected algorithms).
type PartSum struct {
	SHA1   []byte
	MD5    []byte
	SHA256 []byte
}

// Hasher owns one gh_ctx. Safe for concurrent use; calls are serialised inside the library.
type Hasher struct{ ctx *C.gh_ctx }

// ErrNoDevice is returned by New when no CUDA device is usable. There is deliberately no CPU
// fallback in the library; a caller that wants one keeps using crypto/* itself.
var ErrNoDevice = errors.New("gpuhash: no usable CUDA device")

func status(ctx *C.gh_ctx, rc C.int, what string) error {
	if rc == C.GH_OK {
		return nil
	}
	if rc == C.GH_ENODEVICE {
		return ErrNoDevice
	}
	detail := ""
	if ctx != nil {
		detail = C.GoString(C.gh_last_error(ctx))
	}
	return fmt.Errorf("gpuhash: %s: %s (%s)", what, C.GoString(C.gh_strerror(rc)), detail)
}

// New creates a hasher over the given CUDA ordinals (nil = every visible device).
func New(devices []int) (*Hasher, error) {
	var cfg C.gh_config
	var devs []C.int
	if len(devices) > 0 {
		devs = make([]C.int, len(devices))
		for i, d := range devices {
			devs[i] = C.int(d)
		}
		cfg.n_devices = C.int(len(devs))
		cfg.devices = &devs[0]
	}
	h := &Hasher{}
	rc := C.gh_create(&h.ctx, &cfg)
	runtime.KeepAlive(devs)
	if err := status(nil, rc, "gh_create"); err != nil {
		return nil, err
	}
	runtime.SetFinalizer(h, (*Hasher).Close)
	return h, nil
}

// Close releases the context and its staging ring.
func (h *Hasher) Close() {
	if h.ctx != nil {
		C.gh_destroy(h.ctx)
		h.ctx = nil
	}
}

// spans builds the C span array in C memory: Go pointers may not be stored in Go memory that is
// passed to C, and the library never retains them past the call.
func spans(files []FileSpan) (*C.gh_span, C.size_t, func()) {
	n := len(files)
	if n == 0 {
		return nil, 0, func() {}
	}
	arr := (*C.gh_span)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(C.gh_span{}))))
	sl := unsafe.Slice(arr, n)
	var pin runtime.Pinner
	for i, f := range files {
		if len(f) > 0 {
			pin.Pin(&f[0])
			sl[i].host_ptr = unsafe.Pointer(&f[0])
		} else {
			sl[i].host_ptr = nil
		}
		sl[i].len = C.uint64_t(len(f))
	}
	return arr, C.size_t(n), func() { pin.Unpin(); C.free(unsafe.Pointer(arr)) }
}

// VerifyPieces hashes every piece of the stream with SHA-1 and compares it with expected
// (20 bytes per piece, info.Pieces). bitmap bit i is set iff piece i passed.
func (h *Hasher) VerifyPieces(files []FileSpan, pieceLen int64, expected []byte) (bitmap []uint64, digests []byte, bad uint64, err error) {
	var total int64
	for _, f := range files {
		total += int64(len(f))
	}
	if pieceLen <= 0 {
		return nil, nil, 0, errors.New("gpuhash: piece length must be > 0")
	}
	n := (total + pieceLen - 1) / pieceLen
	if expected != nil && int64(len(expected)) != 20*n {
		return nil, nil, 0, errors.New("gpuhash: expected must hold 20 bytes per piece")
	}
	bitmap = make([]uint64, (n+63)/64)
	digests = make([]byte, 20*n)
	arr, cnt, free := spans(files)
	defer free()
	var exp, dig *C.uint8_t
	var bm *C.uint64_t
	if len(expected) > 0 {
		exp = (*C.uint8_t)(unsafe.Pointer(&expected[0]))
	}
	if n > 0 {
		dig = (*C.uint8_t)(unsafe.Pointer(&digests[0]))
		bm = (*C.uint64_t)(unsafe.Pointer(&bitmap[0]))
	}
	var nbad C.uint64_t
	rc := C.gh_verify_pieces(h.ctx, arr, cnt, C.uint64_t(pieceLen), exp, dig, bm, &nbad)
	return bitmap, digests, uint64(nbad), status(h.ctx, rc, "gh_verify_pieces")
}

// HashParts returns the selected digests of every part of the stream.
func (h *Hasher) HashParts(files []FileSpan, partLen int64, algos Algo) ([]PartSum, error) {
	var total int64
	for _, f := range files {
		total += int64(len(f))
	}
	if partLen <= 0 {
		return nil, errors.New("gpuhash: part length must be > 0")
	}
	n := int((total + partLen - 1) / partLen)
	var s1, m5, s2 []byte
	ptr := func(b []byte) *C.uint8_t {
		if len(b) == 0 {
			return nil
		}
		return (*C.uint8_t)(unsafe.Pointer(&b[0]))
	}
	if algos&SHA1 != 0 {
		s1 = make([]byte, 20*n)
	}
	if algos&MD5 != 0 {
		m5 = make([]byte, 16*n)
	}
	if algos&SHA256 != 0 {
		s2 = make([]byte, 32*n)
	}
	arr, cnt, free := spans(files)
	defer free()
	rc := C.gh_hash_parts(h.ctx, arr, cnt, C.uint64_t(partLen), C.int(algos), ptr(s1), ptr(m5), ptr(s2))
	if err := status(h.ctx, rc, "gh_hash_parts"); err != nil {
		return nil, err
	}
	out := make([]PartSum, n)
	for i := range out {
		if s1 != nil {
			out[i].SHA1 = s1[20*i : 20*i+20]
		}
		if m5 != nil {
			out[i].MD5 = m5[16*i : 16*i+16]
		}
		if s2 != nil {
			out[i].SHA256 = s2[32*i : 32*i+32]
		}
	}
	return out, nil
}

// S3ETag is hex(md5(concat(part md5s))) + "-" + N, the ETag S3 reports for a multipart object.
func (h *Hasher) S3ETag(sums []PartSum) (string, error) {
	cat := make([]byte, 0, 16*len(sums))
	for _, s := range sums {
		cat = append(cat, s.MD5...)
	}
	if len(cat) == 0 {
		return "", errors.New("gpuhash: no parts")
	}
	buf := make([]byte, 64)
	rc := C.gh_s3_etag(h.ctx, (*C.uint8_t)(unsafe.Pointer(&cat[0])), C.uint64_t(len(sums)), (*C.char)(unsafe.Pointer(&buf[0])), 64)
	if err := status(h.ctx, rc, "gh_s3_etag"); err != nil {
		return "", err
	}
	return C.GoString((*C.char)(unsafe.Pointer(&buf[0]))), nil
}

// PinnedAlloc returns page-locked memory (spans inside it are DMA'd without the staging copy).
func PinnedAlloc(n int) ([]byte, error) {
	var p unsafe.Pointer
	if rc := C.gh_host_alloc(&p, C.size_t(n)); rc != C.GH_OK {
		return nil, status(nil, rc, "gh_host_alloc")
	}
	return unsafe.Slice((*byte)(p), n), nil
}

// PinnedFree releases memory obtained from PinnedAlloc.
func PinnedFree(b []byte) {
	if len(b) > 0 {
		C.gh_host_free(unsafe.Pointer(&b[0]))
	}
}
