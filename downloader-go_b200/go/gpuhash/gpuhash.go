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
	"time"
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

// ---------------------------------------------------------------------------------------------
// File-backed sources, metainfo and the incremental feed (gh_verify_files, gh_hash_file_parts,
// gh_metainfo_*, gh_verify_torrent, gh_stream_*). Same caveat as above: source only.

// File names one file of the payload and the length the metainfo (or stat) promises for it.
type File struct {
	Path string
	Len  int64
}

func cfiles(files []File) (*C.gh_file, C.size_t, func()) {
	n := len(files)
	if n == 0 {
		return nil, 0, func() {}
	}
	arr := (*C.gh_file)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(C.gh_file{}))))
	sl := unsafe.Slice(arr, n)
	for i, f := range files {
		sl[i].path = C.CString(f.Path)
		sl[i].len = C.uint64_t(f.Len)
	}
	return arr, C.size_t(n), func() {
		for i := range sl {
			C.free(unsafe.Pointer(sl[i].path))
		}
		C.free(unsafe.Pointer(arr))
	}
}

// VerifyFiles is VerifyPieces for payload on disk: the library preads the files into its pinned
// ring itself, so no payload byte crosses cgo.
func (h *Hasher) VerifyFiles(files []File, pieceLen int64, expected []byte) (bitmap []uint64, digests []byte, bad uint64, err error) {
	var total int64
	for _, f := range files {
		total += f.Len
	}
	if pieceLen <= 0 {
		return nil, nil, 0, errors.New("gpuhash: piece length must be > 0")
	}
	n := (total + pieceLen - 1) / pieceLen
	bitmap = make([]uint64, (n+63)/64)
	digests = make([]byte, 20*n)
	arr, cnt, free := cfiles(files)
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
	rc := C.gh_verify_files(h.ctx, arr, cnt, C.uint64_t(pieceLen), exp, dig, bm, &nbad)
	return bitmap, digests, uint64(nbad), status(h.ctx, rc, "gh_verify_files")
}

// HashFileParts is HashParts for files on disk (what uploader.go:73-89 hands to minio).
func (h *Hasher) HashFileParts(files []File, partLen int64, algos Algo) ([]PartSum, error) {
	var total int64
	for _, f := range files {
		total += f.Len
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
	arr, cnt, free := cfiles(files)
	defer free()
	rc := C.gh_hash_file_parts(h.ctx, arr, cnt, C.uint64_t(partLen), C.int(algos), ptr(s1), ptr(m5), ptr(s2))
	if err := status(h.ctx, rc, "gh_hash_file_parts"); err != nil {
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

// HashFilePartsPerFile hashes the parts of every file of an upload job in ONE batch; parts restart at
// every file (each file is its own S3 object). sums[i] holds file i's parts.
func (h *Hasher) HashFilePartsPerFile(files []File, partLen int64, algos Algo) (sums [][]PartSum, err error) {
	if partLen <= 0 {
		return nil, errors.New("gpuhash: part length must be > 0")
	}
	counts := make([]int, len(files))
	n := 0
	for i, f := range files {
		counts[i] = int((f.Len + partLen - 1) / partLen)
		n += counts[i]
	}
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
	arr, cnt, free := cfiles(files)
	defer free()
	rc := C.gh_hash_file_parts_per_file(h.ctx, arr, cnt, C.uint64_t(partLen), C.int(algos), ptr(s1), ptr(m5), ptr(s2))
	if err := status(h.ctx, rc, "gh_hash_file_parts_per_file"); err != nil {
		return nil, err
	}
	sums = make([][]PartSum, len(files))
	k := 0
	for i, c := range counts {
		sums[i] = make([]PartSum, c)
		for j := 0; j < c; j, k = j+1, k+1 {
			if s1 != nil {
				sums[i][j].SHA1 = s1[20*k : 20*k+20]
			}
			if m5 != nil {
				sums[i][j].MD5 = m5[16*k : 16*k+16]
			}
			if s2 != nil {
				sums[i][j].SHA256 = s2[32*k : 32*k+32]
			}
		}
	}
	return sums, nil
}

// Metainfo is a parsed BEP 3 info dictionary.
type Metainfo struct{ m *C.gh_metainfo }

// ParseMetainfo accepts a whole .torrent file or a bare info dictionary.
func ParseMetainfo(b []byte) (*Metainfo, error) {
	if len(b) == 0 {
		return nil, errors.New("gpuhash: empty metainfo")
	}
	var m *C.gh_metainfo
	errbuf := make([]byte, 256)
	rc := C.gh_metainfo_parse((*C.uint8_t)(unsafe.Pointer(&b[0])), C.size_t(len(b)), &m, (*C.char)(unsafe.Pointer(&errbuf[0])), 256)
	if rc != C.GH_OK {
		return nil, fmt.Errorf("gpuhash: bad metainfo: %s", C.GoString((*C.char)(unsafe.Pointer(&errbuf[0]))))
	}
	mi := &Metainfo{m}
	runtime.SetFinalizer(mi, func(x *Metainfo) { C.gh_metainfo_free(x.m) })
	return mi, nil
}

func (m *Metainfo) PieceLength() int64 { return int64(C.gh_metainfo_piece_length(m.m)) }
func (m *Metainfo) NumPieces() int     { return int(C.gh_metainfo_n_pieces(m.m)) }
func (m *Metainfo) Pieces() []byte {
	return C.GoBytes(unsafe.Pointer(C.gh_metainfo_pieces(m.m)), C.int(20*m.NumPieces()))
}

// VerifyTorrent verifies the files laid out under baseDir (anacrolix file-storage layout) against
// the metainfo's piece hashes in one batched call.
func (h *Hasher) VerifyTorrent(m *Metainfo, baseDir string) (bitmap []uint64, digests []byte, bad uint64, err error) {
	n := m.NumPieces()
	bitmap = make([]uint64, (n+63)/64)
	digests = make([]byte, 20*n)
	dir := C.CString(baseDir)
	defer C.free(unsafe.Pointer(dir))
	var dig *C.uint8_t
	var bm *C.uint64_t
	if n > 0 {
		dig = (*C.uint8_t)(unsafe.Pointer(&digests[0]))
		bm = (*C.uint64_t)(unsafe.Pointer(&bitmap[0]))
	}
	var nbad C.uint64_t
	rc := C.gh_verify_torrent(h.ctx, m.m, dir, dig, bm, &nbad)
	return bitmap, digests, uint64(nbad), status(h.ctx, rc, "gh_verify_torrent")
}

// Stream feeds a payload window by window into the library's pinned ring.
type Stream struct {
	h  *Hasher
	st *C.gh_stream
	n  int64
	al Algo
}

func (h *Hasher) NewStream(algos Algo, msgLen, total int64, expected []byte) (*Stream, error) {
	var exp *C.uint8_t
	if len(expected) > 0 {
		exp = (*C.uint8_t)(unsafe.Pointer(&expected[0]))
	}
	s := &Stream{h: h, al: algos, n: (total + msgLen - 1) / msgLen}
	rc := C.gh_stream_open(h.ctx, C.int(algos), C.uint64_t(msgLen), C.uint64_t(total), exp, &s.st)
	return s, status(h.ctx, rc, "gh_stream_open")
}

// Acquire returns the next pinned window (C memory, valid until Submit) or nil when all bytes were fed.
func (s *Stream) Acquire() ([]byte, error) {
	var buf unsafe.Pointer
	var cap_ C.uint64_t
	if rc := C.gh_stream_acquire(s.st, &buf, &cap_); rc != C.GH_OK {
		return nil, status(s.h.ctx, rc, "gh_stream_acquire")
	}
	if cap_ == 0 {
		return nil, nil
	}
	return unsafe.Slice((*byte)(buf), int(cap_)), nil
}

func (s *Stream) Submit(window []byte) error {
	return status(s.h.ctx, C.gh_stream_submit(s.st, C.uint64_t(len(window))), "gh_stream_submit")
}

// Finish waits for the device and returns the SHA-1 digests, the pass bitmap and the number of
// failed pieces (for streams opened with SHA1 and expected hashes).
func (s *Stream) Finish() (bitmap []uint64, sha1 []byte, bad uint64, err error) {
	bitmap = make([]uint64, (s.n+63)/64)
	var d *C.uint8_t
	if s.al&SHA1 != 0 && s.n > 0 {
		sha1 = make([]byte, 20*s.n)
		d = (*C.uint8_t)(unsafe.Pointer(&sha1[0]))
	}
	var bm *C.uint64_t
	if s.n > 0 {
		bm = (*C.uint64_t)(unsafe.Pointer(&bitmap[0]))
	}
	var nbad C.uint64_t
	rc := C.gh_stream_finish(s.st, d, nil, nil, bm, &nbad)
	return bitmap, sha1, uint64(nbad), status(s.h.ctx, rc, "gh_stream_finish")
}

// Abort drops a stream that cannot be completed (an error from Acquire / Submit, or a reader that
// failed): the context is usable again afterwards.
func (s *Stream) Abort() {
	if s.st != nil {
		C.gh_stream_abort(s.st)
		s.st = nil
	}
}

// NewFromEnv creates a hasher configured the way the reference configures everything else -- from
// the environment (GPUHASH_DEVICES, GPUHASH_SLOT_MIB, GPUHASH_SLOTS, GPUHASH_STAGE_THREADS; see
// gh_config_from_env). It returns ErrDisabled when GPUHASH_DISABLE is set: the adapters then leave the
// stock CPU paths of anacrolix / minio-go in place.
func NewFromEnv() (*Hasher, error) {
	var cfg C.gh_config
	var devs [64]C.int
	var disabled C.int
	if rc := C.gh_config_from_env(&cfg, &devs[0], 64, &disabled); rc != C.GH_OK {
		return nil, status(nil, rc, "gh_config_from_env")
	}
	if disabled != 0 {
		return nil, ErrDisabled
	}
	h := &Hasher{}
	rc := C.gh_create(&h.ctx, &cfg)
	runtime.KeepAlive(&devs)
	if err := status(nil, rc, "gh_create"); err != nil {
		return nil, err
	}
	runtime.SetFinalizer(h, (*Hasher).Close)
	return h, nil
}

// ErrDisabled: GPUHASH_DISABLE is set in the environment.
var ErrDisabled = fmt.Errorf("gpuhash: disabled by GPUHASH_DISABLE")

// Batcher is gh_batcher_*: many goroutines hand over one piece each and block for its verdict.
type Batcher struct {
	h *Hasher
	b *C.gh_batcher
}

// Slot is a reserved place in the pinned batch buffer.
type Slot struct{ s C.gh_batch_slot }

func (h *Hasher) NewBatcher(pieceLen int64, maxPieces int, maxDelay time.Duration) (*Batcher, error) {
	b := &Batcher{h: h}
	rc := C.gh_batcher_create(h.ctx, C.uint64_t(pieceLen), C.uint32_t(maxPieces), C.uint32_t(maxDelay/time.Millisecond), &b.b)
	return b, status(h.ctx, rc, "gh_batcher_create")
}

// Reserve returns a slot and its bytes (C memory, pinned): read the piece into them, then Commit or Cancel.
func (b *Batcher) Reserve(length int64) (*Slot, []byte, error) {
	var s Slot
	var buf unsafe.Pointer
	if rc := C.gh_batcher_reserve(b.b, C.uint64_t(length), &s.s, &buf); rc != C.GH_OK {
		return nil, nil, status(b.h.ctx, rc, "gh_batcher_reserve")
	}
	return &s, unsafe.Slice((*byte)(buf), int(length)), nil
}

// Commit blocks until the batch has run. A non-nil error means "verdict unknown": hash on the CPU.
func (b *Batcher) Commit(s *Slot, expected [20]byte) (bool, error) {
	var ok C.int
	rc := C.gh_batcher_commit(b.b, &s.s, (*C.uint8_t)(unsafe.Pointer(&expected[0])), &ok, nil)
	return ok == 1, status(b.h.ctx, rc, "gh_batcher_commit")
}

func (b *Batcher) Cancel(s *Slot) { C.gh_batcher_cancel(b.b, &s.s) }

func (b *Batcher) Close() {
	if b.b != nil {
		C.gh_batcher_destroy(b.b)
		b.b = nil
	}
}

// SuggestDevices measures which n of the visible GPUs give the best aggregate host-to-device
// bandwidth (GPUs behind one PCIe switch share its uplink) and returns their ordinals for New.
func SuggestDevices(n int) ([]int, float64, error) {
	devs := make([]C.int, n)
	var gbs C.double
	if rc := C.gh_suggest_devices(C.int(n), &devs[0], &gbs); rc != C.GH_OK {
		return nil, 0, status(nil, rc, "gh_suggest_devices")
	}
	out := make([]int, n)
	for i, d := range devs {
		out[i] = int(d)
	}
	return out, float64(gbs), nil
}
