package gpustorage

// Micro-batcher for LIVE pieces. anacrolix hashes a piece the moment its last chunk arrives
// (Torrent.hashPiece, a few goroutines at a time). With a `replace` of anacrolix/torrent whose
// hashPiece calls Batcher.Verify instead of crypto/sha1, every such goroutine blocks in ONE cgo call
// until its own piece's verdict is known, and the library (gh_batcher_*, include/gpuhash.h) turns the
// concurrent calls into GPU batches: flush on MaxPieces waiting or on MaxDelay. All batching logic --
// arbitrary piece order, the short last piece, one flusher at a time, error propagation -- lives below
// the C ABI where it is tested (tests/test_batcher.py); this file only moves bytes. SOURCE ONLY (no Go
// toolchain in this image).
//
// Reference seam: the goroutines that run between internal/downloader/torrent/torrent.go:79
// (t.DownloadAll) and :106 (client.WaitAll). Raise ClientConfig.PieceHashersPerTorrent to about
// MaxPieces: a batch is only as large as the number of goroutines blocked in Verify.

import (
	"crypto/sha1"
	"io"
	"time"

	"github.com/tritonmedia/downloader-go/internal/gpuhash"
)

// Batcher verifies whole pieces of ONE torrent (uniform piece length; the last piece may be shorter).
type Batcher struct {
	b        *gpuhash.Batcher
	pieceLen int64
}

// NewBatcher: flush when maxPieces are waiting or the oldest has waited maxDelay.
func NewBatcher(h *gpuhash.Hasher, pieceLen int64, maxPieces int, maxDelay time.Duration) (*Batcher, error) {
	b, err := h.NewBatcher(pieceLen, maxPieces, maxDelay)
	if err != nil {
		return nil, err
	}
	return &Batcher{b: b, pieceLen: pieceLen}, nil
}

// Verify reads the piece (length bytes, at most pieceLen) from r straight into the pinned batch
// buffer, blocks until the batch has been hashed, and reports whether the piece matched `expected`.
// If the GPU path fails -- any error from the library -- the piece is hashed here with crypto/sha1,
// exactly what anacrolix would have done: a library failure is never reported as a bad piece.
func (b *Batcher) Verify(r io.ReaderAt, length int64, expected [20]byte) (bool, error) {
	slot, buf, err := b.b.Reserve(length)
	if err != nil {
		return cpuVerify(r, length, expected)
	}
	if _, err := r.ReadAt(buf, 0); err != nil && err != io.EOF {
		b.b.Cancel(slot) // give the slot back: the batch must not wait for it
		return false, err
	}
	ok, err := b.b.Commit(slot, expected)
	if err != nil {
		return cpuVerify(r, length, expected)
	}
	return ok, nil
}

func cpuVerify(r io.ReaderAt, length int64, expected [20]byte) (bool, error) {
	h := sha1.New()
	if _, err := io.Copy(h, io.NewSectionReader(r, 0, length)); err != nil {
		return false, err
	}
	var sum [20]byte
	copy(sum[:], h.Sum(nil))
	return sum == expected, nil
}

// Close flushes what is waiting and releases the pinned buffers.
func (b *Batcher) Close() { b.b.Close() }
