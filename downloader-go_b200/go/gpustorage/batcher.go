package gpustorage

// Micro-batcher for LIVE pieces: anacrolix hashes a piece the moment its last chunk arrives
// (Torrent.hashPiece, at most ~2 at a time). With a `replace` of anacrolix/torrent whose hashPiece
// calls Batcher.Verify instead of crypto/sha1, completed pieces are copied once into the pinned window
// the library hands out (gh_stream_acquire), and a window is submitted when it is full or when the
// oldest waiting piece has waited MaxDelay. Each caller gets its own piece's verdict, which is what
// Torrent.pieceHashed(index, ok) needs. SOURCE ONLY (no Go toolchain in this image).
//
// Reference seam: the goroutines that run between internal/downloader/torrent/torrent.go:79
// (t.DownloadAll) and :106 (client.WaitAll).

import (
	"io"
	"sync"
	"time"

	"github.com/tritonmedia/downloader-go/internal/gpuhash"
)

type request struct {
	index    int
	r        io.Reader // piece.Storage() section reader, exactly pieceLen bytes (the last piece: fewer)
	expected [20]byte
	done     chan bool
}

// Batcher collects whole pieces of ONE torrent (uniform piece length) into GPU-sized batches.
type Batcher struct {
	h        *gpuhash.Hasher
	pieceLen int64
	MaxDelay time.Duration // flush a partly filled window after this long (default 20 ms)
	MaxBatch int           // pieces per flush (default: what one ring window holds)

	mu      sync.Mutex
	pending []*request
	timer   *time.Timer
}

func NewBatcher(h *gpuhash.Hasher, pieceLen int64) *Batcher {
	return &Batcher{h: h, pieceLen: pieceLen, MaxDelay: 20 * time.Millisecond, MaxBatch: int((256 << 20) / pieceLen)}
}

// Verify blocks until the piece has been hashed on the GPU and reports whether it matched.
func (b *Batcher) Verify(index int, r io.Reader, expected [20]byte) bool {
	req := &request{index: index, r: r, expected: expected, done: make(chan bool, 1)}
	b.mu.Lock()
	b.pending = append(b.pending, req)
	full := len(b.pending) >= b.MaxBatch
	if len(b.pending) == 1 && !full {
		b.timer = time.AfterFunc(b.MaxDelay, b.flush)
	}
	b.mu.Unlock()
	if full {
		b.flush()
	}
	return <-req.done
}

func (b *Batcher) flush() {
	b.mu.Lock()
	batch := b.pending
	b.pending = nil
	if b.timer != nil {
		b.timer.Stop()
		b.timer = nil
	}
	b.mu.Unlock()
	if len(batch) == 0 {
		return
	}
	// Only the torrent's final piece may be short, and it must then be the last message of the
	// stream: keep it for a batch of its own (total = its true length).
	expected := make([]byte, 0, 20*len(batch))
	for _, q := range batch {
		expected = append(expected, q.expected[:]...)
	}
	total := int64(len(batch)) * b.pieceLen
	st, err := b.h.NewStream(gpuhash.SHA1, b.pieceLen, total, expected)
	fail := func() {
		for _, q := range batch {
			q.done <- false // treated as a failed hash: anacrolix re-queues the piece (CPU path stays available)
		}
	}
	if err != nil {
		fail()
		return
	}
	next := 0
	for {
		win, err := st.Acquire()
		if err != nil {
			fail()
			return
		}
		if win == nil {
			break
		}
		for off := int64(0); off < int64(len(win)); off += b.pieceLen {
			if _, err := io.ReadFull(batch[next].r, win[off:off+b.pieceLen]); err != nil {
				fail()
				return
			}
			next++
		}
		if err := st.Submit(win); err != nil {
			fail()
			return
		}
	}
	bitmap, _, _, err := st.Finish()
	if err != nil {
		fail()
		return
	}
	for i, q := range batch {
		q.done <- bitmap[i>>6]>>(uint(i)&63)&1 == 1
	}
}
