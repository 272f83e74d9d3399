// Package gpustorage wraps anacrolix/torrent's file storage so that pieces already on disk are
// verified in one batched GPU call when a torrent is (re)opened.
//
// SOURCE ONLY (no Go toolchain in this image). Seam: internal/downloader/torrent/torrent.go:41
//
//	clientConfig.DefaultStorage = gpustorage.Wrap(storage.NewFile(baseDir), hasher, baseDir)
//
// Nothing else in the reference changes: torrent.go:79 (DownloadAll) and :106 (WaitAll) behave as
// before, but pieces whose Completion() is already {Complete: true, Ok: true} are neither
// re-downloaded nor re-hashed by anacrolix's own hashPiece.
package gpustorage

import (
	"bytes"

	"github.com/anacrolix/torrent/bencode"
	"github.com/anacrolix/torrent/metainfo"
	"github.com/anacrolix/torrent/storage"

	"github.com/tritonmedia/downloader-go/internal/gpuhash"
)

type client struct {
	inner   storage.ClientImpl
	hasher  *gpuhash.Hasher
	baseDir string
}

// Wrap returns a storage.ClientImpl that delegates all I/O to inner.
func Wrap(inner storage.ClientImpl, h *gpuhash.Hasher, baseDir string) storage.ClientImpl {
	return &client{inner, h, baseDir}
}

func (c *client) Close() error { return c.inner.Close() }

func (c *client) OpenTorrent(info *metainfo.Info, infoHash metainfo.Hash) (storage.TorrentImpl, error) {
	t, err := c.inner.OpenTorrent(info, infoHash)
	if err != nil {
		return nil, err
	}
	var buf bytes.Buffer
	if err := bencode.NewEncoder(&buf).Encode(info); err != nil {
		return t, nil // cannot describe the torrent to the library: fall through to the stock path
	}
	mi, err := gpuhash.ParseMetainfo(buf.Bytes())
	if err != nil {
		return t, nil
	}
	bitmap, _, _, err := c.hasher.VerifyTorrent(mi, c.baseDir)
	if err != nil {
		return t, nil // GPU unavailable or failed: anacrolix verifies on the CPU as it always did
	}
	return &torrentImpl{TorrentImpl: t, ok: bitmap}, nil
}

type torrentImpl struct {
	storage.TorrentImpl
	ok []uint64
}

func (t *torrentImpl) Piece(p metainfo.Piece) storage.PieceImpl {
	return &pieceImpl{PieceImpl: t.TorrentImpl.Piece(p), ok: t.ok, index: p.Index()}
}

type pieceImpl struct {
	storage.PieceImpl
	ok    []uint64
	index int
}

// Completion reports a piece as complete when the batched verification proved it; everything else
// is left to the inner storage (and therefore to anacrolix's normal download + hash cycle).
func (p *pieceImpl) Completion() storage.Completion {
	if p.ok[p.index>>6]>>(uint(p.index)&63)&1 == 1 {
		return storage.Completion{Complete: true, Ok: true}
	}
	return p.PieceImpl.Completion()
}

func (p *pieceImpl) MarkNotComplete() error {
	p.ok[p.index>>6] &^= 1 << (uint(p.index) & 63)
	return p.PieceImpl.MarkNotComplete()
}
