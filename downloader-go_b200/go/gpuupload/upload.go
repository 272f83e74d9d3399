// Package gpuupload is the uploader-side adapter: per-part MD5 / SHA-256 computed in one batched GPU
// call, then handed to minio's low-level multipart API, which is the one upstream entry point that
// accepts precomputed sums.
//
// SOURCE ONLY (no Go toolchain in this image). Seam: internal/uploader/uploader.go:89, where
//
//	c.s3client.PutObjectWithContext(ctx, c.bucket, key, f, stats.Size(), minio.PutObjectOptions{})
//
// becomes  gpuupload.Put(ctx, core, hasher, c.bucket, key, fileName, stats.Size()).
package gpuupload

import (
	"context"
	"encoding/base64"
	"encoding/hex"
	"fmt"
	"io"
	"os"

	minio "github.com/minio/minio-go/v6"

	"github.com/tritonmedia/downloader-go/internal/gpuhash"
)

// PartSize is the part length the sums are computed over (BASELINE.json config 3: 64 MiB).
const PartSize = 64 << 20

// PutAll uploads every file of a job: the part sums of ALL files come from one batched GPU call
// (parts restart at every file), then each file goes up through the low-level multipart API.
// keys[i] is the object key of paths[i]. Errors are returned per file, like the reference's loop
// (uploader.go:72-92) which logs and continues.
func PutAll(ctx context.Context, core minio.Core, h *gpuhash.Hasher, bucket string, keys, paths []string, sizes []int64) []error {
	errs := make([]error, len(paths))
	files := make([]gpuhash.File, len(paths))
	for i := range paths {
		files[i] = gpuhash.File{Path: paths[i], Len: sizes[i]}
	}
	sums, err := h.HashFilePartsPerFile(files, PartSize, gpuhash.MD5|gpuhash.SHA256)
	for i := range paths {
		if err != nil {
			errs[i] = err
			continue
		}
		errs[i] = putWithSums(ctx, core, h, bucket, keys[i], paths[i], sizes[i], sums[i])
	}
	return errs
}

// Put uploads one file. Files smaller than one part take minio's ordinary single PUT.
func Put(ctx context.Context, core minio.Core, h *gpuhash.Hasher, bucket, key, path string, size int64) error {
	if size < PartSize {
		return putWithSums(ctx, core, h, bucket, key, path, size, nil)
	}
	sums, err := h.HashFileParts([]gpuhash.File{{Path: path, Len: size}}, PartSize, gpuhash.MD5|gpuhash.SHA256)
	if err != nil {
		return err
	}
	return putWithSums(ctx, core, h, bucket, key, path, size, sums)
}

func putWithSums(ctx context.Context, core minio.Core, h *gpuhash.Hasher, bucket, key, path string, size int64, sums []gpuhash.PartSum) error {
	f, err := os.Open(path)
	if err != nil {
		return err
	}
	defer f.Close()
	if size < PartSize {
		_, err = core.Client.PutObjectWithContext(ctx, bucket, key, f, size, minio.PutObjectOptions{})
		return err
	}
	uploadID, err := core.NewMultipartUpload(bucket, key, minio.PutObjectOptions{})
	if err != nil {
		return err
	}
	parts := make([]minio.CompletePart, len(sums))
	for i, s := range sums {
		off := int64(i) * PartSize
		n := size - off
		if n > PartSize {
			n = PartSize
		}
		op, err := core.PutObjectPart(bucket, key, uploadID, i+1, io.NewSectionReader(f, off, n), n,
			base64.StdEncoding.EncodeToString(s.MD5), hex.EncodeToString(s.SHA256), nil)
		if err != nil {
			_ = core.AbortMultipartUpload(bucket, key, uploadID)
			return err
		}
		if op.ETag != hex.EncodeToString(s.MD5) {
			_ = core.AbortMultipartUpload(bucket, key, uploadID)
			return fmt.Errorf("gpuupload: part %d ETag %s differs from the MD5 sent", i+1, op.ETag)
		}
		parts[i] = minio.CompletePart{PartNumber: i + 1, ETag: op.ETag}
	}
	want, err := h.S3ETag(sums)
	if err != nil {
		return err
	}
	etag, err := core.CompleteMultipartUpload(bucket, key, uploadID, parts)
	if err == nil && etag != want {
		err = fmt.Errorf("gpuupload: object ETag %s, expected %s", etag, want)
	}
	return err
}
