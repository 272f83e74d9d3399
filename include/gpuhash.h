/*
 * gpuhash.h -- C ABI of libgpuhash.so, the B200 (sm_100a) batched content hasher.
 *
 * This is the drop-in boundary for the one data-parallel path of tritonmedia/downloader-go:
 *   - SHA-1 verification of torrent pieces, executed by anacrolix/torrent v1.15.2 (go.mod:6) between
 *     reference internal/downloader/torrent/torrent.go:79 (t.DownloadAll) and :106 (client.WaitAll);
 *     the storage seam is torrent.go:41 (clientConfig.DefaultStorage = storage.NewFile(baseDir)).
 *   - per-part MD5 (ETag / Content-MD5) and SHA-256 (SigV4 payload hash), executed by
 *     minio-go/v6 v6.0.55 (go.mod:11) under reference internal/uploader/uploader.go:89
 *     (s3client.PutObjectWithContext).
 * The reference defines no hashing interface of its own, so each entry point
 * below names the third-party call it replaces and the reference line that reaches it.
 *
 * Conventions
 *   - every function returns an int status, GH_OK == 0; gh_strerror() maps it to a static string.
 *   - plain pointers and sizes only; no C++/torch types; no exceptions cross the boundary.
 *   - the caller owns every buffer it passes; one-shot calls are synchronous and never retain a
 *     caller pointer after returning (the cgo pointer rule). Long-lived memory (staging ring,
 *     midstates) is allocated by the library.
 *   - a gh_ctx may be used from several OS threads; calls are serialised per context. The library
 *     calls cudaSetDevice itself on entry (goroutines migrate between OS threads) and gives the
 *     calling thread its previous current device back on return, so a host that shares the CUDA
 *     runtime with the library is not moved to another GPU by a gh_* call.
 *   - there is NO CPU fallback: without a usable CUDA device gh_create fails with GH_ENODEVICE.
 *   - digests are raw bytes in the standard byte order (what Go's hash.Hash.Sum returns).
 *   - bitmaps: bit (i & 63) of word (i >> 6) is 1 iff piece i matched its expected SHA-1.
 */
#ifndef GPUHASH_H
#define GPUHASH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gh_ctx gh_ctx;

enum {
    GH_OK = 0,
    GH_EINVAL = 1,     /* bad argument (NULL where data is required, zero piece length, ...) */
    GH_ENODEVICE = 2,  /* no CUDA device / driver, or requested ordinal absent */
    GH_ENOMEM = 3,     /* host or device allocation failed */
    GH_ECUDA = 4,      /* a CUDA call failed; gh_last_error(ctx) has the text */
    GH_ECTX_LOST = 5,  /* sticky device error: destroy the context */
    GH_ERANGE = 6,     /* size beyond a documented limit */
    GH_EIO = 7         /* a payload file could not be opened or read (EACCES, EMFILE, EIO ...); gh_last_error names it.
                          A file that does not exist (ENOENT) or is shorter than promised is NOT an error: its
                          missing bytes read as zeros and the pieces over them fail verification */
};

enum { GH_SHA1 = 1, GH_MD5 = 2, GH_SHA256 = 4 }; /* algos bitmask; GH_MD5|GH_SHA256 = one fused pass */

#define GH_STATE_WORDS 17 /* uint32 midstate words per message in a d_state buffer (all three algorithms) */

typedef struct {
    int        n_devices;     /* 0 = every visible device */
    const int* devices;       /* CUDA ordinals, NULL = 0..n_devices-1 */
    uint64_t   slot_bytes;    /* device bytes per ring slot, 0 = default (256 MiB), minimum 1 MiB */
    int        slots;         /* ring slots per device, 0 = default (4), minimum 2 */
    int        stage_threads; /* host threads copying pageable spans / reading files into the pinned ring, 0 = default (one per core, 4..16) */
} gh_config;

/* A run of stream bytes in host memory. The logical stream is the concatenation of the spans, in
 * order: exactly BEP 3's "files concatenated in info order" (what anacrolix/torrent/storage ReadAt
 * presents to Torrent.hashPiece), or one file for the uploader (uploader.go:73-89). */
typedef struct {
    const void* host_ptr;
    uint64_t    len;
} gh_span;

typedef struct {
    uint64_t kernel_launches;  /* hashing kernels launched by this context so far */
    uint64_t h2d_bytes;        /* payload bytes copied host -> device */
    uint64_t d2h_bytes;        /* result bytes copied device -> host */
    double   kernel_ms;        /* CUDA-event time of the hashing kernels of the LAST call, summed */
    double   wall_ms;          /* host wall time of the LAST call */
} gh_stats;

/* Configuration from the environment, following the reference's convention (it reads RABBITMQ_ENDPOINT,
 * S3_ENDPOINT, S3_ACCESS_KEY ... from the environment: cmd/downloader/downloader.go:54-58,
 * internal/uploader/uploader.go:25, minio_credential_provider.go:24-25):
 *   GPUHASH_DEVICES        comma-separated CUDA ordinals ("0,2,4,6")   -> cfg->devices (stored in devices_buf)
 *   GPUHASH_SLOT_MIB       MiB per ring slot                           -> cfg->slot_bytes
 *   GPUHASH_SLOTS          ring slots per device (>= 2)                -> cfg->slots
 *   GPUHASH_STAGE_THREADS  host threads staging pageable spans / files -> cfg->stage_threads
 *   GPUHASH_DISABLE        non-empty and not 0/false/no                -> *disabled = 1
 * Unset or empty variables leave the field 0 (= library default). A malformed value is GH_EINVAL (more
 * devices than devices_cap: GH_ERANGE) and nothing half-parsed is used. The library itself never
 * computes on the CPU: `disabled` is for the adapters, which then do not install themselves and leave
 * the stock crypto/sha1 / crypto/md5 paths in place. Host-only; needs no device. */
int gh_config_from_env(gh_config* cfg, int* devices_buf, int devices_cap, int* disabled);

/* Which n of the visible devices to give a context (no reference counterpart; the reference has no
 * devices). The path is bound by the host->device link end to end, and GPUs that share a PCIe uplink
 * share its bandwidth, so the choice is measured: pinned cudaMemcpyAsync on all devices of a candidate
 * set at once (~0.4 s per candidate), for the first n devices and, when at least 2n are visible, for
 * an evenly strided set; the plain first-n choice stays unless the other is clearly (3 %) better.
 * devices_out receives n ordinals for gh_config.devices / GPUHASH_DEVICES; h2d_gbs_out (optional) the
 * aggregate GB/s measured for them. Allocates 128 MiB of pinned and of device memory per probed device
 * for the duration of the call. */
int gh_suggest_devices(int n, int* devices_out, double* h2d_gbs_out);

/* ---- lifecycle ---------------------------------------------------------------------------- */
int         gh_create(gh_ctx** out, const gh_config* cfg /* NULL = defaults */);
void        gh_destroy(gh_ctx* ctx);
const char* gh_strerror(int status);
const char* gh_last_error(gh_ctx* ctx);           /* detail of the last failure on this context; the pointer stays
                                                     valid until the calling thread's next gh_last_error. ctx == NULL:
                                                     the calling thread's last failure in a context-free call
                                                     (gh_host_alloc, gh_host_register) */
int         gh_device_count(gh_ctx* ctx);
int         gh_get_stats(gh_ctx* ctx, gh_stats* out);
const char* gh_version(void);
/* Name of the hashing kernel variant the calling thread launched last ("gh_hash_trio",
 * "gh_hash_lanes<1, 1, 1, 4>", ...; "" before the first launch). Diagnostic: which of the kernels of
 * DESIGN.md section 4 a batch shape was routed to. No reference counterpart. */
const char* gh_last_kernel(void);

/* Error contract of the calls below: a call that fails has first waited for every copy and kernel it
 * had queued, so no DMA is still reading the caller's buffers when it returns, and it leaves the
 * context usable for the next call -- unless the status is GH_ECTX_LOST (a device fault: destroy the
 * context). Outputs of a failed call are unspecified. */

/* Test hook (no reference counterpart): the n-th payload copy (host->device, counted from 0 over
 * the following calls on this context) fails with GH_ECUDA as if the CUDA runtime had refused it;
 * n < 0 disarms. Exists so that the error contract above can be exercised on a healthy GPU. */
int gh_debug_fail_nth_copy(gh_ctx* ctx, int64_t n);

/* ---- host-buffer (end-to-end) entry points -------------------------------------------------- */

/* Batched piece verification. Replaces, for a whole batch of pieces at once, anacrolix/torrent
 * Torrent.hashPiece + pieceHashed (crypto/sha1 over piece.Storage(), compare with
 * info.Pieces[20i:20i+20]) that run under reference torrent.go:79-106.
 * piece i = stream[i*piece_len, min((i+1)*piece_len, total)), only the last piece may be short.
 * expected: 20*n_pieces bytes or NULL (then only digests are produced and every bit is 0).
 * digests_out: 20*n_pieces bytes or NULL. ok_bitmap_out: ceil(n_pieces/64) words or NULL.
 * n_bad_out: number of pieces whose digest differs from expected, or NULL.
 * Pieces are sharded over the context's devices in contiguous, 64-piece-aligned ranges. */
int gh_verify_pieces(gh_ctx* ctx, const gh_span* spans, size_t n_spans, uint64_t piece_len,
                     const uint8_t* expected, uint8_t* digests_out, uint64_t* ok_bitmap_out,
                     uint64_t* n_bad_out);

/* Batched per-part hashing. Replaces minio-go hashMaterials (crypto/md5 + sha256-simd over one
 * buffered part) that runs under reference uploader.go:89; the sums are what
 * minio.Core.PutObjectPart(..., md5Base64, sha256Hex, ...) accepts precomputed.
 * part k = stream[k*part_len, min((k+1)*part_len, total)). algos is a GH_* bitmask; outputs for
 * unselected algorithms may be NULL. sha1_out 20*n, md5_out 16*n, sha256_out 32*n bytes. */
int gh_hash_parts(gh_ctx* ctx, const gh_span* spans, size_t n_spans, uint64_t part_len, int algos,
                  uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out);

/* The same two calls for payload that lives in FILES: the library preads the files into its pinned
 * ring itself (in parallel, page-cache reads), so no payload byte crosses the FFI or needs a
 * caller-side buffer. This is the shape of both reference call sites: anacrolix's file storage under
 * baseDir (torrent.go:41) read back by hashPiece, and the *os.File handed to PutObjectWithContext
 * (uploader.go:73-89). `len` is the length the metainfo / stat promises; a file that is missing or
 * shorter reads as zeros from there on, so exactly the pieces that touch the hole fail. */
typedef struct {
    const char* path;
    uint64_t    len;
} gh_file;
int gh_verify_files(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t piece_len,
                    const uint8_t* expected, uint8_t* digests_out, uint64_t* ok_bitmap_out,
                    uint64_t* n_bad_out);
int gh_hash_file_parts(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t part_len, int algos,
                       uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out);

/* ---- incremental feed ------------------------------------------------------------------------- */
/* For payload that arrives a window at a time (pieces as the swarm completes them, an HTTP body):
 * the caller fills the library's pinned ring buffers in place, so the bytes are written once and
 * DMA'd from where they landed. This is the micro-batching seam for anacrolix's per-piece hashPiece
 * calls (torrent.go:79-106): flush every `cap` bytes instead of hashing piece by piece.
 *   open    -> fixes algos, message length, total length and (optionally) the expected SHA-1s
 *   acquire -> a pinned buffer and the number of bytes it must receive: whole messages, as many as
 *              fit a ring slot (the last window holds the remainder). buf == NULL, cap == 0 when
 *              everything has been fed. Blocks while the ring is full.
 *   submit  -> the buffer holds exactly `cap` bytes: queue its DMA and kernel, return at once
 *   finish  -> wait, return digests / bitmap / n_bad (same meaning as gh_verify_pieces), close
 *   abort   -> drop the stream
 * Messages longer than a ring slot are refused with GH_ERANGE (sequential feeding would leave one
 * chain per window; gh_hash_parts / gh_hash_file_parts interleave such messages instead). A context
 * carries one stream at a time and refuses one-shot calls while it is open. Windows go round the
 * context's devices (window w on device w mod n_devices, each with its own pinned ring), so a feed
 * that outruns one GPU's link spreads over all of them; finish gathers digests and bitmap from the
 * devices that produced them. After an error from submit or finish the stream can only be aborted. */
typedef struct gh_stream gh_stream;
int  gh_stream_open(gh_ctx* ctx, int algos, uint64_t msg_len, uint64_t total_len, const uint8_t* expected,
                    gh_stream** out);
int  gh_stream_acquire(gh_stream* st, void** buf, uint64_t* cap);
int  gh_stream_submit(gh_stream* st, uint64_t n_bytes);
int  gh_stream_finish(gh_stream* st, uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out,
                      uint64_t* ok_bitmap_out, uint64_t* n_bad_out);
void gh_stream_abort(gh_stream* st);

/* ---- micro-batcher for live pieces ----------------------------------------------------------------- */
/* The seam for a LIVE download: anacrolix/torrent calls Torrent.hashPiece for one piece at a time, from
 * up to PieceHashersPerTorrent goroutines, as pieces complete between torrent.go:79 (t.DownloadAll)
 * and :106 (client.WaitAll). gh_batcher_verify is what a replaced hashPiece calls instead of
 * crypto/sha1: it blocks until THIS piece's verdict is known, while the batcher turns the concurrent
 * calls into GPU batches -- a batch runs when max_pieces are waiting or the oldest has waited
 * max_delay_ms. Pieces may come in any order; the torrent's short last piece is just a smaller `len`.
 * Any number of threads may call concurrently (raise PieceHashersPerTorrent to max_pieces or more: a
 * batch is only as large as the number of callers that are blocked in it).
 *
 * Status: GH_OK = the verdict is valid (*ok_out 1 = digest matched expected20, 0 = mismatch;
 * digest_out, optional, receives the 20-byte SHA-1). ANY OTHER status means the verdict is unknown --
 * the GPU batch failed, every caller of that batch gets the same status -- and the caller must fall
 * back to hashing the piece on the CPU. It is never reported as a mismatch, which would make the
 * torrent client re-download good data and penalise the peer that sent it.
 *
 * gh_batcher_reserve / gh_batcher_commit are the same in two steps for callers that can produce the
 * piece in place: reserve returns a slot of the pinned batch buffer, the caller reads the piece
 * straight into it (io.ReadFull from the piece's storage reader), commit blocks for the verdict.
 * A reservation that cannot be filled must be given back with gh_batcher_cancel: a batch does not run
 * before every reserved slot is committed or cancelled.
 *
 * Batches run on all devices of the context (sharded like gh_verify_pieces) and are serialised with
 * the context's other calls. gh_batcher_destroy flushes what is waiting and returns when every
 * blocked caller has its answer; calls made after it has started are refused with GH_EINVAL. */
typedef struct gh_batcher gh_batcher;
typedef struct {
    uint32_t buffer, index; /* opaque: which slot of which batch buffer */
    uint64_t seq;
} gh_batch_slot;
typedef struct {
    uint64_t batches, pieces, flushed_full, flushed_by_deadline, failed_batches;
} gh_batcher_stats;
int  gh_batcher_create(gh_ctx* ctx, uint64_t piece_len, uint32_t max_pieces, uint32_t max_delay_ms, gh_batcher** out);
int  gh_batcher_verify(gh_batcher* b, const void* data, uint64_t len, const uint8_t* expected20, int* ok_out,
                       uint8_t* digest_out);
int  gh_batcher_reserve(gh_batcher* b, uint64_t len, gh_batch_slot* slot, void** buf);
int  gh_batcher_commit(gh_batcher* b, const gh_batch_slot* slot, const uint8_t* expected20, int* ok_out,
                       uint8_t* digest_out);
int  gh_batcher_cancel(gh_batcher* b, const gh_batch_slot* slot);
int  gh_batcher_flush(gh_batcher* b);   /* run what is waiting now, without waiting for count or deadline */
int  gh_batcher_get_stats(gh_batcher* b, gh_batcher_stats* out);
void gh_batcher_destroy(gh_batcher* b);

/* ---- metainfo (BEP 3 .torrent / info dictionary) ------------------------------------------------- */
/* The data format on the input side of piece verification. In the reference the info dictionary
 * arrives after client.AddMagnet (torrent.go:58) and is awaited at torrent.go:69 (<-t.GotInfo());
 * anacrolix's metainfo package turns it into piece length, piece hashes and the file list that the
 * file storage of torrent.go:41 lays out under baseDir. gh_metainfo_parse accepts a whole .torrent
 * (top-level dictionary with "info") or a bare info dictionary; paths with "..", empty or absolute
 * components are rejected. Host-only: needs no device. err (optional) receives a reason. */
typedef struct gh_metainfo gh_metainfo;
int            gh_metainfo_parse(const uint8_t* data, size_t len, gh_metainfo** out, char* err, size_t err_len);
void           gh_metainfo_free(gh_metainfo* m);
uint64_t       gh_metainfo_piece_length(const gh_metainfo* m);
uint64_t       gh_metainfo_n_pieces(const gh_metainfo* m);
uint64_t       gh_metainfo_total_length(const gh_metainfo* m);
const uint8_t* gh_metainfo_pieces(const gh_metainfo* m);      /* 20 * n_pieces bytes, owned by m */
const char*    gh_metainfo_name(const gh_metainfo* m);
size_t         gh_metainfo_n_files(const gh_metainfo* m);
/* file i in info order; rel_path is relative to the storage base dir (name/, then the path list, for
 * multi-file torrents; name for single-file ones -- anacrolix file storage's layout) */
int            gh_metainfo_file(const gh_metainfo* m, size_t i, const char** rel_path, uint64_t* len);
/* info-hash = SHA-1 of the bencoded info dictionary (the magnet link's btih), computed on the device */
int            gh_metainfo_info_hash(gh_ctx* ctx, const gh_metainfo* m, uint8_t out[20]);
/* Verify a downloaded torrent in place: files under base_dir as laid out above, expected hashes
 * from the metainfo. The batched form of what runs between torrent.go:79 and :106. */
int            gh_verify_torrent(gh_ctx* ctx, const gh_metainfo* m, const char* base_dir,
                                 uint8_t* digests_out, uint64_t* ok_bitmap_out, uint64_t* n_bad_out);

/* Per-file parts: every file is its own S3 object (uploader.go:72 loops over the files of a job, one
 * PutObject each), so parts restart at every file boundary and each file's last part may be short --
 * but ALL files of the job are hashed in ONE batch. More files mean more independent chains, which
 * is what a GPU needs when parts are 64 MiB long (hard part 2): 24 files of
 * 2 GiB give 768 chains instead of 32 at a time. Outputs hold file 0's parts, then file 1's, ...;
 * file i contributes ceil(len_i / part_len) parts (none if empty). */
int gh_hash_parts_per_file(gh_ctx* ctx, const gh_span* files, size_t n_files, uint64_t part_len, int algos,
                           uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out);
int gh_hash_file_parts_per_file(gh_ctx* ctx, const gh_file* files, size_t n_files, uint64_t part_len,
                                int algos, uint8_t* sha1_out, uint8_t* md5_out, uint8_t* sha256_out);

/* S3 multipart ETag of n binary part MD5s: hex(md5(concat)) + "-" + n (what the server returns for
 * the upload started at uploader.go:89; computed on the device like any other message).
 * out needs at least 56 bytes. */
int gh_s3_etag(gh_ctx* ctx, const uint8_t* part_md5s, uint64_t n_parts, char* out, size_t out_len);

/* Lowercase hex of n digests (digest_len bytes each) in order, 2*digest_len characters per digest, no
 * separators, NUL-terminated: the form minio-go's signer uses for X-Amz-Content-Sha256 and, per 64 KiB
 * chunk, in the streaming-SigV4 string-to-sign that the reference's plain-HTTP uploads produce
 * (uploader.go:32-35, :89). out_len >= n*2*digest_len + 1, else GH_ERANGE. Host-only. */
int gh_hex_digests(const uint8_t* digests, uint64_t n, int digest_len, char* out, size_t out_len);

/* ---- device-resident entry points (payload already in HBM; "kernel-only" numbers) ------------ */

/* Same contract as gh_verify_pieces / gh_hash_parts but every pointer is a DEVICE pointer on
 * device `dev_index` (index into the context's device list). d_ok_words: ceil(n/32) uint32 words
 * (the uint64 host bitmap viewed as little-endian uint32 pairs). cuda_stream: a cudaStream_t cast to
 * void*; the launch is asynchronous on it. NULL = the context's own stream and the call returns
 * after the kernel has finished. d_data should be 16-byte aligned and msg_len a multiple of 16 for
 * the staged (coalesced) path; anything else takes the byte-granular path. */
int gh_verify_pieces_device(gh_ctx* ctx, int dev_index, const void* d_data, uint64_t total,
                            uint64_t piece_len, const uint8_t* d_expected, uint8_t* d_digests,
                            uint32_t* d_ok_words, void* cuda_stream);
int gh_hash_parts_device(gh_ctx* ctx, int dev_index, const void* d_data, uint64_t total,
                         uint64_t part_len, int algos, uint8_t* d_sha1, uint8_t* d_md5,
                         uint8_t* d_sha256, void* cuda_stream);

/* Resumable form for messages longer than a staging slot (midstates stay on the device).
 * Absorbs chunk_len bytes of each of n_msgs messages: message m's bytes for this call start at
 * d_data + m*stride. prior_bytes (a multiple of 64) is what earlier calls absorbed of every message;
 * msg_len / last_msg_len are the full message lengths. A message whose end falls inside this call is
 * padded and its digest written; d_state (n_msgs * GH_STATE_WORDS uint32, library layout) carries the midstates
 * between calls and must be the same buffer throughout. chunk_len must be a multiple of 64 unless
 * every message ends in this call. */
int gh_hash_update_device(gh_ctx* ctx, int dev_index, int algos, const void* d_data, uint64_t stride,
                          uint64_t chunk_len, uint64_t prior_bytes, uint64_t n_msgs, uint64_t msg_len,
                          uint64_t last_msg_len, uint32_t* d_state, uint8_t* d_sha1, uint8_t* d_md5,
                          uint8_t* d_sha256, void* cuda_stream);

/* gh_hash_update_device for callers that cannot allocate device memory. It has no counterpart in the
 * reference (which has no device code); it exists because of who calls this library. A download
 * worker sizes its payload buffers and staging ring to the HBM it has been given, so by the time a
 * job reaches hashing there may be nothing left for a midstate buffer and a cudaMalloc would fail
 * after the expensive part (the download) is already done. Other callers sit on a latency-sensitive
 * thread that cannot afford the device-wide synchronisation cudaMalloc/cudaFree imply. (It still
 * makes CUDA calls -- cudaSetDevice, a memset, kernel launches -- so it is not for places where no
 * CUDA call is allowed at all, such as a cudaLaunchHostFunc callback.)
 *
 * Same arguments and results as gh_hash_update_device, except: there is no context (cuda_device is
 * a CUDA ordinal, made current by the call) and no d_state -- the midstates live in a fixed array
 * inside the library image, GH_NOALLOC_MAX_MSGS * GH_STATE_WORDS * 4 bytes on every GPU that loads
 * the library, held until the process exits. n_msgs > GH_NOALLOC_MAX_MSGS is refused with GH_ERANGE
 * before anything is launched. The call makes no cudaMalloc/cudaFree and uses no allocating
 * routine; inputs and outputs remain the caller's buffers.
 *
 * The first call on a device looks up the array's address on that device and sets kernel
 * attributes; make that first call before opening any stream capture. All calls on a device share
 * the one array: a batch's first call (prior_bytes == 0) clears the part it will use, later calls of
 * the batch continue from what the previous call left, and two batches in flight on the same GPU at
 * once (different streams or threads) overwrite each other's midstates -- run them one after another.
 * cuda_stream NULL = the device's default stream, and the call returns after the work is done. */
#define GH_NOALLOC_MAX_MSGS 16384
int gh_hash_update_device_noalloc(int cuda_device, int algos, const void* d_data, uint64_t stride,
                                  uint64_t chunk_len, uint64_t prior_bytes, uint64_t n_msgs,
                                  uint64_t msg_len, uint64_t last_msg_len, uint8_t* d_sha1,
                                  uint8_t* d_md5, uint8_t* d_sha256, void* cuda_stream);

/* Synthetic stream of SURVEY.md Appendix C, generated on the device: byte k of stream(seed) for
 * k in [offset, offset+len) written to d_out. (Test / benchmark input; no reference counterpart.) */
int gh_generate_device(gh_ctx* ctx, int dev_index, void* d_out, uint64_t seed, uint64_t offset,
                       uint64_t len, void* cuda_stream);

/* ---- memory helpers ------------------------------------------------------------------------- */
/* Pinned host memory lets spans go host->device by DMA without the staging copy. */
int gh_host_alloc(void** out, size_t bytes);
int gh_host_free(void* p);
int gh_host_register(void* p, size_t bytes);
int gh_host_unregister(void* p);
int gh_device_alloc(gh_ctx* ctx, int dev_index, void** out, size_t bytes);
int gh_device_free(gh_ctx* ctx, int dev_index, void* p);
int gh_memcpy_d2h(gh_ctx* ctx, int dev_index, void* dst_host, const void* src_dev, size_t bytes);
int gh_memcpy_h2d(gh_ctx* ctx, int dev_index, void* dst_dev, const void* src_host, size_t bytes);

/* Contiguous 64-aligned range of messages that shard `shard` of `n_shards` owns:
 * [*begin, *end). Used for the in-context multi-GPU split and by one-process-per-GPU callers. */
void gh_shard_range(uint64_t n_msgs, int n_shards, int shard, uint64_t* begin, uint64_t* end);

#ifdef __cplusplus
}
#endif
#endif /* GPUHASH_H */
