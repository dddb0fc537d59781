/*
 * xetgpu.h — C ABI of the B200-native Xet content-addressing library (libxetgpu.so).
 *
 * The library computes, on one NVIDIA B200 (sm_100a), what OME reaches through
 * huggingface/xet-core when model shards are content-addressed: gear content-defined
 * chunk boundaries, keyed-BLAKE3 chunk hashes, the merkle root and the Xet file hash.
 * Results are bit-exact with xet-core ("Parity").
 *
 * It sits BESIDE the reference's libxet (pkg/xet/xet.h) and follows its conventions
 * (reference file:line in brackets):
 *   - every fallible call returns XetError* — NULL means success; the caller frees a
 *     non-NULL error with xetgpu_free_error            [pkg/xet/xet.h:86-90,132; src/error.rs:55]
 *   - error codes are the reference's XetErrorCode values      [pkg/xet/src/error.rs:12-23]
 *   - results come back through out-pointers, are owned by the library until the
 *     matching xetgpu_free_* call, and every free function accepts NULL
 *                                                       [pkg/xet/xet.h:132-134; src/ffi.rs:118]
 *   - inputs are borrowed for the duration of the call; strings are NUL-terminated UTF-8
 *   - calls are synchronous (the reference blocks on its runtime) [pkg/xet/src/ffi.rs:190]
 *   - a context may be used from any thread; calls on one context are serialised
 *     internally; it must not be used during xetgpu_context_free  [pkg/xet/README.md:76,88]
 *   - progress / cancellation use the reference's callback shapes   [pkg/xet/xet.h:22-40]; handlers run on
 *     library threads with no library lock held, but the call that invokes them is still in progress: a handler
 *     may call xetgpu_context_set_progress_callback / xetgpu_last_timings, while a hashing call made from
 *     inside a handler (on any handle of this library) is refused with code 1 (it could wait for itself)
 *   - a no-op versioned symbol lets the cgo preamble fail at link time on a mismatch
 *                                                       [pkg/xet/xet.h:94; pkg/xet/xet.go:11-13]
 * No type in this header depends on CUDA or PyTorch; device pointers travel as void*.
 */
#ifndef XETGPU_H
#define XETGPU_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- shared with pkg/xet/xet.h (identical layout; guarded so both headers can be included) */
#ifndef XET_H
typedef struct {
    int32_t code;
    char* message;
    char* details;
} XetError; /* [pkg/xet/xet.h:86-90] */

typedef enum {
    XET_PROGRESS_PHASE_SCANNING = 0,
    XET_PROGRESS_PHASE_DOWNLOADING = 1,
    XET_PROGRESS_PHASE_FINALIZING = 2,
} XetProgressPhase; /* [pkg/xet/xet.h:15-19] */

typedef struct {
    XetProgressPhase phase;
    uint64_t total_bytes;
    uint64_t completed_bytes;
    uint32_t total_files;
    uint32_t completed_files;
    const char* current_file;
    uint64_t current_file_completed_bytes;
    uint64_t current_file_total_bytes;
} XetProgressUpdate; /* [pkg/xet/xet.h:21-30] */

typedef void (*XetProgressCallback)(const XetProgressUpdate* update, void* user_data); /* [xet.h:32] */
typedef bool (*XetCancellationCallback)(void* user_data);                               /* [xet.h:34] */
typedef struct {
    XetCancellationCallback callback;
    void* user_data;
} XetCancellationToken; /* [pkg/xet/xet.h:36-39] */
#endif

/* Error codes: the reference's XetErrorCode [pkg/xet/src/error.rs:12-23]. */
enum {
    XETGPU_OK = 0,
    XETGPU_ERR_INVALID_CONFIG = 1,    /* NULL / malformed argument [src/ffi.rs:140-146] */
    XETGPU_ERR_NOT_FOUND = 4,         /* file does not exist */
    XETGPU_ERR_CHECKSUM_MISMATCH = 6, /* verify: computed hash != expected [src/error.rs:19] */
    XETGPU_ERR_CANCELLED = 7,
    XETGPU_ERR_IO = 8,
    XETGPU_ERR_UNKNOWN = 99,          /* CUDA runtime failures land here, message = cudaGetErrorString */
};

/* Chunking constants of xet-core's deduplication crate. */
#define XETGPU_TARGET_CHUNK 65536u
#define XETGPU_MIN_CHUNK 8192u
#define XETGPU_MAX_CHUNK 131072u

typedef struct XetGpuContext XetGpuContext; /* opaque; one per (process, GPU) */

/* Replaces nothing in xet.h: mirrors XetConfig's role [pkg/xet/xet.h:42-48] for the GPU path. */
typedef struct {
    int32_t device;             /* CUDA device ordinal; -1 = current device */
    void* stream;               /* cudaStream_t to launch on; NULL = a stream the context owns */
    uint64_t staging_bytes;     /* size of EACH of the two pinned host staging buffers used to copy
                                   pageable host memory; 0 = 64 MiB.  xetgpu_hash_files reads through a ring
                                   of 16 pinned slots of min(staging_bytes, 64 MiB) each (16 MiB when 0) */
    uint64_t max_batch_bytes;   /* xetgpu_hash_files: bytes resident on the device per batch; 0 = 8 GiB.
                                   A single file larger than this streams through the device in pieces of this
                                   size (byte ranges with the carried cut, combined at the end), so device memory
                                   stays bounded whatever the file size; xorbs are not formed for such a file */
    uint32_t flags;             /* XETGPU_CTX_* */
} XetGpuConfig;

#define XETGPU_CTX_STAGE_TIMING 1u /* record CUDA events around every stage (xetgpu_last_timings) */
#define XETGPU_CTX_NO_OVERLAP 2u   /* run file groups one after the other on one stream (per-kernel timing) */

/* Where the bytes passed to xetgpu_hash_buffers live. */
enum {
    XETGPU_MEM_AUTO = 0,   /* ask cudaPointerGetAttributes */
    XETGPU_MEM_HOST = 1,   /* host memory (pageable or pinned) */
    XETGPU_MEM_DEVICE = 2, /* device memory on the context's GPU, 16-byte aligned */
};
/* Stream-ordering rule for device inputs: the library reads them on the context's stream (XetGpuConfig.stream, or
 * a non-blocking stream of its own, which takes no implicit ordering from the legacy default stream or from any
 * other stream).  The bytes must therefore be COMPLETE with respect to that stream when the call is made: either
 * create the context on the stream that produces them, or synchronise the producing stream first.  The Python
 * binding does the latter for torch tensors (ome_b200/hasher.py, _order_after_torch). */

#define XETGPU_WANT_CHUNKS 1u /* also return per-chunk end offsets and hashes */
#define XETGPU_WANT_XORBS 2u  /* also form xorbs and return one hash per xorb (see XetGpuXorb) */

/* One file's result.  Hash bytes are in RAW DIGEST ORDER; the Xet hex form
 * (MerkleHash::hex, four LE u64 words as %016x) is produced by xetgpu_hash_to_hex. */
typedef struct {
    uint8_t file_hash[32];   /* merklehash::file_hash of the chunk list */
    uint8_t merkle_root[32]; /* merklehash::aggregated_node_hash (== xorb_hash over all chunks) */
    uint64_t size;           /* bytes */
    uint64_t n_chunks;
    uint64_t first_chunk;    /* index of this file's first chunk in XetGpuResult.chunk_* */
} XetGpuFileResult;

/* One xorb: the unit xet-core's upload side stores in the CAS.  A file's chunks are packed, in
 * order, into xorbs of at most 64 MiB of chunk bytes and at most 8192 chunks; `hash` is merklehash::xorb_hash =
 * aggregated_node_hash over the xorb's (chunk hash, length) list.  The hash FUNCTION is pinned through the file hash;
 * the GROUPING is parity-unpinned (hf_xet.hash_files does not expose xorbs, and a real upload also drops chunks
 * that deduplicate against earlier xorbs): this is the no-dedup, one-file-at-a-time policy. */
typedef struct {
    uint8_t hash[32];        /* raw digest order */
    uint64_t file;           /* index of the file in this result */
    uint64_t first_chunk;    /* index of the xorb's first chunk WITHIN its file */
    uint64_t n_chunks;
    uint64_t bytes;          /* sum of the chunk lengths */
} XetGpuXorb;

typedef struct {
    size_t n_files;
    XetGpuFileResult* files;
    uint64_t n_chunks;       /* 0 unless XETGPU_WANT_CHUNKS */
    uint64_t* chunk_ends;    /* per chunk: exclusive end offset, relative to its own file */
    uint8_t* chunk_hashes;   /* 32 bytes per chunk, raw digest order */
    uint64_t n_xorbs;        /* 0 unless XETGPU_WANT_XORBS (fields added in 0.2.0, at the end of the struct) */
    XetGpuXorb* xorbs;       /* sorted by file, then by first_chunk */
} XetGpuResult;

/* Device time of each stage of the last hash call, in milliseconds (CUDA events on the
 * context's stream; only filled when the context has XETGPU_CTX_STAGE_TIMING). */
typedef struct {
    float h2d_ms;
    float gear_ms;      /* K1: gear window hash + candidate compaction */
    float resolve_ms;   /* K2: min/max boundary resolution */
    float blake3_ms;    /* K3: leaf compression + chunk tree */
    float merkle_ms;    /* K4: merkle levels + file hash */
    float d2h_ms;
    float total_ms;
    float gear_kernel_ms;   /* gear_scan_kernel alone (the K1 hot kernel) */
    float leaf_kernel_ms;   /* leaf_group_kernel alone (the K3 hot kernel, dominant overall) */
    uint64_t bytes;
    uint64_t n_chunks;
    uint64_t n_candidates;
    uint32_t kernel_launches;
} XetGpuTimings;

/* Link-time version guard, same idea as xet_version_1_0_0 [pkg/xet/xet.h:94]. */
extern void xetgpu_version_0_2_0(void);

/* Context lifecycle — shaped like xet_client_new / xet_client_free [pkg/xet/xet.h:97-98],
 * but returns an error object instead of a bare NULL so CUDA failures are reportable. */
XetError* xetgpu_context_new(const XetGpuConfig* config, XetGpuContext** out_ctx);
void xetgpu_context_free(XetGpuContext* ctx);

/* Same callback plumbing as xet_client_set_progress_callback [pkg/xet/xet.h:99-104]. */
XetError* xetgpu_context_set_progress_callback(XetGpuContext* ctx, XetProgressCallback callback,
                                               void* user_data, uint32_t throttle_ms);

/*
 * Hash n_files byte strings packed back to back in `data`: file f is
 * data[offsets[f] .. offsets[f+1]).  offsets has n_files+1 ascending entries, offsets[0]
 * may be non-zero.  This is the entry point for bytes already in memory (the replica
 * agent's checksum switch [internal/ome-agent/replica/replicator/helper.go:86-109], or a
 * download buffer).  Stands where xet-core's clean path would run
 * (deduplication::Chunker + merklehash::{compute_data_hash,file_hash}).
 */
XetError* xetgpu_hash_buffers(XetGpuContext* ctx, const void* data, const uint64_t* offsets,
                              size_t n_files, int32_t memory, uint32_t flags,
                              XetGpuResult** out_result);

/*
 * Hash files on disk: read through the pinned staging buffers and copied to the GPU
 * while earlier bytes are being hashed.  Sizes come from stat() when the call is planned and are re-checked with
 * fstat() on the descriptor each piece is read from: a file that is replaced, grows or shrinks while the call runs
 * fails with XETGPU_ERR_IO instead of being hashed over stale extents.  Files are opened piece by piece, so a batch
 * may name more files than the process may hold open.  The verify-after-download entry point for
 * model-agent [pkg/modelagent/gopher.go:1130-1149 has no verification today; the OCI
 * branch's MD5 pass at gopher.go:843-860 is the shape it mirrors].
 */
XetError* xetgpu_hash_files(XetGpuContext* ctx, const char* const* paths, size_t n_files,
                            uint32_t flags, const XetCancellationToken* cancel_token,
                            XetGpuResult** out_result);

/*
 * Verify: compute the Xet file hash and compare with the 64-digit hex the Hub advertises
 * (X-Xet-Hash header / tree API xetHash) [pkg/xet/src/xet_integration.rs:41-44;
 * pkg/xet/src/hf_adapter.rs:36-44].  NULL = match; code 6 (ChecksumMismatch) with both
 * hashes in `details` otherwise.  Shape precedent: hub.VerifyChecksum
 * [pkg/hfutil/hub/utils.go:139-158].
 */
XetError* xetgpu_verify_file(XetGpuContext* ctx, const char* path, const char* expected_hex);
XetError* xetgpu_verify_buffer(XetGpuContext* ctx, const void* data, uint64_t len, int32_t memory,
                               const char* expected_hex);

/*
 * Verify a whole downloaded snapshot in one call: n files, each against its advertised hash
 * (the list the Hub's tree API returns, HfTreeItem.xet_hash [pkg/xet/src/hf_adapter.rs:36-44]).
 * All files are read and hashed as one batch.  out_ok (optional, n bytes) receives 1/0 per file.
 * NULL when every file matches; otherwise code 6 with the mismatching paths and both hashes in
 * `details` (one per line).  An entry of expected_hex may be NULL or "" to skip that file
 * (files the Hub does not store through Xet carry no xet hash).
 */
XetError* xetgpu_verify_files(XetGpuContext* ctx, const char* const* paths, const char* const* expected_hex,
                              size_t n_files, const XetCancellationToken* cancel_token, uint8_t* out_ok);

/* MerkleHash::hex / MerkleHash::from_hex [used at pkg/xet/src/xet_downloader.rs:124-126]. */
void xetgpu_hash_to_hex(const uint8_t hash[32], char out_hex[65]);
XetError* xetgpu_hash_from_hex(const char* hex, uint8_t out_hash[32]);

/* ---- stage-level entry points (parity tests, benchmarks, multi-GPU range partitioning) ---- */

/*
 * K1 only: indices i (relative to `data`) whose 64-byte gear window hash has its 16 top
 * bits clear (gearhash::Hasher::next_match semantics evaluated at every offset).  `data`
 * is a device or host buffer of `len` bytes.  Returns a malloc'd ascending array
 * (xetgpu_free_u64).
 */
XetError* xetgpu_gear_candidates(XetGpuContext* ctx, const void* data, uint64_t len, int32_t memory,
                                 uint64_t** out_indices, uint64_t* out_count);

/*
 * Byte-range form for a file split across GPUs (BASELINE.json config 4).  `data` holds
 * bytes [range_begin - lead, range_end) of ONE file of `file_len` bytes, where
 * lead = min(range_begin, XETGPU_RANGE_LEAD).  `entry_cut` is the last chunk boundary
 * <= range_begin as resolved by the previous range's owner (0 for the first range).
 * Produces every chunk whose END lies in (range_begin, range_end] — ends are file-relative —
 * plus `exit_cut`, the value to hand to the next range's owner.  No file hash is computed;
 * rank 0 combines the per-range chunk lists with xetgpu_merkle_from_chunks.
 */
#define XETGPU_RANGE_LEAD (XETGPU_MAX_CHUNK + 64u)
XetError* xetgpu_hash_range(XetGpuContext* ctx, const void* data, int32_t memory, uint64_t file_len,
                            uint64_t range_begin, uint64_t range_end, uint64_t entry_cut,
                            uint64_t* out_exit_cut, XetGpuResult** out_result);

/*
 * Whole files and byte-range items in ONE batched call — what a rank of a byte-range partitioned job holds: the
 * files that lie inside its range plus its pieces of the files it shares with its neighbours (BASELINE config 4).
 * Item i is data[offsets[i] .. offsets[i+1]) and is described by specs[i]; a whole file is {len, 0, len, 0}, a
 * byte-range item has the meaning of xetgpu_hash_range's arguments (its data starts `lead` bytes before
 * range_begin).  Per item the result holds, for whole files, root and file hash (for byte-range items those two
 * fields cover only the item's own chunks and are not a file's) and, for byte-range items, the chunk list (ends
 * relative to the FILE; files[i].first_chunk / n_chunks index it) — with XETGPU_WANT_CHUNKS the whole files' chunk
 * lists are returned as well.  out_exit_cuts[i] (optional) receives the cut to hand to the next range's owner.
 */
typedef struct {
    uint64_t file_len;
    uint64_t range_begin, range_end;
    uint64_t entry_cut;
} XetGpuRangeSpec;
XetError* xetgpu_hash_buffers_ranges(XetGpuContext* ctx, const void* data, const uint64_t* offsets, size_t n_items,
                                     int32_t memory, uint32_t flags, const XetGpuRangeSpec* specs,
                                     uint64_t* out_exit_cuts, XetGpuResult** out_result);

/*
 * Boundary resolution is a carry chain across ranges; to keep ranks from waiting on one
 * another, a range owner may first ask for the cut it WOULD hand on if its own entry were
 * `assumed_entry_cut`, forward that, and only re-run when the true entry differs.
 */
XetError* xetgpu_range_exit_cut(XetGpuContext* ctx, const void* data, int32_t memory, uint64_t file_len,
                                uint64_t range_begin, uint64_t range_end, uint64_t entry_cut,
                                uint64_t* out_exit_cut);

/* K4 only: merkle root + file hash of a list of (hash, size) pairs, on the GPU
 * (merklehash::aggregated_node_hash / file_hash).  The two input arrays may be host or device memory. */
XetError* xetgpu_merkle_from_chunks(XetGpuContext* ctx, const uint8_t* chunk_hashes,
                                    const uint64_t* chunk_sizes, uint64_t n_chunks,
                                    uint8_t out_merkle_root[32], uint8_t out_file_hash[32]);

/*
 * Same result as xetgpu_merkle_from_chunks, for callers that cannot allocate device memory.
 * No counterpart in the reference: OME has no GPU code, so it never meets this constraint.
 * A process whose GPU memory is wholly owned by another allocator, a latency path that cannot
 * take the implicit synchronisation of cudaMalloc/cudaFree, or a combiner that holds no
 * XetGpuContext can call this instead: it needs no context and allocates nothing.  Its scratch
 * is a fixed array inside the library image (about 2.2 MB per GPU that loads the library, held
 * until exit), which caps the list at XETGPU_STATIC_MERKLE_MAX_CHUNKS entries; longer lists get
 * error code 1.  Inputs and outputs are DEVICE pointers owned by the caller (hashes: 32 bytes
 * per chunk, raw digest order; sizes: u64; outputs: 32 bytes each; hash and output pointers 4-byte aligned,
 * the size array 8-byte aligned — anything else is refused with code 1).  One kernel is launched on
 * `stream` (a cudaStream_t, NULL = default stream) and the call returns without waiting.
 * The first call on each device also looks up the array's device address — make that call
 * before starting any stream capture.  All calls share the array: calls that can overlap on
 * the GPU must be serialised by the caller.
 */
#define XETGPU_STATIC_MERKLE_MAX_CHUNKS 65536u
XetError* xetgpu_merkle_from_chunks_static(const uint8_t* device_chunk_hashes,
                                           const uint64_t* device_chunk_sizes, uint64_t n_chunks,
                                           uint8_t* device_out_merkle_root, uint8_t* device_out_file_hash,
                                           void* stream);

/* ---- several GPUs from one process -----------------------------------------------------------------
 * "context new(const int* devices, n, ...)".  model-agent is one process per node
 * [cmd/model-agent/main.go:243-271]; a multi-device context lets that one process use every GPU of the node.
 * It owns one XetGpuContext per listed device and runs one host thread per device per call.  The files of a call
 * are laid end to end and cut into one equal byte range per device: a file inside a range is hashed whole by that
 * device, a file that straddles a boundary is shared — each owner hashes its byte range (the XETGPU_RANGE_LEAD bytes
 * in front of it come from the file), the 8-byte carried cut walks down the owners, an owner whose assumed entry
 * was wrong re-hashes a 4 MiB window and splices at the first common cut, and the first owner's device combines the
 * chunk lists (xetgpu_merkle_from_chunks).  Results are identical to the single-device calls.
 * devices == NULL or n_devices == 0 selects every visible device.  config->device is ignored and config->stream
 * must be NULL.  Small jobs use fewer devices (at least 64 MiB of work per device). */
typedef struct XetGpuMulti XetGpuMulti;
XetError* xetgpu_multi_new(const int32_t* devices, size_t n_devices, const XetGpuConfig* config, XetGpuMulti** out_ctx);
void xetgpu_multi_free(XetGpuMulti* ctx);
size_t xetgpu_multi_device_count(const XetGpuMulti* ctx);
XetError* xetgpu_multi_set_progress_callback(XetGpuMulti* ctx, XetProgressCallback callback, void* user_data,
                                             uint32_t throttle_ms);
/* Same contracts as xetgpu_hash_files / xetgpu_verify_files. */
XetError* xetgpu_multi_hash_files(XetGpuMulti* ctx, const char* const* paths, size_t n_files, uint32_t flags,
                                  const XetCancellationToken* cancel_token, XetGpuResult** out_result);
XetError* xetgpu_multi_verify_files(XetGpuMulti* ctx, const char* const* paths, const char* const* expected_hex,
                                    size_t n_files, const XetCancellationToken* cancel_token, uint8_t* out_ok);

/* ---- safetensors-aware ingestion ---------------------------------------------------
 * The shard header as OME's model-config parser reads it [pkg/hfutil/modelconfig/safetensors.go:67-141]: u64-LE
 * header length (at most 10 MB), JSON object of tensor name -> {dtype, shape, data_offsets}, "__metadata__" skipped.
 * parameter_count is what modelconfig.ParseSafetensors returns (product of every non-empty shape, same overflow and
 * invalid-dimension errors, same message texts); the tensor table adds what a partitioner needs: each tensor's
 * absolute byte extent in the file, ascending.  Errors: code 4 / 8 for open / short-read failures, code 1 for a
 * malformed header. */
typedef struct {
    const char* name;      /* owned by the index */
    const char* dtype;
    uint64_t begin, end;   /* byte extent in the file: 8 + header_len + data_offsets */
    uint64_t n_elements;   /* product of the shape; 0 for an empty shape */
    uint32_t rank;
} XetGpuTensorInfo;

typedef struct {
    uint64_t file_size;    /* 0 if unknown (xetgpu_safetensors_parse without a size) */
    uint64_t header_len;
    uint64_t data_start;   /* 8 + header_len */
    uint64_t n_tensors;
    XetGpuTensorInfo* tensors;  /* ascending by begin */
    int64_t parameter_count;
} XetGpuSafetensorsIndex;

XetError* xetgpu_safetensors_open(const char* path, XetGpuSafetensorsIndex** out_index);
/* `head` = the first head_len bytes of a shard (at least 8 + header length of them). */
XetError* xetgpu_safetensors_parse(const void* head, uint64_t head_len, uint64_t file_size,
                                   XetGpuSafetensorsIndex** out_index);
void xetgpu_safetensors_free(XetGpuSafetensorsIndex* index);
/* Byte ranges for n_ranges owners of one shard with interior bounds on tensor boundaries (nearest tensor begin / end
 * within a quarter range of the equal split; the 4096-aligned equal split itself when one tensor is larger than
 * that).  out_bounds: n_ranges + 1 ascending offsets, [0] = 0, [n_ranges] = file_size; every non-empty interior
 * range keeps at least XETGPU_RANGE_LEAD + 4096 bytes, so the ranges are valid inputs to xetgpu_hash_range. */
XetError* xetgpu_safetensors_split(const XetGpuSafetensorsIndex* index, uint64_t file_size, size_t n_ranges,
                                   uint64_t* out_bounds);
/* Flag for xetgpu_multi_hash_files / _verify_files is implied: a *.safetensors file that is shared between devices
 * has its in-file boundaries moved to the nearest tensor boundary (same quarter-range rule) unless
 * XETGPU_NO_TENSOR_SNAP is passed.  Hashes do not depend on where the boundaries fall. */
#define XETGPU_NO_TENSOR_SNAP 4u
/* The byte ranges xetgpu_multi_hash_files would give each device for these files: up to `cap` rows of four u64
 * (device index, file index, begin, end) in device order; *out_count = rows the plan has.  Tooling and tests. */
XetError* xetgpu_multi_plan(XetGpuMulti* ctx, const char* const* paths, size_t n_files, uint32_t flags,
                            uint64_t* out_pieces, size_t cap, size_t* out_count);

/* K0: fill device memory with the SURVEY.md §8(d) counter-based splitmix64 stream:
 * dst[i] = byte (byte_off + i) of stream `seed`.  Benchmarks and tests only. */
XetError* xetgpu_fill_splitmix(XetGpuContext* ctx, void* device_dst, uint64_t seed, uint64_t byte_off,
                               uint64_t len);

/* Stage timings of the most recent hash call on this context. */
XetError* xetgpu_last_timings(XetGpuContext* ctx, XetGpuTimings* out);

/* Memory management — all NULL-safe [pkg/xet/xet.h:132-134]. */
void xetgpu_free_error(XetError* err);
void xetgpu_free_result(XetGpuResult* result);
void xetgpu_free_u64(uint64_t* array);

#ifdef __cplusplus
}
#endif

#endif /* XETGPU_H */
