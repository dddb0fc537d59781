// Package xet — GPU content-addressing additions (drop into github.com/sgl-project/ome/pkg/xet).
//
// This file is ADDITIVE: nothing in xet.go / hf_compat.go / config.go / module.go changes.
// It binds libxetgpu.so (include/xetgpu.h of the xet-b200 repository) the same way xet.go binds
// libxet: cgo preamble with LDFLAGS, a link-time version guard, C callback bridges, and errors
// converted to *XetError.  NOT COMPILED in the xet-b200 build image (no Go toolchain there); the
// C ABI it calls is exercised by that repository's tests through ctypes.
package xet

/*
#cgo linux LDFLAGS: -L${SRCDIR}/gpu -lxetgpu -lstdc++ -ldl -lrt -lpthread
#include <stdlib.h>
#include "xet.h"
#include "xetgpu.h"

// Link-time version check (same idiom as xet_version_1_0_0 in xet.go)
extern void xetgpu_version_0_2_0(void);
static void (*xetgpu_version_check)(void) = &xetgpu_version_0_2_0;

// Existing Go callback bridges exported from xet.go
extern void goXetProgressCallback(XetProgressUpdate* update, void* user_data);
extern bool goXetShouldCancel(void* user_data);
*/
import "C"

import (
	"context"
	"errors"
	"runtime/cgo"
	"time"
	"unsafe"
)

// XetChecksumAlgorithm is the name the replica agent's ChecksumConfig uses for this digest
// (next to "md5" and "sha256" in internal/ome-agent/replica/replicator/helper.go).
const XetChecksumAlgorithm = "xet"

// ErrorCodeChecksumMismatch is XetErrorCode::ChecksumMismatch (pkg/xet/src/error.rs).
const ErrorCodeChecksumMismatch = 6

// FileDigest is one file's result.
type FileDigest struct {
	Hash       string // 64 hex digits, the form the Hub sends as X-Xet-Hash
	MerkleRoot string
	Size       uint64
	Chunks     uint64
}

// GpuHasher owns one libxetgpu context on one GPU.  Like Client it is safe for concurrent use
// (calls serialise inside the library) and must be closed exactly once.
type GpuHasher struct {
	ctx *C.XetGpuContext
}

// NewGpuHasher creates a context on the given CUDA device (-1 = current device).
func NewGpuHasher(device int) (*GpuHasher, error) {
	cfg := C.XetGpuConfig{device: C.int32_t(device)}
	var ctx *C.XetGpuContext
	if cerr := C.xetgpu_context_new(&cfg, &ctx); cerr != nil {
		return nil, convertGpuError(cerr)
	}
	return &GpuHasher{ctx: ctx}, nil
}

// Close releases the context.
func (h *GpuHasher) Close() error {
	if h.ctx != nil {
		C.xetgpu_context_free(h.ctx)
		h.ctx = nil
	}
	return nil
}

// SetProgressHandler mirrors (*Client).SetProgressHandler.
func (h *GpuHasher) SetProgressHandler(handler ProgressHandler, throttle time.Duration) (func(), error) {
	if h.ctx == nil {
		return nil, errors.New("xet: hasher is closed")
	}
	handle := cgo.NewHandle(handler)
	cerr := C.xetgpu_context_set_progress_callback(h.ctx, (C.XetProgressCallback)(C.goXetProgressCallback),
		unsafe.Pointer(handle), C.uint32_t(throttle.Milliseconds()))
	if cerr != nil {
		handle.Delete()
		return nil, convertGpuError(cerr)
	}
	return func() {
		if h.ctx != nil {
			if cerr := C.xetgpu_context_set_progress_callback(h.ctx, nil, nil, 0); cerr != nil {
				C.xetgpu_free_error(cerr)
			}
		}
		handle.Delete()
	}, nil
}

// HashFiles computes the Xet file hash of every path (read through pinned staging buffers,
// hashed on the GPU).  ctx cancellation is polled between files.
func (h *GpuHasher) HashFiles(ctx context.Context, paths []string) ([]FileDigest, error) {
	if h.ctx == nil {
		return nil, errors.New("xet: hasher is closed")
	}
	if len(paths) == 0 {
		return nil, nil
	}
	cpaths := make([]*C.char, len(paths))
	for i, p := range paths {
		cpaths[i] = C.CString(p)
	}
	defer func() {
		for _, p := range cpaths {
			C.free(unsafe.Pointer(p))
		}
	}()
	tok, release := cancelToken(ctx)
	defer release()
	var res *C.XetGpuResult
	if cerr := C.xetgpu_hash_files(h.ctx, &cpaths[0], C.size_t(len(paths)), 0, tok, &res); cerr != nil {
		return nil, convertGpuError(cerr)
	}
	defer C.xetgpu_free_result(res)
	files := unsafe.Slice(res.files, len(paths))
	out := make([]FileDigest, len(paths))
	var hex [65]C.char
	for i := range files {
		C.xetgpu_hash_to_hex(&files[i].file_hash[0], &hex[0])
		out[i].Hash = C.GoString(&hex[0])
		C.xetgpu_hash_to_hex(&files[i].merkle_root[0], &hex[0])
		out[i].MerkleRoot = C.GoString(&hex[0])
		out[i].Size = uint64(files[i].size)
		out[i].Chunks = uint64(files[i].n_chunks)
	}
	return out, nil
}

// HashBytes hashes one in-memory buffer.
func (h *GpuHasher) HashBytes(b []byte) (FileDigest, error) {
	if h.ctx == nil {
		return FileDigest{}, errors.New("xet: hasher is closed")
	}
	offsets := [2]C.uint64_t{0, C.uint64_t(len(b))}
	var p unsafe.Pointer
	if len(b) > 0 {
		p = unsafe.Pointer(&b[0])
	}
	var res *C.XetGpuResult
	if cerr := C.xetgpu_hash_buffers(h.ctx, p, &offsets[0], 1, C.XETGPU_MEM_HOST, 0, &res); cerr != nil {
		return FileDigest{}, convertGpuError(cerr)
	}
	defer C.xetgpu_free_result(res)
	var hex [65]C.char
	C.xetgpu_hash_to_hex(&res.files.file_hash[0], &hex[0])
	return FileDigest{Hash: C.GoString(&hex[0]), Size: uint64(len(b)), Chunks: uint64(res.files.n_chunks)}, nil
}

// VerifyFile returns nil when the file's Xet hash equals expectedHex, and a *XetError with
// Code == ErrorCodeChecksumMismatch (expected and computed hashes in Details) when it does not.
// Shape precedent: hub.VerifyChecksum (pkg/hfutil/hub/utils.go).
func (h *GpuHasher) VerifyFile(path, expectedHex string) error {
	if h.ctx == nil {
		return errors.New("xet: hasher is closed")
	}
	cp, ce := C.CString(path), C.CString(expectedHex)
	defer C.free(unsafe.Pointer(cp))
	defer C.free(unsafe.Pointer(ce))
	return convertGpuError(C.xetgpu_verify_file(h.ctx, cp, ce))
}

// VerifyFiles verifies a downloaded snapshot in one batch (paths[i] against expectedHex[i]; an empty
// expected hash skips the file).  It returns the per-file verdicts; err is a *XetError with
// Code == ErrorCodeChecksumMismatch when any file differs.
func (h *GpuHasher) VerifyFiles(ctx context.Context, paths, expectedHex []string) ([]bool, error) {
	if h.ctx == nil {
		return nil, errors.New("xet: hasher is closed")
	}
	n := len(paths)
	if len(expectedHex) != n {
		return nil, &XetError{Code: 1, Message: "paths and expectedHex differ in length"}
	}
	if n == 0 {
		return nil, nil
	}
	cp := make([]*C.char, n)
	ch := make([]*C.char, n)
	for i := range paths {
		cp[i] = C.CString(paths[i])
		if expectedHex[i] != "" {
			ch[i] = C.CString(expectedHex[i])
		}
	}
	defer func() {
		for i := range cp {
			C.free(unsafe.Pointer(cp[i]))
			if ch[i] != nil {
				C.free(unsafe.Pointer(ch[i]))
			}
		}
	}()
	tok, release := cancelToken(ctx)
	defer release()
	ok := make([]C.uint8_t, n)
	err := convertGpuError(C.xetgpu_verify_files(h.ctx, &cp[0], &ch[0], C.size_t(n), tok, &ok[0]))
	out := make([]bool, n)
	for i := range ok {
		out[i] = ok[i] != 0
	}
	return out, err
}

// cancelToken wraps ctx the way xet.go does for xet_download_file (xet.go:366-375): the reused bridge
// goXetShouldCancel type-asserts *cancellationBridge, so the handle must carry exactly that type.
// A nil or never-cancelled context yields a nil token.
func cancelToken(ctx context.Context) (*C.XetCancellationToken, func()) {
	if ctx == nil || ctx.Done() == nil {
		return nil, func() {}
	}
	handle := cgo.NewHandle(&cancellationBridge{ctx: ctx})
	tok := (*C.XetCancellationToken)(C.malloc(C.size_t(unsafe.Sizeof(C.XetCancellationToken{}))))
	tok.callback = (C.XetCancellationCallback)(C.goXetShouldCancel)
	tok.user_data = unsafe.Pointer(handle)
	return tok, func() {
		C.free(unsafe.Pointer(tok))
		handle.Delete()
	}
}

// XetChecksumMetadataKey is the object-metadata key the replica agent would store the digest under, next to
// OCIObjectMD5MetadataKey / OCIObjectSHA256MetadataKey (internal/ome-agent/replica/replicator/helper.go:28-31).
const XetChecksumMetadataKey = "opc-meta-xet"

// XetFileChecksum has the shape of replicator.GetFileChecksum(filePath, algorithm) for algorithm "xet"
// (internal/ome-agent/replica/replicator/helper.go:86-109): path in, digest string out.  md5/sha256 are
// base64 of the raw digest there; the Xet digest is returned in the form every other Xet tool prints and
// the Hub advertises (64 hex digits, MerkleHash::hex), so that it can be compared with X-Xet-Hash directly.
func (h *GpuHasher) XetFileChecksum(filePath string) (string, error) {
	d, err := h.HashFiles(context.Background(), []string{filePath})
	if err != nil {
		return "", err
	}
	return d[0].Hash, nil
}

// GpuMultiHasher owns one libxetgpu context per listed GPU (xetgpu_multi_*): the form a single
// model-agent process (cmd/model-agent/main.go:243-271) uses to reach every GPU of its node.
type GpuMultiHasher struct {
	ctx *C.XetGpuMulti
}

// NewGpuMultiHasher creates contexts on the given CUDA devices (nil = every visible device).
func NewGpuMultiHasher(devices []int) (*GpuMultiHasher, error) {
	var ctx *C.XetGpuMulti
	var dp *C.int32_t
	cdev := make([]C.int32_t, len(devices))
	for i, d := range devices {
		cdev[i] = C.int32_t(d)
	}
	if len(cdev) > 0 {
		dp = &cdev[0]
	}
	if cerr := C.xetgpu_multi_new(dp, C.size_t(len(cdev)), nil, &ctx); cerr != nil {
		return nil, convertGpuError(cerr)
	}
	return &GpuMultiHasher{ctx: ctx}, nil
}

// Close releases every per-device context.
func (m *GpuMultiHasher) Close() error {
	if m.ctx != nil {
		C.xetgpu_multi_free(m.ctx)
		m.ctx = nil
	}
	return nil
}

// VerifyFiles is (*GpuHasher).VerifyFiles over all devices of the node.
func (m *GpuMultiHasher) VerifyFiles(ctx context.Context, paths, expectedHex []string) ([]bool, error) {
	if m.ctx == nil {
		return nil, errors.New("xet: hasher is closed")
	}
	n := len(paths)
	if len(expectedHex) != n {
		return nil, &XetError{Code: 1, Message: "paths and expectedHex differ in length"}
	}
	if n == 0 {
		return nil, nil
	}
	cp := make([]*C.char, n)
	ch := make([]*C.char, n)
	for i := range paths {
		cp[i] = C.CString(paths[i])
		if expectedHex[i] != "" {
			ch[i] = C.CString(expectedHex[i])
		}
	}
	defer func() {
		for i := range cp {
			C.free(unsafe.Pointer(cp[i]))
			if ch[i] != nil {
				C.free(unsafe.Pointer(ch[i]))
			}
		}
	}()
	tok, release := cancelToken(ctx)
	defer release()
	ok := make([]C.uint8_t, n)
	err := convertGpuError(C.xetgpu_multi_verify_files(m.ctx, &cp[0], &ch[0], C.size_t(n), tok, &ok[0]))
	out := make([]bool, n)
	for i := range ok {
		out[i] = ok[i] != 0
	}
	return out, err
}

// IsChecksumMismatch reports whether err is a hash mismatch from VerifyFile.
func IsChecksumMismatch(err error) bool {
	var xe *XetError
	return errors.As(err, &xe) && xe.Code == ErrorCodeChecksumMismatch
}

func convertGpuError(err *C.XetError) error {
	if err == nil {
		return nil
	}
	defer C.xetgpu_free_error(err)
	xetErr := &XetError{Code: int(err.code), Message: C.GoString(err.message)}
	if err.details != nil {
		xetErr.Details = C.GoString(err.details)
	}
	return xetErr
}
