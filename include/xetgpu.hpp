// xetgpu.hpp — C++ host-side mirror of the Go surface OME's pkg/xet would gain (header only).
//
// The reference's host language is Go (pkg/xet/xet.go over cgo); there is no Go toolchain in this build
// image, so the same surface is given here in C++ over the C ABI of xetgpu.h, name for name with
// integration/go/pkg/xet/xetgpu.go:
//     NewGpuHasher / Close            <-> xet::GpuHasher ctor / Close        (Client lifecycle, xet.go:241-308)
//     HashFiles, HashBytes            <-> same names
//     VerifyFile, VerifyFiles         <-> same names; a mismatch is an XetError with Code 6
//     SetProgressHandler              <-> same name                           (xet.go:93-151)
//     XetError{Code, Message, Details}<-> xet::XetError                       (xet.go:227-238, 507-522)
//     XetFileChecksum, XetChecksumAlgorithm, XetChecksumMetadataKey <-> same names: the "xet" case of the replica
//                                        agent's GetFileChecksum (internal/ome-agent/replica/replicator/helper.go:86-130)
//     NewGpuMultiHasher / VerifyFiles  <-> xet::GpuMultiHasher: every GPU of the node from one process
// Error behaviour follows the reference: every failure carries the reference's XetErrorCode
// (pkg/xet/src/error.rs:12-23) plus message and details; nothing is computed on the CPU.
#pragma once
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "xetgpu.h"

namespace xet {

constexpr int ErrorCodeChecksumMismatch = XETGPU_ERR_CHECKSUM_MISMATCH;
// Names the replica agent's ChecksumConfig / object metadata would use, next to "md5" / "sha256" and
// "opc-meta-md5" / "opc-meta-sha256" (internal/ome-agent/replica/replicator/helper.go:28-31).
constexpr const char* XetChecksumAlgorithm = "xet";
constexpr const char* XetChecksumMetadataKey = "opc-meta-xet";

struct XetError : std::runtime_error {
    int Code;
    std::string Message, Details;
    XetError(int code, std::string message, std::string details)
        : std::runtime_error("xet error " + std::to_string(code) + ": " + message +
                             (details.empty() ? "" : " (" + details + ")")),
          Code(code), Message(std::move(message)), Details(std::move(details)) {}
};

inline bool IsChecksumMismatch(const XetError& e) { return e.Code == ErrorCodeChecksumMismatch; }

struct FileDigest {
    std::string Hash;        // 64 hex digits, the X-Xet-Hash form
    std::string MerkleRoot;
    uint64_t Size = 0;
    uint64_t Chunks = 0;
};

struct ProgressUpdate {      // ProgressUpdate in xet.go
    int Phase = 0;
    uint64_t TotalBytes = 0, CompletedBytes = 0;
    uint32_t TotalFiles = 0, CompletedFiles = 0;
    std::string CurrentFile;
};
using ProgressHandler = std::function<void(const ProgressUpdate&)>;

namespace detail {
inline void check(::XetError* err) {  // convertGpuError
    if (!err) return;
    XetError e(err->code, err->message ? err->message : "", err->details ? err->details : "");
    xetgpu_free_error(err);
    throw e;
}
inline std::string hex(const uint8_t h[32]) {
    char buf[65];
    xetgpu_hash_to_hex(h, buf);
    return buf;
}
}  // namespace detail

class GpuHasher {
  public:
    explicit GpuHasher(int device = -1) {  // NewGpuHasher
        xetgpu_version_0_2_0();
        XetGpuConfig cfg;
        std::memset(&cfg, 0, sizeof cfg);
        cfg.device = device;
        detail::check(xetgpu_context_new(&cfg, &ctx_));
    }
    GpuHasher(const GpuHasher&) = delete;
    GpuHasher& operator=(const GpuHasher&) = delete;
    ~GpuHasher() { Close(); }

    void Close() {
        if (ctx_) xetgpu_context_free(ctx_);
        ctx_ = nullptr;
    }

    void SetProgressHandler(ProgressHandler handler, uint32_t throttle_ms = 200) {
        handler_ = std::move(handler);
        detail::check(xetgpu_context_set_progress_callback(ctx_, handler_ ? &GpuHasher::on_progress : nullptr, this,
                                                           throttle_ms));
    }

    // cancelled() is polled while files are read; returning true aborts with Code 7.
    std::vector<FileDigest> HashFiles(const std::vector<std::string>& paths, std::function<bool()> cancelled = {}) {
        std::vector<const char*> cp;
        for (const auto& p : paths) cp.push_back(p.c_str());
        XetCancellationToken tok{&GpuHasher::on_cancel, &cancelled};
        XetGpuResult* res = nullptr;
        detail::check(xetgpu_hash_files(ctx_, cp.data(), cp.size(), 0, cancelled ? &tok : nullptr, &res));
        std::vector<FileDigest> out(paths.size());
        for (size_t i = 0; i < paths.size(); i++)
            out[i] = FileDigest{detail::hex(res->files[i].file_hash), detail::hex(res->files[i].merkle_root),
                                res->files[i].size, res->files[i].n_chunks};
        xetgpu_free_result(res);
        return out;
    }

    FileDigest HashBytes(const void* data, size_t len) {
        const uint64_t offsets[2] = {0, len};
        XetGpuResult* res = nullptr;
        detail::check(xetgpu_hash_buffers(ctx_, data, offsets, 1, XETGPU_MEM_HOST, 0, &res));
        FileDigest d{detail::hex(res->files[0].file_hash), detail::hex(res->files[0].merkle_root), res->files[0].size,
                     res->files[0].n_chunks};
        xetgpu_free_result(res);
        return d;
    }

    // GetFileChecksum(filePath, "xet") of the replica agent (helper.go:86-109): path in, digest string out.
    // The digest is the 64-hex-digit form the Hub advertises (md5/sha256 are base64 there; a Xet hash is
    // compared with X-Xet-Hash, which is hex).
    std::string XetFileChecksum(const std::string& filePath) { return HashFiles({filePath})[0].Hash; }

    // Returns normally on a match; throws XetError{Code 6} (expected and computed hash in Details) otherwise.
    void VerifyFile(const std::string& path, const std::string& expectedHex) {
        detail::check(xetgpu_verify_file(ctx_, path.c_str(), expectedHex.c_str()));
    }

    // Snapshot verification in one batch.  An empty expected hash skips that file.  Returns the per-file
    // verdicts when all checked files match; throws XetError{Code 6} listing the mismatches otherwise
    // (verdicts are then available through LastVerdicts()).
    std::vector<bool> VerifyFiles(const std::vector<std::string>& paths, const std::vector<std::string>& expectedHex) {
        if (paths.size() != expectedHex.size()) throw XetError(XETGPU_ERR_INVALID_CONFIG, "paths and hashes differ in length", "");
        std::vector<const char*> cp, ch;
        for (size_t i = 0; i < paths.size(); i++) {
            cp.push_back(paths[i].c_str());
            ch.push_back(expectedHex[i].empty() ? nullptr : expectedHex[i].c_str());
        }
        std::vector<uint8_t> ok(paths.size() ? paths.size() : 1);
        ::XetError* err = xetgpu_verify_files(ctx_, cp.data(), ch.data(), paths.size(), nullptr, ok.data());
        verdicts_.assign(paths.size(), false);
        for (size_t i = 0; i < paths.size(); i++) verdicts_[i] = ok[i] != 0;
        detail::check(err);
        return verdicts_;
    }
    const std::vector<bool>& LastVerdicts() const { return verdicts_; }

  private:
    static void on_progress(const XetProgressUpdate* u, void* self) {
        auto* h = static_cast<GpuHasher*>(self);
        if (!h->handler_ || !u) return;
        ProgressUpdate p;
        p.Phase = (int)u->phase;
        p.TotalBytes = u->total_bytes;
        p.CompletedBytes = u->completed_bytes;
        p.TotalFiles = u->total_files;
        p.CompletedFiles = u->completed_files;
        if (u->current_file) p.CurrentFile = u->current_file;
        h->handler_(p);
    }
    static bool on_cancel(void* fn) {
        auto* f = static_cast<std::function<bool()>*>(fn);
        return *f && (*f)();
    }

    XetGpuContext* ctx_ = nullptr;
    ProgressHandler handler_;
    std::vector<bool> verdicts_;
};

// GetFileChecksum with the reference's signature and error behaviour for the algorithm this library adds:
// any other algorithm is "unsupported checksum algorithm: <name>" as in helper.go:95.
inline std::string GetFileChecksum(GpuHasher& h, const std::string& filePath, const std::string& algorithm) {
    if (algorithm != XetChecksumAlgorithm) throw std::invalid_argument("unsupported checksum algorithm: " + algorithm);
    return h.XetFileChecksum(filePath);
}

// Every GPU of the node from one process (xetgpu_multi_*).
class GpuMultiHasher {
  public:
    explicit GpuMultiHasher(const std::vector<int>& devices = {}) {  // NewGpuMultiHasher; empty = all visible devices
        xetgpu_version_0_2_0();
        std::vector<int32_t> d(devices.begin(), devices.end());
        detail::check(xetgpu_multi_new(d.empty() ? nullptr : d.data(), d.size(), nullptr, &ctx_));
    }
    GpuMultiHasher(const GpuMultiHasher&) = delete;
    GpuMultiHasher& operator=(const GpuMultiHasher&) = delete;
    ~GpuMultiHasher() { Close(); }
    void Close() {
        if (ctx_) xetgpu_multi_free(ctx_);
        ctx_ = nullptr;
    }
    size_t Devices() const { return xetgpu_multi_device_count(ctx_); }

    std::vector<FileDigest> HashFiles(const std::vector<std::string>& paths) {
        std::vector<const char*> cp;
        for (const auto& p : paths) cp.push_back(p.c_str());
        XetGpuResult* res = nullptr;
        detail::check(xetgpu_multi_hash_files(ctx_, cp.data(), cp.size(), 0, nullptr, &res));
        std::vector<FileDigest> out(paths.size());
        for (size_t i = 0; i < paths.size(); i++)
            out[i] = FileDigest{detail::hex(res->files[i].file_hash), detail::hex(res->files[i].merkle_root),
                                res->files[i].size, res->files[i].n_chunks};
        xetgpu_free_result(res);
        return out;
    }

    // As GpuHasher::VerifyFiles.
    std::vector<bool> VerifyFiles(const std::vector<std::string>& paths, const std::vector<std::string>& expectedHex) {
        if (paths.size() != expectedHex.size()) throw XetError(XETGPU_ERR_INVALID_CONFIG, "paths and hashes differ in length", "");
        std::vector<const char*> cp, ch;
        for (size_t i = 0; i < paths.size(); i++) {
            cp.push_back(paths[i].c_str());
            ch.push_back(expectedHex[i].empty() ? nullptr : expectedHex[i].c_str());
        }
        std::vector<uint8_t> ok(paths.size() ? paths.size() : 1);
        ::XetError* err = xetgpu_multi_verify_files(ctx_, cp.data(), ch.data(), paths.size(), nullptr, ok.data());
        verdicts_.assign(paths.size(), false);
        for (size_t i = 0; i < paths.size(); i++) verdicts_[i] = ok[i] != 0;
        detail::check(err);
        return verdicts_;
    }
    const std::vector<bool>& LastVerdicts() const { return verdicts_; }

  private:
    XetGpuMulti* ctx_ = nullptr;
    std::vector<bool> verdicts_;
};

}  // namespace xet
