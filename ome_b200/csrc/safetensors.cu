// Safetensors-aware ingestion: read a shard's header the way OME's model-config parser does
// (pkg/hfutil/modelconfig/safetensors.go:67-141: u64-LE header length, at most 10 MB of JSON, per-tensor
// dtype / shape / data_offsets) and use the tensor table to place byte-range boundaries on tensor boundaries.
// Host code only; it is part of libxetgpu because the multi-device file call uses it to plan its ranges.
//
// Where a boundary falls never changes a hash (the byte-range protocol is exact for any cut); aligning cuts with
// tensors keeps each device's share a set of whole tensors, which is what a loader that maps tensors to devices,
// or a per-tensor dedup statistic, wants to see.
#include <ctype.h>
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/xetgpu.h"

namespace {

char* dup_str(const std::string& s) {
    char* p = (char*)malloc(s.size() + 1);
    if (p) memcpy(p, s.c_str(), s.size() + 1);
    return p;
}

XetError* st_error(int32_t code, const std::string& message, const std::string& details = "") {
    XetError* e = (XetError*)malloc(sizeof(XetError));
    if (!e) return nullptr;
    e->code = code;
    e->message = dup_str(message);
    e->details = details.empty() ? nullptr : dup_str(details);
    return e;
}

constexpr uint64_t kMaxHeaderSize = 10ull * 1024 * 1024;  // safetensors.go:85 (maxHeaderSize)

// ---- a JSON reader just large enough for a safetensors header ------------------------------------------
// Grammar: RFC 8259 values.  Only what the header needs is kept: per top-level key an object whose "dtype" is a
// string, "shape" an array of integers and "data_offsets" an array of two integers; everything else is skipped
// structurally (so "__metadata__" with arbitrary string values parses, as encoding/json accepts it).
struct Json {
    const char* p;
    const char* end;
    bool ok = true;

    void ws() {
        while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++;
    }
    bool eat(char c) {
        ws();
        if (p < end && *p == c) {
            p++;
            return true;
        }
        return false;
    }
    bool string(std::string* out) {
        ws();
        if (p >= end || *p != '"') return ok = false;
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (p + 1 >= end) return ok = false;
                p++;
                char c = *p++;
                switch (c) {
                    case 'n': c = '\n'; break;
                    case 't': c = '\t'; break;
                    case 'r': c = '\r'; break;
                    case 'b': c = '\b'; break;
                    case 'f': c = '\f'; break;
                    case 'u': {  // keep the escape as written: tensor names are compared bytewise nowhere
                        if (p + 4 > end) return ok = false;
                        if (out) out->append("\\u").append(p, 4);
                        p += 4;
                        continue;
                    }
                    default: break;  // \" \\ \/
                }
                if (out) out->push_back(c);
            } else {
                if (out) out->push_back(*p);
                p++;
            }
        }
        if (p >= end) return ok = false;
        p++;
        return true;
    }
    // integer without fraction / exponent; anything else is "not an integer"
    bool integer(int64_t* out) {
        ws();
        const char* s = p;
        bool neg = false;
        if (p < end && *p == '-') {
            neg = true;
            p++;
        }
        if (p >= end || *p < '0' || *p > '9') return ok = false;
        unsigned __int128 v = 0;
        while (p < end && *p >= '0' && *p <= '9') {
            v = v * 10 + (unsigned)(*p - '0');
            if (v > (unsigned __int128)INT64_MAX + 1) return ok = false;
            p++;
        }
        if (p < end && (*p == '.' || *p == 'e' || *p == 'E')) {
            p = s;
            return ok = false;
        }
        if (!neg && v > (unsigned __int128)INT64_MAX) return ok = false;
        *out = neg ? -(int64_t)v : (int64_t)v;
        return true;
    }
    bool skip_value() {
        ws();
        if (p >= end) return ok = false;
        if (*p == '"') return string(nullptr);
        if (*p == '{') {
            p++;
            if (eat('}')) return true;
            do {
                if (!string(nullptr) || !eat(':') || !skip_value()) return ok = false;
            } while (eat(','));
            return eat('}') ? true : (ok = false);
        }
        if (*p == '[') {
            p++;
            if (eat(']')) return true;
            do {
                if (!skip_value()) return ok = false;
            } while (eat(','));
            return eat(']') ? true : (ok = false);
        }
        // number / true / false / null
        const char* s = p;
        while (p < end && (isalnum((unsigned char)*p) || *p == '-' || *p == '+' || *p == '.')) p++;
        return p > s ? true : (ok = false);
    }
};

struct TensorRow {
    std::string name, dtype;
    std::vector<int64_t> shape;
    bool has_shape = false;
    int64_t off[2] = {0, 0};
    bool has_offsets = false;
};

bool parse_tensor_object(Json& j, TensorRow* t) {
    if (!j.eat('{')) return j.ok = false;
    if (j.eat('}')) return true;
    do {
        std::string key;
        if (!j.string(&key) || !j.eat(':')) return j.ok = false;
        if (key == "dtype") {
            j.ws();
            if (j.p < j.end && *j.p == '"') {
                if (!j.string(&t->dtype)) return false;
            } else if (!j.skip_value()) return false;
        } else if (key == "shape" || key == "data_offsets") {
            j.ws();
            if (j.p >= j.end || *j.p != '[') {  // e.g. null: encoding/json leaves the field at its zero value
                if (!j.skip_value()) return false;
                continue;
            }
            j.p++;
            std::vector<int64_t> v;
            if (!j.eat(']')) {
                do {
                    int64_t x;
                    if (!j.integer(&x)) return false;
                    v.push_back(x);
                } while (j.eat(','));
                if (!j.eat(']')) return j.ok = false;
            }
            if (key == "shape") {
                t->shape = v;
                t->has_shape = true;
            } else {
                if (v.size() != 2) return j.ok = false;  // the partitioner needs both ends (the reference never reads them)
                t->off[0] = v[0];
                t->off[1] = v[1];
                t->has_offsets = true;
            }
        } else if (!j.skip_value()) {
            return false;
        }
    } while (j.eat(','));
    return j.eat('}') ? true : (j.ok = false);
}

}  // namespace

// ---- index ----------------------------------------------------------------------------------------------
static XetError* build_index(const uint8_t* head, uint64_t head_len, uint64_t file_size, const char* what,
                             XetGpuSafetensorsIndex** out) {
    const std::string where = what ? what : "<memory>";
    if (head_len < 8) return st_error(XETGPU_ERR_IO, "failed to read header length from '" + where + "'", "unexpected EOF");
    uint64_t header_len = 0;
    for (int i = 7; i >= 0; i--) header_len = (header_len << 8) | head[i];
    if (header_len > kMaxHeaderSize)
        return st_error(XETGPU_ERR_INVALID_CONFIG,
                        "header length " + std::to_string(header_len) + " in '" + where +
                            "' exceeds maximum allowed size of " + std::to_string(kMaxHeaderSize) + " bytes");
    if (head_len - 8 < header_len)
        return st_error(XETGPU_ERR_IO, "failed to read JSON header from '" + where + "'", "unexpected EOF");
    Json j{(const char*)head + 8, (const char*)head + 8 + header_len};
    std::vector<TensorRow> rows;
    bool good = j.eat('{');
    if (good && !j.eat('}')) {
        do {
            TensorRow t;
            if (!j.string(&t.name) || !j.eat(':') || !parse_tensor_object(j, &t)) {
                good = false;
                break;
            }
            rows.push_back(std::move(t));
        } while (j.eat(','));
        good = good && j.eat('}');
    }
    j.ws();
    if (!good || !j.ok || j.p != j.end)
        return st_error(XETGPU_ERR_INVALID_CONFIG, "failed to parse JSON header from '" + where + "'",
                        "at byte " + std::to_string((uint64_t)(j.p - ((const char*)head + 8))) + " of the header");
    const uint64_t data_start = 8 + header_len;
    int64_t total = 0;
    std::vector<const TensorRow*> tensors;
    for (const auto& t : rows) {
        if (t.name == "__metadata__") continue;
        // parameter count exactly as modelconfig.ParseSafetensors (safetensors.go:111-141)
        if (!t.shape.empty()) {
            int64_t count = 1;
            for (size_t i = 0; i < t.shape.size(); i++) {
                const int64_t dim = t.shape[i];
                if (dim <= 0)
                    return st_error(XETGPU_ERR_INVALID_CONFIG, "tensor '" + t.name + "' in '" + where + "' has invalid dimension " +
                                                                   std::to_string(dim) + " at index " + std::to_string(i));
                if (dim > INT64_MAX / count)
                    return st_error(XETGPU_ERR_INVALID_CONFIG, "dimension overflow for tensor '" + t.name + "' in '" + where +
                                                                   "': multiplication would exceed maximum value");
                count *= dim;
            }
            if (total > INT64_MAX - count)
                return st_error(XETGPU_ERR_INVALID_CONFIG, "parameter count overflow in '" + where + "': total would exceed maximum value");
            total += count;
        }
        // the partitioner needs sane extents (the reference never looks at data_offsets)
        if (t.has_offsets) {
            if (t.off[0] < 0 || t.off[1] < t.off[0] ||
                (file_size && (uint64_t)t.off[1] > file_size - std::min(file_size, data_start)))
                return st_error(XETGPU_ERR_INVALID_CONFIG, "tensor '" + t.name + "' in '" + where + "' has data_offsets outside the file",
                                "[" + std::to_string(t.off[0]) + ", " + std::to_string(t.off[1]) + "]");
            tensors.push_back(&t);
        }
    }
    std::stable_sort(tensors.begin(), tensors.end(), [](const TensorRow* a, const TensorRow* b) { return a->off[0] < b->off[0]; });
    XetGpuSafetensorsIndex* ix = (XetGpuSafetensorsIndex*)calloc(1, sizeof(XetGpuSafetensorsIndex));
    if (!ix) return st_error(XETGPU_ERR_UNKNOWN, "out of host memory");
    ix->file_size = file_size;
    ix->header_len = header_len;
    ix->data_start = data_start;
    ix->parameter_count = total;
    ix->n_tensors = tensors.size();
    ix->tensors = (XetGpuTensorInfo*)calloc(tensors.size() ? tensors.size() : 1, sizeof(XetGpuTensorInfo));
    if (!ix->tensors) {
        free(ix);
        return st_error(XETGPU_ERR_UNKNOWN, "out of host memory");
    }
    for (size_t i = 0; i < tensors.size(); i++) {
        const TensorRow& t = *tensors[i];
        XetGpuTensorInfo& o = ix->tensors[i];
        o.name = dup_str(t.name);
        o.dtype = dup_str(t.dtype);
        o.begin = data_start + (uint64_t)t.off[0];
        o.end = data_start + (uint64_t)t.off[1];
        o.rank = (uint32_t)t.shape.size();
        uint64_t n = t.shape.empty() ? 0 : 1;
        for (int64_t d : t.shape) n *= (uint64_t)d;
        o.n_elements = n;
    }
    *out = ix;
    return nullptr;
}

extern "C" XetError* xetgpu_safetensors_parse(const void* head, uint64_t head_len, uint64_t file_size,
                                              XetGpuSafetensorsIndex** out_index) {
    if (!out_index || (!head && head_len)) return st_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_safetensors_parse");
    *out_index = nullptr;
    return build_index((const uint8_t*)head, head_len, file_size, nullptr, out_index);
}

extern "C" XetError* xetgpu_safetensors_open(const char* path, XetGpuSafetensorsIndex** out_index) {
    if (!out_index) return st_error(XETGPU_ERR_INVALID_CONFIG, "out_index is NULL");
    *out_index = nullptr;
    if (!path || !*path) return st_error(XETGPU_ERR_INVALID_CONFIG, "safetensors file path cannot be empty");
    int fd = open(path, O_RDONLY | O_CLOEXEC);
    if (fd < 0)
        return st_error(errno == ENOENT ? XETGPU_ERR_NOT_FOUND : XETGPU_ERR_IO,
                        std::string("failed to open safetensors file '") + path + "'", strerror(errno));
    struct stat st;
    uint64_t file_size = fstat(fd, &st) == 0 ? (uint64_t)st.st_size : 0;
    auto read_full = [&](uint8_t* dst, uint64_t n, uint64_t off) -> uint64_t {
        uint64_t got = 0;
        while (got < n) {
            ssize_t r = pread(fd, dst + got, n - got, (off_t)(off + got));
            if (r < 0 && errno == EINTR) continue;
            if (r <= 0) break;
            got += (uint64_t)r;
        }
        return got;
    };
    std::vector<uint8_t> head(8);
    uint64_t got = read_full(head.data(), 8, 0);
    if (got == 8) {
        uint64_t header_len = 0;
        for (int i = 7; i >= 0; i--) header_len = (header_len << 8) | head[i];
        if (header_len <= kMaxHeaderSize) {  // (a larger value is reported by build_index, without allocating for it)
            head.resize(8 + header_len);
            got = 8 + read_full(head.data() + 8, header_len, 8);
        }
    }
    close(fd);
    return build_index(head.data(), got, file_size, path, out_index);
}

extern "C" void xetgpu_safetensors_free(XetGpuSafetensorsIndex* ix) {
    if (!ix) return;
    for (uint64_t i = 0; i < ix->n_tensors; i++) {
        free((void*)ix->tensors[i].name);
        free((void*)ix->tensors[i].dtype);
    }
    free(ix->tensors);
    free(ix);
}

// Cut [0, file_size) into n_ranges byte ranges whose interior bounds sit on tensor boundaries: for the r-th equal-size
// target the nearest tensor begin (or the end of the last tensor) is taken, provided both neighbouring ranges keep at
// least XETGPU_RANGE_LEAD + 4096 bytes (what the halo needs) and the boundary lies within a quarter range of the target;
// otherwise (one tensor larger than a range) the target itself, aligned down to 4096, is used.  Ranges may be empty
// for files too small to share.  out_bounds has n_ranges + 1 ascending entries, [0] = 0, [n_ranges] = file_size.
extern "C" XetError* xetgpu_safetensors_split(const XetGpuSafetensorsIndex* ix, uint64_t file_size, size_t n_ranges,
                                              uint64_t* out_bounds) {
    if (!ix || !out_bounds || n_ranges == 0) return st_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_safetensors_split");
    const uint64_t min_piece = XETGPU_RANGE_LEAD + 4096;
    std::vector<uint64_t> marks;  // candidate boundaries, ascending
    for (uint64_t i = 0; i < ix->n_tensors; i++) {
        marks.push_back(ix->tensors[i].begin);
        marks.push_back(ix->tensors[i].end);
    }
    std::sort(marks.begin(), marks.end());
    marks.erase(std::unique(marks.begin(), marks.end()), marks.end());
    out_bounds[0] = 0;
    for (size_t r = 1; r < n_ranges; r++) {
        const uint64_t target = (uint64_t)((unsigned __int128)file_size * r / n_ranges);
        const uint64_t slack = file_size / n_ranges / 4;
        uint64_t b = target / 4096 * 4096;
        auto it = std::lower_bound(marks.begin(), marks.end(), target);
        uint64_t best = UINT64_MAX, best_d = UINT64_MAX;
        for (int k = 0; k < 2; k++) {
            if (k == 0 && it == marks.end()) continue;
            if (k == 1 && it == marks.begin()) continue;
            const uint64_t m = k == 0 ? *it : *(it - 1);
            const uint64_t d = m > target ? m - target : target - m;
            if (d < best_d) {
                best_d = d;
                best = m;
            }
        }
        if (best != UINT64_MAX && best_d <= slack) b = best;
        const uint64_t prev = out_bounds[r - 1];
        if (b < prev + min_piece) b = prev;                      // too close to the previous bound: empty range
        if (file_size - std::min(file_size, b) < min_piece) b = file_size;  // sliver at the end: fold it in
        if (b != prev && b != file_size && b < min_piece) b = prev;
        out_bounds[r] = std::max(b, prev);
    }
    out_bounds[n_ranges] = file_size;
    return nullptr;
}
