// C ABI of libxetgpu.so (include/xetgpu.h): context, workspace, host<->device ingestion and the
// stage pipeline.  Conventions follow the reference's libxet FFI (pkg/xet/src/ffi.rs, error.rs):
// NULL == success, heap-allocated XetError otherwise, out-pointers for results, NULL-safe frees.
//
// There is no CPU fallback anywhere in this file: if the device or a kernel launch fails the call
// returns an error.
#include <cuda_runtime.h>
#include <errno.h>
#include <fcntl.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/xetgpu.h"
#include "xetgpu_internal.h"

using namespace xetgpu;

namespace xetgpu {
void init_chunk_hash_kernels();  // chunk_hash.cu
void init_gear_kernels();        // gear.cu
}  // namespace xetgpu

// ------------------------------------------------------------------------------------ errors
static char* dup_cstr(const std::string& s) {
    char* p = (char*)malloc(s.size() + 1);
    if (p) memcpy(p, s.c_str(), s.size() + 1);
    return p;
}

static XetError* make_error(int32_t code, const std::string& message, const std::string& details = "") {
    XetError* e = (XetError*)malloc(sizeof(XetError));
    if (!e) return nullptr;
    e->code = code;
    e->message = dup_cstr(message);
    e->details = details.empty() ? nullptr : dup_cstr(details);
    return e;
}

static XetError* cuda_error(cudaError_t err, const char* what) {
    return make_error(XETGPU_ERR_UNKNOWN, std::string("CUDA error in ") + what + ": " + cudaGetErrorString(err),
                      cudaGetErrorName(err));
}

#define CU(expr)                                         \
    do {                                                 \
        cudaError_t _e = (expr);                         \
        if (_e != cudaSuccess) return cuda_error(_e, #expr); \
    } while (0)

extern "C" void xetgpu_free_error(XetError* err) {
    if (!err) return;
    free(err->message);
    free(err->details);
    free(err);
}

extern "C" void xetgpu_version_0_2_0(void) {}

// ----------------------------------------------------------------------------------- context
struct Arena {
    uint8_t* base = nullptr;
    uint64_t cap = 0;
    uint64_t used = 0;
    template <typename T>
    T* take(uint64_t count) {
        uint64_t bytes = (count * sizeof(T) + 255) & ~255ull;
        T* p = reinterpret_cast<T*>(base + used);
        used += bytes;
        return p;
    }
};

// A lane is a workspace + pinned metadata + timing events; lanes alternate between TWO streams.  A batch of several
// file groups gives every group its own lane, so the host enqueues all groups without waiting for any of them and
// collects the results at the end (round 1 had two lanes and had to drain group g before it could launch g + 2).
constexpr int kMaxLanes = 8;
constexpr size_t kReadSlotsDefault = 16;  // pinned 16 MiB slots of the file reader ring (XETGPU_READ_SLOTS overrides; measured on B200:
                                          // 16, 32 and 64 slots x 16 or 32 threads all read tmpfs at 42-45 GB/s — the host memory system is the bound)
constexpr size_t kReadThreadsMax = 16;

struct XetGpuContext {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint32_t flags = 0;
    uint32_t sm_count = 148;
    uint64_t staging_bytes = 64ull << 20;
    uint64_t max_batch_bytes = 8ull << 30;
    std::mutex mu;                            // serialises the hashing entry points of this context
    std::atomic<std::thread::id> owner{};     // thread inside a hashing call (re-entrancy from a callback is refused)
    std::mutex cb_mu;                         // guards the progress fields and `last` (never held while a callback runs)
    // Two lanes: a lane is a stream with its own grow-only workspace, pinned metadata area and timing
    // events.  Lane 0 runs on the context's stream; batches of several files alternate between the lanes
    // so that one group's copies and gear pass overlap the previous group's hashing.
    struct Lane {
        cudaStream_t stream = nullptr;
        Arena ws;
        uint8_t* h_meta = nullptr;
        uint64_t h_meta_cap = 0;
        uint8_t* h_chunks = nullptr;  // pinned bounce buffer for chunk-list readback (pageable D2H runs at ~2.5 GB/s)
        uint64_t h_chunks_cap = 0;
        cudaEvent_t ev[8] = {};
        cudaEvent_t ev_kernel[4] = {};
    } lanes[kMaxLanes];
    cudaStream_t second_stream = nullptr;  // odd lanes launch here, even lanes on `stream`
    cudaEvent_t ev_fork = nullptr;
    cudaEvent_t ev_gear_done[kMaxLanes] = {};
    cudaEvent_t ev_head_done[kMaxLanes] = {};   // staged schedule: K2 + work lists of a group are in place
    cudaEvent_t ev_leaf_done[kMaxLanes] = {};   // staged schedule: the group's leaf kernel has finished
    int force_staged = -1;                      // XETGPU_STAGED=0/1 overrides the staged-schedule policy
    int force_lanes = -1;  // XETGPU_TWO_LANES=0/1 overrides the two-lane policy (experiments)
    uint64_t min_group_bytes = 256ull << 20;  // smallest file group worth its own lane (XETGPU_MIN_GROUP_BYTES)
    uint8_t* d_input = nullptr;  // device copy of host inputs
    uint64_t d_input_cap = 0;
    uint8_t* h_stage[2] = {nullptr, nullptr};  // pinned ingestion buffers
    cudaEvent_t stage_free[2] = {nullptr, nullptr};
    // file ingestion: ring of pinned read slots filled by reader threads (xetgpu_hash_files)
    std::vector<uint8_t*> read_slots;
    std::vector<cudaEvent_t> read_slot_ev;
    uint64_t read_slot_bytes = 16ull << 20;
    size_t read_threads = 8;
    size_t n_read_slots = kReadSlotsDefault;
    uint64_t cand_cap_hint = 0;
    XetGpuTimings last = {};
    // progress
    XetProgressCallback progress = nullptr;
    void* progress_user = nullptr;
    uint32_t throttle_ms = 0;
    std::chrono::steady_clock::time_point last_progress;
};

// Entry-point lock.  Calls on one context serialise on c->mu.  A progress / cancellation callback runs on a thread
// that is inside such a call; if it calls a hashing entry point of the SAME context it would wait for itself, so
// that case is detected and refused with XETGPU_ERR_INVALID_CONFIG instead of deadlocking.
// Set while a user callback (progress or cancellation) runs on this thread: hashing entry points refuse to start
// from inside one, whichever handle they are given (a multi-device call runs callbacks on its worker threads).
static thread_local int tl_in_callback = 0;
struct CallbackScope {
    CallbackScope() { tl_in_callback++; }
    ~CallbackScope() { tl_in_callback--; }
};

struct CtxGuard {
    XetGpuContext* c;
    bool ok;
    explicit CtxGuard(XetGpuContext* ctx) : c(ctx), ok(true) {
        if (tl_in_callback || c->owner.load() == std::this_thread::get_id()) {
            ok = false;
            return;
        }
        c->mu.lock();
        c->owner.store(std::this_thread::get_id());
    }
    ~CtxGuard() {
        if (!ok) return;
        c->owner.store(std::thread::id());
        c->mu.unlock();
    }
};
#define LOCK_CTX(c)                                                                                     \
    CtxGuard _ctx_guard(c);                                                                             \
    if (!_ctx_guard.ok)                                                                                 \
    return make_error(XETGPU_ERR_INVALID_CONFIG, "re-entrant call: a progress or cancellation callback " \
                                                 "called a hashing entry point of the context that is running it")



typedef XetGpuContext::Lane Lane;

static XetError* ensure_meta(Lane* l, uint64_t bytes) {
    if (bytes <= l->h_meta_cap) return nullptr;
    if (l->h_meta) cudaFreeHost(l->h_meta);
    l->h_meta = nullptr;
    l->h_meta_cap = 0;
    uint64_t cap = std::max<uint64_t>(bytes * 2, 1 << 20);
    CU(cudaMallocHost((void**)&l->h_meta, cap));
    l->h_meta_cap = cap;
    return nullptr;
}

static XetError* ensure_input(XetGpuContext* c, uint64_t bytes) {
    bytes += 256;  // slack: kernels read whole 16-byte units
    if (bytes <= c->d_input_cap) return nullptr;
    if (c->d_input) cudaFree(c->d_input);
    c->d_input = nullptr;
    c->d_input_cap = 0;
    CU(cudaMalloc((void**)&c->d_input, bytes));
    c->d_input_cap = bytes;
    return nullptr;
}

extern "C" void xetgpu_context_free(XetGpuContext* c);

extern "C" XetError* xetgpu_context_new(const XetGpuConfig* config, XetGpuContext** out_ctx) {
    if (!out_ctx) return make_error(XETGPU_ERR_INVALID_CONFIG, "out_ctx is NULL");
    *out_ctx = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return make_error(XETGPU_ERR_UNKNOWN, "no CUDA device available (libxetgpu has no CPU path)",
                          e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    int dev = config ? config->device : -1;
    if (dev < 0) CU(cudaGetDevice(&dev));
    if (dev >= ndev) return make_error(XETGPU_ERR_INVALID_CONFIG, "device ordinal out of range");
    CU(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10)
        return make_error(XETGPU_ERR_UNKNOWN, "libxetgpu is built for sm_100a (B200) only", prop.name);
    XetGpuContext* c = new XetGpuContext();
    c->device = dev;
    c->sm_count = (uint32_t)prop.multiProcessorCount;
    if (config) {
        c->flags = config->flags;
        if (config->staging_bytes) c->staging_bytes = (config->staging_bytes + 4095) & ~4095ull;
        if (config->staging_bytes) c->read_slot_bytes = std::min<uint64_t>(c->staging_bytes, 64ull << 20);
        if (config->max_batch_bytes) c->max_batch_bytes = config->max_batch_bytes;
        c->stream = (cudaStream_t)config->stream;
    }
    // From here on every failure tears the context down through xetgpu_context_free, which is NULL-tolerant per member.
    cudaError_t ce = cudaSuccess;
    auto ok = [&](cudaError_t r) {
        if (ce == cudaSuccess && r != cudaSuccess) ce = r;
        return ce == cudaSuccess;
    };
    if (!c->stream) {
        if (ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking))) c->own_stream = true;
    }
    ok(cudaStreamCreateWithFlags(&c->second_stream, cudaStreamNonBlocking));
    for (int k = 0; k < kMaxLanes; k++) c->lanes[k].stream = (k & 1) ? c->second_stream : c->stream;
    for (auto& l : c->lanes) {
        for (auto& ev : l.ev) ok(cudaEventCreate(&ev));
        for (auto& ev : l.ev_kernel) ok(cudaEventCreate(&ev));
    }
    ok(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    for (auto& ev : c->ev_gear_done) ok(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (auto& ev : c->ev_head_done) ok(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    for (auto& ev : c->ev_leaf_done) ok(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    if (const char* v = getenv("XETGPU_STAGED")) c->force_staged = atoi(v) ? 1 : 0;
    for (auto& ev : c->stage_free) ok(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    if (const char* v = getenv("XETGPU_TWO_LANES")) c->force_lanes = atoi(v) ? 1 : 0;
    if (const char* v = getenv("XETGPU_MIN_GROUP_BYTES")) {  // test knob: exercise the two-lane path on small inputs
        unsigned long long n = strtoull(v, nullptr, 10);
        if (n) c->min_group_bytes = n;
    }
    if (ce == cudaSuccess) {
        init_gear_kernels();
        init_chunk_hash_kernels();
        ce = cudaGetLastError();
    }
    if (ce != cudaSuccess) {
        xetgpu_context_free(c);
        return cuda_error(ce, "context set-up (streams, events, kernel attributes)");
    }
    {
        unsigned hc = std::thread::hardware_concurrency();
        c->read_threads = hc ? std::min<size_t>(kReadThreadsMax, hc) : 4;
        if (const char* v = getenv("XETGPU_READ_SLOTS")) {
            unsigned long n = strtoul(v, nullptr, 10);
            if (n) c->n_read_slots = std::min<size_t>(std::max<size_t>(n, 2), 256);
        }
        if (const char* v = getenv("XETGPU_READ_THREADS")) {
            unsigned long n = strtoul(v, nullptr, 10);
            if (n) c->read_threads = std::min<size_t>(n, 64);
        }
    }
    *out_ctx = c;
    return nullptr;
}

extern "C" void xetgpu_context_free(XetGpuContext* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto& l : c->lanes) {
        if (l.stream) cudaStreamSynchronize(l.stream);
        if (l.ws.base) cudaFree(l.ws.base);
        if (l.h_meta) cudaFreeHost(l.h_meta);
        if (l.h_chunks) cudaFreeHost(l.h_chunks);
        for (auto& ev : l.ev)
            if (ev) cudaEventDestroy(ev);
        for (auto& ev : l.ev_kernel)
            if (ev) cudaEventDestroy(ev);
    }
    if (c->second_stream) {
        cudaStreamSynchronize(c->second_stream);
        cudaStreamDestroy(c->second_stream);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (auto& ev : c->ev_gear_done)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : c->ev_head_done)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : c->ev_leaf_done)
        if (ev) cudaEventDestroy(ev);
    if (c->d_input) cudaFree(c->d_input);
    for (auto& p : c->h_stage)
        if (p) cudaFreeHost(p);
    for (auto& p : c->read_slots) cudaFreeHost(p);
    for (auto& ev : c->read_slot_ev) cudaEventDestroy(ev);
    for (auto& ev : c->stage_free)
        if (ev) cudaEventDestroy(ev);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" XetError* xetgpu_context_set_progress_callback(XetGpuContext* c, XetProgressCallback callback,
                                                          void* user_data, uint32_t throttle_ms) {
    if (!c) return make_error(XETGPU_ERR_INVALID_CONFIG, "ctx is NULL");
    std::lock_guard<std::mutex> lk(c->cb_mu);
    c->progress = callback;
    c->progress_user = user_data;
    c->throttle_ms = throttle_ms;
    return nullptr;
}

// The handler is copied under cb_mu and invoked with no library lock held, so it may call
// xetgpu_context_set_progress_callback / xetgpu_last_timings on the same context.
static void report_progress(XetGpuContext* c, const XetProgressUpdate& u, bool force) {
    XetProgressCallback cb;
    void* user;
    {
        std::lock_guard<std::mutex> lk(c->cb_mu);
        if (!c->progress) return;
        auto now = std::chrono::steady_clock::now();
        if (!force && c->throttle_ms &&
            std::chrono::duration_cast<std::chrono::milliseconds>(now - c->last_progress).count() < c->throttle_ms)
            return;
        c->last_progress = now;
        cb = c->progress;
        user = c->progress_user;
    }
    CallbackScope scope;
    cb(&u, user);
}

// --------------------------------------------------------------------------------- hex helpers
extern "C" void xetgpu_hash_to_hex(const uint8_t hash[32], char out_hex[65]) {
    static const char* d = "0123456789abcdef";
    for (int w = 0; w < 4; w++)
        for (int b = 0; b < 8; b++) {
            uint8_t v = hash[8 * w + (7 - b)];  // LE u64 word printed most-significant byte first
            out_hex[16 * w + 2 * b] = d[v >> 4];
            out_hex[16 * w + 2 * b + 1] = d[v & 15];
        }
    out_hex[64] = 0;
}

extern "C" XetError* xetgpu_hash_from_hex(const char* hex, uint8_t out_hash[32]) {
    if (!hex || !out_hash) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument");
    if (strlen(hex) != 64) return make_error(XETGPU_ERR_INVALID_CONFIG, "a Xet hash is 64 hex digits", hex);
    for (int w = 0; w < 4; w++)
        for (int b = 0; b < 8; b++) {
            int v = 0;
            for (int k = 0; k < 2; k++) {
                char ch = hex[16 * w + 2 * b + k];
                int x = (ch >= '0' && ch <= '9')   ? ch - '0'
                        : (ch >= 'a' && ch <= 'f') ? ch - 'a' + 10
                        : (ch >= 'A' && ch <= 'F') ? ch - 'A' + 10
                                                   : -1;
                if (x < 0) return make_error(XETGPU_ERR_INVALID_CONFIG, "invalid hex digit in Xet hash", hex);
                v = v * 16 + x;
            }
            out_hash[8 * w + (7 - b)] = (uint8_t)v;
        }
    return nullptr;
}

// ---------------------------------------------------------------------------------- workspace
struct Plan {
    uint64_t total = 0;
    uint32_t n_files = 0;
    uint64_t n_slots = 0;
    uint32_t n_tiles = 0;
    uint64_t n_groups_cap = 0;
    uint64_t cand_cap = 0;
    uint64_t max_nodes = 0;
    std::vector<uint64_t> slot_off;  // host prefix
    std::vector<uint64_t> seg_off;   // host prefix of per-file segment counts (K2)
    uint64_t n_segs = 0;
    bool has_items = false;          // the batch carries per-file byte-range descriptors (RunOptions::items)
    std::vector<uint64_t> xorb_off;  // host prefix of per-file xorb upper bounds (only with want_xorbs)
    uint64_t n_xorb_slots = 0;
};

static uint32_t bit_length(uint64_t v) {
    uint32_t n = 0;
    while (v) {
        n++;
        v >>= 1;
    }
    return n;
}

static Plan make_plan(const uint64_t* offsets, size_t n_files, uint64_t total, uint64_t cand_cap_hint,
                      bool want_xorbs = false) {
    Plan p;
    if (want_xorbs) {  // a xorb other than a file's last holds more than 64 MiB - 128 KiB
        p.xorb_off.resize(n_files + 1);
        uint64_t x = 0;
        for (size_t f = 0; f < n_files; f++) {
            p.xorb_off[f] = x;
            x += (offsets[f + 1] - offsets[f]) / (kMaxXorbBytes - kMaxChunk) + 2;
        }
        p.xorb_off[n_files] = x;
        p.n_xorb_slots = x;
    }
    p.total = total;
    p.n_files = (uint32_t)n_files;
    p.slot_off.resize(n_files + 1);
    p.seg_off.resize(n_files + 1);
    uint64_t s = 0, max_slots = 0, g = 0;
    for (size_t f = 0; f < n_files; f++) {
        p.slot_off[f] = s;
        p.seg_off[f] = g;
        uint64_t len = offsets[f + 1] - offsets[f];
        g += (len + kSegBytes - 1) / kSegBytes;
        uint64_t ub = len / kMinChunk + 2;  // every chunk but the last is >= 8192 bytes (+1 sentinel slot)
        max_slots = std::max(max_slots, ub);
        s += ub;
    }
    p.slot_off[n_files] = s;
    p.n_slots = s;
    p.seg_off[n_files] = g;
    p.n_segs = g;
    p.n_tiles = (uint32_t)((total + kGearTile - 1) / kGearTile);
    p.n_groups_cap = total / (kLeavesPerGroup * 1024ull) + s;
    p.cand_cap = std::max<uint64_t>(std::max<uint64_t>(total / 1024 + 4096, cand_cap_hint), 1);
    p.max_nodes = max_slots;
    return p;
}

static XetError* carve(Lane* l, const Plan& p, PipelineBuffers* b, bool dry, uint64_t* need) {
    Arena a = l->ws;
    a.used = 0;
    if (dry) a.base = nullptr;
    const uint64_t F = p.n_files, S = p.n_slots, T = p.n_tiles;
    b->n_files = p.n_files;
    b->n_slots = S;
    b->n_tiles = p.n_tiles;
    b->n_groups_cap = p.n_groups_cap;
    b->cand_cap = p.cand_cap;
    b->max_nodes = p.max_nodes;
    b->total_bytes = p.total;
    b->file_off = a.take<uint64_t>(F + 1);
    uint64_t* walk = a.take<uint64_t>(F + 1);
    b->file_walk = p.has_items ? walk : nullptr;
    b->slot_off = a.take<uint64_t>(F + 1);
    b->tile_count = a.take<uint32_t>(T + 1);
    b->tile_stage = a.take<uint32_t>(T * (uint64_t)kGearCap + 1);
    b->tile_off = a.take<uint64_t>(T + 2);
    b->cand = a.take<uint64_t>(p.cand_cap);
    b->slot_end = a.take<uint64_t>(S + 1);
    const uint64_t G = p.n_segs;
    b->n_segs = G;
    b->seg_off = a.take<uint64_t>(F + 1);
    b->seg_spec = a.take<uint64_t>(G * kSegCap + 1);
    b->seg_pre = a.take<uint64_t>(G * kSegCap + 1);
    b->seg_spec_n = a.take<uint32_t>(G + 1);
    b->seg_pre_n = a.take<uint32_t>(G + 1);
    b->seg_merge_idx = a.take<uint32_t>(G + 1);
    b->seg_final_n = a.take<uint32_t>(G + 1);
    b->seg_exit_spec = a.take<uint64_t>(G + 1);
    b->seg_exit_true = a.take<uint64_t>(G + 1);
    b->seg_base = a.take<uint64_t>(G + 2);
    b->file_count = a.take<uint32_t>(F + 1);
    b->first_start = a.take<uint64_t>(F + 1);
    b->file_chunk_base = a.take<uint64_t>(F + 2);
    b->chunk_start = a.take<uint64_t>(S + 1);
    b->chunk_len = a.take<uint32_t>(S + 1);
    b->chunk_file = a.take<uint32_t>(S + 1);
    b->slot_groups = a.take<uint32_t>(S + 1);
    b->group_off = a.take<uint64_t>(S + 2);
    b->group_chunk = a.take<uint32_t>(p.n_groups_cap + 1);
    b->group_cv = a.take<uint32_t>((p.n_groups_cap + 1) * 8);
    b->chunk_hash = a.take<uint8_t>((S + 1) * 32);
    b->node_hash[0] = b->chunk_hash;
    b->node_hash[1] = a.take<uint8_t>((S + 1) * 32);
    b->node_size[0] = a.take<uint64_t>(S + 1);
    b->node_size[1] = a.take<uint64_t>(S + 1);
    b->level_count[0] = a.take<uint32_t>(F + 1);
    b->level_count[1] = a.take<uint32_t>(F + 1);
    b->group_start = a.take<uint32_t>(S + 2);
    b->mk_bits = a.take<uint64_t>(S + 1);
    b->mk_out = a.take<uint64_t>(S + 1);
    b->mk_cnt = a.take<uint64_t>(S + 1);
    b->mk_entry = a.take<uint32_t>(S + 1);
    b->mk_base = a.take<uint32_t>(S + 1);
    b->file_root = a.take<uint8_t>((F + 1) * 32);
    b->file_hash = a.take<uint8_t>((F + 1) * 32);
    b->out_end = a.take<uint64_t>(S + 1);
    b->out_hash = a.take<uint8_t>((S + 1) * 32);
    b->n_xorb_slots = p.n_xorb_slots;
    if (p.n_xorb_slots) {
        const uint64_t X = p.n_xorb_slots;
        b->xorb_slot_off = a.take<uint64_t>(F + 1);
        b->xorb_count = a.take<uint32_t>(F + 1);
        b->xorb_node_off = a.take<uint64_t>(X + 1);
        b->xorb_n = a.take<uint32_t>(X + 1);
        b->xorb_bytes = a.take<uint64_t>(X + 1);
        b->xorb_root = a.take<uint8_t>((X + 1) * 32);
        b->xorb_scratch = a.take<uint8_t>((X + 1) * 32);
        b->xorb_level[0] = a.take<uint32_t>(X + 1);
        b->xorb_level[1] = a.take<uint32_t>(X + 1);
    }
    b->scan_partial = a.take<uint64_t>(scan_blocks(std::max<uint64_t>(std::max<uint64_t>(std::max<uint64_t>(T, S), F), G) + 2) + 1);
    b->scalars = a.take<uint64_t>(8);
    *need = a.used;
    return nullptr;
}

static XetError* ensure_workspace(Lane* l, const Plan& p, PipelineBuffers* b) {
    uint64_t need = 0;
    carve(l, p, b, true, &need);
    if (need > l->ws.cap) {
        if (l->ws.base) {
            cudaStreamSynchronize(l->stream);
            cudaFree(l->ws.base);
        }
        l->ws.base = nullptr;
        l->ws.cap = 0;
        uint64_t cap = need + need / 8;
        CU(cudaMalloc((void**)&l->ws.base, cap));
        l->ws.cap = cap;
    }
    return carve(l, p, b, false, &need);
}

// ---------------------------------------------------------------------------------- ingestion
// Copy host bytes to the device through the two pinned staging buffers: the CPU fills one buffer
// while the DMA engine drains the other.  Pinned callers (cudaHostAlloc / cudaHostRegister) are
// copied directly in staging-sized pieces.
// File ingestion.  A single read() loop moves ~5 GB/s out of the page cache — no faster than the CPU
// reference hashes — so a batch is cut into pieces that a small pool of reader threads pread() into a
// ring of pinned slots; the thread that filled a slot enqueues its host->device copy itself (the
// copies land in disjoint device ranges, so their order on the stream does not matter) and the slot
// is reused once that copy's event has fired.
struct ReadPiece {
    int fd;               // >= 0: a piece of file paths[file], opened by the worker for this piece only (a batch may
                          //       hold more files than the process may keep open); -1: a piece of host memory
    const uint8_t* mem;   // the host memory to memcpy when fd < 0
    uint64_t file_off, len, dev_off;
    size_t file;          // index into paths
    uint64_t file_size;   // size the batch was planned with; the file must still have it when the piece is read
};

struct ReaderPool {
    XetGpuContext* c = nullptr;
    cudaStream_t stream = nullptr;   // stream the host->device copies are enqueued on
    uint8_t* d_base = nullptr;       // device destination of dev_off 0
    const char* const* paths = nullptr;
    std::vector<ReadPiece> pieces;
    std::vector<uint8_t*> slots;
    std::vector<cudaEvent_t> slot_ev;
    std::vector<long> slot_issued;  // index of the last piece whose copy was enqueued from this slot (-1: none)
    std::mutex mu;
    std::condition_variable cv;
    std::atomic<size_t> next{0};
    std::atomic<uint64_t> bytes_done{0};
    std::atomic<size_t> last_file{0};   // file of the piece read most recently (progress: current_file)
    std::atomic<bool> stop{false};
    XetError* err = nullptr;

    void fail(XetError* e) {
        std::lock_guard<std::mutex> lk(mu);
        if (!err) err = e;
        else xetgpu_free_error(e);
        stop = true;
        cv.notify_all();
    }

    void worker() {
        cudaSetDevice(c->device);
        const size_t S = slots.size();
        while (!stop) {
            const size_t i = next.fetch_add(1);
            if (i >= pieces.size()) return;
            const size_t s = i % S;
            {   // the slot is free once the piece that used it last (i - S) has enqueued its copy ...
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || slot_issued[s] == (long)i - (long)S; });
                if (stop) return;
            }
            cudaEventSynchronize(slot_ev[s]);  // ... and the last copy out of this slot (any stream, any call) has drained
            const ReadPiece& p = pieces[i];
            uint64_t got = 0;
            int fd = -1;
            if (p.fd < 0) {
                memcpy(slots[s], p.mem, p.len);
                got = p.len;
            } else {
                // open, then fstat the descriptor: the size is checked on the very object that is read, so a file
                // replaced or resized since the batch was planned is an I/O error, never a hash of stale extents
                fd = open(paths[p.file], O_RDONLY | O_CLOEXEC);
                struct stat st;
                if (fd < 0 || fstat(fd, &st) != 0) {
                    XetError* e = make_error(errno == ENOENT ? XETGPU_ERR_NOT_FOUND : XETGPU_ERR_IO,
                                             std::string("cannot open file: ") + strerror(errno), paths[p.file]);
                    if (fd >= 0) close(fd);
                    fail(e);
                    return;
                }
                if ((uint64_t)st.st_size != p.file_size) {
                    close(fd);
                    fail(make_error(XETGPU_ERR_IO, "file changed size while being hashed", paths[p.file]));
                    return;
                }
            }
            while (got < p.len) {
                ssize_t r = pread(fd, slots[s] + got, p.len - got, (off_t)(p.file_off + got));
                if (r < 0 && errno == EINTR) continue;
                if (r <= 0) {
                    XetError* e = make_error(XETGPU_ERR_IO, r < 0 ? std::string("read failed: ") + strerror(errno)
                                                                  : std::string("file shrank while being read"),
                                             paths ? paths[p.file] : "");
                    close(fd);
                    fail(e);
                    return;
                }
                got += (uint64_t)r;
            }
            if (fd >= 0) {
                close(fd);
                last_file.store(p.file);
            }
            cudaError_t ce = cudaMemcpyAsync(d_base + p.dev_off, slots[s], p.len, cudaMemcpyHostToDevice, stream);
            if (ce == cudaSuccess) ce = cudaEventRecord(slot_ev[s], stream);
            if (ce != cudaSuccess) {
                fail(cuda_error(ce, "cudaMemcpyAsync(H2D)"));
                return;
            }
            bytes_done += p.len;
            {
                std::lock_guard<std::mutex> lk(mu);
                slot_issued[s] = (long)i;
            }
            cv.notify_all();
        }
    }
};

static XetError* ensure_read_slots(XetGpuContext* c) {
    if (c->read_slots.size() == c->n_read_slots) return nullptr;
    for (size_t i = c->read_slots.size(); i < c->n_read_slots; i++) {
        uint8_t* p = nullptr;
        CU(cudaMallocHost((void**)&p, c->read_slot_bytes));
        cudaEvent_t ev;
        CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        c->read_slots.push_back(p);
        c->read_slot_ev.push_back(ev);
    }
    return nullptr;
}

// Runs the pool to completion (all copies enqueued); `poll` is called every few milliseconds on this thread.
template <typename Poll>
static XetError* run_reader_pool(XetGpuContext* c, ReaderPool& pool, Poll poll) {
    if (pool.pieces.empty()) return nullptr;
    pool.c = c;
    pool.slots = c->read_slots;
    pool.slot_ev = c->read_slot_ev;
    const size_t S = c->n_read_slots;
    pool.slot_issued.assign(S, -1);
    for (size_t s = 0; s < S; s++) pool.slot_issued[s] = (long)s - (long)S;
    const size_t nthreads = std::min<size_t>(std::min<size_t>(c->read_threads, pool.pieces.size()), S);
    std::vector<std::thread> threads;
    for (size_t t = 0; t < nthreads; t++) threads.emplace_back([&pool] { pool.worker(); });
    while (pool.next.load() < pool.pieces.size() && !pool.stop) {
        std::this_thread::sleep_for(std::chrono::milliseconds(2));
        poll(pool);
    }
    for (auto& t : threads) t.join();
    return pool.err;
}

static XetError* ensure_staging(XetGpuContext* c) {
    for (int i = 0; i < 2; i++)
        if (!c->h_stage[i]) CU(cudaMallocHost((void**)&c->h_stage[i], c->staging_bytes));
    return nullptr;
}

static bool host_ptr_is_pinned(const void* p) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
}

static XetError* ingest_host(XetGpuContext* c, cudaStream_t stream, const uint8_t* src, uint64_t bytes, uint8_t* d_dst) {
    if (bytes == 0) return nullptr;
    if (host_ptr_is_pinned(src)) {
        for (uint64_t o = 0; o < bytes; o += c->staging_bytes) {
            uint64_t n = std::min<uint64_t>(c->staging_bytes, bytes - o);
            CU(cudaMemcpyAsync(d_dst + o, src + o, n, cudaMemcpyHostToDevice, stream));
        }
        return nullptr;
    }
    if (bytes > c->read_slot_bytes) {  // pageable memory: several threads copy into the pinned slot ring
        if (XetError* e = ensure_read_slots(c)) return e;
        ReaderPool pool;
        pool.stream = stream;
        pool.d_base = d_dst;
        for (uint64_t o = 0; o < bytes; o += c->read_slot_bytes)
            pool.pieces.push_back(ReadPiece{-1, src + o, 0, std::min<uint64_t>(c->read_slot_bytes, bytes - o), o, 0, 0});
        return run_reader_pool(c, pool, [](ReaderPool&) {});
    }
    if (XetError* e = ensure_staging(c)) return e;
    int i = 0;
    for (uint64_t o = 0; o < bytes; o += c->staging_bytes, i ^= 1) {
        uint64_t n = std::min<uint64_t>(c->staging_bytes, bytes - o);
        CU(cudaEventSynchronize(c->stage_free[i]));  // the copy that last used this buffer has drained
        memcpy(c->h_stage[i], src + o, n);
        CU(cudaMemcpyAsync(d_dst + o, c->h_stage[i], n, cudaMemcpyHostToDevice, stream));
        CU(cudaEventRecord(c->stage_free[i], stream));
    }
    return nullptr;
}

// ----------------------------------------------------------------------------------- pipeline
extern "C" void xetgpu_free_result(XetGpuResult* r) {
    if (!r) return;
    free(r->files);
    free(r->chunk_ends);
    free(r->chunk_hashes);
    free(r->xorbs);
    free(r);
}

extern "C" void xetgpu_free_u64(uint64_t* a) { free(a); }

struct RunOptions {
    const RangeWalk* range = nullptr;  // byte-range mode (one file)
    uint64_t end_bias = 0;             // added to every returned chunk end (range mode: file position of data[0])
    bool want_chunks = false;
    bool want_xorbs = false;           // also form xorbs (greedy <= 64 MiB / <= 8192 chunks) and hash each one
    bool resolve_only = false;         // stop after K2 (range exit-cut query)
    uint64_t* exit_cut_abs = nullptr;  // resolve_only / range: last cut (absolute in data), or walk_start if none
    bool skip_merkle = false;
    // A batch that mixes whole files and byte-range items (xetgpu_hash_buffers_ranges): per item, the position in its
    // own data of the entry cut, whether the data ends at the end of the file, and the file position of data[0].
    struct Item {
        uint64_t walk_rel;   // entry cut, relative to the item's first data byte
        bool is_eof;
        uint64_t bias;       // file position of the item's first data byte (added to the chunk ends returned)
        bool partial;        // a byte-range item (its chunk list is always returned) rather than a whole file
    };
    const std::vector<Item>* items = nullptr;
    std::vector<uint64_t>* item_exit_cuts = nullptr;   // per item: last cut (file position), or the entry cut if none
};

// One group of files in flight on a lane: everything launch_group enqueued and finish_group reads.
struct GroupState {
    Plan plan;
    PipelineBuffers b{};
    uint64_t total = 0;
    size_t n_files = 0;
    int launches = 0;
    uint64_t meta_bytes = 0;
    bool had_h2d = false;
    // pinned readback area
    uint64_t* h_scalars = nullptr;
    uint64_t* h_base = nullptr;
    uint32_t* h_count = nullptr;
    uint8_t* h_hash = nullptr;
    uint8_t* h_root = nullptr;
    // xorb readback (want_xorbs)
    uint32_t* h_xcount = nullptr;
    uint64_t* h_xoff = nullptr;
    uint64_t* h_xbytes = nullptr;
    uint32_t* h_xn = nullptr;
    uint8_t* h_xroot = nullptr;
};

// Enqueue the whole pipeline for one group on `lane` (no host synchronisation).
// d_data: device bytes; offsets: host, relative to d_data.
// Enqueue stages of the pipeline for one group (no host synchronisation).  `stages` is a bit mask:
//   1  A: plan, workspace, metadata upload, K1 (gear candidates)
//   2  H: K2 (boundary resolution) + the K3 work lists
//   4  K: the K3 leaf kernel
//   8  T: chunk tree, gather, K4 (+ xorbs) and the result readback
// All on `st`; a caller that spreads the stages over streams orders them with events.  Stage A must come first.
// d_data: device bytes; offsets: host, relative to d_data.
static XetError* launch_group_stages(XetGpuContext* c, Lane* lane, const uint8_t* d_data, const uint64_t* offsets,
                                     size_t n_files, const RunOptions& opt, GroupState* g, int stages, cudaStream_t st,
                                     cudaEvent_t gear_after = nullptr, cudaEvent_t gear_done = nullptr) {
    const bool timing = (c->flags & XETGPU_CTX_STAGE_TIMING) != 0;
    const uint64_t F = n_files;
    const bool xorbs = opt.want_xorbs && !opt.resolve_only && !opt.skip_merkle;
    PipelineBuffers& b = g->b;
    if (stages & 1) {
        const uint64_t total = n_files ? offsets[n_files] : 0;
        g->total = total;
        g->n_files = n_files;
        g->launches = 0;
        g->plan = make_plan(offsets, n_files, total, c->cand_cap_hint, xorbs);
        g->plan.has_items = opt.items != nullptr;
        const Plan& plan = g->plan;
        b = PipelineBuffers{};
        if (XetError* e = ensure_workspace(lane, plan, &b)) return e;
        b.data = d_data;
        if (timing)
            for (int i = 0; i < 4; i++) b.ev_kernel[i] = lane->ev_kernel[i];
        // metadata upload through pinned memory
        const uint64_t X = plan.n_xorb_slots;
        g->meta_bytes = (F + 1) * 32 + 64 + (F + 2) * 8 + (F + 1) * 4 + (F + 1) * 64 + 256 +
                        (xorbs ? (F + 1) * 8 + (F + 2) * 4 + (X + 1) * (8 + 8 + 4 + 32) + 64 : 0);
        if (XetError* e = ensure_meta(lane, g->meta_bytes)) return e;
        uint64_t* h_off = reinterpret_cast<uint64_t*>(lane->h_meta);
        uint64_t* h_slot = h_off + (F + 1);
        uint64_t* h_seg = h_slot + (F + 1);
        uint64_t* h_walk = h_seg + (F + 1);
        if (opt.items)
            for (size_t f = 0; f < F; f++)
                h_walk[f] = (offsets[f] + (*opt.items)[f].walk_rel) | ((*opt.items)[f].is_eof ? 0ull : 1ull << 63);
        memcpy(h_off, offsets, (F + 1) * 8);
        memcpy(h_slot, plan.slot_off.data(), (F + 1) * 8);
        memcpy(h_seg, plan.seg_off.data(), (F + 1) * 8);
        if (timing) cudaEventRecord(lane->ev[0], st);
        CU(cudaMemcpyAsync((void*)b.file_off, h_off, (F + 1) * 8, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync((void*)b.slot_off, h_slot, (F + 1) * 8, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync((void*)b.seg_off, h_seg, (F + 1) * 8, cudaMemcpyHostToDevice, st));
        if (opt.items && F) CU(cudaMemcpyAsync((void*)b.file_walk, h_walk, F * 8, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(b.scalars, 0, 64, st));
        b.sm_count = c->sm_count;
        cudaError_t gear_err = cudaSuccess;
        // Two gear passes side by side only fight for shared-memory bandwidth: a group's gear pass starts when the
        // previous group's is done.
        if (gear_after) CU(cudaStreamWaitEvent(st, gear_after, 0));
        g->launches += launch_gear(b, st, &gear_err);
        if (gear_err != cudaSuccess) return cuda_error(gear_err, "tensor map setup for the gear kernel");
        if (gear_done) CU(cudaEventRecord(gear_done, st));
        if (timing) cudaEventRecord(lane->ev[1], st);
    }
    const Plan& plan = g->plan;
    const uint64_t X = plan.n_xorb_slots;
    if (stages & 2) {
        g->launches += launch_resolve(b, opt.range, st);
        if (timing) cudaEventRecord(lane->ev[2], st);
        if (!opt.resolve_only) g->launches += launch_chunk_hash_stages(b, st, 1);
    }
    if ((stages & 4) && !opt.resolve_only) g->launches += launch_chunk_hash_stages(b, st, 2);
    if (!(stages & 8)) {
        CU(cudaGetLastError());
        return nullptr;
    }
    if (!opt.resolve_only) {
        g->launches += launch_chunk_hash_stages(b, st, 4);
        g->launches += launch_gather(b, st);
        if (timing) cudaEventRecord(lane->ev[3], st);
        if (!opt.skip_merkle) g->launches += launch_merkle(b, st);
        if (xorbs && X) {
            uint64_t* h_xs = reinterpret_cast<uint64_t*>(lane->h_meta + g->meta_bytes - (F + 1) * 8 - 8);  // upload area (tail of the block)
            memcpy(h_xs, plan.xorb_off.data(), (F + 1) * 8);
            CU(cudaMemcpyAsync((void*)b.xorb_slot_off, h_xs, (F + 1) * 8, cudaMemcpyHostToDevice, st));
            g->launches += launch_xorbs(b, st);
        }
        if (timing) cudaEventRecord(lane->ev[4], st);
    } else {
        g->launches += launch_scan_u32(b.file_count, b.file_chunk_base, b.n_files, b.scan_partial,
                                       b.scalars + kScalarChunks, st);
        if (timing) {
            cudaEventRecord(lane->ev[3], st);
            cudaEventRecord(lane->ev[4], st);
        }
    }
    CU(cudaGetLastError());
    // readback: scalars, per-file counts, bases, roots, hashes
    uint8_t* h_rb = lane->h_meta + (F + 1) * 32;
    g->h_scalars = reinterpret_cast<uint64_t*>(h_rb);
    g->h_base = g->h_scalars + 8;
    g->h_count = reinterpret_cast<uint32_t*>(g->h_base + (F + 2));
    g->h_hash = reinterpret_cast<uint8_t*>(g->h_count + (F + 1 + ((F + 1) & 1)));
    g->h_root = g->h_hash + (F + 1) * 32;
    CU(cudaMemcpyAsync(g->h_scalars, b.scalars, 64, cudaMemcpyDeviceToHost, st));
    if (F) {
        CU(cudaMemcpyAsync(g->h_base, b.file_chunk_base, (F + 1) * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(g->h_count, b.file_count, F * 4, cudaMemcpyDeviceToHost, st));
        if (!opt.resolve_only && !opt.skip_merkle) {
            CU(cudaMemcpyAsync(g->h_hash, b.file_hash, F * 32, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(g->h_root, b.file_root, F * 32, cudaMemcpyDeviceToHost, st));
        }
    }
    if (xorbs && X) {
        uint8_t* p = g->h_root + (F + 1) * 32;
        p += (8 - ((uintptr_t)p & 7)) & 7;
        g->h_xoff = reinterpret_cast<uint64_t*>(p);
        g->h_xbytes = g->h_xoff + (X + 1);
        g->h_xn = reinterpret_cast<uint32_t*>(g->h_xbytes + (X + 1));
        g->h_xcount = g->h_xn + (X + 1);
        g->h_xroot = reinterpret_cast<uint8_t*>(g->h_xcount + (F + 2));
        CU(cudaMemcpyAsync(g->h_xoff, b.xorb_node_off, X * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(g->h_xbytes, b.xorb_bytes, X * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(g->h_xn, b.xorb_n, X * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(g->h_xcount, b.xorb_count, F * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(g->h_xroot, b.xorb_root, X * 32, cudaMemcpyDeviceToHost, st));
    }
    if (timing) cudaEventRecord(lane->ev[5], st);
    return nullptr;
}

// The whole pipeline of one group on its lane's stream.
static XetError* launch_group(XetGpuContext* c, Lane* lane, const uint8_t* d_data, const uint64_t* offsets,
                              size_t n_files, const RunOptions& opt, GroupState* g,
                              cudaEvent_t gear_after = nullptr, cudaEvent_t gear_done = nullptr) {
    return launch_group_stages(c, lane, d_data, offsets, n_files, opt, g, 15, lane->stream, gear_after, gear_done);
}

// Result accumulator shared by all groups of one call.
struct ResultBuilder {
    std::vector<XetGpuFileResult> files;
    std::vector<uint64_t> ends;
    std::vector<uint8_t> hashes;
    std::vector<XetGpuXorb> xorbs;
    uint64_t n_cand = 0;
    int launches = 0;
    float gear_ms = 0, resolve_ms = 0, blake3_ms = 0, merkle_ms = 0, d2h_ms = 0, gear_kernel_ms = 0, leaf_kernel_ms = 0,
          h2d_ms = 0;
};

// Wait for a group and collect its results.  *overflow = true when the candidate list was too small
// (dense data): nothing is collected and the caller re-runs the group with the grown list.
static XetError* finish_group(XetGpuContext* c, Lane* lane, GroupState* g, const uint64_t* offsets,
                              const RunOptions& opt, ResultBuilder* rb, bool* overflow) {
    const bool timing = (c->flags & XETGPU_CTX_STAGE_TIMING) != 0;
    cudaStream_t st = lane->stream;
    CU(cudaStreamSynchronize(st));
    *overflow = false;
    const Plan& plan = g->plan;
    PipelineBuffers& b = g->b;
    const size_t F = g->n_files;
    const uint64_t n_cand = g->h_scalars[kScalarCand];
    if (n_cand > plan.cand_cap) {
        c->cand_cap_hint = std::max<uint64_t>(c->cand_cap_hint, n_cand + n_cand / 8 + 1024);
        *overflow = true;
        return nullptr;
    }
    const uint64_t n_chunks = F ? g->h_base[F] : 0;
    if (opt.exit_cut_abs) {
        uint64_t last = opt.range ? opt.range->walk_start : 0;
        if (n_chunks)
            CU(cudaMemcpy(&last, b.slot_end + (plan.slot_off[F - 1] + g->h_count[F - 1] - 1), 8, cudaMemcpyDeviceToHost));
        *opt.exit_cut_abs = last;
    }
    const uint64_t chunk0 = rb->ends.size();
    for (size_t f = 0; f < F; f++) {
        XetGpuFileResult fr{};
        if (!opt.resolve_only && !opt.skip_merkle) {
            memcpy(fr.file_hash, g->h_hash + f * 32, 32);
            memcpy(fr.merkle_root, g->h_root + f * 32, 32);
        }
        fr.size = offsets[f + 1] - offsets[f];
        fr.n_chunks = g->h_count[f];
        fr.first_chunk = (opt.want_chunks ? chunk0 : 0) + g->h_base[f];   // (rewritten below for selective readback)
        rb->files.push_back(fr);
    }
    if (g->h_xcount) {
        const size_t file0 = rb->files.size() - F;
        for (size_t f = 0; f < F; f++)
            for (uint32_t k = 0; k < g->h_xcount[f]; k++) {
                const uint64_t sx = plan.xorb_off[f] + k;
                XetGpuXorb xo{};
                memcpy(xo.hash, g->h_xroot + sx * 32, 32);
                xo.file = file0 + f;
                xo.first_chunk = g->h_xoff[sx] - g->h_base[f];  // file-relative chunk index
                xo.n_chunks = g->h_xn[sx];
                xo.bytes = g->h_xbytes[sx];
                rb->xorbs.push_back(xo);
            }
    }
    if (opt.want_chunks && n_chunks) {
        rb->ends.resize(chunk0 + n_chunks);
        rb->hashes.resize((chunk0 + n_chunks) * 32);
        if (n_chunks * 40 > lane->h_chunks_cap) {
            if (lane->h_chunks) cudaFreeHost(lane->h_chunks);
            lane->h_chunks = nullptr;
            lane->h_chunks_cap = 0;
            const uint64_t cap = n_chunks * 40 + n_chunks * 5 + 4096;
            CU(cudaMallocHost((void**)&lane->h_chunks, cap));
            lane->h_chunks_cap = cap;
        }
        uint8_t* pe = lane->h_chunks;
        uint8_t* ph = lane->h_chunks + n_chunks * 8;
        if (opt.resolve_only) {  // ends only: gather was skipped; single file in this mode
            CU(cudaMemcpyAsync(pe, b.slot_end, n_chunks * 8, cudaMemcpyDeviceToHost, st));
        } else {
            CU(cudaMemcpyAsync(pe, b.out_end, n_chunks * 8, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(ph, b.out_hash, n_chunks * 32, cudaMemcpyDeviceToHost, st));
        }
        CU(cudaStreamSynchronize(st));
        memcpy(rb->ends.data() + chunk0, pe, n_chunks * 8);
        if (!opt.resolve_only) memcpy(rb->hashes.data() + chunk0 * 32, ph, n_chunks * 32);
        if (opt.end_bias)
            for (uint64_t i = 0; i < n_chunks; i++) rb->ends[chunk0 + i] += opt.end_bias;
        if (opt.items)
            for (size_t f = 0; f < F; f++) {
                const uint64_t bias = (*opt.items)[f].bias;
                const uint64_t c0 = chunk0 + g->h_base[f], c1 = c0 + g->h_count[f];
                for (uint64_t i = c0; i < c1; i++) rb->ends[i] += bias;
            }
    }
    if (opt.items && !opt.want_chunks && !opt.resolve_only) {
        // selective readback: chunk lists of the byte-range items only (a rank's whole shards would be ~50 MB of
        // records nobody reads); one pinned bounce of the few items' slices
        uint64_t need = 0;
        for (size_t f = 0; f < F; f++)
            if ((*opt.items)[f].partial) need += g->h_count[f];
        if (need * 40 > lane->h_chunks_cap) {
            if (lane->h_chunks) cudaFreeHost(lane->h_chunks);
            lane->h_chunks = nullptr;
            lane->h_chunks_cap = 0;
            const uint64_t cap = need * 45 + 4096;
            CU(cudaMallocHost((void**)&lane->h_chunks, cap));
            lane->h_chunks_cap = cap;
        }
        uint64_t w = 0;
        for (size_t f = 0; f < F; f++) {
            if (!(*opt.items)[f].partial || !g->h_count[f]) continue;
            CU(cudaMemcpyAsync(lane->h_chunks + w * 8, b.out_end + g->h_base[f], g->h_count[f] * 8ull, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(lane->h_chunks + need * 8 + w * 32, b.out_hash + g->h_base[f] * 32, g->h_count[f] * 32ull,
                               cudaMemcpyDeviceToHost, st));
            w += g->h_count[f];
        }
        if (need) CU(cudaStreamSynchronize(st));
        rb->ends.resize(chunk0 + need);
        rb->hashes.resize((chunk0 + need) * 32);
        if (need) {
            memcpy(rb->ends.data() + chunk0, lane->h_chunks, need * 8);
            memcpy(rb->hashes.data() + chunk0 * 32, lane->h_chunks + need * 8, need * 32);
        }
        w = 0;
        const size_t file0 = rb->files.size() - F;
        for (size_t f = 0; f < F; f++) {
            XetGpuFileResult& fr = rb->files[file0 + f];
            fr.first_chunk = chunk0 + w;
            if (!(*opt.items)[f].partial) continue;
            const uint64_t bias = (*opt.items)[f].bias;
            for (uint64_t i = 0; i < g->h_count[f]; i++) rb->ends[chunk0 + w + i] += bias;
            w += g->h_count[f];
        }
    }
    if (opt.items && opt.item_exit_cuts) {
        opt.item_exit_cuts->assign(F, 0);
        const size_t file0 = rb->files.size() - F;
        for (size_t f = 0; f < F; f++) {
            const RunOptions::Item& it = (*opt.items)[f];
            const bool have = g->h_count[f] && (opt.want_chunks || it.partial);
            const uint64_t last = rb->files[file0 + f].first_chunk + g->h_count[f];
            (*opt.item_exit_cuts)[f] = have ? rb->ends[last - 1] : it.partial ? it.bias + it.walk_rel : it.bias + (offsets[f + 1] - offsets[f]);
        }
    }
    if (timing) {
        cudaEventRecord(lane->ev[6], st);
        cudaEventSynchronize(lane->ev[6]);
        float t = 0;
        cudaEventElapsedTime(&t, lane->ev[0], lane->ev[1]);
        rb->gear_ms += t;
        cudaEventElapsedTime(&t, lane->ev[1], lane->ev[2]);
        rb->resolve_ms += t;
        cudaEventElapsedTime(&t, lane->ev[2], lane->ev[3]);
        rb->blake3_ms += t;
        cudaEventElapsedTime(&t, lane->ev[3], lane->ev[4]);
        rb->merkle_ms += t;
        cudaEventElapsedTime(&t, lane->ev[4], lane->ev[6]);
        rb->d2h_ms += t;
        if (plan.n_tiles) {
            cudaEventElapsedTime(&t, lane->ev_kernel[0], lane->ev_kernel[1]);
            rb->gear_kernel_ms += t;
        }
        if (!opt.resolve_only && plan.n_slots && plan.n_groups_cap) {
            cudaEventElapsedTime(&t, lane->ev_kernel[2], lane->ev_kernel[3]);
            rb->leaf_kernel_ms += t;
        }
        if (g->had_h2d) {
            cudaEventElapsedTime(&t, lane->ev[7], lane->ev[0]);
            rb->h2d_ms += t;
        }
    }
    rb->n_cand += n_cand;
    rb->launches += g->launches;
    return nullptr;
}

static XetError* build_result(const ResultBuilder& rb, bool want_chunks, XetGpuResult** out_result) {
    if (!out_result) return nullptr;
    XetGpuResult* r = (XetGpuResult*)calloc(1, sizeof(XetGpuResult));
    if (!r) return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
    r->n_files = rb.files.size();
    r->files = (XetGpuFileResult*)calloc(rb.files.size() ? rb.files.size() : 1, sizeof(XetGpuFileResult));
    if (!rb.files.empty()) memcpy(r->files, rb.files.data(), rb.files.size() * sizeof(XetGpuFileResult));
    if (want_chunks && !rb.ends.empty()) {
        r->n_chunks = rb.ends.size();
        r->chunk_ends = (uint64_t*)malloc(rb.ends.size() * 8);
        r->chunk_hashes = (uint8_t*)malloc(rb.hashes.size());
        if (!r->chunk_ends || !r->chunk_hashes) {
            xetgpu_free_result(r);
            return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
        }
        memcpy(r->chunk_ends, rb.ends.data(), rb.ends.size() * 8);
        memcpy(r->chunk_hashes, rb.hashes.data(), rb.hashes.size());
    }
    if (!rb.xorbs.empty()) {
        r->n_xorbs = rb.xorbs.size();
        r->xorbs = (XetGpuXorb*)malloc(rb.xorbs.size() * sizeof(XetGpuXorb));
        if (!r->xorbs) {
            xetgpu_free_result(r);
            return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
        }
        memcpy(r->xorbs, rb.xorbs.data(), rb.xorbs.size() * sizeof(XetGpuXorb));
    }
    *out_result = r;
    return nullptr;
}

static void publish_timings(XetGpuContext* c, const ResultBuilder& rb, uint64_t bytes, float wall_ms) {
    std::lock_guard<std::mutex> lk(c->cb_mu);
    XetGpuTimings& t = c->last;
    t.h2d_ms = rb.h2d_ms;
    t.gear_ms = rb.gear_ms;
    t.resolve_ms = rb.resolve_ms;
    t.blake3_ms = rb.blake3_ms;
    t.merkle_ms = rb.merkle_ms;
    t.d2h_ms = rb.d2h_ms;
    t.gear_kernel_ms = rb.gear_kernel_ms;
    t.leaf_kernel_ms = rb.leaf_kernel_ms;
    t.total_ms = wall_ms;
    t.bytes = bytes;
    t.n_chunks = 0;
    for (const auto& f : rb.files) t.n_chunks += f.n_chunks;
    t.n_candidates = rb.n_cand;
    t.kernel_launches = (uint32_t)rb.launches;
}

// One group, synchronously, on lane 0 (also the retry path when the candidate list overflowed).
static XetError* run_group_sync(XetGpuContext* c, Lane* lane, const uint8_t* d_data, const uint64_t* offsets,
                                size_t n_files, const RunOptions& opt, ResultBuilder* rb) {
    for (int attempt = 0; attempt < 3; attempt++) {
        GroupState g;
        if (XetError* e = launch_group(c, lane, d_data, offsets, n_files, opt, &g)) return e;
        bool overflow = false;
        if (XetError* e = finish_group(c, lane, &g, offsets, opt, rb, &overflow)) return e;
        if (!overflow) return nullptr;
    }
    return make_error(XETGPU_ERR_UNKNOWN, "candidate list kept overflowing");
}

static XetError* run_pipeline(XetGpuContext* c, const uint8_t* d_data, const uint64_t* offsets, size_t n_files,
                              const RunOptions& opt, XetGpuResult** out_result) {
    auto t0 = std::chrono::steady_clock::now();
    ResultBuilder rb;
    if (XetError* e = run_group_sync(c, &c->lanes[0], d_data, offsets, n_files, opt, &rb)) return e;
    float wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    publish_timings(c, rb, n_files ? offsets[n_files] : 0, wall);
    return build_result(rb, opt.want_chunks || opt.items != nullptr, out_result);
}

// ---- file groups on two lanes -------------------------------------------------------------------
// A batch of several files can be cut into groups of whole files (>= 1/8 of the batch and >= 256 MiB each).
// Group g runs on lane g % 2, so its host->device copy and its many small kernels overlap the previous
// group's bulk kernels.
struct HostSource {
    const uint8_t* base = nullptr;  // host bytes of the batch; nullptr = bytes are already on the device
};

static XetError* run_batch(XetGpuContext* c, const uint8_t* d_data, const HostSource& host, const uint64_t* offsets,
                           size_t n_files, const RunOptions& opt, XetGpuResult** out_result) {
    const uint64_t total = n_files ? offsets[n_files] : 0;
    const bool timing = (c->flags & XETGPU_CTX_STAGE_TIMING) != 0;
    auto t0 = std::chrono::steady_clock::now();
    // group boundaries (file indices)
    std::vector<size_t> cut{0};
    const uint64_t target = std::max<uint64_t>(total / 8, c->min_group_bytes);
    // Measured on B200: two lanes pay off when there is a host->device copy to hide (+4 % end to end) or many
    // files (+8 % on 10,000 files: the short latency-bound kernels of one group hide behind the other group's
    // bulk kernels); a few large device-resident files run fastest as one group (the gear and BLAKE3 kernels
    // each fill the SMs' shared memory, so they cannot be co-resident anyway).
    bool two_lanes = !(c->flags & XETGPU_CTX_NO_OVERLAP) && !opt.range && (host.base != nullptr || n_files >= 16);
    if (c->force_lanes >= 0 && !opt.range) two_lanes = c->force_lanes == 1;
    if (opt.items) two_lanes = false;   // per-item descriptors are indexed by position in the batch: one group
    // The staged schedule below (bulk kernels strictly one after the other on one stream, every group's short
    // latency-bound kernels on the second stream behind the next bulk kernel) is OFF unless XETGPU_STAGED=1: measured
    // on B200 with config 2 it runs 14.6 ms against 13.7 ms for one group — the small kernels are not free to hide,
    // they stretch the gear and leaf kernels they run beside by about what they cost alone (4.21 -> 4.46 ms and
    // 8.31 -> 8.96 ms) and the per-group dependencies add gaps (profiles/r2_summary.md).
    bool staged = false;
    if (c->force_staged == 1 && !(c->flags & XETGPU_CTX_NO_OVERLAP) && !opt.range && !opt.items && !opt.resolve_only &&
        !host.base && n_files >= 2)
        staged = true;
    if (staged) two_lanes = true;   // same grouping rule
    if (two_lanes) {
        uint64_t acc = 0;
        for (size_t f = 0; f < n_files; f++) {
            acc += offsets[f + 1] - offsets[f];
            if (acc >= target && f + 1 < n_files && total - offsets[f + 1] >= c->min_group_bytes / 2) {
                cut.push_back(f + 1);
                acc = 0;
            }
        }
    }
    cut.push_back(n_files);
    const size_t G = cut.size() - 1;
    ResultBuilder rb;
    auto ingest = [&](Lane* lane, uint64_t lo, uint64_t hi, GroupState* g) -> XetError* {
        if (!host.base || hi == lo) return nullptr;
        if (timing) cudaEventRecord(lane->ev[7], lane->stream);
        g->had_h2d = true;
        return ingest_host(c, lane->stream, host.base + lo, hi - lo, const_cast<uint8_t*>(d_data) + lo);
    };
    if (staged && G >= 2 && G <= (size_t)kMaxLanes) {
        // Order on the bulk stream B: A0 A1 K0 A2 K1 ... K(G-1)  (A = K1 gear pass, K = K3 leaf kernel: the two kernels
        // that each fill the SMs never overlap one another); on the small stream S: H0 H1 T0 H2 T1 ... T(G-1)
        // (H = K2 + work lists, T = chunk tree + gather + K4 + readback), tied to B by events.
        cudaStream_t B = c->stream, S = c->second_stream;
        CU(cudaEventRecord(c->ev_fork, B));
        CU(cudaStreamWaitEvent(S, c->ev_fork, 0));
        std::vector<GroupState> state(G);
        std::vector<std::vector<uint64_t>> rel(G);
        std::vector<uint64_t> group_base(G, 0);
        auto stage = [&](size_t gi, int stages, cudaStream_t st) -> XetError* {
            const size_t f0 = cut[gi], f1 = cut[gi + 1];
            return launch_group_stages(c, &c->lanes[gi], d_data + group_base[gi], rel[gi].data(), f1 - f0, opt, &state[gi], stages,
                                       st, nullptr, stages == 1 ? c->ev_gear_done[gi] : nullptr);
        };
        auto leaf_and_tail = [&](size_t gi) -> XetError* {
            CU(cudaStreamWaitEvent(B, c->ev_head_done[gi], 0));
            if (XetError* e = stage(gi, 4, B)) return e;
            CU(cudaEventRecord(c->ev_leaf_done[gi], B));
            CU(cudaStreamWaitEvent(S, c->ev_leaf_done[gi], 0));
            return stage(gi, 8, S);
        };
        for (size_t gi = 0; gi < G; gi++) {
            const size_t f0 = cut[gi], f1 = cut[gi + 1];
            const uint64_t base_lo = offsets[f0] & ~15ull;   // 16-byte aligned group base (TMA), as below
            group_base[gi] = base_lo;
            rel[gi].assign(f1 - f0 + 1, 0);
            for (size_t f = f0; f <= f1; f++) rel[gi][f - f0] = offsets[f] - base_lo;
            if (XetError* e = stage(gi, 1, B)) return e;
            // the head of group gi goes on S before the tail of gi - 1: it is what the bulk stream waits for next
            CU(cudaStreamWaitEvent(S, c->ev_gear_done[gi], 0));
            if (XetError* e = stage(gi, 2, S)) return e;
            CU(cudaEventRecord(c->ev_head_done[gi], S));
            if (gi)
                if (XetError* e = leaf_and_tail(gi - 1)) return e;
        }
        if (XetError* e = leaf_and_tail(G - 1)) return e;
        // the call returns with everything complete on both streams, and B ordered after S for the caller
        CU(cudaEventRecord(c->ev_fork, S));
        CU(cudaStreamWaitEvent(B, c->ev_fork, 0));
        CU(cudaStreamSynchronize(S));
        CU(cudaStreamSynchronize(B));
        for (size_t gi = 0; gi < G; gi++) {
            bool overflow = false;
            if (XetError* e = finish_group(c, &c->lanes[gi], &state[gi], rel[gi].data(), opt, &rb, &overflow)) return e;
            if (overflow)
                if (XetError* e = run_group_sync(c, &c->lanes[gi], d_data + group_base[gi], rel[gi].data(), cut[gi + 1] - cut[gi],
                                                 opt, &rb))
                    return e;
        }
    } else if (G <= 1) {
        GroupState dummy;
        if (XetError* e = ingest(&c->lanes[0], 0, total, &dummy)) return e;
        // the copy is on lane 0's stream, in front of the pipeline: timing of it is taken there
        for (int attempt = 0; attempt < 3; attempt++) {
            GroupState g;
            g.had_h2d = dummy.had_h2d && attempt == 0;
            if (XetError* e = launch_group(c, &c->lanes[0], d_data, offsets, n_files, opt, &g)) return e;
            bool overflow = false;
            if (XetError* e = finish_group(c, &c->lanes[0], &g, offsets, opt, &rb, &overflow)) return e;
            if (!overflow) break;
            if (attempt == 2) return make_error(XETGPU_ERR_UNKNOWN, "candidate list kept overflowing");
        }
    } else {
        // the second stream must see what the caller enqueued on the context's stream before this call
        CU(cudaEventRecord(c->ev_fork, c->stream));
        CU(cudaStreamWaitEvent(c->second_stream, c->ev_fork, 0));
        // Group gi runs on lane gi % kMaxLanes (its own workspace and pinned metadata; streams alternate), so every
        // group is enqueued without waiting for an earlier one; a lane is drained only when a batch has more groups
        // than lanes.  Results are collected in group order at the end.
        std::vector<GroupState> state(kMaxLanes);
        std::vector<std::vector<uint64_t>> rel(kMaxLanes);
        uint64_t group_base[kMaxLanes] = {};
        long pending[kMaxLanes];
        for (auto& x : pending) x = -1;
        auto finish = [&](int li) -> XetError* {
            if (pending[li] < 0) return nullptr;
            const size_t gi = (size_t)pending[li];
            pending[li] = -1;
            bool overflow = false;
            if (XetError* e = finish_group(c, &c->lanes[li], &state[li], rel[li].data(), opt, &rb, &overflow)) return e;
            if (overflow)  // dense candidates: redo this group alone with the grown list
                return run_group_sync(c, &c->lanes[li], d_data + group_base[li], rel[li].data(),
                                      cut[gi + 1] - cut[gi], opt, &rb);
            return nullptr;
        };
        size_t next_finish = 0;   // groups finish in order: results are appended in group order
        for (size_t gi = 0; gi < G; gi++) {
            const int li = (int)(gi % kMaxLanes);
            while (gi - next_finish >= (size_t)kMaxLanes) {
                if (XetError* e = finish((int)(next_finish % kMaxLanes))) return e;
                next_finish++;
            }
            const size_t f0 = cut[gi], f1 = cut[gi + 1];
            // the group's device base must stay 16-byte aligned (TMA): start it at the aligned address at or
            // below its first file; the few bytes of the previous file in front are ignored by every kernel
            const uint64_t base_lo = offsets[f0] & ~15ull;
            group_base[li] = base_lo;
            rel[li].assign(f1 - f0 + 1, 0);
            for (size_t f = f0; f <= f1; f++) rel[li][f - f0] = offsets[f] - base_lo;
            state[li] = GroupState{};
            if (XetError* e = ingest(&c->lanes[li], offsets[f0], offsets[f1], &state[li])) return e;
            const int prev = (int)((gi + kMaxLanes - 1) % kMaxLanes);
            if (XetError* e = launch_group(c, &c->lanes[li], d_data + base_lo, rel[li].data(), f1 - f0, opt, &state[li],
                                           gi ? c->ev_gear_done[prev] : nullptr, c->ev_gear_done[li]))
                return e;
            pending[li] = (long)gi;
        }
        for (; next_finish < G; next_finish++)
            if (XetError* e = finish((int)(next_finish % kMaxLanes))) return e;
    }
    float wall = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    publish_timings(c, rb, total, wall);
    return build_result(rb, opt.want_chunks || opt.items != nullptr, out_result);
}

static XetError* resolve_memory_kind(const void* data, int32_t memory, bool* is_device) {
    if (memory == XETGPU_MEM_DEVICE) {
        *is_device = true;
    } else if (memory == XETGPU_MEM_HOST) {
        *is_device = false;
    } else if (memory == XETGPU_MEM_AUTO) {
        cudaPointerAttributes attr;
        cudaError_t e = cudaPointerGetAttributes(&attr, data);
        if (e != cudaSuccess) {
            cudaGetLastError();
            *is_device = false;
        } else {
            *is_device = attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
        }
    } else {
        return make_error(XETGPU_ERR_INVALID_CONFIG, "memory must be XETGPU_MEM_AUTO, _HOST or _DEVICE");
    }
    return nullptr;
}

// Device inputs: returns a 16-byte aligned device pointer to data[lo] (copying once, device to device,
// if the caller's pointer is not aligned).  Host inputs: returns the context's device input buffer,
// sized for hi - lo bytes; the bytes are copied later, group by group (run_batch).
static XetError* stage_input(XetGpuContext* c, const void* data, uint64_t lo, uint64_t hi, int32_t memory,
                             const uint8_t** d_out, bool* is_host) {
    bool is_device = false;
    if (XetError* e = resolve_memory_kind(data, memory, &is_device)) return e;
    const uint8_t* p = (const uint8_t*)data + lo;
    *is_host = !is_device;
    if (is_device) {
        if (((uintptr_t)p & 15u) == 0) {
            *d_out = p;
            return nullptr;
        }
        if (XetError* e = ensure_input(c, hi - lo)) return e;
        CU(cudaMemcpyAsync(c->d_input, p, hi - lo, cudaMemcpyDeviceToDevice, c->stream));
        *d_out = c->d_input;
        return nullptr;
    }
    if (XetError* e = ensure_input(c, hi - lo)) return e;
    *d_out = c->d_input;
    return nullptr;
}

// stage_input + the host->device copy on the context's stream (single-group callers).
static XetError* stage_input_now(XetGpuContext* c, const void* data, uint64_t lo, uint64_t hi, int32_t memory,
                                 const uint8_t** d_out) {
    bool is_host = false;
    if (XetError* e = stage_input(c, data, lo, hi, memory, d_out, &is_host)) return e;
    if (is_host) return ingest_host(c, c->stream, (const uint8_t*)data + lo, hi - lo, c->d_input);
    return nullptr;
}

extern "C" XetError* xetgpu_hash_buffers(XetGpuContext* c, const void* data, const uint64_t* offsets,
                                         size_t n_files, int32_t memory, uint32_t flags,
                                         XetGpuResult** out_result) {
    if (!c || !out_result || (n_files && !offsets))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_hash_buffers");
    *out_result = nullptr;
    for (size_t f = 0; f < n_files; f++)
        if (offsets[f + 1] < offsets[f])
            return make_error(XETGPU_ERR_INVALID_CONFIG, "offsets must be ascending");
    if (n_files && offsets[n_files] > offsets[0] && !data)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "data is NULL");
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    std::vector<uint64_t> rel(n_files + 1, 0);
    const uint64_t lo = n_files ? offsets[0] : 0;
    for (size_t f = 0; f <= n_files && n_files; f++) rel[f] = offsets[f] - lo;
    const uint8_t* d = nullptr;
    HostSource host;
    if (n_files && rel[n_files]) {
        bool is_host = false;
        if (XetError* e = stage_input(c, data, lo, offsets[n_files], memory, &d, &is_host)) return e;
        if (is_host) host.base = (const uint8_t*)data + lo;
    }
    RunOptions opt;
    opt.want_chunks = (flags & XETGPU_WANT_CHUNKS) != 0;
    opt.want_xorbs = (flags & XETGPU_WANT_XORBS) != 0;
    return run_batch(c, d, host, rel.data(), n_files, opt, out_result);
}

// Whole files and byte-range items in ONE batch (the config-4 rank: its whole shards plus its pieces of the two
// shards it shares).  Item i is data[offsets[i] .. offsets[i+1]) = bytes [range_begin - lead, range_end) of a file of
// file_len bytes, lead = min(range_begin, XETGPU_RANGE_LEAD); a whole file is {len, 0, len, 0}.
extern "C" XetError* xetgpu_hash_buffers_ranges(XetGpuContext* c, const void* data, const uint64_t* offsets, size_t n_items,
                                                int32_t memory, uint32_t flags, const XetGpuRangeSpec* specs,
                                                uint64_t* out_exit_cuts, XetGpuResult** out_result) {
    if (!c || !out_result || (n_items && (!offsets || !specs)))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_hash_buffers_ranges");
    *out_result = nullptr;
    std::vector<RunOptions::Item> items(n_items);
    for (size_t i = 0; i < n_items; i++) {
        const XetGpuRangeSpec& sp = specs[i];
        if (offsets[i + 1] < offsets[i]) return make_error(XETGPU_ERR_INVALID_CONFIG, "offsets must be ascending");
        if (sp.range_begin > sp.range_end || sp.range_end > sp.file_len || sp.entry_cut > sp.range_begin)
            return make_error(XETGPU_ERR_INVALID_CONFIG, "need entry_cut <= range_begin <= range_end <= file_len");
        const uint64_t lead = std::min<uint64_t>(sp.range_begin, XETGPU_RANGE_LEAD);
        const uint64_t data_pos = sp.range_begin - lead;
        if (sp.entry_cut < data_pos || sp.range_begin - sp.entry_cut > XETGPU_MAX_CHUNK)
            return make_error(XETGPU_ERR_INVALID_CONFIG, "entry_cut must lie within one maximum chunk before range_begin");
        if (offsets[i + 1] - offsets[i] != sp.range_end - data_pos)
            return make_error(XETGPU_ERR_INVALID_CONFIG, "an item's data must be bytes [range_begin - lead, range_end) of its file");
        items[i] = RunOptions::Item{sp.entry_cut - data_pos, sp.range_end == sp.file_len, data_pos,
                                    !(sp.range_begin == 0 && sp.range_end == sp.file_len)};
    }
    if (n_items && offsets[n_items] > offsets[0] && !data) return make_error(XETGPU_ERR_INVALID_CONFIG, "data is NULL");
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    std::vector<uint64_t> rel(n_items + 1, 0);
    const uint64_t lo = n_items ? offsets[0] : 0;
    for (size_t f = 0; f <= n_items && n_items; f++) rel[f] = offsets[f] - lo;
    const uint8_t* d = nullptr;
    HostSource host;
    if (n_items && rel[n_items]) {
        bool is_host = false;
        if (XetError* e = stage_input(c, data, lo, offsets[n_items], memory, &d, &is_host)) return e;
        if (is_host) host.base = (const uint8_t*)data + lo;
    }
    RunOptions opt;
    opt.want_chunks = (flags & XETGPU_WANT_CHUNKS) != 0;   // without it only the byte-range items carry chunk lists
    opt.items = &items;
    std::vector<uint64_t> exits;
    opt.item_exit_cuts = &exits;
    if (XetError* e = run_batch(c, d, host, rel.data(), n_items, opt, out_result)) return e;
    if (out_exit_cuts)
        for (size_t i = 0; i < n_items; i++) out_exit_cuts[i] = exits[i];
    return nullptr;
}

// -------------------------------------------------------------------------------------- files
static XetError* stat_regular(const char* path, uint64_t* size) {
    struct stat st;
    if (stat(path, &st) != 0)
        return make_error(errno == ENOENT ? XETGPU_ERR_NOT_FOUND : XETGPU_ERR_IO,
                          std::string("cannot stat file: ") + strerror(errno), path);
    if (!S_ISREG(st.st_mode)) return make_error(XETGPU_ERR_IO, "not a regular file", path);
    *size = (uint64_t)st.st_size;
    return nullptr;
}

static bool cancelled(const XetCancellationToken* cancel) {
    if (!cancel || !cancel->callback) return false;
    CallbackScope scope;
    return cancel->callback(cancel->user_data);
}

// Progress of one xetgpu_hash_files call (or of one device's share of a multi-device call).
struct FileProgress {
    XetGpuContext* report_to = nullptr;   // context whose handler receives the updates
    std::atomic<uint64_t>* shared_bytes = nullptr;  // multi-device: bytes completed by all devices together
    std::atomic<uint32_t>* shared_files = nullptr;  // multi-device: files completed by all devices together
    XetProgressUpdate pu{};
    uint64_t base_bytes = 0;              // single-device: bytes of the batches / pieces already read
    uint64_t skip_bytes = 0;              // bytes of the current read that are a range's lead (not counted as progress)
};

// Reads byte ranges of files into c->d_input through the reader pool.  Caller holds the context.
struct FileSpan {
    size_t file;            // index into paths
    uint64_t off, len;      // range of the file
    uint64_t file_size;
};

static XetError* read_spans_to_device(XetGpuContext* c, const char* const* paths, const std::vector<FileSpan>& spans,
                                      std::vector<uint64_t>* rel, const XetCancellationToken* cancel,
                                      FileProgress* prog) {
    uint64_t bytes = 0;
    for (const auto& sp : spans) bytes += sp.len;
    if (XetError* e = ensure_input(c, bytes)) return e;
    if (XetError* e = ensure_read_slots(c)) return e;
    ReaderPool pool;
    pool.paths = paths;
    pool.stream = c->stream;
    pool.d_base = c->d_input;
    rel->assign(spans.size() + 1, 0);
    uint64_t dev_off = 0;
    for (size_t k = 0; k < spans.size(); k++) {
        const FileSpan& sp = spans[k];
        (*rel)[k] = dev_off;
        for (uint64_t o = 0; o < sp.len; o += c->read_slot_bytes)
            pool.pieces.push_back(ReadPiece{0, nullptr, sp.off + o, std::min<uint64_t>(c->read_slot_bytes, sp.len - o),
                                            dev_off + o, sp.file, sp.file_size});
        dev_off += sp.len;
    }
    (*rel)[spans.size()] = dev_off;
    // this thread reports progress and polls for cancellation while the readers work
    uint64_t reported = 0;
    return run_reader_pool(c, pool, [&](ReaderPool& pl) {
        if (prog && prog->report_to) {
            const uint64_t raw = pl.bytes_done.load();
            const uint64_t done = raw > prog->skip_bytes ? raw - prog->skip_bytes : 0;
            if (prog->shared_bytes) {
                prog->pu.completed_bytes = prog->shared_bytes->fetch_add(done - reported) + (done - reported);
                reported = done;
            } else {
                prog->pu.completed_bytes = prog->base_bytes + done;
            }
            prog->pu.current_file = paths[pl.last_file.load()];
            report_progress(prog->report_to, prog->pu, false);
        }
        if (cancelled(cancel)) pl.fail(make_error(XETGPU_ERR_CANCELLED, "cancelled"));
    });
}

struct MultiPiece {
    size_t file;
    uint64_t begin, end, file_len;
    bool whole() const { return begin == 0 && end == file_len; }
};

// Port of sharded.plan_concatenated: per device, its pieces in file order.
// One owner's chunk list for its range of a shared file.
struct RangePart {
    std::vector<uint64_t> ends;     // file-relative
    std::vector<uint8_t> hashes;
    uint64_t exit_cut = 0;
};

static XetError* hash_file_range_locked(XetGpuContext* c, const char* const* paths, const MultiPiece& p, uint64_t end,
                                        uint64_t entry_cut, const XetCancellationToken* cancel, FileProgress* prog,
                                        RangePart* out);
static XetError* merkle_from_chunks_locked(XetGpuContext* c, const uint8_t* chunk_hashes, const uint64_t* chunk_sizes,
                                           uint64_t n_chunks, uint8_t out_merkle_root[32], uint8_t out_file_hash[32]);
struct FileChunks {
    std::vector<uint64_t> ends;
    std::vector<uint8_t> hashes;
};

// A file larger than max_batch_bytes never sits on the device whole: it streams through in pieces of that size.
// Piece k holds bytes [begin_k - RANGE_LEAD, end_k) and is hashed as a byte range whose entry cut is the exit cut
// of piece k - 1 (exact, no speculation: the pieces run in order), and the chunk lists are combined at the end.
static XetError* hash_large_file_streaming(XetGpuContext* c, const char* const* paths, size_t f, uint64_t size,
                                           const XetCancellationToken* cancel, FileProgress* prog,
                                           XetGpuFileResult* out, std::vector<uint64_t>* ends_out,
                                           std::vector<uint8_t>* hashes_out) {
    const uint64_t piece = std::max<uint64_t>(c->max_batch_bytes / 4096 * 4096, 1ull << 20);
    std::vector<uint64_t> ends;
    std::vector<uint8_t> hashes;
    uint64_t entry = 0;
    for (uint64_t begin = 0; begin < size; begin += piece) {
        const uint64_t end = std::min<uint64_t>(size, begin + piece);
        RangePart part;
        if (XetError* e = hash_file_range_locked(c, paths, MultiPiece{f, begin, end, size}, end, entry, cancel, prog, &part))
            return e;
        ends.insert(ends.end(), part.ends.begin(), part.ends.end());
        hashes.insert(hashes.end(), part.hashes.begin(), part.hashes.end());
        entry = part.exit_cut;
        if (prog) prog->base_bytes += end - begin;
    }
    std::vector<uint64_t> csz(ends.size());
    for (size_t i = 0; i < csz.size(); i++) csz[i] = ends[i] - (i ? ends[i - 1] : 0);
    *out = XetGpuFileResult{};
    out->size = size;
    out->n_chunks = ends.size();
    if (XetError* e = merkle_from_chunks_locked(c, hashes.data(), csz.data(), csz.size(), out->merkle_root, out->file_hash))
        return e;
    if (ends_out) *ends_out = std::move(ends);
    if (hashes_out) *hashes_out = std::move(hashes);
    return nullptr;
}

// Hash whole files (paths[idx[k]] for k in [0, n)) on context c, batch by batch.  Results land in files_out[idx[k]]
// and, with want_chunks, in the per-file chunk vectors.  Caller holds the context.

static XetError* hash_whole_files(XetGpuContext* c, const char* const* paths, const std::vector<size_t>& idx,
                                  const std::vector<uint64_t>& sizes, bool want_chunks, const XetCancellationToken* cancel,
                                  FileProgress* prog, XetGpuFileResult* files_out, std::vector<FileChunks>* chunks_out,
                                  std::vector<XetGpuXorb>* xorbs_out = nullptr) {
    size_t k0 = 0;
    while (k0 < idx.size()) {
        // batch = as many whole files as fit in max_batch_bytes (at least one)
        size_t k1 = k0;
        uint64_t bytes = 0;
        while (k1 < idx.size() && (k1 == k0 || bytes + sizes[idx[k1]] <= c->max_batch_bytes)) bytes += sizes[idx[k1++]];
        if (cancelled(cancel)) return make_error(XETGPU_ERR_CANCELLED, "cancelled");
        if (k1 == k0 + 1 && sizes[idx[k0]] > c->max_batch_bytes && sizes[idx[k0]] > (1ull << 20)) {
            // one file beyond the device budget: stream it (xorbs are not formed for such a file)
            const size_t f = idx[k0];
            FileChunks* fc = want_chunks && chunks_out ? &(*chunks_out)[f] : nullptr;
            if (XetError* e = hash_large_file_streaming(c, paths, f, sizes[f], cancel, prog, &files_out[f],
                                                        fc ? &fc->ends : nullptr, fc ? &fc->hashes : nullptr))
                return e;
            if (prog && prog->report_to && !prog->shared_bytes) {
                prog->pu.completed_bytes = prog->base_bytes;
                prog->pu.completed_files = (uint32_t)k1;
                report_progress(prog->report_to, prog->pu, false);
            } else if (prog && prog->shared_files) {
                prog->shared_files->fetch_add(1);
            }
            k0 = k1;
            continue;
        }
        std::vector<FileSpan> spans;
        for (size_t k = k0; k < k1; k++) spans.push_back(FileSpan{idx[k], 0, sizes[idx[k]], sizes[idx[k]]});
        std::vector<uint64_t> rel;
        if (XetError* e = read_spans_to_device(c, paths, spans, &rel, cancel, prog)) return e;
        if (prog) prog->base_bytes += bytes;
        if (prog && prog->report_to) {
            if (!prog->shared_bytes) {
                prog->pu.completed_bytes = prog->base_bytes;
                prog->pu.completed_files = (uint32_t)k1;
            } else if (prog->shared_files) {
                prog->pu.completed_files = prog->shared_files->fetch_add((uint32_t)(k1 - k0)) + (uint32_t)(k1 - k0);
            }
            report_progress(prog->report_to, prog->pu, false);
        }
        RunOptions opt;
        opt.want_chunks = want_chunks;
        opt.want_xorbs = xorbs_out != nullptr;
        XetGpuResult* part = nullptr;
        if (XetError* e = run_pipeline(c, c->d_input, rel.data(), k1 - k0, opt, &part)) return e;
        for (uint64_t x = 0; xorbs_out && x < part->n_xorbs; x++) {
            XetGpuXorb xo = part->xorbs[x];
            xo.file = idx[k0 + xo.file];
            xorbs_out->push_back(xo);
        }
        for (size_t k = k0; k < k1; k++) {
            const XetGpuFileResult& fr = part->files[k - k0];
            files_out[idx[k]] = fr;
            if (want_chunks && chunks_out && fr.n_chunks) {
                FileChunks& fc = (*chunks_out)[idx[k]];
                fc.ends.assign(part->chunk_ends + fr.first_chunk, part->chunk_ends + fr.first_chunk + fr.n_chunks);
                fc.hashes.assign(part->chunk_hashes + fr.first_chunk * 32, part->chunk_hashes + (fr.first_chunk + fr.n_chunks) * 32);
            }
        }
        xetgpu_free_result(part);
        k0 = k1;
    }
    return nullptr;
}

// Per-file results + per-file chunk vectors -> one XetGpuResult (files in caller order, chunks file by file).
static XetError* assemble_result(std::vector<XetGpuFileResult>& files, std::vector<FileChunks>& chunks, bool want_chunks,
                                 XetGpuResult** out_result, const std::vector<XetGpuXorb>* xorbs = nullptr) {
    XetGpuResult* total = (XetGpuResult*)calloc(1, sizeof(XetGpuResult));
    if (!total) return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
    total->n_files = files.size();
    total->files = (XetGpuFileResult*)calloc(files.size() ? files.size() : 1, sizeof(XetGpuFileResult));
    if (!total->files) {
        free(total);
        return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
    }
    uint64_t n_chunks = 0;
    for (size_t f = 0; f < files.size(); f++) {
        files[f].first_chunk = want_chunks ? n_chunks : 0;
        if (want_chunks) n_chunks += chunks[f].ends.size();
        total->files[f] = files[f];
    }
    if (want_chunks && n_chunks) {
        total->n_chunks = n_chunks;
        total->chunk_ends = (uint64_t*)malloc(n_chunks * 8);
        total->chunk_hashes = (uint8_t*)malloc(n_chunks * 32);
        if (!total->chunk_ends || !total->chunk_hashes) {
            xetgpu_free_result(total);
            return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
        }
        uint64_t w = 0;
        for (size_t f = 0; f < files.size(); f++) {
            const FileChunks& fc = chunks[f];
            if (fc.ends.empty()) continue;
            memcpy(total->chunk_ends + w, fc.ends.data(), fc.ends.size() * 8);
            memcpy(total->chunk_hashes + w * 32, fc.hashes.data(), fc.hashes.size());
            w += fc.ends.size();
        }
    }
    if (xorbs && !xorbs->empty()) {
        total->n_xorbs = xorbs->size();
        total->xorbs = (XetGpuXorb*)malloc(xorbs->size() * sizeof(XetGpuXorb));
        if (!total->xorbs) {
            xetgpu_free_result(total);
            return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
        }
        memcpy(total->xorbs, xorbs->data(), xorbs->size() * sizeof(XetGpuXorb));
        // batches complete in caller order, so the list is already sorted by file
    }
    *out_result = total;
    return nullptr;
}

extern "C" XetError* xetgpu_hash_files(XetGpuContext* c, const char* const* paths, size_t n_files, uint32_t flags,
                                       const XetCancellationToken* cancel, XetGpuResult** out_result) {
    if (!c || !out_result || (n_files && !paths))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_hash_files");
    *out_result = nullptr;
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    std::vector<uint64_t> sizes(n_files);
    uint64_t all_bytes = 0;
    for (size_t f = 0; f < n_files; f++) {
        if (!paths[f]) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL path");
        if (XetError* e = stat_regular(paths[f], &sizes[f])) return e;
        all_bytes += sizes[f];
    }
    const bool want_chunks = (flags & XETGPU_WANT_CHUNKS) != 0;
    FileProgress prog;
    prog.report_to = c;
    prog.pu.phase = XET_PROGRESS_PHASE_SCANNING;
    prog.pu.total_bytes = all_bytes;
    prog.pu.total_files = (uint32_t)n_files;
    std::vector<size_t> idx(n_files);
    for (size_t f = 0; f < n_files; f++) idx[f] = f;
    std::vector<XetGpuFileResult> files(n_files);
    std::vector<FileChunks> chunks(want_chunks ? n_files : 0);
    const bool want_xorbs = (flags & XETGPU_WANT_XORBS) != 0;
    std::vector<XetGpuXorb> xorbs;
    if (XetError* e = hash_whole_files(c, paths, idx, sizes, want_chunks, cancel, &prog, files.data(),
                                       want_chunks ? &chunks : nullptr, want_xorbs ? &xorbs : nullptr)) {
        cudaStreamSynchronize(c->stream);
        return e;
    }
    if (XetError* e = assemble_result(files, chunks, want_chunks, out_result, &xorbs)) return e;
    prog.pu.phase = XET_PROGRESS_PHASE_FINALIZING;
    prog.pu.completed_bytes = all_bytes;
    prog.pu.completed_files = (uint32_t)n_files;
    prog.pu.current_file = nullptr;
    report_progress(c, prog.pu, true);
    return nullptr;
}

// ------------------------------------------------------------------------------------- verify
static XetError* verify_against(const XetGpuResult* r, const char* const* paths, const char* const* expected_hex,
                                const std::vector<std::vector<uint8_t>>& want, size_t n_files, uint8_t* out_ok);
static XetError* parse_expected(const char* const* paths, const char* const* expected_hex, size_t n_files,
                                std::vector<std::vector<uint8_t>>* want);
static XetError* compare_hash(const uint8_t got[32], const char* expected_hex, const char* what) {
    uint8_t want[32];
    if (XetError* e = xetgpu_hash_from_hex(expected_hex, want)) return e;
    if (memcmp(got, want, 32) == 0) return nullptr;
    char hex[65];
    xetgpu_hash_to_hex(got, hex);
    return make_error(XETGPU_ERR_CHECKSUM_MISMATCH, std::string("xet hash mismatch for ") + what,
                      std::string("expected ") + expected_hex + " computed " + hex);
}

extern "C" XetError* xetgpu_verify_file(XetGpuContext* c, const char* path, const char* expected_hex) {
    if (!c || !path || !expected_hex) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_verify_file");
    uint8_t want[32];
    if (XetError* e = xetgpu_hash_from_hex(expected_hex, want)) return e;
    XetGpuResult* r = nullptr;
    const char* paths[1] = {path};
    if (XetError* e = xetgpu_hash_files(c, paths, 1, 0, nullptr, &r)) return e;
    XetError* e = compare_hash(r->files[0].file_hash, expected_hex, path);
    xetgpu_free_result(r);
    return e;
}

extern "C" XetError* xetgpu_verify_files(XetGpuContext* c, const char* const* paths, const char* const* expected_hex,
                                         size_t n_files, const XetCancellationToken* cancel, uint8_t* out_ok) {
    if (!c || (n_files && (!paths || !expected_hex)))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_verify_files");
    std::vector<std::vector<uint8_t>> want;
    if (XetError* e = parse_expected(paths, expected_hex, n_files, &want)) return e;
    XetGpuResult* r = nullptr;
    if (XetError* e = xetgpu_hash_files(c, paths, n_files, 0, cancel, &r)) return e;
    XetError* e = verify_against(r, paths, expected_hex, want, n_files, out_ok);
    xetgpu_free_result(r);
    return e;
}

extern "C" XetError* xetgpu_verify_buffer(XetGpuContext* c, const void* data, uint64_t len, int32_t memory,
                                          const char* expected_hex) {
    if (!c || !expected_hex || (len && !data))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_verify_buffer");
    uint8_t want[32];
    if (XetError* e = xetgpu_hash_from_hex(expected_hex, want)) return e;
    XetGpuResult* r = nullptr;
    uint64_t offsets[2] = {0, len};
    if (XetError* e = xetgpu_hash_buffers(c, data, offsets, 1, memory, 0, &r)) return e;
    XetError* e = compare_hash(r->files[0].file_hash, expected_hex, "buffer");
    xetgpu_free_result(r);
    return e;
}

// ------------------------------------------------------------------------- stage-level entries
extern "C" XetError* xetgpu_gear_candidates(XetGpuContext* c, const void* data, uint64_t len, int32_t memory,
                                            uint64_t** out_indices, uint64_t* out_count) {
    if (!c || !out_indices || !out_count || (len && !data))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_gear_candidates");
    *out_indices = nullptr;
    *out_count = 0;
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    if (len == 0) return nullptr;
    const uint8_t* d = nullptr;
    if (XetError* e = stage_input_now(c, data, 0, len, memory, &d)) return e;
    uint64_t offsets[2] = {0, len};
    for (int attempt = 0; attempt < 3; attempt++) {
        Plan plan = make_plan(offsets, 1, len, c->cand_cap_hint);
        PipelineBuffers b{};
        if (XetError* e = ensure_workspace(&c->lanes[0], plan, &b)) return e;
        b.data = d;
        CU(cudaMemsetAsync(b.scalars, 0, 64, c->stream));
        b.sm_count = c->sm_count;
        cudaError_t gear_err = cudaSuccess;
        launch_gear(b, c->stream, &gear_err);
        if (gear_err != cudaSuccess) return cuda_error(gear_err, "tensor map setup for the gear kernel");
        CU(cudaGetLastError());
        uint64_t n = 0;
        CU(cudaMemcpyAsync(&n, b.scalars + kScalarCand, 8, cudaMemcpyDeviceToHost, c->stream));
        CU(cudaStreamSynchronize(c->stream));
        if (n > plan.cand_cap) {
            c->cand_cap_hint = n + n / 8 + 1024;
            continue;
        }
        uint64_t* out = (uint64_t*)malloc((n ? n : 1) * 8);
        if (!out) return make_error(XETGPU_ERR_UNKNOWN, "out of host memory");
        if (n) CU(cudaMemcpy(out, b.cand, n * 8, cudaMemcpyDeviceToHost));
        *out_indices = out;
        *out_count = n;
        return nullptr;
    }
    return make_error(XETGPU_ERR_UNKNOWN, "candidate list kept overflowing");
}

static XetError* range_common(XetGpuContext* c, const void* data, int32_t memory, uint64_t file_len,
                              uint64_t range_begin, uint64_t range_end, uint64_t entry_cut, bool resolve_only,
                              uint64_t* out_exit_cut, XetGpuResult** out_result) {
    if (!c || !out_exit_cut) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to range call");
    if (range_begin > range_end || range_end > file_len || entry_cut > range_begin)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "need entry_cut <= range_begin <= range_end <= file_len");
    const uint64_t lead = std::min<uint64_t>(range_begin, XETGPU_RANGE_LEAD);
    const uint64_t data_pos = range_begin - lead;  // file position of data[0]
    if (entry_cut < data_pos || range_begin - entry_cut > XETGPU_MAX_CHUNK)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "entry_cut must lie within one maximum chunk before range_begin");
    if (out_result) *out_result = nullptr;
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    const uint64_t len = range_end - data_pos;
    if (len && !data) return make_error(XETGPU_ERR_INVALID_CONFIG, "data is NULL");
    const uint8_t* d = nullptr;
    if (len)
        if (XetError* e = stage_input_now(c, data, 0, len, memory, &d)) return e;
    uint64_t offsets[2] = {0, len};
    RangeWalk rw{entry_cut - data_pos, range_end == file_len ? 1u : 0u};
    RunOptions opt;
    opt.range = &rw;
    opt.end_bias = data_pos;
    opt.want_chunks = !resolve_only;
    opt.resolve_only = resolve_only;
    opt.skip_merkle = true;
    uint64_t exit_abs = 0;
    opt.exit_cut_abs = &exit_abs;
    if (XetError* e = run_pipeline(c, d, offsets, 1, opt, out_result)) return e;
    *out_exit_cut = exit_abs + data_pos;
    return nullptr;
}

extern "C" XetError* xetgpu_hash_range(XetGpuContext* c, const void* data, int32_t memory, uint64_t file_len,
                                       uint64_t range_begin, uint64_t range_end, uint64_t entry_cut,
                                       uint64_t* out_exit_cut, XetGpuResult** out_result) {
    if (!out_result) return make_error(XETGPU_ERR_INVALID_CONFIG, "out_result is NULL");
    return range_common(c, data, memory, file_len, range_begin, range_end, entry_cut, false, out_exit_cut, out_result);
}

extern "C" XetError* xetgpu_range_exit_cut(XetGpuContext* c, const void* data, int32_t memory, uint64_t file_len,
                                           uint64_t range_begin, uint64_t range_end, uint64_t entry_cut,
                                           uint64_t* out_exit_cut) {
    return range_common(c, data, memory, file_len, range_begin, range_end, entry_cut, true, out_exit_cut, nullptr);
}

extern "C" XetError* xetgpu_merkle_from_chunks(XetGpuContext* c, const uint8_t* chunk_hashes,
                                               const uint64_t* chunk_sizes, uint64_t n_chunks,
                                               uint8_t out_merkle_root[32], uint8_t out_file_hash[32]) {
    if (!c || !out_merkle_root || !out_file_hash || (n_chunks && (!chunk_hashes || !chunk_sizes)))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_merkle_from_chunks");
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    return merkle_from_chunks_locked(c, chunk_hashes, chunk_sizes, n_chunks, out_merkle_root, out_file_hash);
}

static XetError* merkle_from_chunks_locked(XetGpuContext* c, const uint8_t* chunk_hashes, const uint64_t* chunk_sizes,
                                           uint64_t n_chunks, uint8_t out_merkle_root[32], uint8_t out_file_hash[32]) {
    // a one-file plan whose slot range holds exactly n_chunks nodes
    Plan plan;
    plan.total = 0;
    plan.n_files = 1;
    plan.slot_off = {0, n_chunks + 1};
    plan.n_slots = n_chunks + 1;
    plan.n_tiles = 0;
    plan.n_groups_cap = 1;
    plan.cand_cap = 1;
    plan.max_nodes = n_chunks + 1;
    PipelineBuffers b{};
    if (XetError* e = ensure_workspace(&c->lanes[0], plan, &b)) return e;
    cudaStream_t st = c->stream;
    uint64_t h_slot[2] = {0, n_chunks + 1};
    uint32_t cnt = (uint32_t)n_chunks;
    CU(cudaMemcpyAsync((void*)b.slot_off, h_slot, 16, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(b.file_count, &cnt, 4, cudaMemcpyHostToDevice, st));
    if (n_chunks) {
        // cudaMemcpyDefault: the lists may live in host memory or (rank-0 combine of a sharded file) on the device
        CU(cudaMemcpyAsync(b.chunk_hash, chunk_hashes, n_chunks * 32, cudaMemcpyDefault, st));
        // sizes may exceed 32 bits here (callers combine arbitrary (hash, size) lists): upload u64
        // sizes straight into the level-0 size array and skip the u32 chunk_len initialisation
        CU(cudaMemcpyAsync(b.node_size[0], chunk_sizes, n_chunks * 8, cudaMemcpyDefault, st));
    }
    CU(cudaStreamSynchronize(st));  // the host sources above are pageable stack/heap memory
    PipelineBuffers m = b;
    m.chunk_len = nullptr;  // launch_merkle: sizes already in place
    launch_merkle(m, st);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_merkle_root, b.file_root, 32, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out_file_hash, b.file_hash, 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return nullptr;
}

extern "C" XetError* xetgpu_fill_splitmix(XetGpuContext* c, void* device_dst, uint64_t seed, uint64_t byte_off,
                                          uint64_t len) {
    if (!c || (len && !device_dst)) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_fill_splitmix");
    LOCK_CTX(c);
    CU(cudaSetDevice(c->device));
    launch_fill_splitmix((uint8_t*)device_dst, seed, byte_off, len, c->stream);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(c->stream));
    return nullptr;
}

extern "C" XetError* xetgpu_last_timings(XetGpuContext* c, XetGpuTimings* out) {
    if (!c || !out) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_last_timings");
    std::lock_guard<std::mutex> lk(c->cb_mu);
    *out = c->last;
    return nullptr;
}

// ------------------------------------------------------------------------------- several GPUs, one process
// model-agent is ONE process per node (cmd/model-agent/main.go:243-271), so the node's GPUs are reached from one
// address space: a multi-device context owns one XetGpuContext per device and one host thread per device per call.
// Work is laid out like BASELINE config 4: the files end to end, cut into one equal byte range per device.  A file
// inside a range is hashed whole by that device; a file that straddles a range boundary is shared: every owner
// hashes its range speculatively (entry cut assumed at the range begin) with the RANGE_LEAD bytes in front of it
// read from the file itself, then the true 8-byte carry walks down the owners on the host and an owner whose
// assumption was wrong re-hashes a small window and splices at the first common cut (same protocol as
// ome_b200/sharded.py, which does it across processes).  The merkle combine runs on the first owner's device.
struct XetGpuMulti {
    std::vector<XetGpuContext*> ctx;
    std::mutex mu;
    uint64_t min_bytes_per_device = 64ull << 20;  // below this much work per device, fewer devices are used
};

// Nearest tensor boundary of a *.safetensors shard to `off`, if one lies within a quarter range and leaves both
// neighbours room for the halo; `off` itself otherwise (other file types, unparsable headers, giant tensors).
static uint64_t snap_to_tensor(const char* path, uint64_t size, uint64_t off, uint64_t range_bytes) {
    const size_t len = strlen(path);
    if (len < 12 || strcmp(path + len - 12, ".safetensors") != 0) return off;
    XetGpuSafetensorsIndex* ix = nullptr;
    if (XetError* e = xetgpu_safetensors_open(path, &ix)) {
        xetgpu_free_error(e);
        return off;
    }
    uint64_t best = off, best_d = range_bytes / 4 + 1;
    for (uint64_t i = 0; i < ix->n_tensors; i++)
        for (uint64_t m : {ix->tensors[i].begin, ix->tensors[i].end}) {
            const uint64_t d = m > off ? m - off : off - m;
            if (d < best_d && m >= XETGPU_RANGE_LEAD + 4096 && size - std::min(size, m) >= XETGPU_RANGE_LEAD + 4096) {
                best_d = d;
                best = m;
            }
        }
    xetgpu_safetensors_free(ix);
    return best;
}

// snap(f, off, range_bytes): optional policy hook that may move an in-file boundary (tensor boundaries of a shard).
typedef std::function<uint64_t(size_t, uint64_t, uint64_t)> SnapFn;

static std::vector<std::vector<MultiPiece>> plan_concatenated(const std::vector<uint64_t>& sizes, size_t world,
                                                              const SnapFn& snap = nullptr, uint64_t align = 4096) {
    std::vector<uint64_t> starts(sizes.size() + 1, 0);
    for (size_t f = 0; f < sizes.size(); f++) starts[f + 1] = starts[f] + sizes[f];
    const uint64_t total = starts[sizes.size()];
    const uint64_t min_piece = XETGPU_RANGE_LEAD + align;
    std::vector<uint64_t> cuts{0};
    for (size_t r = 1; r < world; r++) {
        const uint64_t pos = (uint64_t)((unsigned __int128)total * r / world);
        size_t f = (size_t)(std::upper_bound(starts.begin(), starts.end(), pos) - starts.begin());
        f = f ? f - 1 : 0;
        if (f >= sizes.size()) f = sizes.size() - 1;
        uint64_t off = (pos - starts[f]) / align * align;
        if (snap && off >= min_piece && sizes[f] - off >= min_piece) off = snap(f, off, total / world);
        if (off < min_piece) off = 0;
        else if (sizes[f] - off < min_piece) off = sizes[f];
        cuts.push_back(std::max(starts[f] + off, cuts.back()));
    }
    cuts.push_back(total);
    std::vector<std::vector<MultiPiece>> plan(world);
    for (size_t r = 0; r < world; r++) {
        const uint64_t lo = cuts[r], hi = cuts[r + 1];
        for (size_t f = 0; f < sizes.size(); f++) {
            const uint64_t fb = starts[f], fe = starts[f] + sizes[f];
            if (sizes[f] == 0) {
                if ((lo <= fb && fb < hi) || (r == world - 1 && fb == total)) plan[r].push_back(MultiPiece{f, 0, 0, 0});
                continue;
            }
            const uint64_t b = std::max(lo, fb), e = std::min(hi, fe);
            if (b < e) plan[r].push_back(MultiPiece{f, b - fb, e - fb, sizes[f]});
        }
    }
    return plan;
}

extern "C" XetError* xetgpu_multi_new(const int32_t* devices, size_t n_devices, const XetGpuConfig* config,
                                      XetGpuMulti** out) {
    if (!out) return make_error(XETGPU_ERR_INVALID_CONFIG, "out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return make_error(XETGPU_ERR_UNKNOWN, "no CUDA device available (libxetgpu has no CPU path)",
                          e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    std::vector<int32_t> devs;
    if (devices && n_devices) devs.assign(devices, devices + n_devices);
    else
        for (int d = 0; d < ndev; d++) devs.push_back(d);  // NULL / 0 = every visible device
    // (a device may be listed more than once: it then runs that many workers — the single-GPU tests use this to
    //  exercise the shared-file protocol)
    if (config && config->stream)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "a multi-device context owns its streams: config->stream must be NULL");
    XetGpuMulti* m = new XetGpuMulti();
    if (const char* v = getenv("XETGPU_MULTI_MIN_BYTES")) {  // test knob: spread small jobs over the devices too
        unsigned long long n = strtoull(v, nullptr, 10);
        if (n) m->min_bytes_per_device = n;
    }
    int prev = 0;
    cudaGetDevice(&prev);
    for (int32_t d : devs) {
        XetGpuConfig cfg{};
        if (config) cfg = *config;
        cfg.device = d;
        cfg.stream = nullptr;
        XetGpuContext* c = nullptr;
        if (XetError* err = xetgpu_context_new(&cfg, &c)) {
            for (auto* x : m->ctx) xetgpu_context_free(x);
            delete m;
            cudaSetDevice(prev);
            return err;
        }
        m->ctx.push_back(c);
    }
    cudaSetDevice(prev);
    *out = m;
    return nullptr;
}

extern "C" void xetgpu_multi_free(XetGpuMulti* m) {
    if (!m) return;
    for (auto* c : m->ctx) xetgpu_context_free(c);
    delete m;
}

extern "C" size_t xetgpu_multi_device_count(const XetGpuMulti* m) { return m ? m->ctx.size() : 0; }

extern "C" XetError* xetgpu_multi_set_progress_callback(XetGpuMulti* m, XetProgressCallback callback, void* user_data,
                                                        uint32_t throttle_ms) {
    if (!m) return make_error(XETGPU_ERR_INVALID_CONFIG, "ctx is NULL");
    return xetgpu_context_set_progress_callback(m->ctx[0], callback, user_data, throttle_ms);  // one handler, one throttle
}

// Caller holds c.  Reads [begin - lead, end) of the file and hashes the range from entry_cut.
static XetError* hash_file_range_locked(XetGpuContext* c, const char* const* paths, const MultiPiece& p, uint64_t end,
                                        uint64_t entry_cut, const XetCancellationToken* cancel, FileProgress* prog,
                                        RangePart* out) {
    const uint64_t lead = std::min<uint64_t>(p.begin, XETGPU_RANGE_LEAD);
    const uint64_t data_pos = p.begin - lead;
    std::vector<FileSpan> spans{FileSpan{p.file, data_pos, end - data_pos, p.file_len}};
    std::vector<uint64_t> rel;
    if (prog) prog->skip_bytes = lead;
    XetError* re = read_spans_to_device(c, paths, spans, &rel, cancel, prog);
    if (prog) prog->skip_bytes = 0;
    if (re) return re;
    uint64_t offsets[2] = {0, end - data_pos};
    RangeWalk rw{entry_cut - data_pos, end == p.file_len ? 1u : 0u};
    RunOptions opt;
    opt.range = &rw;
    opt.end_bias = data_pos;
    opt.want_chunks = true;
    opt.skip_merkle = true;
    uint64_t exit_abs = 0;
    opt.exit_cut_abs = &exit_abs;
    XetGpuResult* r = nullptr;
    if (XetError* e = run_pipeline(c, c->d_input, offsets, 1, opt, &r)) return e;
    out->exit_cut = exit_abs + data_pos;
    out->ends.assign(r->chunk_ends, r->chunk_ends + r->n_chunks);
    out->hashes.assign(r->chunk_hashes, r->chunk_hashes + r->n_chunks * 32);
    xetgpu_free_result(r);
    return nullptr;
}

// First cut of the re-synchronised chain `fix` that is also a cut of the speculative chain `spec`: the chunk that
// ends there comes from `fix` (it starts elsewhere in the two chains), everything after it from `spec`.
static bool splice_at_common_cut(const RangePart& fix, RangePart* spec) {
    size_t i = 0, j = 0;
    while (i < fix.ends.size() && j < spec->ends.size()) {
        if (fix.ends[i] == spec->ends[j]) break;
        if (fix.ends[i] < spec->ends[j]) i++;
        else j++;
    }
    if (i >= fix.ends.size() || j >= spec->ends.size()) return false;
    RangePart merged;
    merged.ends.assign(fix.ends.begin(), fix.ends.begin() + i + 1);
    merged.ends.insert(merged.ends.end(), spec->ends.begin() + j + 1, spec->ends.end());
    merged.hashes.assign(fix.hashes.begin(), fix.hashes.begin() + (i + 1) * 32);
    merged.hashes.insert(merged.hashes.end(), spec->hashes.begin() + (j + 1) * 32, spec->hashes.end());
    merged.exit_cut = spec->exit_cut;
    *spec = std::move(merged);
    return true;
}

extern "C" XetError* xetgpu_multi_hash_files(XetGpuMulti* m, const char* const* paths, size_t n_files, uint32_t flags,
                                             const XetCancellationToken* cancel, XetGpuResult** out_result) {
    if (!m || !out_result || (n_files && !paths))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_multi_hash_files");
    *out_result = nullptr;
    if (tl_in_callback)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "re-entrant call: a progress or cancellation callback called a hashing "
                                                     "entry point of the library");
    std::lock_guard<std::mutex> lk(m->mu);
    std::vector<uint64_t> sizes(n_files);
    uint64_t all_bytes = 0;
    for (size_t f = 0; f < n_files; f++) {
        if (!paths[f]) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL path");
        if (XetError* e = stat_regular(paths[f], &sizes[f])) return e;
        all_bytes += sizes[f];
    }
    if (flags & XETGPU_WANT_XORBS)
        return make_error(XETGPU_ERR_INVALID_CONFIG, "XETGPU_WANT_XORBS is not offered by the multi-device call: a shared "
                                                     "file's chunk list is assembled on the host; form xorbs per device");
    const bool want_chunks = (flags & XETGPU_WANT_CHUNKS) != 0;
    const size_t world = (size_t)std::min<uint64_t>(m->ctx.size(), std::max<uint64_t>(1, all_bytes / m->min_bytes_per_device));
    // Boundaries inside a *.safetensors shard move to the nearest tensor boundary (safetensors.cu); a file whose
    // header does not parse is simply cut where the equal split falls.
    SnapFn snap = nullptr;
    if (!(flags & XETGPU_NO_TENSOR_SNAP))
        snap = [&](size_t f, uint64_t off, uint64_t range_bytes) -> uint64_t { return snap_to_tensor(paths[f], sizes[f], off, range_bytes); };
    const auto plan = plan_concatenated(sizes, world, snap);
    std::vector<std::vector<size_t>> owners(n_files);   // devices that hold a piece of each file, ascending
    for (size_t d = 0; d < world; d++)
        for (const auto& p : plan[d]) owners[p.file].push_back(d);

    std::vector<XetGpuFileResult> files(n_files);
    std::vector<FileChunks> chunks(n_files);   // shared files need their lists whatever the flags say
    std::vector<std::vector<RangePart>> parts(n_files);
    for (size_t f = 0; f < n_files; f++) {
        files[f].size = sizes[f];
        if (owners[f].size() > 1) parts[f].resize(owners[f].size());
    }
    std::atomic<uint64_t> bytes_done{0};
    std::atomic<uint32_t> files_done{0};
    std::atomic<bool> stop{false};
    // any device's failure cancels the others through the same token shape
    struct Cancel {
        const XetCancellationToken* user;
        std::atomic<bool>* stop;
    } cstate{cancel, &stop};
    XetCancellationToken tok{[](void* u) -> bool {
                                 Cancel* s = (Cancel*)u;
                                 return s->stop->load() || cancelled(s->user);
                             },
                             &cstate};
    std::vector<XetError*> errs(world, nullptr);
    auto device_work = [&](size_t d) {
        XetGpuContext* c = m->ctx[d];
        CtxGuard g(c);
        if (!g.ok) {
            errs[d] = make_error(XETGPU_ERR_INVALID_CONFIG, "re-entrant call from a callback");
            return;
        }
        if (cudaSetDevice(c->device) != cudaSuccess) {
            errs[d] = make_error(XETGPU_ERR_UNKNOWN, "cudaSetDevice failed");
            return;
        }
        FileProgress prog;
        prog.report_to = m->ctx[0];
        prog.shared_bytes = &bytes_done;
        prog.shared_files = &files_done;
        prog.pu.phase = XET_PROGRESS_PHASE_SCANNING;
        prog.pu.total_bytes = all_bytes;
        prog.pu.total_files = (uint32_t)n_files;
        std::vector<size_t> whole;
        for (const auto& p : plan[d])
            if (p.whole()) whole.push_back(p.file);
        errs[d] = hash_whole_files(c, paths, whole, sizes, want_chunks, &tok, &prog, files.data(), want_chunks ? &chunks : nullptr);
        for (const auto& p : plan[d]) {
            if (errs[d]) break;
            if (p.whole()) continue;
            size_t k = 0;
            while (owners[p.file][k] != d) k++;
            errs[d] = hash_file_range_locked(c, paths, p, p.end, p.begin /* assumed entry; exact for begin == 0 */, &tok,
                                             &prog, &parts[p.file][k]);
        }
        if (errs[d]) stop.store(true);
        cudaStreamSynchronize(c->stream);
    };
    std::vector<std::thread> threads;
    for (size_t d = 1; d < world; d++) threads.emplace_back(device_work, d);
    device_work(0);
    for (auto& t : threads) t.join();
    XetError* first = nullptr;
    for (auto*& e : errs) {
        // report the root cause, not the cancellations it triggered on the other devices
        if (e && (!first || (first->code == XETGPU_ERR_CANCELLED && e->code != XETGPU_ERR_CANCELLED))) std::swap(first, e);
        if (e) xetgpu_free_error(e);
    }
    if (first) return first;

    // carry chain + merkle combine of the shared files (host side, a few small launches per boundary)
    for (size_t f = 0; f < n_files; f++) {
        if (owners[f].size() < 2) continue;
        uint64_t entry = 0;
        FileChunks& fc = chunks[f];
        fc.ends.clear();
        fc.hashes.clear();
        for (size_t k = 0; k < owners[f].size(); k++) {
            const size_t d = owners[f][k];
            const MultiPiece* p = nullptr;
            for (const auto& q : plan[d])
                if (q.file == f) p = &q;
            RangePart& spec = parts[f][k];
            if (k > 0 && entry != p->begin) {
                XetGpuContext* c = m->ctx[d];
                CtxGuard g(c);
                if (!g.ok) return make_error(XETGPU_ERR_INVALID_CONFIG, "re-entrant call from a callback");
                CU(cudaSetDevice(c->device));
                uint64_t window = 4ull << 20;
                while (true) {
                    const uint64_t w_end = std::min<uint64_t>(p->end, p->begin + window);
                    RangePart fix;
                    if (XetError* e = hash_file_range_locked(c, paths, *p, w_end, entry, cancel, nullptr, &fix)) return e;
                    if (splice_at_common_cut(fix, &spec)) break;
                    if (w_end == p->end) {   // the chains never met inside this range: the re-hash is the answer
                        spec = std::move(fix);
                        break;
                    }
                    window *= 4;
                }
            }
            fc.ends.insert(fc.ends.end(), spec.ends.begin(), spec.ends.end());
            fc.hashes.insert(fc.hashes.end(), spec.hashes.begin(), spec.hashes.end());
            entry = spec.exit_cut;
        }
        std::vector<uint64_t> csz(fc.ends.size());
        for (size_t i = 0; i < csz.size(); i++) csz[i] = fc.ends[i] - (i ? fc.ends[i - 1] : 0);
        if (XetError* e = xetgpu_merkle_from_chunks(m->ctx[owners[f][0]], fc.hashes.data(), csz.data(), csz.size(),
                                                    files[f].merkle_root, files[f].file_hash))
            return e;
        files[f].n_chunks = csz.size();
    }
    if (XetError* e = assemble_result(files, chunks, want_chunks, out_result)) return e;
    XetProgressUpdate pu{};
    pu.phase = XET_PROGRESS_PHASE_FINALIZING;
    pu.total_bytes = pu.completed_bytes = all_bytes;
    pu.total_files = pu.completed_files = (uint32_t)n_files;
    report_progress(m->ctx[0], pu, true);
    return nullptr;
}

static XetError* verify_against(const XetGpuResult* r, const char* const* paths, const char* const* expected_hex,
                                const std::vector<std::vector<uint8_t>>& want, size_t n_files, uint8_t* out_ok) {
    std::string bad;
    size_t n_bad = 0;
    for (size_t f = 0; f < n_files; f++) {
        const bool ok = want[f].empty() || memcmp(r->files[f].file_hash, want[f].data(), 32) == 0;
        if (out_ok) out_ok[f] = ok ? 1 : 0;
        if (!ok) {
            char hex[65];
            xetgpu_hash_to_hex(r->files[f].file_hash, hex);
            if (n_bad < 16) bad += std::string(paths[f]) + ": expected " + expected_hex[f] + " computed " + hex + "\n";
            n_bad++;
        }
    }
    if (!n_bad) return nullptr;
    return make_error(XETGPU_ERR_CHECKSUM_MISMATCH,
                      "xet hash mismatch for " + std::to_string(n_bad) + " of " + std::to_string(n_files) + " files", bad);
}

static XetError* parse_expected(const char* const* paths, const char* const* expected_hex, size_t n_files,
                                std::vector<std::vector<uint8_t>>* want) {
    want->assign(n_files, {});
    for (size_t f = 0; f < n_files; f++) {
        if (!paths[f]) return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL path");
        if (expected_hex[f] && expected_hex[f][0]) {
            (*want)[f].resize(32);
            if (XetError* e = xetgpu_hash_from_hex(expected_hex[f], (*want)[f].data())) return e;
        }
    }
    return nullptr;
}

extern "C" XetError* xetgpu_multi_verify_files(XetGpuMulti* m, const char* const* paths, const char* const* expected_hex,
                                               size_t n_files, const XetCancellationToken* cancel, uint8_t* out_ok) {
    if (!m || (n_files && (!paths || !expected_hex)))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_multi_verify_files");
    std::vector<std::vector<uint8_t>> want;
    if (XetError* e = parse_expected(paths, expected_hex, n_files, &want)) return e;
    XetGpuResult* r = nullptr;
    if (XetError* e = xetgpu_multi_hash_files(m, paths, n_files, 0, cancel, &r)) return e;
    XetError* e = verify_against(r, paths, expected_hex, want, n_files, out_ok);
    xetgpu_free_result(r);
    return e;
}

// Test / tooling hook: the byte ranges xetgpu_multi_hash_files would give each device for these files.
// out_pieces receives up to cap entries of (device, file, begin, end), in device order; returns the count.
extern "C" XetError* xetgpu_multi_plan(XetGpuMulti* m, const char* const* paths, size_t n_files, uint32_t flags,
                                       uint64_t* out_pieces, size_t cap, size_t* out_count) {
    if (!m || !out_count || (n_files && !paths) || (cap && !out_pieces))
        return make_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_multi_plan");
    std::vector<uint64_t> sizes(n_files);
    uint64_t all_bytes = 0;
    for (size_t f = 0; f < n_files; f++) {
        if (XetError* e = stat_regular(paths[f], &sizes[f])) return e;
        all_bytes += sizes[f];
    }
    const size_t world = (size_t)std::min<uint64_t>(m->ctx.size(), std::max<uint64_t>(1, all_bytes / m->min_bytes_per_device));
    SnapFn snap = nullptr;
    if (!(flags & XETGPU_NO_TENSOR_SNAP))
        snap = [&](size_t f, uint64_t off, uint64_t range_bytes) -> uint64_t { return snap_to_tensor(paths[f], sizes[f], off, range_bytes); };
    const auto plan = plan_concatenated(sizes, world, snap);
    size_t n = 0;
    for (size_t d = 0; d < plan.size(); d++)
        for (const auto& p : plan[d]) {
            if (n < cap) {
                out_pieces[4 * n] = d;
                out_pieces[4 * n + 1] = p.file;
                out_pieces[4 * n + 2] = p.begin;
                out_pieces[4 * n + 3] = p.end;
            }
            n++;
        }
    *out_count = n;
    return nullptr;
}
