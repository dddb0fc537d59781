// xetgpu_merkle_from_chunks_static — merkle root + file hash of a chunk list WITHOUT allocating.
//
// Who this is for.  xetgpu_merkle_from_chunks (api.cu) carves its level buffers out of a
// context workspace that it grows with cudaMalloc on demand.  Some callers must not trigger a
// device allocation at all: code that runs while another framework's allocator owns the whole
// GPU (a serving process whose memory pool was sized to the last byte), code on a latency path
// where an implicit device synchronisation inside cudaMalloc/cudaFree is unacceptable, and code
// that holds no XetGpuContext because it only ever combines a handful of chunk lists (the rank-0
// merge of a byte-range partitioned shard).  For them this entry point keeps its scratch in a
// file-scope __device__ array that is part of the loaded library image: every GPU that loads
// libxetgpu.so holds those ~2.2 MB until the process exits, and nothing is allocated per call.
// The caller still owns the inputs and outputs, which are DEVICE pointers.
//
// Limits and rules.
//   * n_chunks <= XETGPU_STATIC_MERKLE_MAX_CHUNKS (65,536: any file up to 512 MiB, ~4 GiB of
//     typical 64 KiB chunks); larger lists are rejected with code 1 and must use the
//     context entry point.
//   * The first call on each device looks the array's device address up once
//     (cudaGetSymbolAddress, with that device current).  Make that first call outside any stream
//     capture; later calls on the device only launch one kernel.
//   * All calls on a device share the one array and it keeps whatever the previous call left in
//     it.  The kernel never depends on that: every scratch word it reads was written earlier in
//     the same launch, and the per-level group count lives in shared memory and is reset at the
//     top of every level.  Two calls whose kernels overlap on the GPU (different streams, or
//     different host threads on different streams) race on the array — serialise them.
//   * The launch is asynchronous on `stream`; outputs are valid once the stream has drained.
//   * No cudaMalloc, cudaFree, cudaMallocAsync or library routine that allocates device memory is
//     called here (host-side error objects come from malloc, as everywhere in this ABI).
//
// The arithmetic is merklehash::aggregated_node_hash / file_hash exactly as in merkle.cu
// (SURVEY.md A.5); this file is a single-CTA variant of those level kernels.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/xetgpu.h"
#include "blake3_dev.cuh"

namespace {

using namespace xetgpu;

constexpr uint32_t kMaxN = XETGPU_STATIC_MERKLE_MAX_CHUNKS;
constexpr uint32_t kThreads = 1024;

// scratch carve-up (byte offsets, each a multiple of 256)
constexpr size_t align256(size_t x) { return (x + 255) & ~size_t(255); }
constexpr size_t kOffGroup = 0;                                               // u32 [kMaxN + 1]
constexpr size_t kOffHashA = kOffGroup + align256((kMaxN + 1) * 4);            // level 1, 3, ...: <= n/2 + 1 nodes
constexpr size_t kOffSizeA = kOffHashA + align256((kMaxN / 2 + 1) * 32);
constexpr size_t kOffHashB = kOffSizeA + align256((kMaxN / 2 + 1) * 8);        // level 2, 4, ...: <= n/4 + 2 nodes
constexpr size_t kOffSizeB = kOffHashB + align256((kMaxN / 4 + 2) * 32);
constexpr size_t kScratchBytes = kOffSizeB + align256((kMaxN / 4 + 2) * 8);

// The scratch itself: lives in the library image on every device that loads it.
static __device__ __align__(256) unsigned char g_merkle_scratch[kScratchBytes];

__device__ __forceinline__ uint32_t hex_digit(uint32_t v) { return v < 10 ? '0' + v : 'a' + (v - 10); }

// One node: keyed BLAKE3 (INTERNAL_NODE_KEY) of "{hex(hash)} : {size}\n" for children [c0, c1).
__device__ void hash_node(const uint8_t* __restrict__ cur_hash, const uint64_t* __restrict__ cur_size, uint32_t c0,
                          uint32_t c1, uint8_t* __restrict__ out_hash, uint64_t* __restrict__ out_size) {
    uint32_t w[208];  // 832 bytes >= 9 children * 89
#pragma unroll 1
    for (int i = 0; i < 208; i++) w[i] = 0u;
    uint32_t len = 0;
    auto put = [&](uint32_t byte) {
        w[len >> 2] |= byte << (8u * (len & 3u));
        len++;
    };
    uint64_t total = 0;
    for (uint32_t c = c0; c < c1; c++) {
        const uint8_t* h = cur_hash + (size_t)c * 32;
        for (int wd = 0; wd < 4; wd++)  // MerkleHash::hex: four LE u64 words as %016x
            for (int b = 7; b >= 0; b--) {
                const uint32_t v = h[8 * wd + b];
                put(hex_digit(v >> 4));
                put(hex_digit(v & 15u));
            }
        put(' ');
        put(':');
        put(' ');
        const uint64_t sz = cur_size[c];
        total += sz;
        char dec[20];
        int nd = 0;
        uint64_t t = sz;
        do {
            dec[nd++] = (char)('0' + (t % 10));
            t /= 10;
        } while (t);
        while (nd) put((uint32_t)dec[--nd]);
        put('\n');
    }
    const uint32_t key[8] = XG_NODE_KEY_WORDS;
    uint32_t cv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) cv[i] = key[i];
    const uint32_t nb = len <= 64 ? 1u : (len + 63u) / 64u;
    for (uint32_t b = 0; b < nb; b++) {
        uint32_t m[16];
#pragma unroll
        for (int i = 0; i < 16; i++) m[i] = w[b * 16 + i];
        uint32_t flags = B3_KEYED_HASH | (b == 0 ? B3_CHUNK_START : 0u);
        uint32_t bl = 64;
        if (b == nb - 1) {
            flags |= B3_CHUNK_END | B3_ROOT;
            bl = len - b * 64;
        }
        b3_compress(cv, m, 0u, 0u, bl, flags);
    }
    uint32_t* o = reinterpret_cast<uint32_t*>(out_hash);
#pragma unroll
    for (int i = 0; i < 8; i++) o[i] = cv[i];
    *out_size = total;
}

__global__ void __launch_bounds__(kThreads)
    merkle_static_kernel(const uint8_t* __restrict__ in_hash, const uint64_t* __restrict__ in_size, uint32_t n,
                         uint8_t* __restrict__ out_root, uint8_t* __restrict__ out_file_hash,
                         unsigned char* __restrict__ scratch) {
    __shared__ uint32_t s_flags[kMaxN / 32];  // bit i: node i ends a group (hash % 4 == 0)
    __shared__ uint32_t s_ng;
    const uint32_t tid = threadIdx.x;
    if (n == 0) {  // merklehash::file_hash of an empty list is the zero hash
        if (tid < 32) out_root[tid] = 0, out_file_hash[tid] = 0;
        return;
    }
    uint32_t* group_start = reinterpret_cast<uint32_t*>(scratch + kOffGroup);
    uint8_t* lvl_hash[2] = {scratch + kOffHashA, scratch + kOffHashB};
    uint64_t* lvl_size[2] = {reinterpret_cast<uint64_t*>(scratch + kOffSizeA),
                             reinterpret_cast<uint64_t*>(scratch + kOffSizeB)};
    const uint8_t* cur_hash = in_hash;
    const uint64_t* cur_size = in_size;
    int nxt = 0;
    while (n > 1) {
        if (tid == 0) s_ng = 0;  // reset every level: never carried over from an earlier level or call
        for (uint32_t wi = tid; wi < (n + 31) / 32; wi += kThreads) {
            uint32_t bits = 0;
            for (uint32_t b = 0; b < 32 && wi * 32 + b < n; b++)
                bits |= ((cur_hash[(size_t)(wi * 32 + b) * 32 + 24] & 3u) == 0u ? 1u : 0u) << b;
            s_flags[wi] = bits;
        }
        __syncthreads();
        if (tid == 0) {  // greedy grouping: 2..9 children, cut after a flagged child at index >= 2
            uint32_t r = 0, ng = 0;
            while (r < n) {
                const uint32_t rest = n - r;
                uint32_t k;
                if (rest <= 2) {
                    k = rest;
                } else {
                    const uint32_t end = rest < 9 ? rest : 9;
                    k = end;
                    for (uint32_t i = 2; i < end; i++) {
                        const uint32_t p = r + i;
                        if ((s_flags[p >> 5] >> (p & 31)) & 1u) {
                            k = i + 1;
                            break;
                        }
                    }
                }
                group_start[ng++] = r;
                r += k;
            }
            group_start[ng] = n;
            s_ng = ng;
        }
        __syncthreads();
        const uint32_t ng = s_ng;
        for (uint32_t g = tid; g < ng; g += kThreads)
            hash_node(cur_hash, cur_size, group_start[g], group_start[g + 1], lvl_hash[nxt] + (size_t)g * 32,
                      lvl_size[nxt] + g);
        __syncthreads();
        cur_hash = lvl_hash[nxt];
        cur_size = lvl_size[nxt];
        nxt ^= 1;
        n = ng;
    }
    if (tid == 0) {
        const uint32_t* r = reinterpret_cast<const uint32_t*>(cur_hash);
        uint32_t m[16] = {0};
        uint32_t* o = reinterpret_cast<uint32_t*>(out_root);
        for (int i = 0; i < 8; i++) {
            m[i] = r[i];
            o[i] = r[i];
        }
        uint32_t cv[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // file_hash: keyed by the all-zero salt
        b3_compress(cv, m, 0u, 0u, 32u, B3_KEYED_HASH | B3_CHUNK_START | B3_CHUNK_END | B3_ROOT);
        uint32_t* f = reinterpret_cast<uint32_t*>(out_file_hash);
        for (int i = 0; i < 8; i++) f[i] = cv[i];
    }
}

XetError* static_error(int32_t code, const char* message, const char* details) {
    XetError* e = (XetError*)malloc(sizeof(XetError));
    if (!e) return nullptr;
    e->code = code;
    e->message = strdup(message);
    e->details = details ? strdup(details) : nullptr;
    return e;
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {};  // device address of g_merkle_scratch, per device

}  // namespace

extern "C" XetError* xetgpu_merkle_from_chunks_static(const uint8_t* device_chunk_hashes,
                                                      const uint64_t* device_chunk_sizes, uint64_t n_chunks,
                                                      uint8_t* device_out_merkle_root, uint8_t* device_out_file_hash,
                                                      void* stream) {
    if (!device_out_merkle_root || !device_out_file_hash || (n_chunks && (!device_chunk_hashes || !device_chunk_sizes)))
        return static_error(XETGPU_ERR_INVALID_CONFIG, "NULL argument to xetgpu_merkle_from_chunks_static", nullptr);
    if (n_chunks > kMaxN) {
        char buf[160];
        snprintf(buf, sizeof buf, "n_chunks = %llu, limit XETGPU_STATIC_MERKLE_MAX_CHUNKS = %u",
                 (unsigned long long)n_chunks, kMaxN);
        return static_error(XETGPU_ERR_INVALID_CONFIG,
                            "chunk list too long for the static-scratch entry point; use xetgpu_merkle_from_chunks", buf);
    }
    // The kernel moves hashes as 32-bit words and sizes as 64-bit words; a pointer into the middle of a packed
    // record would fault on the device, and such a fault is sticky for the whole process, so refuse it here.
    if (((uintptr_t)device_chunk_hashes | (uintptr_t)device_out_merkle_root | (uintptr_t)device_out_file_hash) & 3u ||
        ((uintptr_t)device_chunk_sizes & 7u))
        return static_error(XETGPU_ERR_INVALID_CONFIG,
                            "misaligned pointer: hashes and outputs must be 4-byte aligned, sizes 8-byte aligned", nullptr);
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_error(XETGPU_ERR_UNKNOWN, "cudaGetDevice failed", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices)
        return static_error(XETGPU_ERR_INVALID_CONFIG, "device ordinal beyond the static-scratch table", nullptr);
    unsigned char* scratch = nullptr;
    {
        // Once per device, with that device current: in host code the array's name is only a
        // stand-in for the symbol, not a device address, so ask the runtime for the real one.
        std::lock_guard<std::mutex> lk(g_setup_mu);
        if (!g_scratch_addr[dev]) {
            unsigned char* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_merkle_scratch);
            if (e != cudaSuccess || !p)
                return static_error(XETGPU_ERR_UNKNOWN, "cudaGetSymbolAddress(g_merkle_scratch) failed",
                                    cudaGetErrorString(e));
            g_scratch_addr[dev] = p;
        }
        scratch = g_scratch_addr[dev];
    }
    merkle_static_kernel<<<1, kThreads, 0, (cudaStream_t)stream>>>(device_chunk_hashes, device_chunk_sizes,
                                                                  (uint32_t)n_chunks, device_out_merkle_root,
                                                                  device_out_file_hash, scratch);
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_error(XETGPU_ERR_UNKNOWN, "merkle_static_kernel launch failed", cudaGetErrorString(e));
    return nullptr;
}
