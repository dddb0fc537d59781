// K4: merkle aggregation and file hash.
//
// Restates merklehash::aggregated_node_hash / file_hash (xet-core merklehash 0.14.5, pinned at
// /root/reference pkg/xet/Cargo.lock:1651-1663) per SURVEY.md A.5:
//   - a level is cut greedily into groups of 2..9 children; after the first two children a group
//     ends with the first child whose hash (as LE u64 word 3) is 0 mod 4, or at nine children;
//     a tail of <= 2 children forms the last group
//   - node = BLAKE3_keyed(INTERNAL_NODE_KEY, concat("{hex(hash)} : {size}\n")), size = sum of sizes
//   - levels repeat until one node remains; file hash = BLAKE3_keyed(0^32, root)
// A node message is <= 9 * 89 = 801 bytes, i.e. always a single BLAKE3 leaf (<= 13 blocks).
//
// Every file in the batch advances one level per round of launches: the grouping kernels below,
// then one thread per new node that formats and hashes its lines.  This stage touches 40 bytes
// per ~64 KiB chunk, so it is launch/latency-bound, not bandwidth-bound.
#include "blake3_dev.cuh"
#include "xetgpu_internal.h"

namespace xetgpu {

// ---- grouping -----------------------------------------------------------------------------
// Cutting a level into groups is a greedy chain (next start = start + k(start), k in 1..9), but the
// jump is bounded, so the chain through a block of 32 nodes is a function of just the offset
// (0..8) at which it enters the block.  Per level:
//   A  one thread per 32-node block tabulates, for each of the 9 possible entry offsets, the exit
//      offset and the number of group starts inside the block                    (parallel)
//   B  one thread per file composes the block tables left to right: a shift and a mask per block
//   C  one thread per block replays its 32 nodes from its now-known entry offset and writes the
//      group starts                                                             (parallel)
// Level bookkeeping for file f lives in the file's slot range [slot_off[f], slot_off[f+1]).

// Children taken by a group starting at block-relative node r, with `rest` nodes left in the
// level; bits holds the "hash % 4 == 0" flags of block nodes 0..39 (8 nodes of lookahead).
__device__ __forceinline__ uint32_t group_len(uint64_t bits, uint32_t r, uint32_t rest) {
    if (rest <= 2) return rest;
    const uint32_t end = rest < 9 ? rest : 9;
    const uint32_t m = (uint32_t)(bits >> (r + 2)) & ((1u << (end - 2)) - 1u);  // children 2 .. end-1
    return m ? (uint32_t)__ffs(m) + 2u : end;
}

__device__ __forceinline__ uint32_t file_of_slot_g(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t s) {
    uint32_t lo = 0, hi = n_files;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (slot_off[mid] <= s) lo = mid;
        else hi = mid;
    }
    return lo;
}

__global__ void merkle_block_map_kernel(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t n_slots,
                                        const uint8_t* __restrict__ cur_hash, const uint32_t* __restrict__ cur_count,
                                        uint64_t* __restrict__ block_bits, uint64_t* __restrict__ map_out,
                                        uint64_t* __restrict__ map_cnt) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t f = file_of_slot_g(slot_off, n_files, s);
    const uint64_t base = slot_off[f];
    const uint32_t n = cur_count[f];
    const uint64_t b = s - base;
    if (n <= 1 || b * 32 >= n) return;
    const uint32_t first = (uint32_t)b * 32;
    uint64_t bits = 0;
    for (uint32_t i = 0; i < 40 && first + i < n; i++)
        // MerkleHash % 4 == word 3 (LE u64 at bytes 24..31) % 4 == low two bits of byte 24
        bits |= (uint64_t)((cur_hash[(base + first + i) * 32 + 24] & 3u) == 0u) << i;
    uint64_t mo = 0, mc = 0;
    for (uint32_t o = 0; o < 9; o++) {
        uint32_t r = o, cnt = 0;
        while (r < 32 && first + r < n) {
            r += group_len(bits, r, n - (first + r));
            cnt++;
        }
        const uint32_t out = r >= 32 ? r - 32 : 0;
        mo |= (uint64_t)out << (4 * o);
        mc |= (uint64_t)cnt << (5 * o);
    }
    block_bits[s] = bits;
    map_out[s] = mo;
    map_cnt[s] = mc;
}

__global__ void merkle_compose_kernel(const uint64_t* __restrict__ slot_off, uint32_t n_files,
                                      const uint8_t* __restrict__ cur_hash, const uint32_t* __restrict__ cur_count,
                                      const uint64_t* __restrict__ map_out, const uint64_t* __restrict__ map_cnt,
                                      uint32_t* __restrict__ block_entry, uint32_t* __restrict__ block_base,
                                      uint32_t* __restrict__ next_count, uint32_t* __restrict__ group_start,
                                      uint8_t* __restrict__ file_root) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_files) return;
    const uint32_t n = cur_count[f];
    const uint64_t base = slot_off[f];
    if (n <= 1) {
        if (n == 1) {  // finished: latch the root
            const uint4* src = reinterpret_cast<const uint4*>(cur_hash + base * 32);
            uint4* dst = reinterpret_cast<uint4*>(file_root + (uint64_t)f * 32);
            dst[0] = src[0];
            dst[1] = src[1];
        }
        next_count[f] = 0;
        return;
    }
    const uint32_t nb = (n + 31) / 32;
    uint32_t o = 0, total = 0;
#pragma unroll 8
    for (uint32_t b = 0; b < nb; b++) {
        const uint64_t mo = map_out[base + b], mc = map_cnt[base + b];
        block_entry[base + b] = o;
        block_base[base + b] = total;
        total += (uint32_t)(mc >> (5 * o)) & 31u;
        o = (uint32_t)(mo >> (4 * o)) & 15u;
    }
    group_start[base + total] = n;  // sentinel; total <= (n + 2) / 3 < n for n >= 2
    next_count[f] = total;
}

__global__ void merkle_emit_kernel(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t n_slots,
                                   const uint32_t* __restrict__ cur_count, const uint64_t* __restrict__ block_bits,
                                   const uint32_t* __restrict__ block_entry, const uint32_t* __restrict__ block_base,
                                   uint32_t* __restrict__ group_start) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t f = file_of_slot_g(slot_off, n_files, s);
    const uint64_t base = slot_off[f];
    const uint32_t n = cur_count[f];
    const uint64_t b = s - base;
    if (n <= 1 || b * 32 >= n) return;
    const uint32_t first = (uint32_t)b * 32;
    const uint64_t bits = block_bits[s];
    uint32_t r = block_entry[s];
    uint64_t w = base + block_base[s];
    while (r < 32 && first + r < n) {
        group_start[w++] = first + r;
        r += group_len(bits, r, n - (first + r));
    }
}

__device__ __forceinline__ uint32_t file_of_slot_m(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t s) {
    uint32_t lo = 0, hi = n_files;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (slot_off[mid] <= s) lo = mid;
        else hi = mid;
    }
    return lo;
}

struct MsgBuf {
    uint32_t w[208];  // 832 bytes >= 801
    uint32_t len;
    __device__ __forceinline__ void put(uint32_t byte) {
        w[len >> 2] |= byte << (8u * (len & 3u));
        len++;
    }
};

__device__ __forceinline__ uint32_t hex_digit(uint32_t v) { return v < 10 ? '0' + v : 'a' + (v - 10); }

__global__ void merkle_hash_kernel(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t n_slots,
                                   const uint8_t* __restrict__ cur_hash, const uint64_t* __restrict__ cur_size,
                                   const uint32_t* __restrict__ next_count, const uint32_t* __restrict__ group_start,
                                   uint8_t* __restrict__ next_hash, uint64_t* __restrict__ next_size) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t f = file_of_slot_m(slot_off, n_files, s);
    const uint64_t base = slot_off[f];
    const uint32_t gi = (uint32_t)(s - base);
    if (gi >= next_count[f]) return;
    const uint32_t c0 = group_start[base + gi], c1 = group_start[base + gi + 1];

    MsgBuf msg;
#pragma unroll 1
    for (int i = 0; i < 208; i++) msg.w[i] = 0u;
    msg.len = 0;
    uint64_t total = 0;
    for (uint32_t c = c0; c < c1; c++) {
        const uint8_t* h = cur_hash + (base + c) * 32;
        // MerkleHash::hex: four LE u64 words, each printed as 16 hex digits
        for (int wd = 0; wd < 4; wd++)
            for (int b = 7; b >= 0; b--) {
                const uint32_t v = h[8 * wd + b];
                msg.put(hex_digit(v >> 4));
                msg.put(hex_digit(v & 15u));
            }
        msg.put(' ');
        msg.put(':');
        msg.put(' ');
        const uint64_t sz = cur_size[base + c];
        total += sz;
        char dec[20];
        int nd = 0;
        uint64_t t = sz;
        do {
            dec[nd++] = (char)('0' + (t % 10));
            t /= 10;
        } while (t);
        while (nd) msg.put((uint32_t)dec[--nd]);
        msg.put('\n');
    }
    // single-leaf keyed BLAKE3 of the message
    const uint32_t key[8] = XG_NODE_KEY_WORDS;
    uint32_t cv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) cv[i] = key[i];
    const uint32_t nb = msg.len <= 64 ? 1u : (msg.len + 63u) / 64u;
    for (uint32_t b = 0; b < nb; b++) {
        uint32_t m[16];
#pragma unroll
        for (int i = 0; i < 16; i++) m[i] = msg.w[b * 16 + i];
        uint32_t flags = B3_KEYED_HASH;
        if (b == 0) flags |= B3_CHUNK_START;
        uint32_t bl = 64;
        if (b == nb - 1) {
            flags |= B3_CHUNK_END | B3_ROOT;
            bl = msg.len - b * 64;
        }
        b3_compress(cv, m, 0u, 0u, bl, flags);
    }
    uint4* dst = reinterpret_cast<uint4*>(next_hash + s * 32);
    dst[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
    dst[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
    next_size[s] = total;
}

// Level-0 node sizes are the chunk lengths.
__global__ void merkle_init_kernel(uint64_t n_slots, const uint32_t* __restrict__ chunk_len,
                                   uint64_t* __restrict__ node_size) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n_slots) node_size[s] = chunk_len[s];
}

// file_hash = keyed BLAKE3 of the 32-byte root under the all-zero salt; empty file -> zero hash.
__global__ void file_hash_kernel(uint32_t n_files, const uint32_t* __restrict__ file_count,
                                 uint8_t* __restrict__ file_root, uint8_t* __restrict__ file_hash) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_files) return;
    uint4* root = reinterpret_cast<uint4*>(file_root + (uint64_t)f * 32);
    uint4* out = reinterpret_cast<uint4*>(file_hash + (uint64_t)f * 32);
    if (file_count[f] == 0) {
        root[0] = root[1] = out[0] = out[1] = make_uint4(0, 0, 0, 0);
        return;
    }
    const uint4 r0 = root[0], r1 = root[1];
    uint32_t m[16] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, 0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t cv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    b3_compress(cv, m, 0u, 0u, 32u, B3_KEYED_HASH | B3_CHUNK_START | B3_CHUNK_END | B3_ROOT);
    out[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
    out[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
}

int launch_merkle(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_files == 0) return 0;
    int n = 0;
    const unsigned fblocks = (b.n_files + 127) / 128;
    const unsigned sblocks = (unsigned)((b.n_slots + 127) / 128);
    // level 0: hashes = chunk hashes (node_hash[0] aliases chunk_hash), sizes = chunk lengths
    if (b.n_slots && b.chunk_len) {
        merkle_init_kernel<<<(unsigned)((b.n_slots + 255) / 256), 256, 0, st>>>(b.n_slots, b.chunk_len, b.node_size[0]);
        n++;
    }
    cudaMemcpyAsync(b.level_count[0], b.file_count, sizeof(uint32_t) * b.n_files, cudaMemcpyDeviceToDevice, st);
    int cur = 0;
    for (uint32_t level = 0; level <= b.max_levels; level++) {
        if (b.n_slots) {
            merkle_block_map_kernel<<<sblocks, 128, 0, st>>>(b.slot_off, b.n_files, b.n_slots, b.node_hash[cur],
                                                             b.level_count[cur], b.mk_bits, b.mk_out, b.mk_cnt);
            n++;
        }
        merkle_compose_kernel<<<fblocks, 128, 0, st>>>(b.slot_off, b.n_files, b.node_hash[cur], b.level_count[cur],
                                                       b.mk_out, b.mk_cnt, b.mk_entry, b.mk_base,
                                                       b.level_count[cur ^ 1], b.group_start, b.file_root);
        n++;
        if (level == b.max_levels || b.n_slots == 0) break;
        merkle_emit_kernel<<<sblocks, 128, 0, st>>>(b.slot_off, b.n_files, b.n_slots, b.level_count[cur], b.mk_bits,
                                                    b.mk_entry, b.mk_base, b.group_start);
        merkle_hash_kernel<<<sblocks, 128, 0, st>>>(b.slot_off, b.n_files, b.n_slots, b.node_hash[cur],
                                                    b.node_size[cur], b.level_count[cur ^ 1], b.group_start,
                                                    b.node_hash[cur ^ 1], b.node_size[cur ^ 1]);
        n += 2;
        cur ^= 1;
    }
    file_hash_kernel<<<fblocks, 128, 0, st>>>(b.n_files, b.file_count, b.file_root, b.file_hash);
    return n + 1;
}

}  // namespace xetgpu
