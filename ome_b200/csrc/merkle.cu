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
#include <stdio.h>

#include <algorithm>

#include "blake3_dev.cuh"
#include "xetgpu_internal.h"

namespace xetgpu {

// ---- grouping -----------------------------------------------------------------------------
// Cutting a level into groups is a greedy chain (next start = start + k(start), k in 1..9), but the
// jump is bounded, so the chain through a block of 32 nodes is a function of just the offset
// (0..8) at which it enters the block.  Per level:
//   A  one thread per 32-node block tabulates, for each of the 9 possible entry offsets, the exit
//      offset and the number of group starts inside the block                    (parallel)
//   B  one warp per file composes the block tables left to right: a shift and a mask per block
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

// A: table of one 32-node block (slot s of its file's range).
__device__ __forceinline__ void block_map_item(uint64_t s, const uint64_t* __restrict__ slot_off, uint32_t n_files,
                                               const uint8_t* __restrict__ cur_hash, const uint32_t* __restrict__ cur_count,
                                               uint64_t* __restrict__ block_bits, uint64_t* __restrict__ map_out,
                                               uint64_t* __restrict__ map_cnt) {
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

// B: one WARP per file: the block tables of 32 consecutive blocks are loaded coalesced, one per lane, and the chain
// is then advanced through them with register shuffles only (no memory latency on the sequential path); every lane
// runs the same uniform recurrence and keeps the state at which the chain entered its own block.
// Returns the file's node count at the next level (0 for a finished or empty file).
__device__ __forceinline__ uint32_t compose_file(uint32_t f, uint32_t lane, const uint64_t* __restrict__ slot_off,
                                                 const uint8_t* __restrict__ cur_hash, const uint32_t* __restrict__ cur_count,
                                                 const uint64_t* __restrict__ map_out, const uint64_t* __restrict__ map_cnt,
                                                 uint32_t* __restrict__ block_entry, uint32_t* __restrict__ block_base,
                                                 uint32_t* __restrict__ next_count, uint32_t* __restrict__ group_start,
                                                 uint8_t* __restrict__ file_root) {
    const uint32_t n = cur_count[f];
    const uint64_t base = slot_off[f];
    if (n <= 1) {
        if (n == 1 && lane < 2)  // finished: latch the root
            reinterpret_cast<uint4*>(file_root + (uint64_t)f * 32)[lane] =
                reinterpret_cast<const uint4*>(cur_hash + base * 32)[lane];
        if (lane == 0) next_count[f] = 0;
        return 0;
    }
    const uint32_t nb = (n + 31) / 32;
    uint32_t o = 0, total = 0;
    for (uint32_t b0 = 0; b0 < nb; b0 += 32) {
        const uint32_t mine = b0 + lane;
        const uint64_t my_out = mine < nb ? map_out[base + mine] : 0ull;
        const uint64_t my_cnt = mine < nb ? map_cnt[base + mine] : 0ull;
        uint32_t my_entry = 0, my_base = 0;
        const uint32_t m = nb - b0 < 32 ? nb - b0 : 32;
        for (uint32_t k = 0; k < m; k++) {
            const uint64_t mo = __shfl_sync(0xffffffffu, my_out, k);
            const uint64_t mc = __shfl_sync(0xffffffffu, my_cnt, k);
            if (lane == k) {
                my_entry = o;
                my_base = total;
            }
            total += (uint32_t)(mc >> (5 * o)) & 31u;
            o = (uint32_t)(mo >> (4 * o)) & 15u;
        }
        if (mine < nb) {
            block_entry[base + mine] = my_entry;
            block_base[base + mine] = my_base;
        }
    }
    if (lane == 0) {
        group_start[base + total] = n;  // sentinel; total <= (n + 2) / 3 < n for n >= 2
        next_count[f] = total;
    }
    return total;
}

// C: replay one block from its known entry offset and write the group starts.
__device__ __forceinline__ void emit_item(uint64_t s, const uint64_t* __restrict__ slot_off, uint32_t n_files,
                                          const uint32_t* __restrict__ cur_count, const uint64_t* __restrict__ block_bits,
                                          const uint32_t* __restrict__ block_entry, const uint32_t* __restrict__ block_base,
                                          uint32_t* __restrict__ group_start) {
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

// ---- node message formatting ------------------------------------------------------------------
// A node's message is the concatenation of "{hex(hash)} : {size}\n" over its children, hex being
// MerkleHash::hex (four LE u64 words as %016x).  It is produced four characters at a
// time: two hash bytes are spread into four nibbles with two byte permutes, turned into ASCII with
// a SIMD-within-register adjust, and appended through a 64-bit packer that emits whole 32-bit
// message words — so the buffer needs no zero initialisation and no read-modify-write.
struct Packer {
    uint32_t* w;
    uint32_t n;     // words emitted
    uint64_t acc;   // pending bytes, little endian
    uint32_t pend;  // number of pending bytes (0..3)
    __device__ __forceinline__ void put(uint32_t bytes_le, uint32_t k) {  // k = 1..4 valid bytes, rest zero
        acc |= (uint64_t)bytes_le << (8u * pend);
        pend += k;
        if (pend >= 4u) {
            w[n++] = (uint32_t)acc;
            acc >>= 32;
            pend -= 4u;
        }
    }
    __device__ __forceinline__ uint32_t bytes() const { return n * 4u + pend; }
};

// Four hex digits of the 16-bit value t (most significant nibble first) as four ASCII bytes.
__device__ __forceinline__ uint32_t hex4(uint32_t t) {
    const uint32_t z = __byte_perm(t, 0u, 0x4140);                 // [t.b0, 0, t.b1, 0]
    const uint32_t nib = ((z << 4) | z) & 0x0F0F0F0Fu;             // [lo&15, lo>>4, hi&15, hi>>4]
    const uint32_t x = __byte_perm(nib, 0u, 0x0123);               // [hi>>4, hi&15, lo>>4, lo&15]
    return x + 0x30303030u + (((x + 0x06060606u) >> 4) & 0x01010101u) * 0x27u;  // 0-9 -> '0'.., 10-15 -> 'a'..
}

// Formats children [c0, c1) of the level at (cur_hash, cur_size) into msg (>= 208 words); returns
// the message length in bytes, zero-pads the last 64-byte block, and accumulates the sizes.
__device__ __forceinline__ uint32_t format_node(const uint8_t* __restrict__ cur_hash,
                                                const uint64_t* __restrict__ cur_size, uint64_t first_child,
                                                uint32_t n_children, uint32_t* msg, uint64_t* total_out) {
    Packer pk{msg, 0u, 0ull, 0u};
    uint64_t total = 0;
    for (uint32_t c = 0; c < n_children; c++) {
        const uint4* hp = reinterpret_cast<const uint4*>(cur_hash + (first_child + c) * 32);
        const uint4 h0 = hp[0], h1 = hp[1];
        const uint32_t hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
#pragma unroll
        for (int wd = 0; wd < 4; wd++) {  // LE u64 word = (hw[2wd+1] << 32) | hw[2wd], printed high byte first
            pk.put(hex4(hw[2 * wd + 1] >> 16), 4);
            pk.put(hex4(hw[2 * wd + 1] & 0xFFFFu), 4);
            pk.put(hex4(hw[2 * wd] >> 16), 4);
            pk.put(hex4(hw[2 * wd] & 0xFFFFu), 4);
        }
        pk.put(0x203A20u, 3);  // " : "
        const uint64_t sz = cur_size[first_child + c];
        total += sz;
        // decimal, most significant digit first; 32-bit arithmetic whenever the size allows
        char dec[20];
        int nd = 0;
        if (sz >> 32) {
            uint64_t t = sz;
            do {
                dec[nd++] = (char)('0' + (uint32_t)(t % 10u));
                t /= 10u;
            } while (t);
        } else {
            uint32_t t = (uint32_t)sz;
            do {
                dec[nd++] = (char)('0' + t % 10u);
                t /= 10u;
            } while (t);
        }
        while (nd) pk.put((uint32_t)dec[--nd], 1);
        pk.put('\n', 1);
    }
    const uint32_t len = pk.bytes();
    if (pk.pend) msg[pk.n++] = (uint32_t)pk.acc;
    while (pk.n & 15u) msg[pk.n++] = 0u;  // zero padding of the final block
    *total_out = total;
    return len;
}

// Single-leaf keyed BLAKE3 (INTERNAL_NODE_KEY) of a formatted message.
__device__ __forceinline__ void hash_node_message(const uint32_t* msg, uint32_t len, uint32_t (&cv)[8]) {
    const uint32_t key[8] = XG_NODE_KEY_WORDS;
#pragma unroll
    for (int i = 0; i < 8; i++) cv[i] = key[i];
    const uint32_t nb = len <= 64 ? 1u : (len + 63u) / 64u;
    for (uint32_t b = 0; b < nb; b++) {
        uint32_t m[16];
#pragma unroll
        for (int i = 0; i < 16; i++) m[i] = msg[b * 16 + i];
        uint32_t flags = B3_KEYED_HASH;
        if (b == 0) flags |= B3_CHUNK_START;
        uint32_t bl = 64;
        if (b == nb - 1) {
            flags |= B3_CHUNK_END | B3_ROOT;
            bl = len - b * 64;
        }
        b3_compress(cv, m, 0u, 0u, bl, flags);
    }
}

// D: one new node: format its children's lines and hash them.
__device__ __forceinline__ void hash_node_item(const uint8_t* __restrict__ cur_hash, const uint64_t* __restrict__ cur_size,
                                               uint64_t first_child, uint32_t n_children, uint8_t* __restrict__ dst_hash,
                                               uint64_t* __restrict__ dst_size) {
    uint32_t msg[224];
    uint64_t total;
    const uint32_t len = format_node(cur_hash, cur_size, first_child, n_children, msg, &total);
    uint32_t cv[8];
    hash_node_message(msg, len, cv);
    uint4* dst = reinterpret_cast<uint4*>(dst_hash);
    dst[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
    dst[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
    *dst_size = total;
}

// ---- the whole stage in ONE cooperative launch ----------------------------------------------------------
// K4 is latency-bound (40 bytes of input per ~64 KiB chunk), so what it costs is launches and the gaps between them:
// as separate kernels it took four launches per level plus three more, and the number of level rounds had to come
// from a host-side upper bound (chunk slots, ~8x the real chunk count), i.e. two or three empty rounds per call.
// Here one cooperative grid runs every phase with grid-wide barriers in between and decides ON THE DEVICE when the
// level rounds are over:
//   0   level-0 sizes from the chunk lengths, level counts from the per-file chunk counts
//   repeat while some file holds more than kFinishNodes nodes:   A block tables | B compose | C emit | D hash
//   then one CTA per file (CTAs stride over the files) finishes its tree in shared memory and hashes the root
//   under the zero salt (file hash).
constexpr uint32_t kFinishNodes = 8192;
constexpr uint32_t kFinishBlocks = kFinishNodes / 32;   // 32-node blocks per finishing file
constexpr int kMerkleThreads = 512;

struct MerkleArgs {
    const uint64_t* slot_off;
    uint32_t n_files;
    uint64_t n_slots;
    const uint32_t* chunk_len;     // nullptr: level-0 sizes are already in node_size[0]
    const uint32_t* file_count;
    uint8_t* node_hash[2];
    uint64_t* node_size[2];
    uint32_t* level_count[2];
    uint32_t* group_start;
    uint64_t* mk_bits;
    uint64_t* mk_out;
    uint64_t* mk_cnt;
    uint32_t* mk_entry;
    uint32_t* mk_base;
    uint8_t* file_root;
    uint8_t* file_hash;
    uint32_t* round_max;           // [2] device words: largest next-level count seen in a round
};

__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t& epoch, uint32_t n_blocks);

__global__ void __launch_bounds__(kMerkleThreads) merkle_coop_kernel(MerkleArgs a, uint32_t* barrier_word) {
    __shared__ uint64_t s_bits[kFinishBlocks], s_out[kFinishBlocks], s_cnt[kFinishBlocks];
    __shared__ uint32_t s_entry[kFinishBlocks], s_base[kFinishBlocks];
    __shared__ uint32_t s_start[kFinishNodes / 3 + 4];
    __shared__ uint32_t s_ng;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint64_t gtid = (uint64_t)blockIdx.x * blockDim.x + tid;
    const uint64_t gthreads = (uint64_t)gridDim.x * blockDim.x;
    const uint32_t gwarp = (uint32_t)(gtid >> 5), gwarps = (uint32_t)(gthreads >> 5);
    uint32_t epoch = 0;
#ifdef XG_MERKLE_TRACE
    unsigned long long tr[40];
    int ntr = 0;
#define XG_TR() do { if (gtid == 0 && ntr < 40) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tr[ntr++] = t_; } } while (0)
#else
#define XG_TR() do {} while (0)
#endif
    XG_TR();

    // Slot loops hand consecutive groups of 32 slots to consecutive CTAs: the live slots of a level are the first
    // few of every file's range, and a plain grid-stride loop would give all of them to a handful of CTAs.
    const uint32_t wpb = blockDim.x >> 5, lw = tid >> 5;
#define XG_FOR_SLOTS(s)                                                                                          \
    for (uint64_t _sw = (uint64_t)lw * gridDim.x + blockIdx.x, s = _sw * 32 + lane; _sw * 32 < a.n_slots;        \
         _sw += (uint64_t)wpb * gridDim.x, s = _sw * 32 + lane)                                                  \
        if (s < a.n_slots)

    // phase 0
    if (a.chunk_len)
        for (uint64_t s = gtid; s < a.n_slots; s += gthreads) a.node_size[0][s] = a.chunk_len[s];
    uint32_t local_max = 0;
    for (uint64_t f = gtid; f < a.n_files; f += gthreads) {
        const uint32_t c = a.file_count[f];
        a.level_count[0][f] = c;
        local_max = max(local_max, c);
    }
    if (gtid == 0) a.round_max[1] = 0;
    if (local_max) atomicMax(&a.round_max[0], local_max);   // round_max[0] was zeroed by the host before the launch
    grid_barrier(barrier_word, epoch, gridDim.x);
    XG_TR();

    int cur = 0;
    uint32_t r = 0;   // round_max[r & 1] holds the largest count of the level the grid is about to process
    while (*(volatile uint32_t*)&a.round_max[r & 1] > kFinishNodes) {
        const uint8_t* cur_hash = a.node_hash[cur];
        const uint64_t* cur_size = a.node_size[cur];
        const uint32_t* cur_count = a.level_count[cur];
        uint32_t* next_count = a.level_count[cur ^ 1];
        XG_FOR_SLOTS(s) block_map_item(s, a.slot_off, a.n_files, cur_hash, cur_count, a.mk_bits, a.mk_out, a.mk_cnt);
        XG_TR();
        grid_barrier(barrier_word, epoch, gridDim.x);
        XG_TR();
        uint32_t mx = 0;
        for (uint32_t f = gwarp; f < a.n_files; f += gwarps)
            mx = max(mx, compose_file(f, lane, a.slot_off, cur_hash, cur_count, a.mk_out, a.mk_cnt, a.mk_entry, a.mk_base,
                                      next_count, a.group_start, a.file_root));
        if (lane == 0 && mx) atomicMax(&a.round_max[(r + 1) & 1], mx);
        XG_TR();
        grid_barrier(barrier_word, epoch, gridDim.x);
        XG_TR();
        XG_FOR_SLOTS(s) emit_item(s, a.slot_off, a.n_files, cur_count, a.mk_bits, a.mk_entry, a.mk_base, a.group_start);
        XG_TR();
        grid_barrier(barrier_word, epoch, gridDim.x);
        XG_TR();
        XG_FOR_SLOTS(s) {
            const uint32_t f = file_of_slot_g(a.slot_off, a.n_files, s);
            const uint64_t base = a.slot_off[f];
            const uint32_t gi = (uint32_t)(s - base);
            if (gi >= next_count[f]) continue;
            const uint32_t c0 = a.group_start[base + gi], c1 = a.group_start[base + gi + 1];
            hash_node_item(cur_hash, cur_size, base + c0, c1 - c0, a.node_hash[cur ^ 1] + s * 32, a.node_size[cur ^ 1] + s);
        }
        if (gtid == 0) a.round_max[r & 1] = 0;   // read by everyone before the barrier after phase A; free for round r + 2
        XG_TR();
        grid_barrier(barrier_word, epoch, gridDim.x);
        XG_TR();
        cur ^= 1;
        r++;
    }

#undef XG_FOR_SLOTS
    // finish: one CTA per file, all remaining levels in shared memory
    for (uint32_t f = blockIdx.x; f < a.n_files; f += gridDim.x) {
        uint32_t n = a.level_count[cur][f];
        const uint64_t base = a.slot_off[f];
        uint8_t* cur_hash = a.node_hash[cur] + base * 32;
        uint64_t* cur_size = a.node_size[cur] + base;
        uint8_t* nxt_hash = a.node_hash[cur ^ 1] + base * 32;
        uint64_t* nxt_size = a.node_size[cur ^ 1] + base;
        const bool run = n >= 1 && n <= kFinishNodes;   // 0: empty file, or root latched during the level rounds
        while (run && n > 1) {
            const uint32_t nb = (n + 31) / 32;
            if (tid < nb) {  // A: block table
                const uint32_t first = tid * 32;
                uint64_t bits = 0;
                for (uint32_t i = 0; i < 40 && first + i < n; i++)
                    bits |= (uint64_t)((cur_hash[(uint64_t)(first + i) * 32 + 24] & 3u) == 0u) << i;
                uint64_t mo = 0, mc = 0;
                for (uint32_t o = 0; o < 9; o++) {
                    uint32_t q = o, cnt = 0;
                    while (q < 32 && first + q < n) {
                        q += group_len(bits, q, n - (first + q));
                        cnt++;
                    }
                    mo |= (uint64_t)(q >= 32 ? q - 32 : 0) << (4 * o);
                    mc |= (uint64_t)cnt << (5 * o);
                }
                s_bits[tid] = bits;
                s_out[tid] = mo;
                s_cnt[tid] = mc;
            }
            __syncthreads();
            if (tid == 0) {  // B: compose
                uint32_t o = 0, total = 0;
                for (uint32_t b = 0; b < nb; b++) {
                    s_entry[b] = o;
                    s_base[b] = total;
                    total += (uint32_t)(s_cnt[b] >> (5 * o)) & 31u;
                    o = (uint32_t)(s_out[b] >> (4 * o)) & 15u;
                }
                s_start[total] = n;
                s_ng = total;
            }
            __syncthreads();
            if (tid < nb) {  // C: emit group starts
                const uint32_t first = tid * 32;
                const uint64_t bits = s_bits[tid];
                uint32_t q = s_entry[tid], w = s_base[tid];
                while (q < 32 && first + q < n) {
                    s_start[w++] = first + q;
                    q += group_len(bits, q, n - (first + q));
                }
            }
            __syncthreads();
            const uint32_t ng = s_ng;
            for (uint32_t g = tid; g < ng; g += blockDim.x)  // D: hash the new nodes
                hash_node_item(cur_hash, cur_size, s_start[g], s_start[g + 1] - s_start[g], nxt_hash + (uint64_t)g * 32,
                               nxt_size + g);
            __syncthreads();
            uint8_t* th = cur_hash;
            cur_hash = nxt_hash;
            nxt_hash = th;
            uint64_t* ts = cur_size;
            cur_size = nxt_size;
            nxt_size = ts;
            n = ng;
        }
        // root + file hash (keyed BLAKE3 of the 32-byte root under the all-zero salt; empty file -> zero hash)
        if (tid == 0) {
            uint4* root = reinterpret_cast<uint4*>(a.file_root + (uint64_t)f * 32);
            uint4* out = reinterpret_cast<uint4*>(a.file_hash + (uint64_t)f * 32);
            if (a.file_count[f] == 0) {
                root[0] = root[1] = out[0] = out[1] = make_uint4(0, 0, 0, 0);
            } else {
                if (run) {
                    root[0] = reinterpret_cast<const uint4*>(cur_hash)[0];
                    root[1] = reinterpret_cast<const uint4*>(cur_hash)[1];
                }
                const uint4 r0 = root[0], r1 = root[1];
                uint32_t m[16] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, 0, 0, 0, 0, 0, 0, 0, 0};
                uint32_t cv[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                b3_compress(cv, m, 0u, 0u, 32u, B3_KEYED_HASH | B3_CHUNK_START | B3_CHUNK_END | B3_ROOT);
                out[0] = make_uint4(cv[0], cv[1], cv[2], cv[3]);
                out[1] = make_uint4(cv[4], cv[5], cv[6], cv[7]);
            }
        }
        __syncthreads();
        XG_TR();
    }
#ifdef XG_MERKLE_TRACE
    if (gtid == 0) {
        printf("merkle trace (ns since start):");
        for (int i = 1; i < ntr; i++) printf(" %llu", tr[i] - tr[0]);
        printf("\n");
    }
#endif
}

// Grid-wide barrier for a cooperative launch (every CTA is resident, which cudaLaunchCooperativeKernel guarantees):
// a monotonically growing arrival counter; CTA-local epochs keep the word reusable without a reset.
__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t& epoch, uint32_t n_blocks) {
    __syncthreads();
    epoch++;
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
        const uint32_t target = epoch * n_blocks;
        while (*(volatile uint32_t*)counter < target) __nanosleep(32);
        __threadfence();
    }
    __syncthreads();
}

static int g_merkle_blocks_per_sm[64] = {};

int launch_merkle(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_files == 0) return 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && g_merkle_blocks_per_sm[dev] == 0) {
        int nb = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, merkle_coop_kernel, kMerkleThreads, 0);
        g_merkle_blocks_per_sm[dev] = nb < 1 ? 1 : (nb > 2 ? 2 : nb);
    }
    const unsigned per_sm = dev >= 0 && dev < 64 ? g_merkle_blocks_per_sm[dev] : 1;
    MerkleArgs a{};
    a.slot_off = b.slot_off;
    a.n_files = b.n_files;
    a.n_slots = b.n_slots;
    a.chunk_len = b.chunk_len;
    a.file_count = b.file_count;
    for (int i = 0; i < 2; i++) {
        a.node_hash[i] = b.node_hash[i];
        a.node_size[i] = b.node_size[i];
        a.level_count[i] = b.level_count[i];
    }
    a.group_start = b.group_start;
    a.mk_bits = b.mk_bits;
    a.mk_out = b.mk_out;
    a.mk_cnt = b.mk_cnt;
    a.mk_entry = b.mk_entry;
    a.mk_base = b.mk_base;
    a.file_root = b.file_root;
    a.file_hash = b.file_hash;
    // three device words of the per-call scalar block: [0,1] round maxima, [2] barrier counter
    uint32_t* words = reinterpret_cast<uint32_t*>(b.scalars + kScalarMerkle);
    cudaMemsetAsync(words, 0, 16, st);
    a.round_max = words;
    uint32_t* barrier_word = words + 2;
    unsigned sms = b.sm_count;
    if (!sms) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        sms = v > 0 ? (unsigned)v : 1u;
    }
    unsigned grid = sms * per_sm;
    // a grid larger than the work only adds barrier arrivals
    const uint64_t want = std::max<uint64_t>((b.n_slots + kMerkleThreads - 1) / kMerkleThreads, b.n_files);
    if (grid > want) grid = (unsigned)std::max<uint64_t>(want, 1);
    void* args[] = {&a, &barrier_word};
    launch_cooperative_gated((const void*)merkle_coop_kernel, dim3(grid), dim3(kMerkleThreads), args, 0, st);
    return 1;
}

// ---- xorb formation -------------------------------------------------------------
// xet-core's upload side packs a file's chunks, in order, into xorbs of at most 64 MiB of chunk bytes and at most
// 8192 chunks; a xorb's hash is merklehash::xorb_hash = aggregated_node_hash over its (chunk hash, length) list —
// the very function K4 computes per file.  So: one thread per file walks the greedy packing over the dense chunk
// ends (a binary search per xorb, ~N / 64 MiB steps), then the K4 finish kernel runs with "file" = xorb (every
// xorb holds <= 8192 nodes, which is what one finish CTA handles).  The grouping policy is parity-unpinned: the
// reference binary (hf_xet.hash_files) does not expose xorbs; the tests compare with an independent CPU twin.
__global__ void xorb_group_kernel(uint32_t n_files, const uint64_t* __restrict__ file_chunk_base,
                                  const uint32_t* __restrict__ file_count, const uint64_t* __restrict__ out_end,
                                  const uint64_t* __restrict__ xorb_slot_off, uint32_t* __restrict__ xorb_count,
                                  uint64_t* __restrict__ xorb_node_off, uint32_t* __restrict__ xorb_n,
                                  uint64_t* __restrict__ xorb_bytes) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_files) return;
    const uint64_t base = file_chunk_base[f];
    const uint32_t n = file_count[f];
    const uint64_t s0 = xorb_slot_off[f], cap = xorb_slot_off[f + 1] - s0;
    const uint64_t* ends = out_end + base;   // file-relative exclusive ends
    uint32_t i = 0;
    uint64_t x = 0;
    while (i < n) {
        const uint64_t start = i ? ends[i - 1] : 0;
        // largest j in (i, min(n, i + kMaxXorbChunks)] with ends[j-1] - start <= kMaxXorbBytes; j = i + 1 always fits
        uint32_t lo = i + 1, hi = n - i > kMaxXorbChunks ? i + kMaxXorbChunks : n;
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo + 1) / 2;
            if (ends[mid - 1] - start <= kMaxXorbBytes) lo = mid;
            else hi = mid - 1;
        }
        if (x < cap) {
            xorb_node_off[s0 + x] = base + i;
            xorb_n[s0 + x] = lo - i;
            xorb_bytes[s0 + x] = ends[lo - 1] - start;
        }
        x++;
        i = lo;
    }
    xorb_count[f] = (uint32_t)(x < cap ? x : cap);
    for (uint64_t k = x; k < cap; k++) {  // unused slots: empty "files" for the merkle pass
        xorb_node_off[s0 + k] = 0;
        xorb_n[s0 + k] = 0;
        xorb_bytes[s0 + k] = 0;
    }
}

// Level 0 of the xorb trees: dense chunk hashes and lengths (file-relative ends -> lengths).
__global__ void xorb_prepare_kernel(uint32_t n_files, const uint64_t* __restrict__ scalars,
                                    const uint64_t* __restrict__ file_chunk_base, const uint64_t* __restrict__ out_end,
                                    const uint8_t* __restrict__ out_hash, uint8_t* __restrict__ node_hash,
                                    uint64_t* __restrict__ node_size) {
    const uint64_t d = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= scalars[kScalarChunks]) return;
    uint32_t lo = 0, hi = n_files;  // last f with file_chunk_base[f] <= d
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (file_chunk_base[mid] <= d) lo = mid;
        else hi = mid;
    }
    // (files without chunks share their base with the next file: the search lands on the LAST of them, which is
    //  the one that owns d or an empty one before it — either way d > base means a predecessor in the same file)
    while (lo + 1 < n_files && file_chunk_base[lo + 1] <= d) lo++;
    node_size[d] = out_end[d] - (d == file_chunk_base[lo] ? 0 : out_end[d - 1]);
    const uint4* src = reinterpret_cast<const uint4*>(out_hash + d * 32);
    uint4* dst = reinterpret_cast<uint4*>(node_hash + d * 32);
    dst[0] = src[0];
    dst[1] = src[1];
}

int launch_xorbs(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_files == 0 || b.n_xorb_slots == 0) return 0;
    xorb_group_kernel<<<(b.n_files + 127) / 128, 128, 0, st>>>(b.n_files, b.file_chunk_base, b.file_count, b.out_end,
                                                               b.xorb_slot_off, b.xorb_count, b.xorb_node_off, b.xorb_n,
                                                               b.xorb_bytes);
    int n = 1;
    if (b.n_slots) {
        xorb_prepare_kernel<<<(unsigned)((b.n_slots + 255) / 256), 256, 0, st>>>(
            b.n_files, b.scalars, b.file_chunk_base, b.out_end, b.out_hash, b.node_hash[0], b.node_size[0]);
        n++;
    }
    // the K4 machinery with "file" = xorb slot: node ranges start at xorb_node_off, every tree fits one finish CTA
    PipelineBuffers m = b;
    m.n_files = (uint32_t)b.n_xorb_slots;
    m.slot_off = b.xorb_node_off;
    m.file_count = b.xorb_n;
    m.level_count[0] = b.xorb_level[0];
    m.level_count[1] = b.xorb_level[1];
    m.file_root = b.xorb_root;
    m.file_hash = b.xorb_scratch;
    m.max_nodes = kMaxXorbChunks;
    m.chunk_len = nullptr;  // level-0 sizes are already in node_size[0]
    return n + launch_merkle(m, st);
}

}  // namespace xetgpu
