// K3: keyed-BLAKE3 chunk hashes (merklehash::compute_data_hash = blake3::keyed_hash(DATA_KEY, chunk);
// xet-core merklehash 0.14.5 pinned at /root/reference pkg/xet/Cargo.lock:1651-1663; the reference's
// only direct call site is pkg/xet/src/xet_downloader.rs:217).
//
// Work decomposition.  BLAKE3 splits a chunk into 1 KiB leaves whose chaining values are merged
// by a left-heavy binary tree.  Leaf positions are relative to the (data-dependent) chunk start,
// so the work list is built after K2: a *group* is up to four consecutive leaves of one chunk,
// aligned at a multiple of four leaves from the chunk start, i.e. a complete subtree.  One lane
// hashes one group: 16 serial compressions per leaf with the whole state in registers, then the
// 2-level subtree merge in registers.  All 32 lanes of a warp do identical work except at chunk
// tails.  The few remaining tree levels (<= 32 group values per chunk) are folded by K3b.
//
// Data movement.  A lane's leaf is 1 KiB at an arbitrary byte address.  It is staged half a leaf
// (512 bytes, 8 blocks) at a time: every lane issues ONE bulk TMA copy (cp.async.bulk -> UBLKCP)
// of the 16-byte-aligned span that covers its half leaf into its own 528-byte row of the warp's
// shared-memory stage; completion is signalled on a per-warp mbarrier.  No LSU instruction is
// spent on staging and HBM lines are read once.  The small stage (16.5 KiB per warp) is what lets
// 12 warps share an SM — the compression chains are latency-bound below ~3 warps per scheduler.
// Message words are read with LDS.32 from the row (row stride 528 B = 132 words: lanes l, l+8,
// l+16, l+24 share a bank — a 4-way conflict that costs ~68 of the ~1000 issue cycles of a
// compression) and realigned with one funnel shift per word for the chunk's byte offset.
#include <algorithm>

#include "blake3_dev.cuh"
#include "tma.cuh"
#include "xetgpu_internal.h"

namespace xetgpu {

int launch_chunk_bookkeeping(const PipelineBuffers& b, cudaStream_t st);  // resolve.cu


// The subtree merges are 3 compressions per 64 leaf compressions; keeping them out of line keeps one
// copy of the ~13 KB compression body in the leaf kernel's instruction stream.
__device__ __noinline__ void parent_out_of_line(const uint32_t* l, const uint32_t* r, uint32_t root, uint32_t* out) {
    const uint32_t key[8] = XG_DATA_KEY_WORDS;
    uint32_t lv[8], rv[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        lv[i] = l[i];
        rv[i] = r[i];
    }
    b3_parent(key, lv, rv, root != 0u, o);
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = o[i];
}

constexpr int kLeafThreads = kLeafWarpsPerCta * 32;
constexpr int kLeafStage = 32 * kLeafRow;                       // bytes of shared memory per warp
constexpr int kLeafSmem = kLeafWarpsPerCta * kLeafStage;

// Compress the (up to 8) 64-byte blocks of one staged half leaf into cv.
//   row      : the lane's shared-memory row (16-byte aligned copy of the aligned span)
//   off      : byte offset of the first byte of this half inside the row (0..15)
//   b0       : index of the half's first block inside the leaf (0 or 8)
//   nb       : blocks in the whole leaf;  leaf_len : bytes in the whole leaf
__device__ __forceinline__ void hash_half_leaf(const uint32_t* __restrict__ row, uint32_t off, uint32_t b0, uint32_t nb,
                                               uint32_t leaf_len, uint32_t leaf_index, bool active, bool root_leaf,
                                               uint32_t (&cv)[8]) {
    const uint32_t shift = (off & 3u) * 8u;
    const uint32_t* wrow = row + (off >> 2);
#pragma unroll 1
    for (uint32_t k = 0; k < kLeafHalfBlocks; k++) {
        const uint32_t b = b0 + k;
        const bool go = active && b < nb;
        if (!__any_sync(0xffffffffu, go)) break;
        if (go) {
            uint32_t w[17];
#pragma unroll
            for (int i = 0; i < 17; i++) w[i] = wrow[k * 16 + i];
            uint32_t m[16];
#pragma unroll
            for (int i = 0; i < 16; i++) m[i] = __funnelshift_r(w[i], w[i + 1], shift);
            uint32_t bl = 64;
            if (b == nb - 1) {
                bl = leaf_len - b * 64;  // 1..64 (a leaf is never empty: K2 emits no empty chunk)
                if (bl < 64) {
#pragma unroll
                    for (int i = 0; i < 16; i++) {
                        const int lo = 4 * i;
                        if (bl <= (uint32_t)lo) m[i] = 0u;
                        else if (bl < (uint32_t)lo + 4u) m[i] &= (1u << (8u * (bl - lo))) - 1u;
                    }
                }
            }
            uint32_t flags = B3_KEYED_HASH;
            if (b == 0) flags |= B3_CHUNK_START;
            if (b == nb - 1) flags |= B3_CHUNK_END | (root_leaf ? B3_ROOT : 0u);
            b3_compress(cv, m, leaf_index, 0u, bl, flags);
        }
    }
}

__global__ void __launch_bounds__(kLeafThreads)
    leaf_group_kernel(const uint8_t* __restrict__ data, const uint64_t* __restrict__ scalars,
                      const uint32_t* __restrict__ group_chunk, const uint64_t* __restrict__ group_off,
                      const uint64_t* __restrict__ chunk_start, const uint32_t* __restrict__ chunk_len,
                      uint32_t* __restrict__ group_cv, uint8_t* __restrict__ chunk_hash) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t s_bar[kLeafWarpsPerCta];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint64_t n_groups = scalars[kScalarGroups];
    const uint64_t g = ((uint64_t)blockIdx.x * kLeafWarpsPerCta + wid) * 32u + lane;
    if (g - lane >= n_groups) return;  // whole warp out of range (warps never sync with each other)

    uint64_t* bar = &s_bar[wid];
    if (lane == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncwarp();
    uint8_t* stage = smem + wid * kLeafStage;
    uint8_t* my_row = stage + lane * kLeafRow;

    const bool valid = g < n_groups;
    uint32_t slot = 0, clen = 0, goff = 0, glen = 0, gleaves = 0, cleaves = 0;
    uint64_t cstart = 0;
    if (valid) {
        slot = group_chunk[g];
        const uint32_t j = (uint32_t)(g - group_off[slot]);
        cstart = chunk_start[slot];
        clen = chunk_len[slot];
        cleaves = clen <= 1024u ? 1u : (clen + 1023u) / 1024u;
        goff = j * (kLeavesPerGroup * 1024u);
        glen = clen - goff < kLeavesPerGroup * 1024u ? clen - goff : kLeavesPerGroup * 1024u;
        gleaves = glen <= 1024u ? 1u : (glen + 1023u) / 1024u;
    }
    const uint32_t key[8] = XG_DATA_KEY_WORDS;
    uint32_t a[8], bq[8], cv[8];
    uint32_t parity = 0;

#pragma unroll 1
    for (uint32_t li = 0; li < kLeavesPerGroup; li++) {
        const bool active = valid && li < gleaves;
        if (!__any_sync(0xffffffffu, active)) break;
        const uint32_t llen = active ? (glen - li * 1024u < 1024u ? glen - li * 1024u : 1024u) : 0u;
        const uint32_t nb = llen <= 64u ? 1u : (llen + 63u) / 64u;
#pragma unroll
        for (int i = 0; i < 8; i++) cv[i] = key[i];
#pragma unroll 1
        for (uint32_t half = 0; half < (1024u + kLeafHalfBytes - 1u) / kLeafHalfBytes; half++) {
            // stage this half leaf of all 32 lanes: one bulk copy per lane, one mbarrier phase
            const uint32_t hbeg = half * kLeafHalfBytes;
            const bool hact = active && (hbeg < llen || half == 0);
            if (!__any_sync(0xffffffffu, hact)) break;
            const uint32_t hlen = hact ? (llen - hbeg < kLeafHalfBytes ? llen - hbeg : kLeafHalfBytes) : 0u;
            const uint64_t src = cstart + goff + li * 1024u + hbeg;
            const uint32_t off = (uint32_t)(src & 15u);
            const uint32_t copy = hact ? ((off + hlen + 15u) & ~15u) : 0u;
            const uint32_t tx = __reduce_add_sync(0xffffffffu, copy);
            if (lane == 0) mbar_arrive_expect_tx(bar, tx);
            __syncwarp();
            if (copy) tma_bulk_g2s(my_row, data + (src - off), copy, bar);
            mbar_wait(bar, parity);
            parity ^= 1u;
            hash_half_leaf(reinterpret_cast<const uint32_t*>(my_row), off, half * kLeafHalfBlocks, nb, llen,
                           goff / 1024u + li, hact, /*root_leaf=*/cleaves == 1u, cv);
            __syncwarp();  // all lanes are done reading the stage before the next copy overwrites it
        }

        if (active) {
            if (li == 0) {
#pragma unroll
                for (int i = 0; i < 8; i++) a[i] = cv[i];
            } else if (li == 1) {
                parent_out_of_line(a, cv, cleaves == 2u, a);
            } else if (li == 2) {
#pragma unroll
                for (int i = 0; i < 8; i++) bq[i] = cv[i];
            } else {
                parent_out_of_line(bq, cv, 0u, bq);
            }
        }
    }
    if (!valid) return;
    if (gleaves >= 3) {
        parent_out_of_line(a, bq, cleaves <= (uint32_t)kLeavesPerGroup, a);
    }
    uint4* dst = cleaves <= (uint32_t)kLeavesPerGroup ? reinterpret_cast<uint4*>(chunk_hash + (uint64_t)slot * 32)
                                                       : reinterpret_cast<uint4*>(group_cv + g * 8);
    dst[0] = make_uint4(a[0], a[1], a[2], a[3]);
    dst[1] = make_uint4(a[4], a[5], a[6], a[7]);
}

// K3b: fold a chunk's group values (complete 4-leaf subtrees, the last possibly partial) into the chunk hash.
// Pairing neighbours level by level and promoting an odd tail unchanged reproduces BLAKE3's left-heavy tree.
// A chunk holds at most 32 groups (128 KiB / 4 KiB).  Eight lanes share a chunk: a lane loads four consecutive
// group values (128 contiguous bytes) and folds them in registers (two levels), then three more levels run across
// the eight lanes with warp shuffles — five dependent compressions per chunk instead of up to 31, and four chunks
// per warp, so even a 1 GiB file (16 k chunks) fills the SMs.
constexpr int kTreeLanes = 8;           // lanes per chunk
constexpr int kTreeThreads = 128;

__device__ __noinline__ void tree_parent(uint32_t (&l)[8], const uint32_t (&r)[8], bool root) {
    const uint32_t key[8] = XG_DATA_KEY_WORDS;
    uint32_t o[8];
    b3_parent(key, l, r, root, o);
#pragma unroll
    for (int i = 0; i < 8; i++) l[i] = o[i];
}

__device__ __forceinline__ void chunk_tree_item(uint64_t d, uint64_t n_chunks, uint32_t lane, uint32_t sub, uint32_t n_files,
                                                const uint64_t* __restrict__ file_chunk_base,
                                                const uint64_t* __restrict__ slot_off, const uint32_t* __restrict__ slot_groups,
                                                const uint64_t* __restrict__ group_off, const uint32_t* __restrict__ group_cv,
                                                uint8_t* __restrict__ chunk_hash) {
    uint64_t s = 0;
    uint32_t ng = 0;
    if (d < n_chunks) {
        uint32_t lo = 0, hi = n_files;   // last f with file_chunk_base[f] <= d (owner of chunk d)
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (file_chunk_base[mid] <= d) lo = mid;
            else hi = mid;
        }
        s = slot_off[lo] + (d - file_chunk_base[lo]);
        ng = slot_groups[s];             // <= 1: finalised by the leaf kernel
    }
    const bool live = ng > 1;
    if (!__any_sync(0xffffffffu, live)) return;
    // this lane's items: groups [4 sub, 4 sub + cnt)
    const uint32_t cnt = live && 4u * sub < ng ? min(4u, ng - 4u * sub) : 0u;
    uint32_t a0[8], a1[8], a2[8], a3[8];
    const uint4* src = reinterpret_cast<const uint4*>(group_cv + (live ? (group_off[s] + 4u * sub) * 8 : 0));
#define XG_LOAD(dst, k)                                                     \
    if (cnt > (k)) {                                                        \
        const uint4 x = src[2 * (k)], y = src[2 * (k) + 1];                 \
        dst[0] = x.x; dst[1] = x.y; dst[2] = x.z; dst[3] = x.w;             \
        dst[4] = y.x; dst[5] = y.y; dst[6] = y.z; dst[7] = y.w;             \
    }
    XG_LOAD(a0, 0) XG_LOAD(a1, 1) XG_LOAD(a2, 2) XG_LOAD(a3, 3)
#undef XG_LOAD
    // level 1 (inside the lane): (a0,a1) and (a2,a3); the merge is the root one iff it leaves a single item
    if (__any_sync(0xffffffffu, cnt >= 2)) {
        if (cnt >= 2) tree_parent(a0, a1, ng == 2);
    }
    if (__any_sync(0xffffffffu, cnt == 4)) {
        if (cnt == 4) tree_parent(a2, a3, false);
    }
    // level 2: (a0', a2') — a2 is promoted unchanged when the lane holds three items
    if (__any_sync(0xffffffffu, cnt >= 3)) {
        if (cnt >= 3) tree_parent(a0, a2, ng <= 4);
    }
    // levels 3..5 across the lanes of the chunk: lane i takes lane i + d when that lane holds a value
    const uint32_t nl = (ng + 3u) / 4u;   // lanes of this chunk that hold a value (0 if not live)
#pragma unroll 1
    for (uint32_t d = 1; d < kTreeLanes; d <<= 1) {
        uint32_t r[8];
#pragma unroll
        for (int i = 0; i < 8; i++) r[i] = __shfl_down_sync(0xffffffffu, a0[i], d, kTreeLanes);
        const bool merge = live && (sub & (2 * d - 1)) == 0 && sub + d < nl;
        if (__any_sync(0xffffffffu, merge)) {
            if (merge) tree_parent(a0, r, nl <= 2 * d);
        }
    }
    if (live && sub == 0) {
        uint4* dst = reinterpret_cast<uint4*>(chunk_hash + s * 32);
        dst[0] = make_uint4(a0[0], a0[1], a0[2], a0[3]);
        dst[1] = make_uint4(a0[4], a0[5], a0[6], a0[7]);
    }
}

// Chunks are addressed densely (d-th chunk of the batch, through the scan of the per-file chunk counts): chunk SLOTS
// are per-file upper bounds, about eight times the real count on ordinary data, and the grid is sized on the host.
__global__ void __launch_bounds__(kTreeThreads)
    chunk_tree_kernel(const uint64_t* __restrict__ scalars, uint32_t n_files, const uint64_t* __restrict__ file_chunk_base,
                      const uint64_t* __restrict__ slot_off, const uint32_t* __restrict__ slot_groups,
                      const uint64_t* __restrict__ group_off, const uint32_t* __restrict__ group_cv,
                      uint8_t* __restrict__ chunk_hash) {
    const uint32_t lane = threadIdx.x & 31u, sub = lane & (kTreeLanes - 1);
    const uint64_t n_chunks = scalars[kScalarChunks];
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x / kTreeLanes;
    // (the loop bound is rounded up to a whole warp's worth of chunks so that all 32 lanes stay together)
    for (uint64_t d0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x - lane) / kTreeLanes; d0 < n_chunks; d0 += stride)
        chunk_tree_item(d0 + lane / kTreeLanes, n_chunks, lane, sub, n_files, file_chunk_base, slot_off, slot_groups, group_off,
                        group_cv, chunk_hash);
}

void init_chunk_hash_kernels() {
    cudaFuncSetAttribute(leaf_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kLeafSmem);
}

// The stage in three parts, so that a caller can place them on different streams: `stages` is a bit mask of
// 1 = bookkeeping (work lists), 2 = the leaf kernel, 4 = dense chunk numbering + chunk tree.
int launch_chunk_hash_stages(const PipelineBuffers& b, cudaStream_t st, int stages) {
    int n = 0;
    if (stages & 1) n += launch_chunk_bookkeeping(b, st);
    if (b.n_slots == 0 || b.n_groups_cap == 0) return n;
    if (stages & 2) {
        const uint64_t warps = (b.n_groups_cap + 31) / 32;
        const unsigned blocks = (unsigned)((warps + kLeafWarpsPerCta - 1) / kLeafWarpsPerCta);
        if (b.ev_kernel[2]) cudaEventRecord(b.ev_kernel[2], st);
        leaf_group_kernel<<<blocks, kLeafThreads, kLeafSmem, st>>>(b.data, b.scalars, b.group_chunk, b.group_off,
                                                                   b.chunk_start, b.chunk_len, b.group_cv, b.chunk_hash);
        if (b.ev_kernel[3]) cudaEventRecord(b.ev_kernel[3], st);
        n++;
    }
    if (stages & 4) {
        // dense chunk numbering: exclusive scan of the per-file chunk counts (+ the batch's chunk count), also used
        // by the gather and by the xorb pass
        n += launch_scan_u32(b.file_count, b.file_chunk_base, b.n_files, b.scan_partial, b.scalars + kScalarChunks, st);
        const uint64_t want = (b.n_slots * kTreeLanes + kTreeThreads - 1) / kTreeThreads;
        const unsigned tblocks = (unsigned)std::min<uint64_t>(want, (uint64_t)b.sm_count * 16);
        chunk_tree_kernel<<<tblocks ? tblocks : 1, kTreeThreads, 0, st>>>(b.scalars, b.n_files, b.file_chunk_base,
                                                                          b.slot_off, b.slot_groups, b.group_off,
                                                                          b.group_cv, b.chunk_hash);
        n++;
    }
    return n;
}

int launch_chunk_hash(const PipelineBuffers& b, cudaStream_t st) { return launch_chunk_hash_stages(b, st, 7); }

}  // namespace xetgpu
