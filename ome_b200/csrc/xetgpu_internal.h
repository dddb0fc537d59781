// Internal declarations shared by the kernel translation units and the C-ABI layer.
// Not installed; the public surface is include/xetgpu.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace xetgpu {

// ---- chunking constants (xet-core deduplication crate)
constexpr uint32_t kMinChunk = 8192;
constexpr uint32_t kMaxChunk = 131072;
constexpr uint32_t kTopBits = 16;  // MASK = 0xFFFF000000000000: the 16 top bits must be clear

// ---- K1 geometry (gear.cu).  A tile is the 128 KiB of contiguous bytes one warp scans per step:
// 32 rows of 4 KiB, one row per lane.  Candidates are staged per tile.
constexpr int kGearTile = 131072;
constexpr int kGearCap = 128;     // staged candidates per tile (fast path; expected ~2)

// ---- K2 geometry.  Boundary resolution speculates per segment of kSegBytes (>= kMaxChunk, so every
// full segment holds at least one cut); a segment holds at most kSegBytes/kMinChunk + 1 cuts.
#ifndef XG_SEG_KIB
#define XG_SEG_KIB 2048
#endif
constexpr uint64_t kSegBytes = (uint64_t)XG_SEG_KIB << 10;
constexpr uint32_t kSegCap = (uint32_t)(kSegBytes / kMinChunk) + 2;

// ---- K3 geometry.  A "group" is up to kLeavesPerGroup consecutive 1 KiB BLAKE3 leaves of one
// chunk, hashed by one lane, which also merges them into one subtree chaining value.
constexpr int kLeavesPerGroup = 4;
#ifndef XG_LEAF_PART
#define XG_LEAF_PART 512
#endif
#ifndef XG_LEAF_WARPS
#define XG_LEAF_WARPS 4
#endif
constexpr int kLeafHalfBytes = XG_LEAF_PART;            // bytes of a leaf staged per round
constexpr int kLeafHalfBlocks = kLeafHalfBytes / 64;
constexpr int kLeafRow = kLeafHalfBytes + 16;           // smem bytes per staged half leaf (+16: alignment slack)
constexpr int kLeafWarpsPerCta = XG_LEAF_WARPS;          // 4 x 16.5 KiB = 66 KiB per CTA, 3 CTAs = 12 warps per SM

struct PipelineBuffers {
    // inputs
    const uint8_t* data;        // device, 16-byte aligned
    uint64_t total_bytes;       // bytes addressable from data
    const uint64_t* file_off;   // device [n_files+1], offsets into data
    uint32_t n_files;
    const uint64_t* file_walk;  // optional device [n_files]: per-file walk (byte-range items in a batch): bits 0..62 =
                                // absolute position in data of the entry cut, bit 63 set = the item's end is NOT the
                                // end of its file.  nullptr: every item is a whole file.
    // K1
    uint32_t n_tiles;
    uint32_t* tile_count;       // [n_tiles]
    uint32_t* tile_stage;       // [n_tiles * kGearCap] tile-relative positions
    uint64_t* tile_off;         // [n_tiles+1] exclusive scan of tile_count
    uint64_t* cand;             // [cand_cap] absolute candidate byte indices, ascending
    uint64_t cand_cap;
    // K2
    const uint64_t* slot_off;   // device [n_files+1], host-computed prefix of per-file chunk upper bounds
    uint64_t n_slots;
    uint64_t* slot_end;         // [n_slots] absolute exclusive end offset of chunk in slot
    uint32_t* file_count;       // [n_files] chunks per file
    uint64_t* first_start;      // [n_files] absolute start of the first emitted chunk
    uint64_t* file_chunk_base;  // [n_files+1] exclusive scan of file_count
    // K2 segment speculation (resolve.cu)
    const uint64_t* seg_off;    // device [n_files+1] first segment of each file
    uint64_t n_segs;
    uint64_t* seg_spec;         // [n_segs * kSegCap]
    uint64_t* seg_pre;          // [n_segs * kSegCap]
    uint32_t* seg_spec_n;       // [n_segs] ...
    uint32_t* seg_pre_n;
    uint32_t* seg_merge_idx;
    uint32_t* seg_final_n;
    uint64_t* seg_exit_spec;
    uint64_t* seg_exit_true;
    uint64_t* seg_base;         // [n_segs+1]
    // K3
    uint64_t* chunk_start;      // [n_slots] absolute start
    uint32_t* chunk_len;        // [n_slots] 0 = empty slot
    uint32_t* chunk_file;       // [n_slots]
    uint32_t* slot_groups;      // [n_slots] groups per chunk
    uint64_t* group_off;        // [n_slots+1]
    uint32_t* group_chunk;      // [n_groups_cap]
    uint64_t n_groups_cap;
    uint32_t* group_cv;         // [n_groups_cap * 8]
    uint8_t* chunk_hash;        // [n_slots * 32]
    // K4
    uint8_t* node_hash[2];      // ping-pong [n_slots * 32]; level 0 aliases chunk_hash
    uint64_t* node_size[2];     // ping-pong [n_slots]
    uint32_t* level_count[2];   // [n_files] nodes per file at the current / next level
    uint32_t* group_start;      // [n_slots] group start indices within a file's level
    uint64_t* mk_bits;          // [n_slots] per 32-node block: group-end flags of nodes 0..39
    uint64_t* mk_out;           // [n_slots] per block: exit offset for each of the 9 entry offsets (4 bits each)
    uint64_t* mk_cnt;           // [n_slots] per block: group starts for each entry offset (5 bits each)
    uint32_t* mk_entry;         // [n_slots] per block: entry offset of the true chain
    uint32_t* mk_base;          // [n_slots] per block: group starts before the block
    uint8_t* file_root;         // [n_files * 32]
    uint8_t* file_hash;         // [n_files * 32]
    // dense outputs
    uint64_t* out_end;          // [n_slots] file-relative ends, dense by file_chunk_base
    uint8_t* out_hash;          // [n_slots * 32]
    // optional xorb formation (merkle.cu, launch_xorbs): file f owns xorb slots [xorb_slot_off[f], xorb_slot_off[f+1])
    const uint64_t* xorb_slot_off;  // device [n_files+1], host-computed prefix of per-file upper bounds
    uint64_t n_xorb_slots;
    uint32_t* xorb_count;       // [n_files] xorbs per file
    uint64_t* xorb_node_off;    // [n_xorb_slots+1] dense index (out_end / out_hash) of the xorb's first chunk
    uint32_t* xorb_n;           // [n_xorb_slots] chunks in the xorb (0 = unused slot)
    uint64_t* xorb_bytes;       // [n_xorb_slots]
    uint8_t* xorb_root;         // [n_xorb_slots * 32] xorb_hash = aggregated_node_hash of the xorb's chunks
    uint8_t* xorb_scratch;      // [n_xorb_slots * 32]
    uint32_t* xorb_level[2];    // [n_xorb_slots+1] ping-pong level counts
    // scan scratch
    uint64_t* scan_partial;     // [scan_blocks_cap]
    uint64_t* scalars;          // [8] device scalars: 0 = n_cand, 1 = n_groups, 2 = n_chunks
    uint64_t max_nodes;         // host bound on the largest per-file chunk count (merkle level-0 size)
    uint32_t sm_count;          // SMs of the device (persistent-kernel grid size)
    // optional timing events (nullptr = off): [0,1] around gear_scan_kernel, [2,3] around leaf_group_kernel
    cudaEvent_t ev_kernel[4];
};

enum { kScalarCand = 0, kScalarGroups = 1, kScalarChunks = 2, kScalarFlags = 3, kScalarMerkle = 4 /* 4,5: K4's round maxima + barrier word */ };

// Launchers (all asynchronous on `st`); each returns the number of kernels it launched.
int launch_fill_splitmix(uint8_t* dst, uint64_t seed, uint64_t byte_off, uint64_t len, cudaStream_t st);
int launch_gear(const PipelineBuffers& b, cudaStream_t st, cudaError_t* err);
// Range mode (one "file" that is a byte range of a larger file): the walk starts at the absolute
// position walk_start, the buffer end is NOT the end of the file unless is_eof, and the walk stops at
// the first cut beyond the buffer.  Whole-file mode: walk_start is ignored (range == nullptr).
struct RangeWalk {
    uint64_t walk_start;   // absolute position in data of the entry cut
    uint32_t is_eof;       // buffer end == file end
};
int launch_resolve(const PipelineBuffers& b, const RangeWalk* range, cudaStream_t st);
int launch_chunk_hash(const PipelineBuffers& b, cudaStream_t st);
int launch_chunk_hash_stages(const PipelineBuffers& b, cudaStream_t st, int stages);  // 1 work lists, 2 leaf kernel, 4 tree
int launch_merkle(const PipelineBuffers& b, cudaStream_t st);
int launch_gather(const PipelineBuffers& b, cudaStream_t st);
// Xorb formation over the dense per-file chunk lists (after launch_gather and launch_merkle): greedy packing into
// xorbs of <= kMaxXorbBytes and <= kMaxXorbChunks, one merkle root per xorb.
int launch_xorbs(const PipelineBuffers& b, cudaStream_t st);
constexpr uint64_t kMaxXorbBytes = 64ull << 20;
constexpr uint32_t kMaxXorbChunks = 8192;
int launch_scan_u32(const uint32_t* in, uint64_t* out, uint64_t n, uint64_t* partial, uint64_t* total_out,
                    cudaStream_t st);

// Cooperative kernels (grid-wide barriers) of this library never overlap one another on a device: two grids that
// are each only partly resident would wait for CTAs that the other one's spinning CTAs keep out.  Every cooperative
// launch goes through this gate, which chains it behind the previous one on the same device with an event (any
// stream, any context of this process).  Bulk kernels are not affected.
cudaError_t launch_cooperative_gated(const void* kernel, dim3 grid, dim3 block, void** args, size_t smem, cudaStream_t st);

inline uint64_t scan_blocks(uint64_t n) { return (n + 2047) / 2048; }

}  // namespace xetgpu
