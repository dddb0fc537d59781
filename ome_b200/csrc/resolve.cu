// K2: min/max-size boundary resolution, plus the bookkeeping kernels that turn resolved cuts into
// the work lists of K3 and the dense output arrays.
//
// Restates deduplication::Chunker::next_boundary (xet-core deduplication 0.14.5, pinned at
// /root/reference pkg/xet/Cargo.lock:642-655) in the candidate-list form of SURVEY.md A.2:
//     start = 0
//     while start < N:
//         c   = first candidate index with c + 1 >= start + 8192
//         cut = c + 1  if it exists and c + 1 <= min(N, start + 131072)  else  min(N, start + 131072)
// Chunk starts never influence candidacy (no cut is accepted below 8192 >= 64 bytes), so the
// sequential chunker and this walk over the pre-computed candidate list produce the same cuts.
#include "xetgpu_internal.h"

namespace xetgpu {

struct CandIndex {
    const uint64_t* cand;
    const uint64_t* tile_off;
    uint64_t n_cand;
    uint32_t n_tiles;

    // Index of the first candidate >= x (n_cand if none).  tile_off gives O(1) access to the
    // tile that contains x; inside the tile a binary search (tiles hold ~1 candidate on random
    // data, up to the tile size on adversarial data).
    __device__ __forceinline__ uint64_t lower_bound(uint64_t x) const {
        const uint64_t t = x / kGearTile;
        if (t >= n_tiles) return n_cand;
        uint64_t lo = tile_off[t], hi = tile_off[t + 1];
        if (hi > n_cand) hi = n_cand;
        while (lo < hi) {
            uint64_t mid = (lo + hi) >> 1;
            if (cand[mid] < x) lo = mid + 1;
            else hi = mid;
        }
        return lo;  // == first candidate of a later tile when this tile has none >= x
    }
};

// Walks one file (or one byte range of a file) and writes absolute cut positions.
//   begin/end    : absolute extent of the bytes present in `data`
//   walk_start   : absolute position of the cut the walk starts from (== begin for whole files)
//   is_eof       : `end` is the end of the file (flush the tail); otherwise the walk stops at the
//                  first cut that would lie beyond `end` (it belongs to the next range)
__device__ __forceinline__ uint32_t walk_cuts(const CandIndex& ci, uint64_t walk_start, uint64_t end, bool is_eof,
                                              uint64_t* __restrict__ out, uint64_t out_cap) {
    uint64_t start = walk_start;
    uint64_t idx = 0;
    bool have_idx = false;
    uint32_t n = 0;
    while (true) {
        if (is_eof && start >= end) break;
        uint64_t hi = start + kMaxChunk;
        if (is_eof && hi > end) hi = end;
        uint64_t cut = hi;
        if (start + kMinChunk <= hi) {
            const uint64_t x = start + kMinChunk - 1;  // smallest admissible candidate index
            // the cached lower bound stays valid while it is still >= x (no candidate lies between)
            if (!have_idx || (idx < ci.n_cand && ci.cand[idx] < x)) {
                idx = ci.lower_bound(x);
                have_idx = true;
            }
            if (idx < ci.n_cand) {
                const uint64_t c = ci.cand[idx];
                if (c + 1 <= hi) cut = c + 1;
            }
        }
        if (!is_eof && cut > end) break;
        if (n < out_cap) out[n] = cut;
        n++;
        start = cut;
    }
    return n;
}

__global__ void resolve_files_kernel(const uint64_t* __restrict__ cand, const uint64_t* __restrict__ tile_off,
                                     const uint64_t* __restrict__ scalars, uint64_t cand_cap, uint32_t n_tiles,
                                     const uint64_t* __restrict__ file_off, const uint64_t* __restrict__ slot_off,
                                     uint32_t n_files, uint64_t* __restrict__ slot_end,
                                     uint32_t* __restrict__ file_count, uint64_t* __restrict__ first_start,
                                     RangeWalk range, bool range_mode) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_files) return;
    uint64_t n_cand = scalars[kScalarCand];
    if (n_cand > cand_cap) n_cand = cand_cap;  // host detects the overflow and re-runs with a larger list
    CandIndex ci{cand, tile_off, n_cand, n_tiles};
    const uint64_t begin = file_off[f], end = file_off[f + 1];
    const uint64_t ws = range_mode ? range.walk_start : begin;
    const bool eof = range_mode ? (range.is_eof != 0) : true;
    const uint64_t s0 = slot_off[f];
    file_count[f] = walk_cuts(ci, ws, end, eof, slot_end + s0, slot_off[f + 1] - s0);
    first_start[f] = ws;
}

int launch_resolve(const PipelineBuffers& b, const RangeWalk* range, cudaStream_t st) {
    if (b.n_files == 0) return 0;
    RangeWalk r{0, 1};
    if (range) r = *range;
    const int threads = 32;
    resolve_files_kernel<<<(b.n_files + threads - 1) / threads, threads, 0, st>>>(
        b.cand, b.tile_off, b.scalars, b.cand_cap, b.n_tiles, b.file_off, b.slot_off, b.n_files, b.slot_end,
        b.file_count, b.first_start, r, range != nullptr);
    return 1;
}

// ------------------------------------------------------------------------------- bookkeeping
__device__ __forceinline__ uint32_t file_of_slot(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t s) {
    uint32_t lo = 0, hi = n_files;  // last f with slot_off[f] <= s
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (slot_off[mid] <= s) lo = mid;
        else hi = mid;
    }
    return lo;
}

// Per chunk slot: extent, owning file, and how many leaf groups K3 will hash for it.
__global__ void chunk_prepare_kernel(const uint64_t* __restrict__ slot_off, uint32_t n_files, uint64_t n_slots,
                                     const uint64_t* __restrict__ slot_end, const uint32_t* __restrict__ file_count,
                                     const uint64_t* __restrict__ first_start, uint64_t* __restrict__ chunk_start,
                                     uint32_t* __restrict__ chunk_len, uint32_t* __restrict__ chunk_file,
                                     uint32_t* __restrict__ slot_groups) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t f = file_of_slot(slot_off, n_files, s);
    const uint64_t k = s - slot_off[f];
    uint32_t len = 0, groups = 0;
    uint64_t start = 0;
    if (k < file_count[f]) {
        start = k == 0 ? first_start[f] : slot_end[s - 1];
        len = (uint32_t)(slot_end[s] - start);
        const uint32_t leaves = len <= 1024 ? 1u : (len + 1023u) / 1024u;
        groups = (leaves + kLeavesPerGroup - 1) / kLeavesPerGroup;
    }
    chunk_start[s] = start;
    chunk_len[s] = len;
    chunk_file[s] = f;
    slot_groups[s] = groups;
}

__global__ void group_expand_kernel(uint64_t n_slots, const uint32_t* __restrict__ slot_groups,
                                    const uint64_t* __restrict__ group_off, uint32_t* __restrict__ group_chunk,
                                    uint64_t n_groups_cap) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const uint32_t g = slot_groups[s];
    const uint64_t base = group_off[s];
    for (uint32_t j = 0; j < g; j++)
        if (base + j < n_groups_cap) group_chunk[base + j] = (uint32_t)s;
}

// Dense, per-file-contiguous outputs: file-relative chunk ends and chunk hashes.
__global__ void gather_kernel(const uint64_t* __restrict__ slot_off, const uint64_t* __restrict__ file_off,
                              uint32_t n_files, uint64_t n_slots, const uint64_t* __restrict__ slot_end,
                              const uint32_t* __restrict__ chunk_len, const uint32_t* __restrict__ chunk_file,
                              const uint64_t* __restrict__ file_chunk_base, const uint8_t* __restrict__ chunk_hash,
                              uint64_t* __restrict__ out_end, uint8_t* __restrict__ out_hash) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots || chunk_len[s] == 0) return;
    const uint32_t f = chunk_file[s];
    const uint64_t d = file_chunk_base[f] + (s - slot_off[f]);
    out_end[d] = slot_end[s] - file_off[f];
    const uint4* src = reinterpret_cast<const uint4*>(chunk_hash + s * 32);
    uint4* dst = reinterpret_cast<uint4*>(out_hash + d * 32);
    dst[0] = src[0];
    dst[1] = src[1];
}

int launch_gather(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_files == 0) return 0;
    int n = launch_scan_u32(b.file_count, b.file_chunk_base, b.n_files, b.scan_partial, b.scalars + kScalarChunks, st);
    if (b.n_slots) {
        gather_kernel<<<(unsigned)((b.n_slots + 255) / 256), 256, 0, st>>>(b.slot_off, b.file_off, b.n_files,
                                                                          b.n_slots, b.slot_end, b.chunk_len,
                                                                          b.chunk_file, b.file_chunk_base,
                                                                          b.chunk_hash, b.out_end, b.out_hash);
        n++;
    }
    return n;
}

// Used by chunk_hash.cu's launcher.
int launch_chunk_bookkeeping(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_slots == 0) {
        cudaMemsetAsync(b.scalars + kScalarGroups, 0, sizeof(uint64_t), st);
        return 0;
    }
    const unsigned blocks = (unsigned)((b.n_slots + 255) / 256);
    chunk_prepare_kernel<<<blocks, 256, 0, st>>>(b.slot_off, b.n_files, b.n_slots, b.slot_end, b.file_count,
                                                 b.first_start, b.chunk_start, b.chunk_len, b.chunk_file,
                                                 b.slot_groups);
    int n = 1 + launch_scan_u32(b.slot_groups, b.group_off, b.n_slots, b.scan_partial, b.scalars + kScalarGroups, st);
    group_expand_kernel<<<blocks, 256, 0, st>>>(b.n_slots, b.slot_groups, b.group_off, b.group_chunk, b.n_groups_cap);
    return n + 1;
}

}  // namespace xetgpu
