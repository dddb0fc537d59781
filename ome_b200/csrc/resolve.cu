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

// Greedy walk from the cut `start`.  on_cut(cut) is called for every cut in order and returns
// false to stop.  `end` is the end of the bytes present; is_eof says whether that is the end of
// the file (flush the tail) — otherwise the walk stops at the first cut it cannot decide.
template <typename F>
__device__ __forceinline__ void walk_cuts(const CandIndex& ci, uint64_t start, uint64_t end, bool is_eof, F on_cut) {
    uint64_t idx = 0;
    bool have_idx = false;
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
        if (!is_eof && cut > end) break;  // belongs to the next byte range
        if (!on_cut(cut)) break;
        start = cut;
    }
}

// ---- segment-speculative resolution --------------------------------------------------------
// The greedy walk is a sequential chain (next cut = f(previous cut)), ~N/66 KB steps per file.
// It is parallelised by speculation: a file is split into kSegBytes segments and
//   A  every segment walks from its own begin as if a cut were there          (parallel)
//   A2 every segment j >= 1 re-walks from segment j-1's speculative exit cut until it lands on
//      a cut of its own speculative chain — from a shared cut on, two chains are identical —
//      and records the few cuts before that merge point                        (parallel)
//   B  one warp per file checks, 32 segments at a time, that each segment's assumed entry was
//      the true exit of its predecessor, and re-walks the rare segment where it was not
//   C  an exclusive scan of the per-segment final counts and a copy into the file's slots.
// The result is the sequential chain exactly, for any input: B falls back to walking (worst case:
// chains that never merge, e.g. long runs without any candidate after an unaligned prefix).
struct SegArrays {
    const uint64_t* seg_off;   // [n_files+1] first segment of each file
    uint64_t* spec;            // [n_segs * kSegCap] speculative cuts
    uint64_t* pre;             // [n_segs * kSegCap] cuts before the merge point
    uint32_t* spec_n;          // [n_segs]
    uint32_t* pre_n;           // [n_segs]
    uint32_t* merge_idx;       // [n_segs] first speculative cut that is kept
    uint64_t* exit_spec;       // [n_segs] last cut <= segment end of the speculative chain
    uint64_t* exit_true;       // [n_segs] same for the re-synchronised chain
    uint32_t* final_n;         // [n_segs]
    uint64_t* seg_base;        // [n_segs+1] exclusive scan of final_n
};

struct FileView {
    uint64_t begin, end, walk_start;
    bool is_eof;
    uint32_t f;
    uint64_t seg0, seg1;  // this file's segments
};

__device__ __forceinline__ FileView file_view(uint32_t f, const uint64_t* file_off, const uint64_t* seg_off,
                                              RangeWalk range, bool range_mode, const uint64_t* file_walk) {
    FileView v;
    v.f = f;
    v.begin = file_off[f];
    v.end = file_off[f + 1];
    if (file_walk) {  // a batch that mixes whole files and byte-range items
        const uint64_t w = file_walk[f];
        v.walk_start = w & ~(1ull << 63);
        v.is_eof = (w >> 63) == 0;
    } else {
        v.walk_start = range_mode ? range.walk_start : v.begin;
        v.is_eof = range_mode ? (range.is_eof != 0) : true;
    }
    v.seg0 = seg_off[f];
    v.seg1 = seg_off[f + 1];
    return v;
}

__device__ __forceinline__ uint32_t file_of_seg(const uint64_t* __restrict__ seg_off, uint32_t n_files, uint64_t s) {
    uint32_t lo = 0, hi = n_files;  // last f with seg_off[f] <= s (files without segments are skipped)
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (seg_off[mid] <= s) lo = mid;
        else hi = mid;
    }
    return lo;
}

// Re-synchronise segment s of file v: walk from `entry` until a cut of the speculative chain is hit.
__device__ __forceinline__ void sync_segment(const CandIndex& ci, const SegArrays& sa, const FileView& v, uint64_t s,
                                             uint64_t entry) {
    const uint64_t j = s - v.seg0;
    const uint64_t seg_e = min(v.end, v.begin + (j + 1) * kSegBytes);
    const uint64_t* spec = sa.spec + s * kSegCap;
    uint64_t* pre = sa.pre + s * kSegCap;
    const uint32_t ns = sa.spec_n[s];
    uint32_t i = 0, np = 0;
    bool merged = false;
    uint64_t last = entry;
    walk_cuts(ci, entry, v.end, v.is_eof, [&](uint64_t cut) {
        if (cut > seg_e) return false;
        while (i < ns && spec[i] < cut) i++;
        if (i < ns && spec[i] == cut) {
            merged = true;
            return false;
        }
        if (np < kSegCap) pre[np] = cut;
        np++;
        last = cut;
        return true;
    });
    sa.pre_n[s] = np;
    sa.merge_idx[s] = merged ? i : ns;
    sa.exit_true[s] = merged ? sa.exit_spec[s] : last;
    sa.final_n[s] = np + (merged ? ns - i : 0u);
}

struct ResolveArgs {
    const uint64_t* cand;
    const uint64_t* tile_off;
    const uint64_t* scalars;
    uint64_t cand_cap;
    uint32_t n_tiles;
    const uint64_t* file_off;
    uint32_t n_files;
    uint64_t n_segs;
    RangeWalk range;
    bool range_mode;
    const uint64_t* file_walk;
    SegArrays sa;
};

__device__ __forceinline__ CandIndex cand_index(const ResolveArgs& a) {
    uint64_t n_cand = a.scalars[kScalarCand];
    if (n_cand > a.cand_cap) n_cand = a.cand_cap;  // host detects the overflow and re-runs with a larger list
    return CandIndex{a.cand, a.tile_off, n_cand, a.n_tiles};
}

// A: speculative chain of every segment.
__global__ void resolve_spec_kernel(ResolveArgs a) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= a.n_segs) return;
    const CandIndex ci = cand_index(a);
    const FileView v = file_view(file_of_seg(a.sa.seg_off, a.n_files, s), a.file_off, a.sa.seg_off, a.range, a.range_mode, a.file_walk);
    const uint64_t j = s - v.seg0;
    const uint64_t seg_b = v.begin + j * kSegBytes;
    const uint64_t seg_e = min(v.end, seg_b + kSegBytes);
    const uint64_t start = j == 0 ? v.walk_start : seg_b;
    uint64_t* spec = a.sa.spec + s * kSegCap;
    uint32_t n = 0;
    uint64_t last = start;
    walk_cuts(ci, start, v.end, v.is_eof, [&](uint64_t cut) {
        if (cut > seg_e) return false;
        if (n < kSegCap) spec[n] = cut;
        n++;
        last = cut;
        return true;
    });
    a.sa.spec_n[s] = n;
    a.sa.exit_spec[s] = last;
    if (j == 0) {  // the first segment's entry is the true one
        a.sa.pre_n[s] = 0;
        a.sa.merge_idx[s] = 0;
        a.sa.exit_true[s] = last;
        a.sa.final_n[s] = n;
    }
}

// A2: re-synchronise every later segment from its predecessor's speculative exit.
__global__ void resolve_sync_kernel(ResolveArgs a) {
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= a.n_segs) return;
    const FileView v = file_view(file_of_seg(a.sa.seg_off, a.n_files, s), a.file_off, a.sa.seg_off, a.range, a.range_mode, a.file_walk);
    if (s == v.seg0) return;
    sync_segment(cand_index(a), a.sa, v, s, a.sa.exit_spec[s - 1]);
}

// B: validate the speculation chain, 32 segments per step; repair where the assumed entry was wrong.
__global__ void resolve_fix_kernel(ResolveArgs a) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (f >= a.n_files) return;
    const FileView v = file_view(f, a.file_off, a.sa.seg_off, a.range, a.range_mode, a.file_walk);
    const CandIndex ci = cand_index(a);
    uint64_t s = v.seg0 + 1;
    while (s < v.seg1) {
        const uint64_t mine = s + lane;
        // segment `mine` assumed entry exit_spec[mine-1]; it was right iff that is the predecessor's true exit
        const bool bad = mine < v.seg1 && a.sa.exit_true[mine - 1] != a.sa.exit_spec[mine - 1];
        const uint32_t bad_mask = __ballot_sync(0xffffffffu, bad);
        if (bad_mask == 0) {
            s += 32;
            continue;
        }
        const uint64_t sb = s + (__ffs(bad_mask) - 1);  // everything before sb is final
        if (lane == 0) sync_segment(ci, a.sa, v, sb, a.sa.exit_true[sb - 1]);
        __threadfence_block();
        __syncwarp();
        s = sb + 1;
    }
}

// C: copy the final cuts of each segment into its file's slot range (one warp per segment).
__global__ void resolve_stitch_kernel(ResolveArgs a, const uint64_t* __restrict__ slot_off,
                                      uint64_t* __restrict__ slot_end) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t s = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (s >= a.n_segs) return;
    const uint32_t f = file_of_seg(a.sa.seg_off, a.n_files, s);
    const uint64_t s0 = a.sa.seg_off[f];
    uint64_t* dst = slot_end + slot_off[f] + (a.sa.seg_base[s] - a.sa.seg_base[s0]);
    const uint64_t room = slot_off[f + 1] - slot_off[f] - (a.sa.seg_base[s] - a.sa.seg_base[s0]);
    const uint32_t np = a.sa.pre_n[s], mi = a.sa.merge_idx[s], ns = a.sa.spec_n[s];
    const uint64_t* pre = a.sa.pre + s * kSegCap;
    const uint64_t* spec = a.sa.spec + s * kSegCap;
    for (uint32_t i = lane; i < np; i += 32)
        if (i < room) dst[i] = pre[i];
    for (uint32_t i = lane; mi + i < ns; i += 32)
        if (np + i < room) dst[np + i] = spec[mi + i];
}

__global__ void resolve_file_counts_kernel(ResolveArgs a, uint32_t* __restrict__ file_count,
                                           uint64_t* __restrict__ first_start) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= a.n_files) return;
    file_count[f] = (uint32_t)(a.sa.seg_base[a.sa.seg_off[f + 1]] - a.sa.seg_base[a.sa.seg_off[f]]);
    first_start[f] = a.file_walk ? (a.file_walk[f] & ~(1ull << 63)) : a.range_mode ? a.range.walk_start : a.file_off[f];
}

int launch_resolve(const PipelineBuffers& b, const RangeWalk* range, cudaStream_t st) {
    if (b.n_files == 0) return 0;
    ResolveArgs a{};
    a.cand = b.cand;
    a.tile_off = b.tile_off;
    a.scalars = b.scalars;
    a.cand_cap = b.cand_cap;
    a.n_tiles = b.n_tiles;
    a.file_off = b.file_off;
    a.n_files = b.n_files;
    a.n_segs = b.n_segs;
    a.range = range ? *range : RangeWalk{0, 1};
    a.range_mode = range != nullptr;
    a.file_walk = b.file_walk;
    a.sa = SegArrays{b.seg_off, b.seg_spec, b.seg_pre, b.seg_spec_n, b.seg_pre_n, b.seg_merge_idx,
                     b.seg_exit_spec, b.seg_exit_true, b.seg_final_n, b.seg_base};
    int n = 0;
    if (b.n_segs) {
        const unsigned sblocks = (unsigned)((b.n_segs + 63) / 64);
        resolve_spec_kernel<<<sblocks, 64, 0, st>>>(a);
        resolve_sync_kernel<<<sblocks, 64, 0, st>>>(a);
        resolve_fix_kernel<<<(b.n_files * 32 + 127) / 128, 128, 0, st>>>(a);
        n += 3;
    }
    n += launch_scan_u32(b.seg_final_n, b.seg_base, b.n_segs, b.scan_partial, nullptr, st);
    if (b.n_segs) {
        resolve_stitch_kernel<<<(unsigned)((b.n_segs * 32 + 127) / 128), 128, 0, st>>>(a, b.slot_off, b.slot_end);
        n++;
    }
    resolve_file_counts_kernel<<<(b.n_files + 127) / 128, 128, 0, st>>>(a, b.file_count, b.first_start);
    return n + 1;
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
    int n = 0;   // file_chunk_base / the chunk total were scanned by launch_chunk_hash
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
