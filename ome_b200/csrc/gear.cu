// K0 (synthetic fill), K1 (gear window hash at every byte offset -> boundary candidates) and the
// exclusive-scan utility.  sm_100a only.
//
// K1 restates gearhash::Hasher::next_match (gearhash 0.1.3, pinned at /root/reference
// pkg/xet/Cargo.lock:952-955) in position-parallel form:
//     G[i] = sum_{k=0..63} T[d[i-k]] << k  (mod 2^64),   candidate  <=>  G[i] >> 48 == 0.
// The hash at byte i depends only on the 64 bytes ending at i, so a thread that owns a strip of
// bytes warms its rolling state on the 64 bytes in front of the strip and is exact from the
// strip's first byte on.
//
// Data movement: one CTA owns a 66 KiB tile.  The tile and the 64-byte halo in front of it are
// brought into shared memory by bulk TMA copies (cp.async.bulk, completion on an mbarrier — SASS
// UBLKCP), so no LSU instruction is spent on staging and HBM sees whole, aligned lines exactly once.
// Each lane then walks its own 528-byte strip with 16-byte LDS; 528 = 33*16 is an odd number of
// 16-byte bank groups, which makes the 32 strided lane accesses conflict-free.  The gear table
// lookup is the hot shared-memory access (one 8-byte load per input byte per lane): the table is
// replicated 16x in shared memory, lane l reading replica l%16, so the 16 lanes of each LDS.64
// phase always hit 16 distinct bank pairs whatever the data bytes are.
#include "tma.cuh"
#include "xetgpu_internal.h"
#include "../../include/xet_gear_table.h"

namespace xetgpu {

__constant__ uint64_t c_gear[256] = XET_GEAR_TABLE_INIT;

// ------------------------------------------------------------------------------------------ K0
__device__ __forceinline__ uint64_t splitmix_word(uint64_t seed, uint64_t word_index_1based) {
    uint64_t z = seed + word_index_1based * 0x9E3779B97F4A7C15ULL;
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z;
}

// dst[i] = byte (byte_off + i) of the stream.  Head/tail bytes that do not fill an aligned 8-byte
// destination word are written bytewise by the first threads.
__global__ void fill_splitmix_kernel(uint8_t* __restrict__ dst, uint64_t seed, uint64_t byte_off, uint64_t len) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nthreads = (uint64_t)gridDim.x * blockDim.x;
    // aligned body: destination address and stream position both multiples of 8
    uint64_t head = (8 - ((uintptr_t)dst & 7)) & 7;
    if (head > len) head = len;
    const bool stream_aligned = ((byte_off + head) & 7) == 0;
    uint64_t body_words = stream_aligned ? (len - head) / 8 : 0;
    uint64_t* body = reinterpret_cast<uint64_t*>(dst + head);
    const uint64_t first_word = (byte_off + head) / 8 + 1;
    for (uint64_t w = tid; w < body_words; w += nthreads) body[w] = splitmix_word(seed, first_word + w);
    // the rest, bytewise (at most 14 bytes when the stream is aligned; everything otherwise)
    const uint64_t rest = len - body_words * 8;
    for (uint64_t k = tid; k < rest; k += nthreads) {
        uint64_t i = k < head ? k : body_words * 8 + k;
        uint64_t pos = byte_off + i;
        dst[i] = (uint8_t)(splitmix_word(seed, pos / 8 + 1) >> (8 * (pos & 7)));
    }
}

int launch_fill_splitmix(uint8_t* dst, uint64_t seed, uint64_t byte_off, uint64_t len, cudaStream_t st) {
    if (len == 0) return 0;
    uint64_t want = (len / 8 + 255) / 256;
    int blocks = (int)(want < 148 * 16 ? (want ? want : 1) : 148 * 16);
    fill_splitmix_kernel<<<blocks, 256, 0, st>>>(dst, seed, byte_off, len);
    return 1;
}

// ------------------------------------------------------------------------------------------ K1
constexpr int kGearTableBytes = 256 * kGearTableCopies * 8;                 // 32 KiB
constexpr int kGearTileSmem = kGearHalo + kGearTile + 16;                   // staged bytes (+ tail slack)
constexpr int kGearSmem = kGearTableBytes + kGearTileSmem;
constexpr int kGearPieces = 4;                                              // bulk copies per tile

// Stage tile `tile` (and its halo) in s_tile; s_tile[0] corresponds to absolute byte
// tile_start - kGearHalo.  Returns the number of valid tile bytes.  All threads call it.
__device__ __forceinline__ uint32_t gear_load_tile(const uint8_t* __restrict__ data, uint64_t total, uint32_t tile,
                                                   uint8_t* s_tile, uint64_t* bar, uint32_t parity) {
    const uint64_t tile_start = (uint64_t)tile * kGearTile;
    const uint64_t src_begin = tile ? tile_start - kGearHalo : 0;
    const uint32_t dst_off = tile ? 0 : kGearHalo;
    uint64_t tile_end = tile_start + kGearTile;
    if (tile_end > total) tile_end = total;
    const uint32_t avail = (uint32_t)(tile_end - src_begin);
    const uint32_t bulk = avail & ~15u;
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, bulk);
        constexpr uint32_t piece = (kGearTile / kGearPieces);  // 16,896 B, a multiple of 16
        for (uint32_t o = 0; o < bulk; o += piece) {
            uint32_t n = bulk - o < piece ? bulk - o : piece;
            tma_bulk_g2s(s_tile + dst_off + o, data + src_begin + o, n, bar);
        }
    }
    // bytes that do not fill a 16-byte unit at the end of the buffer, and the zero halo of tile 0
    if (threadIdx.x < avail - bulk) s_tile[dst_off + bulk + threadIdx.x] = data[src_begin + bulk + threadIdx.x];
    if (tile == 0 && threadIdx.x < kGearHalo / 4) reinterpret_cast<uint32_t*>(s_tile)[threadIdx.x] = 0u;
    mbar_wait(bar, parity);
    __syncthreads();
    return (uint32_t)(tile_end - tile_start);
}

__device__ __forceinline__ void gear_load_table(uint64_t* s_table) {
    for (int i = threadIdx.x; i < 256 * kGearTableCopies; i += blockDim.x) s_table[i] = c_gear[i / kGearTableCopies];
}

// Walk one strip: 64 warm-up bytes, then kGearStrip tested bytes.  `on_match(pos)` is called with
// the tile-relative position of every candidate below `valid`.
template <typename F>
__device__ __forceinline__ void gear_strip(const uint8_t* s_tile, const uint64_t* s_table, uint32_t valid, F on_match) {
    const uint32_t strip_first = threadIdx.x * kGearStrip;  // tile-relative position of first tested byte
    if (strip_first >= valid) return;
    const uint64_t* tab = s_table + (threadIdx.x & (kGearTableCopies - 1));
    const uint4* src = reinterpret_cast<const uint4*>(s_tile + strip_first);  // halo makes this strip_first - 64
    uint64_t h = 0;
#pragma unroll 1
    for (int k = 0; k < kGearHalo / 16; k++) {
        uint4 v = src[k];
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 16; j++) {
            uint32_t b = (w[j >> 2] >> (8 * (j & 3))) & 0xffu;
            h = (h << 1) + tab[b * kGearTableCopies];
        }
    }
#pragma unroll 1
    for (int k = 0; k < kGearStrip / 16; k++) {
        uint4 v = src[kGearHalo / 16 + k];
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 16; j++) {
            uint32_t b = (w[j >> 2] >> (8 * (j & 3))) & 0xffu;
            h = (h << 1) + tab[b * kGearTableCopies];
            if ((uint32_t)(h >> 32) < (1u << (32 - kTopBits))) {
                uint32_t pos = strip_first + k * 16 + j;
                if (pos < valid) on_match(pos);
            }
        }
    }
}

// Pass 1 over every tile: count candidates, stage up to kGearCap of them in ascending order.
__global__ void __launch_bounds__(kGearThreads)
    gear_scan_kernel(const uint8_t* __restrict__ data, uint64_t total, uint32_t* __restrict__ tile_count,
                     uint32_t* __restrict__ tile_stage) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* s_table = reinterpret_cast<uint64_t*>(smem);
    uint8_t* s_tile = smem + kGearTableBytes;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_n;
    __shared__ uint32_t s_list[kGearCap];

    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        s_n = 0;
    }
    gear_load_table(s_table);
    __syncthreads();
    const uint32_t tile = blockIdx.x;
    const uint32_t valid = gear_load_tile(data, total, tile, s_tile, &s_bar, 0);

    gear_strip(s_tile, s_table, valid, [&](uint32_t pos) {
        uint32_t slot = atomicAdd(&s_n, 1u);
        if (slot < kGearCap) s_list[slot] = pos;
    });
    __syncthreads();
    const uint32_t n = s_n;
    if (threadIdx.x == 0) tile_count[tile] = n;
    if (n <= kGearCap && threadIdx.x < n) {  // rank sort (positions are distinct)
        const uint32_t mine = s_list[threadIdx.x];
        uint32_t rank = 0;
        for (uint32_t j = 0; j < n; j++) rank += s_list[j] < mine;
        tile_stage[(uint64_t)tile * kGearCap + rank] = mine;
    }
}

// Pass 2: move staged candidates to their final place; tiles that overflowed the stage (dense
// candidates — adversarial data only) are recomputed with an ordered two-sweep emission.
__global__ void __launch_bounds__(kGearThreads)
    gear_compact_kernel(const uint8_t* __restrict__ data, uint64_t total, const uint32_t* __restrict__ tile_count,
                        const uint32_t* __restrict__ tile_stage, const uint64_t* __restrict__ tile_off,
                        uint64_t* __restrict__ cand, uint64_t cand_cap) {
    const uint32_t tile = blockIdx.x;
    const uint32_t n = tile_count[tile];
    if (n == 0) return;
    const uint64_t base = tile_off[tile];
    const uint64_t tile_start = (uint64_t)tile * kGearTile;
    if (n <= kGearCap) {
        if (threadIdx.x < n && base + threadIdx.x < cand_cap)
            cand[base + threadIdx.x] = tile_start + tile_stage[(uint64_t)tile * kGearCap + threadIdx.x];
        return;
    }
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* s_table = reinterpret_cast<uint64_t*>(smem);
    uint8_t* s_tile = smem + kGearTableBytes;
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_cnt[kGearThreads];
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
    }
    gear_load_table(s_table);
    __syncthreads();
    const uint32_t valid = gear_load_tile(data, total, tile, s_tile, &s_bar, 0);
    uint32_t mine = 0;
    gear_strip(s_tile, s_table, valid, [&](uint32_t) { mine++; });
    s_cnt[threadIdx.x] = mine;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t j = 0; j < threadIdx.x; j++) before += s_cnt[j];
    uint64_t w = base + before;
    gear_strip(s_tile, s_table, valid, [&](uint32_t pos) {
        if (w < cand_cap) cand[w] = tile_start + pos;
        w++;
    });
}

// ------------------------------------------------------------------------------ exclusive scan
// out[i] = sum_{j<i} in[j] for i in [0, n]; out has n+1 entries.  Three small kernels: per-block
// sums, a single-CTA scan of the block sums, per-block scan with the block's base.
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;  // 2048 items per block

__device__ __forceinline__ uint64_t block_exclusive_scan(uint64_t v, uint64_t* s_warp, uint64_t* block_total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    uint64_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint64_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    uint64_t wbase = 0, tot = 0;
    for (int i = 0; i < nw; i++) {
        uint64_t s = s_warp[i];
        if (i < wid) wbase += s;
        tot += s;
    }
    __syncthreads();
    *block_total = tot;
    return wbase + inc - v;
}

__global__ void __launch_bounds__(kScanThreads)
    scan_block_sums_kernel(const uint32_t* __restrict__ in, uint64_t n, uint64_t* __restrict__ partial) {
    __shared__ uint64_t s_warp[kScanThreads / 32];
    const uint64_t base = (uint64_t)blockIdx.x * (kScanThreads * kScanItems) + threadIdx.x * kScanItems;
    uint64_t v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++)
        if (base + i < n) v += in[base + i];
    uint64_t tot;
    block_exclusive_scan(v, s_warp, &tot);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(1024) scan_partials_kernel(uint64_t* __restrict__ partial, uint64_t nblocks,
                                                              uint64_t* __restrict__ total_out) {
    __shared__ uint64_t s_warp[32];
    uint64_t carry = 0;
    for (uint64_t start = 0; start < nblocks; start += 1024) {
        uint64_t i = start + threadIdx.x;
        uint64_t v = i < nblocks ? partial[i] : 0;
        uint64_t tot;
        uint64_t ex = block_exclusive_scan(v, s_warp, &tot);
        if (i < nblocks) partial[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

__global__ void __launch_bounds__(kScanThreads)
    scan_apply_kernel(const uint32_t* __restrict__ in, uint64_t n, const uint64_t* __restrict__ partial,
                      uint64_t* __restrict__ out) {
    __shared__ uint64_t s_warp[kScanThreads / 32];
    const uint64_t base = (uint64_t)blockIdx.x * (kScanThreads * kScanItems) + threadIdx.x * kScanItems;
    uint32_t x[kScanItems];
    uint64_t v = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        x[i] = base + i < n ? in[base + i] : 0u;
        v += x[i];
    }
    uint64_t tot;
    uint64_t run = partial[blockIdx.x] + block_exclusive_scan(v, s_warp, &tot);
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        if (base + i <= n) out[base + i] = run;  // entry n receives the grand total
        run += x[i];
    }
}

int launch_scan_u32(const uint32_t* in, uint64_t* out, uint64_t n, uint64_t* partial, uint64_t* total_out,
                    cudaStream_t st) {
    // n+1 outputs: size the grid on n+1 so the last entry is always written
    const uint64_t nblocks = scan_blocks(n + 1);
    scan_block_sums_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(in, n, partial);
    scan_partials_kernel<<<1, 1024, 0, st>>>(partial, nblocks, total_out);
    scan_apply_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(in, n, partial, out);
    return 3;
}

// ------------------------------------------------------------------------------------ launcher
// Called once per context with its device current (function attributes are per device).
void init_gear_kernels() {
    cudaFuncSetAttribute(gear_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGearSmem);
    cudaFuncSetAttribute(gear_compact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGearSmem);
}

int launch_gear(const PipelineBuffers& b, cudaStream_t st) {
    if (b.n_tiles == 0) {
        cudaMemsetAsync(b.tile_off, 0, sizeof(uint64_t), st);
        cudaMemsetAsync(b.scalars + kScalarCand, 0, sizeof(uint64_t), st);
        return 0;
    }
    if (b.ev_kernel[0]) cudaEventRecord(b.ev_kernel[0], st);
    gear_scan_kernel<<<b.n_tiles, kGearThreads, kGearSmem, st>>>(b.data, b.total_bytes, b.tile_count, b.tile_stage);
    if (b.ev_kernel[1]) cudaEventRecord(b.ev_kernel[1], st);
    int n = 1 + launch_scan_u32(b.tile_count, b.tile_off, b.n_tiles, b.scan_partial, b.scalars + kScalarCand, st);
    gear_compact_kernel<<<b.n_tiles, kGearThreads, kGearSmem, st>>>(b.data, b.total_bytes, b.tile_count, b.tile_stage,
                                                                    b.tile_off, b.cand, b.cand_cap);
    return n + 1;
}

}  // namespace xetgpu
