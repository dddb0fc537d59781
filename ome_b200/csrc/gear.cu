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
// Data movement and the table lookup are described at the K1 section below.
#include <cuda.h>
#include <string.h>

#include <mutex>

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
// Geometry: the buffer is viewed as a 2-D tensor of 4 KiB rows.  A warp owns a tile of 32
// consecutive rows (128 KiB of contiguous bytes); lane l walks row l from left to right.  The rows
// stream through shared memory as TMA boxes of {128 bytes x 32 rows} (one cp.async.bulk.tensor per
// box, SASS UTMALDG, 128-byte swizzle), kGearStages boxes in flight per warp behind mbarriers, so
// HBM latency is hidden by the TMA unit and no LSU instruction is spent on staging.  The 64 bytes
// in front of every row (the previous row's tail) arrive as one extra box per tile and only warm
// the rolling hash: 64 / 4096 = 1.6 % overhead.
constexpr int kGearRow = 4096;
static_assert(kGearTile == 32 * kGearRow, "a tile is 32 rows");
#ifndef XG_GEAR_WARPS
#define XG_GEAR_WARPS 16
#endif
#ifndef XG_GEAR_STAGES
#define XG_GEAR_STAGES 2
#endif
constexpr int kGearWarps = XG_GEAR_WARPS;
constexpr int kGearThreads = kGearWarps * 32;
constexpr int kGearStages = XG_GEAR_STAGES;
constexpr int kGearBox = 32 * 128;                                           // bytes per box
constexpr int kGearBoxesPerTile = 1 + kGearRow / 128;                        // halo box + 32
constexpr int kGearTableBytes = 256 * 32 * 8;                                // [byte][lane] u64, 64 KiB
constexpr int kGearOffStages = kGearTableBytes;
constexpr int kGearOffLists = kGearOffStages + kGearWarps * kGearStages * kGearBox;
constexpr int kGearOffBars = kGearOffLists + kGearWarps * kGearCap * 4;
constexpr int kGearOffCnt = kGearOffBars + kGearWarps * kGearStages * 8;
constexpr int kGearSmem = kGearOffCnt + kGearWarps * 4 + 1024;               // + slack to align the base to 1 KiB

__device__ __forceinline__ void tma_box_g2s(void* dst_smem, const CUtensorMap* tmap, int32_t c0, int32_t c1,
                                            uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
            "r"(smem_u32(dst_smem)),
        "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// The table is laid out [byte value][lane] so that one PRMT builds the whole shared-memory offset
// (byte << 8) | (lane * 8), and the 16 lanes of each LDS.64 phase hit 16 distinct bank pairs.
__device__ __forceinline__ void gear_load_table(uint64_t* s_table) {
    for (int i = threadIdx.x; i < 256 * 32; i += blockDim.x) s_table[i] = c_gear[i >> 5];
}

struct GearLane {
    uint32_t tab;     // shared-memory address of the table
    uint32_t lane8;   // lane * 8
    uint64_t h;       // rolling hash
    uint32_t two;     // the constant 2, loaded from constant memory so that ptxas keeps the IMADs below
};

// h' = 2h + t (mod 2^64).  XG_GEAR_ROLL selects the pipe the roll issues on.  Measured on B200 (config 2, 16.06 GB,
// profiles/r2_coresidency_variants.jsonl): 0 -> 4.20 ms, 1 -> 4.45 ms, 2 -> 4.63 ms for the kernel alone — the
// IMAD.WIDE is a half-rate instruction, so moving the roll off the ALU pipe costs more issue slots than it
// frees; mode 1 exists for builds in which an ALU-bound kernel shares the SM.
//   0  ALU:  IADD3 + IADD3.X
//   1  FMA:  IMAD.WIDE.U32 (low word * 2 + t, 64-bit result carries into the high word) + IMAD (high word * 2 + carry)
//   2  split: IMAD.WIDE.U32 on the FMA pipe, high word as a 3-input add on the ALU pipe
#ifndef XG_GEAR_ROLL
#define XG_GEAR_ROLL 0
#endif
static __constant__ uint32_t c_gear_two = 2u;

__device__ __forceinline__ uint64_t gear_roll(uint64_t h, uint64_t t, uint32_t two) {
#if XG_GEAR_ROLL == 0
    return (h << 1) + t;
#else
    const uint32_t hl = (uint32_t)h, hh = (uint32_t)(h >> 32);
    // mad.lo.cc + madc.hi is the PTX form ptxas turns into ONE IMAD.WIDE.U32 with a 64-bit addend
    // ({tl, th} + hl * 2: low word and th + carry); mad.wide.u32 with a 64-bit addend and the C expression are
    // both split into IMAD.WIDE + extra ALU adds.
    const uint32_t tl = (uint32_t)t, th = (uint32_t)(t >> 32);
    uint32_t l, hc;
    asm("mad.lo.cc.u32 %0, %2, %3, %4;\n\tmadc.hi.u32 %1, %2, %3, %5;" : "=r"(l), "=r"(hc) : "r"(hl), "r"(two), "r"(tl), "r"(th));
    const uint64_t lo = ((uint64_t)hc << 32) | l;
    uint32_t nh;
#if XG_GEAR_ROLL == 1
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(nh) : "r"(hh), "r"(two), "r"(hc));
#else
    nh = hh + hh + hc;
#endif
    return ((uint64_t)nh << 32) | (uint32_t)lo;
#endif
}

template <int J>
__device__ __forceinline__ uint64_t gear_lookup(const GearLane& g, uint32_t w) {
    // bytes of the result: [0] = lane*8, [1] = byte J of w, [2] = [3] = 0  ->  (byte << 8) | lane*8
    const uint32_t off = __byte_perm(w, g.lane8, 0x5504 | (J << 4));
    uint64_t v;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(g.tab + off));
    return v;
}

// 16 bytes: roll, and report through on_match every position (base + j) whose hash is a candidate.
template <bool kTest, typename F>
__device__ __forceinline__ void gear_chunk16(GearLane& g, uint4 v, uint32_t base, F on_match) {
    uint64_t t[16];
    t[0] = gear_lookup<0>(g, v.x); t[1] = gear_lookup<1>(g, v.x); t[2] = gear_lookup<2>(g, v.x); t[3] = gear_lookup<3>(g, v.x);
    t[4] = gear_lookup<0>(g, v.y); t[5] = gear_lookup<1>(g, v.y); t[6] = gear_lookup<2>(g, v.y); t[7] = gear_lookup<3>(g, v.y);
    t[8] = gear_lookup<0>(g, v.z); t[9] = gear_lookup<1>(g, v.z); t[10] = gear_lookup<2>(g, v.z); t[11] = gear_lookup<3>(g, v.z);
    t[12] = gear_lookup<0>(g, v.w); t[13] = gear_lookup<1>(g, v.w); t[14] = gear_lookup<2>(g, v.w); t[15] = gear_lookup<3>(g, v.w);
    const uint64_t h0 = g.h;
    uint64_t h = h0;
    uint32_t lowest = 0xffffffffu;  // min over the 16 positions of the hash's high word: one ALU op per byte
#pragma unroll
    for (int j = 0; j < 16; j++) {
        h = gear_roll(h, t[j], g.two);
        if (kTest) lowest = min(lowest, (uint32_t)(h >> 32));
    }
    g.h = h;
    if (kTest && lowest < (1u << (32 - kTopBits))) {  // rare (2^-12 per chunk on random data): replay for positions
        uint64_t r = h0;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            r = gear_roll(r, t[j], g.two);
            if ((uint32_t)(r >> 32) < (1u << (32 - kTopBits))) on_match(base + j);
        }
    }
}

// Bytes that are not part of a full 4 KiB row (the end of the buffer): plain global loads.
template <typename F>
__device__ __forceinline__ void gear_slow_span(const uint8_t* __restrict__ data, const uint64_t* s_table,
                                               uint32_t tab_stride, uint32_t lane, uint64_t begin, uint64_t end,
                                               uint32_t pos0, F on_match) {
    if (begin >= end) return;
    uint64_t h = 0;
    for (uint64_t i = begin >= 64 ? begin - 64 : 0; i < end; i++) {
        h = (h << 1) + s_table[(uint32_t)data[i] * tab_stride + lane];
        if (i >= begin && (uint32_t)(h >> 32) < (1u << (32 - kTopBits))) on_match(pos0 + (uint32_t)(i - begin));
    }
}

// Pass 1: count candidates per tile and stage up to kGearCap of them in ascending order.
__global__ void __launch_bounds__(kGearThreads, 1)
    gear_scan_kernel(const __grid_constant__ CUtensorMap tmap, const uint8_t* __restrict__ data, uint64_t total,
                     uint64_t rows, uint32_t n_tiles, uint32_t* __restrict__ tile_count,
                     uint32_t* __restrict__ tile_stage) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint64_t* s_table = reinterpret_cast<uint64_t*>(smem);
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    uint8_t* stages = smem + kGearOffStages + wid * (kGearStages * kGearBox);
    uint32_t* s_list = reinterpret_cast<uint32_t*>(smem + kGearOffLists) + wid * kGearCap;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kGearOffBars) + wid * kGearStages;
    uint32_t* s_cnt = reinterpret_cast<uint32_t*>(smem + kGearOffCnt) + wid;

    gear_load_table(s_table);
    if (lane == 0) {
        for (int i = 0; i < kGearStages; i++) mbar_init(&bars[i], 1);
        mbar_fence_init();
        *s_cnt = 0;
    }
    __syncthreads();

    GearLane g;
    g.tab = smem_u32(s_table);
    g.lane8 = lane * 8u;
    g.two = c_gear_two;
    auto emit = [&](uint32_t pos) {
        const uint32_t slot = atomicAdd(s_cnt, 1u);
        if (slot < kGearCap) s_list[slot] = pos;
    };
    const uint32_t swz = lane & 7u;
    uint32_t q = 0;  // boxes this warp has consumed so far: stage = q % kGearStages, parity = (q / kGearStages) & 1

    for (uint32_t tile = blockIdx.x * kGearWarps + wid; tile < n_tiles; tile += gridDim.x * kGearWarps) {
        const uint64_t row0 = (uint64_t)tile * 32u;
        const bool row_valid = row0 + lane < rows;
        if (row0 < rows) {
            auto issue = [&](uint32_t bi) {  // box bi of this tile into stage (q0 + bi) % kGearStages
                uint64_t* bar = &bars[(q + bi) % kGearStages];
                mbar_arrive_expect_tx(bar, kGearBox);
                if (bi == 0) tma_box_g2s(stages + ((q + bi) % kGearStages) * kGearBox, &tmap, kGearRow - 128, (int32_t)row0 - 1, bar);
                else tma_box_g2s(stages + ((q + bi) % kGearStages) * kGearBox, &tmap, (int32_t)(bi - 1) * 128, (int32_t)row0, bar);
            };
            if (lane == 0)
                for (uint32_t bi = 0; bi < kGearStages; bi++) issue(bi);
            g.h = 0;
#pragma unroll 1
            for (uint32_t bi = 0; bi < kGearBoxesPerTile; bi++) {
                const uint32_t st = (q + bi) % kGearStages;
                mbar_wait(&bars[st], ((q + bi) / kGearStages) & 1u);
                const uint8_t* rowp = stages + st * kGearBox + lane * 128u;
                if (row_valid) {
                    if (bi == 0) {  // the 64 bytes in front of the row: warm-up only
#pragma unroll
                        for (uint32_t c = 4; c < 8; c++)
                            gear_chunk16<false>(g, *reinterpret_cast<const uint4*>(rowp + ((c ^ swz) << 4)), 0u, emit);
                    } else {
                        const uint32_t pos = lane * kGearRow + (bi - 1) * 128u;
#pragma unroll 2
                        for (uint32_t c = 0; c < 8; c++)
                            gear_chunk16<true>(g, *reinterpret_cast<const uint4*>(rowp + ((c ^ swz) << 4)), pos + c * 16u, emit);
                    }
                }
                __syncwarp();
                if (lane == 0 && bi + kGearStages < kGearBoxesPerTile) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    issue(bi + kGearStages);
                }
            }
            q += kGearBoxesPerTile;
        }
        // the partial last row of the buffer, if it lies in this tile: 128-byte pieces, one per lane
        if (rows >= row0 && rows < row0 + 32u && total > rows * kGearRow) {
            const uint64_t tb = rows * kGearRow;
            const uint64_t b0 = tb + lane * 128u;
            const uint64_t b1 = b0 + 128u < total ? b0 + 128u : total;
            gear_slow_span(data, s_table, 32u, lane, b0, b1, (uint32_t)(rows - row0) * kGearRow + lane * 128u, emit);
        }
        __syncwarp();
        const uint32_t n = *s_cnt;
        if (lane == 0) tile_count[tile] = n;
        if (n <= kGearCap)
            for (uint32_t i = lane; i < n; i += 32) {  // rank sort (positions are distinct)
                const uint32_t mine = s_list[i];
                uint32_t rank = 0;
                for (uint32_t j = 0; j < n; j++) rank += s_list[j] < mine;
                tile_stage[(uint64_t)tile * kGearCap + rank] = mine;
            }
        __syncwarp();
        if (lane == 0) *s_cnt = 0;
        __syncwarp();
    }
}

// Pass 2: move staged candidates to their final place.  A tile that overflowed its stage (dense
// candidates — adversarial data only) is recomputed here with an ordered two-sweep emission, one
// 1 KiB piece per thread, straight from global memory.
constexpr int kCompactThreads = 128;
__global__ void __launch_bounds__(kCompactThreads)
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
    __shared__ uint64_t s_tab[256];
    __shared__ uint32_t s_cnt[kCompactThreads];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_tab[i] = c_gear[i];
    __syncthreads();
    constexpr uint32_t piece = kGearTile / kCompactThreads;  // 1 KiB
    uint64_t b0 = tile_start + threadIdx.x * piece;
    uint64_t b1 = b0 + piece;
    if (b0 > total) b0 = total;
    if (b1 > total) b1 = total;
    uint32_t mine = 0;
    gear_slow_span(data, s_tab, 1u, 0u, b0, b1, 0u, [&](uint32_t) { mine++; });
    s_cnt[threadIdx.x] = mine;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t j = 0; j < threadIdx.x; j++) before += s_cnt[j];
    uint64_t w = base + before;
    gear_slow_span(data, s_tab, 1u, 0u, b0, b1, 0u, [&](uint32_t pos) {
        if (w < cand_cap) cand[w] = b0 + pos;
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

// Small inputs (the per-file and per-segment arrays of small batches): one CTA, one launch.
__global__ void __launch_bounds__(1024) scan_small_kernel(const uint32_t* __restrict__ in, uint64_t n,
                                                           uint64_t* __restrict__ out, uint64_t* __restrict__ total_out) {
    __shared__ uint64_t s_warp[32];
    uint64_t carry = 0;
    for (uint64_t start = 0; start <= n; start += 1024) {
        const uint64_t i = start + threadIdx.x;
        const uint64_t v = i < n ? in[i] : 0;
        uint64_t tot;
        const uint64_t ex = block_exclusive_scan(v, s_warp, &tot);
        if (i <= n) out[i] = carry + ex;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

int launch_scan_u32(const uint32_t* in, uint64_t* out, uint64_t n, uint64_t* partial, uint64_t* total_out,
                    cudaStream_t st) {
    if (n <= 2048) {  // measured: one CTA takes 42 us for 16 k entries, the three-kernel form 12 us
        scan_small_kernel<<<1, 1024, 0, st>>>(in, n, out, total_out);
        return 1;
    }
    // n+1 outputs: size the grid on n+1 so the last entry is always written
    const uint64_t nblocks = scan_blocks(n + 1);
    scan_block_sums_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(in, n, partial);
    scan_partials_kernel<<<1, 1024, 0, st>>>(partial, nblocks, total_out);
    scan_apply_kernel<<<(unsigned)nblocks, kScanThreads, 0, st>>>(in, n, partial, out);
    return 3;
}

// ------------------------------------------------------------------------- cooperative-launch gate
namespace {
struct CoopGate {
    std::mutex mu;
    cudaEvent_t done = nullptr;
};
CoopGate g_coop_gate[64];
}  // namespace

cudaError_t launch_cooperative_gated(const void* kernel, dim3 grid, dim3 block, void** args, size_t smem, cudaStream_t st) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    CoopGate& g = g_coop_gate[dev];
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.done) {
        e = cudaEventCreateWithFlags(&g.done, cudaEventDisableTiming);
        if (e != cudaSuccess) return e;
    } else {
        e = cudaStreamWaitEvent(st, g.done, 0);
        if (e != cudaSuccess) return e;
    }
    e = cudaLaunchCooperativeKernel(kernel, grid, block, args, smem, st);
    if (e != cudaSuccess) return e;
    return cudaEventRecord(g.done, st);
}

// ------------------------------------------------------------------------------------ launcher
// Called once per context with its device current (function attributes are per device).
void init_gear_kernels() {
    cudaFuncSetAttribute(gear_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGearSmem);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault, &qres) ==
                cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

int launch_gear(const PipelineBuffers& b, cudaStream_t st, cudaError_t* err) {
    *err = cudaSuccess;
    if (b.n_tiles == 0) {
        cudaMemsetAsync(b.tile_off, 0, sizeof(uint64_t), st);
        cudaMemsetAsync(b.scalars + kScalarCand, 0, sizeof(uint64_t), st);
        return 0;
    }
    const uint64_t rows = b.total_bytes / kGearRow;
    alignas(64) CUtensorMap tmap;
    memset(&tmap, 0, sizeof tmap);
    if (rows) {
        EncodeTiledFn enc = encode_tiled_fn();
        if (!enc) {
            *err = cudaErrorSymbolNotFound;
            return 0;
        }
        const cuuint64_t gdim[2] = {kGearRow, rows};
        const cuuint64_t gstride[1] = {kGearRow};
        const cuuint32_t box[2] = {128, 32};
        const cuuint32_t estride[2] = {1, 1};
        CUresult r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)b.data, gdim, gstride, box, estride,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            *err = cudaErrorInvalidValue;
            return 0;
        }
    }
    unsigned ctas = (b.n_tiles + kGearWarps - 1) / kGearWarps;
    if (ctas > b.sm_count) ctas = b.sm_count;
    if (b.ev_kernel[0]) cudaEventRecord(b.ev_kernel[0], st);
    gear_scan_kernel<<<ctas, kGearThreads, kGearSmem, st>>>(tmap, b.data, b.total_bytes, rows, b.n_tiles, b.tile_count,
                                                            b.tile_stage);
    if (b.ev_kernel[1]) cudaEventRecord(b.ev_kernel[1], st);
    int n = 1 + launch_scan_u32(b.tile_count, b.tile_off, b.n_tiles, b.scan_partial, b.scalars + kScalarCand, st);
    gear_compact_kernel<<<b.n_tiles, kCompactThreads, 0, st>>>(b.data, b.total_bytes, b.tile_count, b.tile_stage,
                                                               b.tile_off, b.cand, b.cand_cap);
    return n + 1;
}

}  // namespace xetgpu
