// BLAKE3 compression function for sm_100a, one compression per thread.
//
// Follows the BLAKE3 specification (§2.2 compression function, §2.3 flags); the keyed
// mode used by xet-core's merklehash::compute_data_hash (blake3::keyed_hash, crate blake3
// 1.8.2 pinned at /root/reference pkg/xet/Cargo.lock:189-192).
//
// Everything is fully unrolled: the seven rounds index the message through a compile-time
// schedule, so the 16 message words and 16 state words live in registers and no permute
// instruction is ever executed.  rot16/rot8 are byte permutes (PRMT), rot12/rot7 funnel
// shifts (SHF); both issue on the ALU pipe next to IADD3/LOP3.
#pragma once
#include <stdint.h>

namespace xetgpu {

enum : uint32_t {
    B3_CHUNK_START = 1u,
    B3_CHUNK_END = 2u,
    B3_PARENT = 4u,
    B3_ROOT = 8u,
    B3_KEYED_HASH = 16u,
};

#define XG_IV0 0x6A09E667u
#define XG_IV1 0xBB67AE85u
#define XG_IV2 0x3C6EF372u
#define XG_IV3 0xA54FF53Au

__device__ __forceinline__ uint32_t rotr16(uint32_t x) { return __byte_perm(x, 0, 0x1032); }
__device__ __forceinline__ uint32_t rotr8(uint32_t x) { return __byte_perm(x, 0, 0x0321); }
__device__ __forceinline__ uint32_t rotr12(uint32_t x) { return __funnelshift_r(x, x, 12); }
__device__ __forceinline__ uint32_t rotr7(uint32_t x) { return __funnelshift_r(x, x, 7); }

// Pipe balance (measured on B200, profiles/r1_int_pipes.txt): LOP3/SHF/PRMT/IADD3 issue on the ALU
// pipe at 64 lanes/clk/SM, IMAD on the FMA pipe at another 64 lanes/clk/SM, and the two overlap.
// A G function is 4 xors + 4 rotates (ALU only) + 6 additions; writing every addition as
// x * 1 + y with a multiplier the assembler cannot see through (a __constant__ it must load) keeps
// all eight two-input additions on the FMA pipe: 8 ALU + 8 FMA slots per G instead of 10 + 2.
// (a + m) is formed first — it does not wait for b — so the dependent chain per G stays 12 deep.
static __constant__ uint32_t c_b3_one = 1u;

__device__ __forceinline__ uint32_t fma_add(uint32_t x, uint32_t y, uint32_t one) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(one), "r"(y));
    return r;
}

// A rotate can also run on the FMA pipe: x * 2^(32-n) as a 64-bit product holds x >> n in its high word and
// x << (32-n) in its low word; their sum is the rotation.  IMAD.WIDE issues at half rate, so this costs
// three FMA slots instead of one ALU slot — worth it only for as many rotates as balance the two pipes
// (ALU ~533 slots vs FMA ~365 per compression): XG_GF uses it for the rotate by 16.
__device__ __forceinline__ uint32_t rotr_fma(uint32_t x, uint32_t pow2, uint32_t one) {
    uint32_t lo, hi, r;
    asm("{\n\t.reg .u64 t;\n\tmul.wide.u32 t, %2, %3;\n\tmov.b64 {%0, %1}, t;\n\t}" : "=r"(lo), "=r"(hi) : "r"(x), "r"(pow2));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(one), "r"(hi));
    return r;
}

#define XG_GF(a, b, c, d, mx, my)             \
    do {                                      \
        a = fma_add(fma_add(a, (mx), one), b, one); \
        d = rotr_fma(d ^ a, p16, one);        \
        c = fma_add(c, d, one);               \
        b = rotr12(b ^ c);                    \
        a = fma_add(fma_add(a, (my), one), b, one); \
        d = rotr8(d ^ a);                     \
        c = fma_add(c, d, one);               \
        b = rotr7(b ^ c);                     \
    } while (0)

#define XG_G(a, b, c, d, mx, my)              \
    do {                                      \
        a = fma_add(fma_add(a, (mx), one), b, one); \
        d = rotr16(d ^ a);                    \
        c = fma_add(c, d, one);               \
        b = rotr12(b ^ c);                    \
        a = fma_add(fma_add(a, (my), one), b, one); \
        d = rotr8(d ^ a);                     \
        c = fma_add(c, d, one);               \
        b = rotr7(b ^ c);                     \
    } while (0)

// Measured on B200: with 28 of the 448 rotates per compression on the FMA pipe the leaf kernel is 2.5 % SLOWER
// (8.69 -> 8.91 ms per 16 GB): the extra instruction per rotate costs more issue slots than the ALU slot it frees.
// Kept for reference, off by default.
#ifndef XG_FMA_ROT
#define XG_FMA_ROT 0
#endif
#if XG_FMA_ROT
#define XG_G0 XG_GF
#else
#define XG_G0 XG_G
#endif

// One round with message word indices given as literals.
#define XG_ROUND(i0, i1, i2, i3, i4, i5, i6, i7, i8, i9, i10, i11, i12, i13, i14, i15) \
    do {                                                                                \
        XG_G0(v0, v4, v8, v12, m[i0], m[i1]);                                           \
        XG_G(v1, v5, v9, v13, m[i2], m[i3]);                                            \
        XG_G0(v2, v6, v10, v14, m[i4], m[i5]);                                          \
        XG_G(v3, v7, v11, v15, m[i6], m[i7]);                                           \
        XG_G0(v0, v5, v10, v15, m[i8], m[i9]);                                          \
        XG_G(v1, v6, v11, v12, m[i10], m[i11]);                                         \
        XG_G0(v2, v7, v8, v13, m[i12], m[i13]);                                         \
        XG_G(v3, v4, v9, v14, m[i14], m[i15]);                                          \
    } while (0)

// cv: 8-word chaining value (updated in place with the first 8 output words — all this
// path ever needs: chaining values and 32-byte digests).  m: 16 message words.
__device__ __forceinline__ void b3_compress(uint32_t (&cv)[8], const uint32_t (&m)[16],
                                            uint32_t counter_lo, uint32_t counter_hi,
                                            uint32_t block_len, uint32_t flags) {
    uint32_t v0 = cv[0], v1 = cv[1], v2 = cv[2], v3 = cv[3];
    uint32_t v4 = cv[4], v5 = cv[5], v6 = cv[6], v7 = cv[7];
    uint32_t v8 = XG_IV0, v9 = XG_IV1, v10 = XG_IV2, v11 = XG_IV3;
    uint32_t v12 = counter_lo, v13 = counter_hi, v14 = block_len, v15 = flags;
    const uint32_t one = c_b3_one;
    const uint32_t p16 = c_b3_one << 16;  // 2^16, opaque to the assembler like `one`
    // message schedule = repeated application of the BLAKE3 permutation
    XG_ROUND(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
    XG_ROUND(2, 6, 3, 10, 7, 0, 4, 13, 1, 11, 12, 5, 9, 14, 15, 8);
    XG_ROUND(3, 4, 10, 12, 13, 2, 7, 14, 6, 5, 9, 0, 11, 15, 8, 1);
    XG_ROUND(10, 7, 12, 9, 14, 3, 13, 15, 4, 0, 11, 2, 5, 8, 1, 6);
    XG_ROUND(12, 13, 9, 11, 15, 10, 14, 8, 7, 2, 5, 3, 0, 1, 6, 4);
    XG_ROUND(9, 14, 11, 5, 8, 12, 15, 1, 13, 3, 0, 10, 2, 6, 4, 7);
    XG_ROUND(11, 15, 5, 0, 1, 9, 8, 6, 14, 10, 2, 12, 3, 4, 7, 13);
    cv[0] = v0 ^ v8;
    cv[1] = v1 ^ v9;
    cv[2] = v2 ^ v10;
    cv[3] = v3 ^ v11;
    cv[4] = v4 ^ v12;
    cv[5] = v5 ^ v13;
    cv[6] = v6 ^ v14;
    cv[7] = v7 ^ v15;
}

// Parent node: message = left CV || right CV, chaining input = key.
__device__ __forceinline__ void b3_parent(const uint32_t (&key)[8], const uint32_t (&l)[8],
                                          const uint32_t (&r)[8], bool root, uint32_t (&out)[8]) {
    uint32_t m[16];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        m[i] = l[i];
        m[8 + i] = r[i];
    }
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = key[i];
    b3_compress(out, m, 0u, 0u, 64u, B3_KEYED_HASH | B3_PARENT | (root ? B3_ROOT : 0u));
}

// merklehash DATA_KEY / INTERNAL_NODE_KEY as little-endian words.
#define XG_DATA_KEY_WORDS                                                                       \
    {0x77f59766u, 0xde50955bu, 0xaccb3531u, 0x1c1897a5u, 0x1021e49du, 0x582beb9bu, 0x4bb0d0b4u, \
     0x29f2ad93u}
#define XG_NODE_KEY_WORDS                                                                       \
    {0xc7c57e01u, 0x962947a5u, 0x666694fdu, 0xe6028ab4u, 0x6f53dd5du, 0xd26dc737u, 0xe65263f8u, \
     0x3f71534au}

}  // namespace xetgpu
