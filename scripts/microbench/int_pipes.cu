// INT32 issue-roof microbenchmark for sm_100a: lane-ops per clock per SM for the instruction kinds the
// hashing kernels are made of, alone and mixed.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// -o int_pipes int_pipes.cu ; run on the B200.  Results are recorded in profiles/r1_int_pipes.md.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define ITERS 4096
#define CHAINS 8

#define KERNEL(name, BODY)                                                        \
    __global__ void name(uint32_t* out, uint32_t seed) {                          \
        uint32_t x[CHAINS], y = seed ^ threadIdx.x, z = seed * 7 + 1;             \
        _Pragma("unroll") for (int c = 0; c < CHAINS; c++) x[c] = seed + c * 977 + threadIdx.x; \
        for (int i = 0; i < ITERS; i++) {                                         \
            _Pragma("unroll") for (int c = 0; c < CHAINS; c++) { BODY }           \
        }                                                                         \
        uint32_t s = 0;                                                           \
        _Pragma("unroll") for (int c = 0; c < CHAINS; c++) s ^= x[c];             \
        out[blockIdx.x * blockDim.x + threadIdx.x] = s + y + z;                   \
    }

// one op per BODY unless noted
KERNEL(k_iadd3, asm volatile("add.u32 %0, %0, %1;" : "+r"(x[c]) : "r"(y));)
KERNEL(k_iadd3_3in, asm volatile("{.reg .u32 t; add.u32 t, %0, %1; add.u32 %0, t, %2;}" : "+r"(x[c]) : "r"(y), "r"(z));)
KERNEL(k_lop3, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(y), "r"(z));)
KERNEL(k_shf, asm volatile("shf.r.wrap.b32 %0, %0, %0, 7;" : "+r"(x[c]));)
KERNEL(k_prmt, asm volatile("prmt.b32 %0, %0, %1, 0x1032;" : "+r"(x[c]) : "r"(y));)
KERNEL(k_imad, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(y), "r"(z));)
KERNEL(k_imad_imm, asm volatile("mad.lo.u32 %0, %0, 4097, %1;" : "+r"(x[c]) : "r"(z));)
KERNEL(k_imad_iadd, asm volatile("mad.lo.u32 %0, %1, 1, %0;" : "+r"(x[c]) : "r"(y));)
KERNEL(k_imad_shl, asm volatile("mad.lo.u32 %0, %0, 128, %1;" : "+r"(x[c]) : "r"(z));)
KERNEL(k_imad_wide, asm volatile("{.reg .u64 t; .reg .u32 lo, hi; mul.wide.u32 t, %0, %1; mov.b64 {lo, hi}, t; add.u32 %0, lo, hi;}" : "+r"(x[c]) : "r"(y));)  // 2 ops
KERNEL(k_imad_hi, asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[c]) : "r"(y));)
KERNEL(k_rot_via_wide, asm volatile("{.reg .u64 t; .reg .u32 lo, hi; mul.wide.u32 t, %0, 128; mov.b64 {lo, hi}, t; mad.lo.u32 %0, lo, 1, hi;}" : "+r"(x[c]));)  // rotl 7: wide + add
KERNEL(k_mix_lop_imad, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(y), "r"(z));)  // 2 ops
KERNEL(k_mix_lop_imadiadd, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; mad.lo.u32 %0, %1, 1, %0;" : "+r"(x[c]) : "r"(y), "r"(z));)  // 2 ops
KERNEL(k_mix_2lop_imad, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96; shf.r.wrap.b32 %0, %0, %0, 7; mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(y), "r"(z));)  // 3 ops
KERNEL(k_g_alu, asm volatile("{.reg .u32 a; add.u32 a, %0, %1; add.u32 a, a, %2; xor.b32 %0, %0, a; prmt.b32 %0, %0, %0, 0x1032;}" : "+r"(x[c]) : "r"(y), "r"(z));)  // ~4 ops

typedef void (*kern_t)(uint32_t*, uint32_t);
struct Case { const char* name; kern_t k; int ops; };

int main() {
    Case cases[] = {{"add.u32 (IADD3)", k_iadd3, 1}, {"add3 (2x add -> IADD3?)", k_iadd3_3in, 1}, {"lop3", k_lop3, 1},
                    {"shf.r.wrap", k_shf, 1}, {"prmt", k_prmt, 1}, {"mad.lo reg", k_imad, 1}, {"mad.lo imm", k_imad_imm, 1},
                    {"mad.lo x*1+y", k_imad_iadd, 1}, {"mad.lo x*128+y", k_imad_shl, 1}, {"mul.wide + add", k_imad_wide, 2},
                    {"mul.hi", k_imad_hi, 1}, {"rotl via mul.wide+mad", k_rot_via_wide, 2}, {"lop3 + mad.lo", k_mix_lop_imad, 2},
                    {"lop3 + mad x*1+y", k_mix_lop_imadiadd, 2}, {"lop3 + shf + mad.lo", k_mix_2lop_imad, 3},
                    {"add3+xor+prmt", k_g_alu, 4}};
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    int sms = prop.multiProcessorCount;
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const int threads = 512, blocks = sms * 4;
    uint32_t* out;
    cudaMalloc(&out, (size_t)threads * blocks * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    printf("SMs %d, max clock %d kHz; %d blocks x %d threads, %d iters x %d chains\n", sms, clk_khz, blocks, threads, ITERS, CHAINS);
    printf("%-28s %10s %14s %18s\n", "case", "ms", "PTX ops/ns/SM", "PTX ops/clk/SM@max");
    for (auto& c : cases) {
        c.k<<<blocks, threads>>>(out, 1);
        cudaDeviceSynchronize();
        float best = 1e9;
        for (int r = 0; r < 3; r++) {
            cudaEventRecord(e0);
            c.k<<<blocks, threads>>>(out, 1);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        double ops = (double)threads * blocks * ITERS * CHAINS * c.ops;
        double per_ns_sm = ops / (best * 1e6) / sms;
        printf("%-28s %10.3f %14.2f %18.2f\n", c.name, best, per_ns_sm, per_ns_sm / (clk_khz / 1e6));
    }
    return 0;
}
