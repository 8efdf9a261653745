// rt_peaks.cuh - micro-benchmarks that MEASURE the denominators of the render path's rooflines on the device at hand
// ("the builder must measure it with an FMA micro-benchmark and use the measured value as the denominator"):
//   * FP32 FMA throughput of the CUDA cores (the intersection arithmetic: no tensor cores on this path),
//   * warp-instruction issue rate (what actually binds k_shadow / k_hit: 4 schedulers per SM, one warp instruction per cycle each),
//   * L1 data-pipe (LSU) wavefront rate (the 128-bit BVH node loads: a warp-wide LDG.128 is four 128-byte wavefronts).
// Every loop body is inline PTX so the instruction counts are exactly what is written here (checked with cuobjdump -sass).
#pragma once
#include <cuda_runtime.h>

#define PK_BLOCK 256
#define PK_BLOCKS_PER_SM 8          // 2048 resident threads per SM = 16 warps per scheduler

// 128 independent-enough FFMA per loop trip: 8 accumulator chains x 16 rounds (FFMA latency 4 cycles, 16 warps per scheduler)
__global__ void __launch_bounds__(PK_BLOCK, PK_BLOCKS_PER_SM) k_peak_fma(float* out, int trips, float b, float c) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    for (int t = 0; t < trips; t++) {
#pragma unroll
        for (int r = 0; r < 16; r++) {
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a0) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a1) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a2) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a3) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a4) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a5) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a6) : "f"(b), "f"(c));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a7) : "f"(b), "f"(c));
        }
    }
    const float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 123.456f) out[0] = s;          // never true in practice: keeps the chains alive
}
#define PK_FMA_PER_TRIP 128

// issue rate: FFMA (FMA pipe) and integer adds / logic ops (ALU pipe) interleaved 1:1, so neither pipe is the limit (each runs
// at half its rate) and the scheduler's one-instruction-per-cycle issue slot is.  128 instructions per loop trip.
__global__ void __launch_bounds__(PK_BLOCK, PK_BLOCKS_PER_SM) k_peak_issue(float* out, int trips, float b, float c, unsigned m) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    unsigned i0 = threadIdx.x, i1 = i0 + m, i2 = i0 + 2, i3 = i0 ^ m;
    for (int t = 0; t < trips; t++) {
#pragma unroll
        for (int r = 0; r < 16; r++) {
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a0) : "f"(b), "f"(c));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(i0) : "r"(i1));      // add / xor recurrences ptxas cannot fold into one instruction
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a1) : "f"(b), "f"(c));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(i2) : "r"(i3));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a2) : "f"(b), "f"(c));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(i1) : "r"(i0));
            asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a3) : "f"(b), "f"(c));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(i3) : "r"(i2));
        }
    }
    const float s = (a0 + a1) + (a2 + a3) + (float)((i0 ^ i1) + (i2 ^ i3));
    if (s == 123.456f) out[0] = s;
}
#define PK_INST_PER_TRIP 128

// L1 data pipe: every lane reads its own 16 bytes (one warp instruction = 512 contiguous, 128-byte aligned bytes = 4 wavefronts)
// of a 64 KB table that stays in L1.  16 loads per trip at 16 different addresses that also move with the trip number (ptxas
// merges plain loads of one address and hoists loop-invariant ones), none of them dependent on loaded data.
#define PK_LSU_TABLE_BYTES (65536 + 256)
__global__ void __launch_bounds__(PK_BLOCK, PK_BLOCKS_PER_SM) k_peak_lsu(const float4* __restrict__ table, float* out, int trips) {
    const float4* p = table + threadIdx.x;                                               // 8 warps x 512 B = 4 KB window per block
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    for (int t = 0; t < trips; t++) {
        const float4* q = p + ((t & 1) << 3);
#pragma unroll
        for (int r = 0; r < 16; r++) {
            float4 v;
            asm volatile("ld.global.ca.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(q + (r << 8)) : "memory");
            s0 += v.x; s1 += v.y; s2 += v.z; s3 += v.w;
        }
    }
    const float s = (s0 + s1) + (s2 + s3);
    if (s == 123.456f) out[0] = s;
}
#define PK_LOADS_PER_TRIP 16
