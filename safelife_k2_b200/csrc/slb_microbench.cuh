// slb_microbench.cuh — integer-pipe issue-rate microbenchmark: the measured denominator of the
// compute-bound (fused rollout, side effects) roofline, SURVEY.md 8(d).
#pragma once
#include <stdint.h>

namespace slb {

template <int KIND>
__global__ void __launch_bounds__(256) int_peak_kernel(uint32_t *sink, int iters) {
    uint32_t a[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = threadIdx.x * 2654435761u + j * 40503u + blockIdx.x;
    const uint32_t k1 = 0x9E3779B9u ^ blockIdx.x, k2 = 0x85EBCA6Bu + threadIdx.x;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (KIND == 0 || (j & 1) == 0)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(k1), "r"(k2));
            else
                asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(k1), "r"(k2));
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) x ^= a[j];
    if (x == 0x12345u) sink[0] = x;  // practically never: keeps the chains alive
}

}  // namespace slb
