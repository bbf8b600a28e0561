// Common macros for code shared between device kernels and the host-side unit-test harness.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __CUDACC__
#define RFQ_HD __host__ __device__ __forceinline__
#define RFQ_D __device__ __forceinline__
// rare, register-hungry paths (big-integer formatting, exact strtod): kept out of line so they
// do not set the register count of the kernels that call them
#define RFQ_HD_COLD static __host__ __device__ __noinline__
#else
#define RFQ_HD inline
#define RFQ_D inline
#define RFQ_HD_COLD inline
#endif

// Exact IEEE double ops with NO fma contraction (nvcc contracts a*b+c by default, which would
// change the rounding of pandas' `number = number * 10. + digit`).
#if defined(__CUDA_ARCH__)
#define RFQ_DMUL(a, b) __dmul_rn((a), (b))
#define RFQ_DADD(a, b) __dadd_rn((a), (b))
#define RFQ_DDIV(a, b) __ddiv_rn((a), (b))
#else
static inline double rfq_dmul_(double a, double b) { volatile double r = a * b; return r; }
static inline double rfq_dadd_(double a, double b) { volatile double r = a + b; return r; }
static inline double rfq_ddiv_(double a, double b) { volatile double r = a / b; return r; }
#define RFQ_DMUL(a, b) rfq_dmul_((a), (b))
#define RFQ_DADD(a, b) rfq_dadd_((a), (b))
#define RFQ_DDIV(a, b) rfq_ddiv_((a), (b))
#endif

#ifdef __CUDACC__
// largest d in [0, n) with table[d] <= x (table ascending, table[0] <= x), found by the whole warp: 32 probes
// per step, so a table of a million documents takes four dependent loads instead of twenty
__device__ __forceinline__ uint32_t warp_find_doc(const uint32_t* __restrict__ table, uint32_t n, uint32_t x, int lane) {
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    const uint32_t step = (hi - lo + 31u) / 32u;
    const uint32_t idx = lo + (uint32_t)lane * step;
    const uint32_t v = idx < hi ? __ldg(table + idx) : 0xFFFFFFFFu;
    const unsigned le = __ballot_sync(0xffffffffu, v <= x);   // a prefix of the lanes (lane 0 always)
    const uint32_t k = (uint32_t)__popc(le) - 1u;
    lo += k * step;
    hi = min(hi, lo + step);
  }
  return lo;
}
#endif

#define RFQ_TILE 512u          // bytes per warp-tile (32 lanes x 16 B)
#define RFQ_ALIGN 16u          // documents start on 16-byte boundaries inside the arena
