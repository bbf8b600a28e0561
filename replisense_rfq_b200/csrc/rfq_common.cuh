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

#define RFQ_TILE 512u          // bytes per warp-tile (32 lanes x 16 B)
#define RFQ_ALIGN 16u          // documents start on 16-byte boundaries inside the arena
