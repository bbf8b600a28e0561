// Device-wide exclusive scans used by every pipeline stage (tile counts, doc sizes, FSM
// transition vectors).  Reduce-then-scan in three launches: per-block reduce, one-CTA scan of
// the block partials, per-block scan with carry-in.  `Op` may be non-commutative (the FSM
// composition is), so order is preserved everywhere.
#pragma once
#include "rfq_common.cuh"
#include "launch.h"

namespace rfq {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;                       // items per thread
constexpr int SCAN_BLOCK = SCAN_THREADS * SCAN_ITEMS;  // 4096 items per CTA

struct SumU32 {
  typedef uint32_t T;
  __device__ static T identity() { return 0u; }
  __device__ static T op(T a, T b) { return a + b; }  // a happens before b
};
struct SumU64 {
  typedef unsigned long long T;
  __device__ static T identity() { return 0ull; }
  __device__ static T op(T a, T b) { return a + b; }
};

template <class Op>
__device__ typename Op::T warp_incl_scan(typename Op::T v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    typename Op::T o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v = Op::op(o, v);
  }
  return v;
}

// Block-wide exclusive scan of one value per thread; returns the exclusive prefix and writes
// the block total to *total (valid in every thread).
template <class Op>
__device__ typename Op::T block_excl_scan(typename Op::T v, typename Op::T* total) {
  typedef typename Op::T T;
  __shared__ T warp_tot[SCAN_THREADS / 32];
  __shared__ T block_tot;
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T incl = warp_incl_scan<Op>(v, lane);
  if (lane == 31) warp_tot[w] = incl;
  __syncthreads();
  if (w == 0) {
    T x = lane < SCAN_THREADS / 32 ? warp_tot[lane] : Op::identity();
    T xi = warp_incl_scan<Op>(x, lane);
    if (lane < SCAN_THREADS / 32) warp_tot[lane] = xi;
    if (lane == SCAN_THREADS / 32 - 1) block_tot = xi;
  }
  __syncthreads();
  T excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = Op::identity();
  if (w > 0) excl = Op::op(warp_tot[w - 1], excl);
  *total = block_tot;
  __syncthreads();
  return excl;
}

template <class Op>
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(
    const typename Op::T* __restrict__ in, typename Op::T* __restrict__ partial, uint32_t n) {
  typedef typename Op::T T;
  uint32_t base = blockIdx.x * SCAN_BLOCK + threadIdx.x * SCAN_ITEMS;
  T acc = Op::identity();
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i)
    if (base + i < n) acc = Op::op(acc, in[base + i]);
  T tot;
  block_excl_scan<Op>(acc, &tot);
  if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// One CTA scans the partials in place (exclusive); *grand receives the total.
template <class Op>
__global__ void __launch_bounds__(SCAN_THREADS) scan_partials_kernel(
    typename Op::T* __restrict__ partial, uint32_t nb, typename Op::T* __restrict__ grand) {
  typedef typename Op::T T;
  __shared__ T carry_s;
  if (threadIdx.x == 0) carry_s = Op::identity();
  __syncthreads();
  for (uint32_t start = 0; start < nb; start += SCAN_THREADS) {
    uint32_t i = start + threadIdx.x;
    T v = i < nb ? partial[i] : Op::identity();
    T tot;
    T ex = block_excl_scan<Op>(v, &tot);
    T carry = carry_s;
    if (i < nb) partial[i] = Op::op(carry, ex);
    __syncthreads();
    if (threadIdx.x == 0) carry_s = Op::op(carry, tot);
    __syncthreads();
  }
  if (threadIdx.x == 0 && grand) *grand = carry_s;
}

template <class Op>
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(
    const typename Op::T* __restrict__ in, const typename Op::T* __restrict__ partial,
    typename Op::T* __restrict__ out, uint32_t n) {
  typedef typename Op::T T;
  uint32_t base = blockIdx.x * SCAN_BLOCK + threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS];
  T acc = Op::identity();
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    v[i] = base + i < n ? in[base + i] : Op::identity();
    acc = Op::op(acc, v[i]);
  }
  T tot;
  T ex = block_excl_scan<Op>(acc, &tot);
  T run = Op::op(partial[blockIdx.x], ex);
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    if (base + i < n) out[base + i] = run;
    run = Op::op(run, v[i]);
  }
}

// Exclusive scan of in[0..n) into out[0..n); optional grand total to *grand (device pointer).
// `partial` is scratch of at least ceil(n / SCAN_BLOCK) elements. in may alias out.
template <class Op>
inline void exclusive_scan(const typename Op::T* in, typename Op::T* out, uint32_t n,
                           typename Op::T* partial, typename Op::T* grand,
                           cudaStream_t stream) {
  if (n == 0) {
    if (grand) cudaMemsetAsync(grand, 0, sizeof(typename Op::T), stream);
    return;
  }
  uint32_t nb = (n + SCAN_BLOCK - 1) / SCAN_BLOCK;
  RFQ_LAUNCH(scan_reduce_kernel<Op>, nb, SCAN_THREADS, 0, stream, in, partial, n);
  RFQ_LAUNCH(scan_partials_kernel<Op>, 1, SCAN_THREADS, 0, stream, partial, nb, grand);
  RFQ_LAUNCH(scan_apply_kernel<Op>, nb, SCAN_THREADS, 0, stream, in, partial, out, n);
}

}  // namespace rfq
