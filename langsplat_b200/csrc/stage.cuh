// stage.cuh — warp-cooperative, fully coalesced movement of per-Gaussian rows between global and shared memory.
//
// The API fixes the tensors as arrays of structures (means [P,3], SH [P,16,3], gradients likewise), so a thread that
// walks its own row issues 32 different cache lines per load instruction.  Here a warp moves the rows of its 32
// consecutive Gaussians as one contiguous byte range: lane l handles 16-byte (or 8-byte) chunk it*32+l, rows that are
// not wanted (culled Gaussians) are predicated off chunk by chunk, and the shared-memory copy is padded by one chunk
// per row so that each lane can then read or write ITS row with conflict-free 128-bit accesses.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lsr {

// smem chunk index of (row g, part) for rows of ROW chunks padded to ROW+1
template <int ROW>
__device__ __forceinline__ int row_slot(int g, int part) { return g * (ROW + 1) + part; }

// Loads rows [0,32) of ROW float4 each from `src` (16-B aligned, row g at src + g*ROW) into `dst` (padded layout).
// `want` bit g set = row g is loaded.  All 32 lanes must call.
template <int ROW>
__device__ __forceinline__ void warp_load_rows(const float4* __restrict__ src, float4* dst, uint32_t want, int lane) {
#pragma unroll
  for (int it = 0; it < ROW; it++) {
    const int c = it * 32 + lane;
    const int g = c / ROW, part = c - g * ROW;
    if ((want >> g) & 1u) dst[row_slot<ROW>(g, part)] = __ldg(src + c);
  }
}

// ---- asynchronous variants (LDGSTS / cp.async: global -> shared without a register round trip) ----
// Issued early, they let a warp's row traffic fly while it does its per-Gaussian arithmetic, which is what keeps the
// HBM-bound per-Gaussian kernels fed at the modest occupancy their register use allows.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem_dst, const void* gmem_src) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int ROW>
__device__ __forceinline__ void warp_load_rows_async(const float4* __restrict__ src, float4* dst, uint32_t want,
                                                     int lane) {
#pragma unroll
  for (int it = 0; it < ROW; it++) {
    const int c = it * 32 + lane;
    const int g = c / ROW, part = c - g * ROW;
    if ((want >> g) & 1u) cp_async16(dst + row_slot<ROW>(g, part), src + c);
  }
}

// flat array of `nfloats` floats starting at src + first_float (see stage_flat); 16-byte chunks when aligned
template <int THREADS>
__device__ __forceinline__ void stage_flat_async(const float* __restrict__ src, float* dst, size_t first_float,
                                                 int nfloats, int tid) {
  const float* base = src + first_float;
  if ((((uintptr_t)base) & 15) == 0) {
    const int n4 = nfloats >> 2;
    for (int i = tid; i < n4; i += THREADS) cp_async16(reinterpret_cast<float4*>(dst) + i, reinterpret_cast<const float4*>(base) + i);
    for (int i = (n4 << 2) + tid; i < nfloats; i += THREADS) cp_async4(dst + i, base + i);
  } else {
    for (int i = tid; i < nfloats; i += THREADS) cp_async4(dst + i, base + i);
  }
}

// Stores rows from the padded smem layout to `dst`; rows whose `want` bit is clear are written as zeros when
// `zero_others` (bit set in `exists` = the row is inside the array at all).
template <int ROW>
__device__ __forceinline__ void warp_store_rows(float4* __restrict__ dst, const float4* src, uint32_t want,
                                                uint32_t exists, bool zero_others, int lane) {
#pragma unroll
  for (int it = 0; it < ROW; it++) {
    const int c = it * 32 + lane;
    const int g = c / ROW, part = c - g * ROW;
    if ((want >> g) & 1u) dst[c] = src[row_slot<ROW>(g, part)];
    else if (zero_others && ((exists >> g) & 1u)) dst[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// Cooperative, coalesced copy of `count` consecutive rows of K floats (K = 3: xyz triplets) into smem as a flat array.
template <int THREADS>
__device__ __forceinline__ void stage_flat(const float* __restrict__ src, float* dst, size_t first_float, int nfloats,
                                           int tid) {
  const float* base = src + first_float;
  if ((((uintptr_t)base) & 15) == 0) {
    const int n4 = nfloats >> 2;
    for (int i = tid; i < n4; i += THREADS) reinterpret_cast<float4*>(dst)[i] = __ldg(reinterpret_cast<const float4*>(base) + i);
    for (int i = (n4 << 2) + tid; i < nfloats; i += THREADS) dst[i] = __ldg(base + i);
  } else {
    for (int i = tid; i < nfloats; i += THREADS) dst[i] = __ldg(base + i);
  }
}

// Coalesced write-out of a flat smem array (rows of K floats, `count` rows) to global.
template <int THREADS>
__device__ __forceinline__ void unstage_flat(float* __restrict__ dst, const float* src, size_t first_float, int nfloats,
                                             int tid) {
  float* base = dst + first_float;
  if ((((uintptr_t)base) & 15) == 0) {
    const int n4 = nfloats >> 2;
    for (int i = tid; i < n4; i += THREADS) reinterpret_cast<float4*>(base)[i] = reinterpret_cast<const float4*>(src)[i];
    for (int i = (n4 << 2) + tid; i < nfloats; i += THREADS) base[i] = src[i];
  } else {
    for (int i = tid; i < nfloats; i += THREADS) base[i] = src[i];
  }
}

}  // namespace lsr
