// preprocess_fwd.cu — K1': per-Gaussian cull, cov3D, EWA cov2D, conic, radius, tile rect, SH colour.
// Replaces FORWARD::preprocess / preprocessCUDA (RAST/cuda_rasterizer/forward.cu:156-257) and in_frustum
// (RAST/cuda_rasterizer/auxiliary.h:139-164).  HBM-bound: algorithmic bytes 20*P + 291*V.
//
// Layout choices (B200): inputs are the API's AoS tensors; a CTA of 256 threads stages its 256 xyz / scale
// triplets through shared memory with fully coalesced 128-bit loads, reads quaternions as float4, and pulls each
// visible Gaussian's 192-byte SH row with 12 x LDG.128.  All per-instance data the blend kernels need is written
// as one 48-byte GeomRec.
#include "common.cuh"
#include "gs_math.cuh"
#include "stage.cuh"

namespace lsr {

namespace {

constexpr int PRE_THREADS = 256;

__device__ __forceinline__ float3 sh_to_rgb(int deg, const float* __restrict__ sh, float x, float y, float z,
                                            int stride /*floats between coefficients = 3*/) {
  // RAST/cuda_rasterizer/forward.cu:21-72
  float3 res;
  float* rp = &res.x;
#pragma unroll
  for (int ch = 0; ch < 3; ch++) {
    const float* s = sh + ch;
    float v = LSR_SH_C0 * s[0];
    if (deg > 0) {
      v = v - LSR_SH_C1 * y * s[1 * stride] + LSR_SH_C1 * z * s[2 * stride] - LSR_SH_C1 * x * s[3 * stride];
      if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        v = v + LSR_SH_C2_0 * xy * s[4 * stride] + LSR_SH_C2_1 * yz * s[5 * stride] +
            LSR_SH_C2_2 * (2.0f * zz - xx - yy) * s[6 * stride] + LSR_SH_C2_3 * xz * s[7 * stride] +
            LSR_SH_C2_4 * (xx - yy) * s[8 * stride];
        if (deg > 2) {
          v = v + LSR_SH_C3_0 * y * (3.0f * xx - yy) * s[9 * stride] + LSR_SH_C3_1 * xy * z * s[10 * stride] +
              LSR_SH_C3_2 * y * (4.0f * zz - xx - yy) * s[11 * stride] +
              LSR_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * s[12 * stride] +
              LSR_SH_C3_4 * x * (4.0f * zz - xx - yy) * s[13 * stride] + LSR_SH_C3_5 * z * (xx - yy) * s[14 * stride] +
              LSR_SH_C3_6 * x * (xx - 3.0f * yy) * s[15 * stride];
        }
      }
    }
    rp[ch] = v + 0.5f;
  }
  return res;
}

constexpr int SH_ROW = 12;  // float4 chunks per [16,3] SH row
constexpr size_t PRE_SH_SMEM = (size_t)(PRE_THREADS / 32) * 32 * (SH_ROW + 1) * sizeof(float4);

// LSR_PRE_STAGE_ALL = 1 also stages scales / language features through shared memory (6 KB more per CTA: 3 CTAs per
// SM); 0 reads those two 12-byte triplets directly (a warp's 384 contiguous bytes) and fits 4 CTAs per SM.
#ifndef LSR_PRE_STAGE_ALL
#define LSR_PRE_STAGE_ALL 0
#endif
#ifndef LSR_PRE_MINB
#define LSR_PRE_MINB 4
#endif
template <bool SH16>
__global__ void __launch_bounds__(PRE_THREADS, LSR_PRE_MINB)
preprocess_fwd_kernel(int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
                      float scale_modifier, const float* __restrict__ rotations, const float* __restrict__ opacities,
                      const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                      const float* __restrict__ colors_precomp, const float* __restrict__ lang,
                      const float* __restrict__ viewmatrix, const float* __restrict__ projmatrix,
                      const float* __restrict__ campos, int W, int H, float tan_fovx, float tan_fovy, float focal_x,
                      float focal_y, int gx, int gy, int prefiltered, int32_t* __restrict__ radii, GeomRec* __restrict__ rec,
                      float* __restrict__ depths, uint32_t* __restrict__ tiles_touched, float* __restrict__ cov3Ds,
                      uint8_t* __restrict__ clamped, uint32_t* __restrict__ ctrl) {
  extern __shared__ __align__(16) float4 s_dyn[];  // per warp: 32 padded SH rows; reused to stage rec / cov3D
  __shared__ __align__(16) float s_xyz[PRE_THREADS * 3];
#if LSR_PRE_STAGE_ALL
  __shared__ __align__(16) float s_scl[PRE_THREADS * 3];
  __shared__ __align__(16) float s_lng[PRE_THREADS * 3];
#endif
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  __shared__ uint32_t s_negmin, s_max, s_tiles;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int first = blockIdx.x * PRE_THREADS;
  const int count = min(PRE_THREADS, P - first);
  if (tid < 16) {
    s_view[tid] = viewmatrix[tid];
    s_proj[tid] = projmatrix[tid];
  }
  if (tid < 3) s_cam[tid] = campos[tid];
  if (tid == 0) { s_negmin = 0u; s_max = 0u; s_tiles = 0u; }
  stage_flat_async<PRE_THREADS>(means3D, s_xyz, (size_t)first * 3, count * 3, tid);
#if LSR_PRE_STAGE_ALL
  if (scales) stage_flat_async<PRE_THREADS>(scales, s_scl, (size_t)first * 3, count * 3, tid);
  if (lang) stage_flat_async<PRE_THREADS>(lang, s_lng, (size_t)first * 3, count * 3, tid);
#endif
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  const bool valid = tid < count;
  const int idx = first + tid;
  const float mx = valid ? s_xyz[3 * tid] : 0.f, my = valid ? s_xyz[3 * tid + 1] : 0.f,
              mz = valid ? s_xyz[3 * tid + 2] : 0.f;
#if !LSR_PRE_STAGE_ALL
  float sc0 = 0.f, sc1 = 0.f, sc2 = 0.f;  // requested now, used after the projection arithmetic
  if (valid && scales) {
    sc0 = __ldg(scales + 3 * (size_t)idx); sc1 = __ldg(scales + 3 * (size_t)idx + 1); sc2 = __ldg(scales + 3 * (size_t)idx + 2);
  }
#endif
  float4* wsm = s_dyn + (size_t)warp * 32 * (SH_ROW + 1);
  // Speculative prefetch of the SH rows (cp.async), so that they fly during the projection arithmetic below: a
  // Gaussian in front of the camera whose centre projects inside the image (plus a 16-pixel band) is almost surely
  // visible.  Rows guessed wrongly cost a little bandwidth; visible Gaussians the guess missed (large splats centred
  // off-screen) are fetched synchronously afterwards, so the result never depends on the guess.
  uint32_t spec_mask = 0u;
  if (SH16 && colors_precomp == nullptr) {
    bool guess = false;
    if (valid) {
      const float vz = tp_row(s_view, 2, mx, my, mz);
      const float hw = tp_row(s_proj, 3, mx, my, mz) + 0.0000001f;
      const float ax = ((tp_row(s_proj, 0, mx, my, mz) / hw + 1.f) * (float)W - 1.f) * 0.5f;
      const float ay = ((tp_row(s_proj, 1, mx, my, mz) / hw + 1.f) * (float)H - 1.f) * 0.5f;
      guess = vz > 0.2f && ax > -16.f && ax < (float)W + 16.f && ay > -16.f && ay < (float)H + 16.f;
    }
    spec_mask = __ballot_sync(0xffffffffu, guess);
    warp_load_rows_async<SH_ROW>(reinterpret_cast<const float4*>(shs) + (size_t)(first + warp * 32) * SH_ROW, wsm,
                                 spec_mask, lane);
    cp_async_commit();
  }

  int radius_out = 0;
  uint32_t tiles_out = 0;
  float viewz = 0.f, px = 0.f, py = 0.f, conx = 0.f, cony = 0.f, conz = 0.f;
  float cov3D[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (valid) do {
    // in_frustum: the only cull is p_view.z <= 0.2 (auxiliary.h:152-154)
    viewz = tp_row(s_view, 2, mx, my, mz);
    if (viewz <= 0.2f) {
      if (prefiltered) atomicOr(&ctrl[CTRL_ERROR], 1u);
      break;
    }
    const float hx = tp_row(s_proj, 0, mx, my, mz);
    const float hy = tp_row(s_proj, 1, mx, my, mz);
    const float hw = tp_row(s_proj, 3, mx, my, mz);
    const float p_w = __fdiv_rn(1.0f, __fadd_rn(hw, 0.0000001f));
    const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
    if (cov3D_precomp) {
#pragma unroll
      for (int i = 0; i < 6; i++) cov3D[i] = __ldg(cov3D_precomp + (size_t)idx * 6 + i);
    } else {
      const float4 q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
#if LSR_PRE_STAGE_ALL
      compute_cov3D(s_scl[3 * tid], s_scl[3 * tid + 1], s_scl[3 * tid + 2], scale_modifier, q.x, q.y, q.z, q.w,
                    cov3D);
#else
      compute_cov3D(sc0, sc1, sc2, scale_modifier, q.x, q.y, q.z, q.w, cov3D);
#endif
    }
    Cov2D c2;
    compute_cov2D(mx, my, mz, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, s_view, c2);
    const float det = __fmaf_rn(c2.a, c2.c, -__fmul_rn(c2.b, c2.b));
    if (det == 0.0f) break;
    const float det_inv = __fdiv_rn(1.f, det);
    conx = __fmul_rn(c2.c, det_inv); cony = __fmul_rn(-c2.b, det_inv); conz = __fmul_rn(c2.a, det_inv);
    const float mid = __fmul_rn(0.5f, __fadd_rn(c2.a, c2.c));
    const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
    const float lambda1 = __fadd_rn(mid, sq), lambda2 = __fadd_rn(mid, -sq);
    const int my_radius = __float2int_rz(ceilf(__fmul_rn(3.f, __fsqrt_rn(fmaxf(lambda1, lambda2)))));
    px = ndc2pix(projx, W); py = ndc2pix(projy, H);
    uint32_t x0, y0, x1, y1;
    get_rect(px, py, my_radius, gx, gy, x0, y0, x1, y1);
    const uint32_t nt = (x1 - x0) * (y1 - y0);
    if (nt == 0) break;
    radius_out = my_radius;
    tiles_out = nt;
  } while (0);
  const bool vis = radius_out > 0;
  const uint32_t vis_mask = __ballot_sync(0xffffffffu, vis);

  // ---- colour: SH rows of the warp's visible Gaussians (prefetched above; the missed ones are fetched now) ----
  float3 col = make_float3(0.f, 0.f, 0.f);
  uint8_t clamp_bits = 0;
  if (colors_precomp == nullptr) {
    if (SH16) {
      const uint32_t missed = vis_mask & ~spec_mask;
      if (missed)
        warp_load_rows<SH_ROW>(reinterpret_cast<const float4*>(shs) + (size_t)(first + warp * 32) * SH_ROW, wsm, missed,
                               lane);
      cp_async_wait<0>();
      __syncwarp();
    }
    if (vis) {
      float dx = mx - s_cam[0], dy = my - s_cam[1], dz = mz - s_cam[2];
      const float len = sqrtf(dx * dx + dy * dy + dz * dz);
      dx = dx / len; dy = dy / len; dz = dz / len;
      if (SH16) {
        // RGB[ch] = sum_k basis[k] * sh[k][ch] + 0.5, coefficients consumed straight from the staged row (the terms of
        // each channel in the reference's order, forward.cu:31-68), so the 48 floats never sit in registers
        float bs[16];
        sh_basis(D, dx, dy, dz, bs);
        const int nfl = 3 * (D + 1) * (D + 1);  // floats of the row that belong to active coefficients
        float acc[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < SH_ROW; i++) {
          const float4 t = wsm[row_slot<SH_ROW>(lane, i)];
          const float e[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int f = 4 * i + j;
            if (f < nfl) acc[f % 3] = f < 3 ? bs[0] * e[j] : acc[f % 3] + bs[f / 3] * e[j];
          }
        }
        col = make_float3(acc[0] + 0.5f, acc[1] + 0.5f, acc[2] + 0.5f);
      } else {
        col = sh_to_rgb(D, shs + (size_t)idx * M * 3, dx, dy, dz, 3);
      }
      clamp_bits = (col.x < 0.f ? 1 : 0) | (col.y < 0.f ? 2 : 0) | (col.z < 0.f ? 4 : 0);
      col.x = fmaxf(col.x, 0.f); col.y = fmaxf(col.y, 0.f); col.z = fmaxf(col.z, 0.f);
    }
    __syncwarp();
  } else if (vis) {
    col.x = __ldg(colors_precomp + (size_t)idx * 3);
    col.y = __ldg(colors_precomp + (size_t)idx * 3 + 1);
    col.z = __ldg(colors_precomp + (size_t)idx * 3 + 2);
  }

  // ---- outputs: the 48-byte records and cov3D go out through shared memory as coalesced 16-byte chunks ----
  if (vis) {
    float f0 = 0.f, f1 = 0.f, f2 = 0.f;
#if LSR_PRE_STAGE_ALL
    if (lang) { f0 = s_lng[3 * tid]; f1 = s_lng[3 * tid + 1]; f2 = s_lng[3 * tid + 2]; }
#else
    if (lang) { f0 = __ldg(lang + 3 * (size_t)idx); f1 = __ldg(lang + 3 * (size_t)idx + 1); f2 = __ldg(lang + 3 * (size_t)idx + 2); }
#endif
    wsm[row_slot<3>(lane, 0)] = make_float4(px, py, conx, cony);
    wsm[row_slot<3>(lane, 1)] = make_float4(conz, __ldg(opacities + idx), col.x, col.y);
    wsm[row_slot<3>(lane, 2)] = make_float4(col.z, f0, f1, f2);
  }
  __syncwarp();
  warp_store_rows<3>(reinterpret_cast<float4*>(rec + first + warp * 32), wsm, vis_mask, 0u, false, lane);
  __syncwarp();
  if (!cov3D_precomp) {
    float2* w2 = reinterpret_cast<float2*>(wsm);
    if (vis) {
      w2[lane * 3] = make_float2(cov3D[0], cov3D[1]);
      w2[lane * 3 + 1] = make_float2(cov3D[2], cov3D[3]);
      w2[lane * 3 + 2] = make_float2(cov3D[4], cov3D[5]);
    }
    __syncwarp();
    float2* dst = reinterpret_cast<float2*>(cov3Ds + (size_t)(first + warp * 32) * 6);
#pragma unroll
    for (int it = 0; it < 3; it++) {
      const int c = it * 32 + lane;
      if ((vis_mask >> (c / 3)) & 1u) dst[c] = w2[c];
    }
  }
  if (vis) {
    depths[idx] = viewz;
    clamped[idx] = clamp_bits;
  }
  if (valid) {
    radii[idx] = radius_out;
    tiles_touched[idx] = tiles_out;
  }
  // range of the depth word over the visible Gaussians (feeds the sort's digit plan)
  if (vis_mask) {
    const uint32_t bits = __float_as_uint(viewz);
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, vis ? bits : 0u);
    const uint32_t wneg = __reduce_max_sync(0xffffffffu, vis ? ~bits : 0u);
    const uint32_t wtiles = __reduce_add_sync(0xffffffffu, tiles_out);
    if (lane == 0) { atomicMax(&s_max, wmax); atomicMax(&s_negmin, wneg); atomicAdd(&s_tiles, wtiles); }
  }
  __syncthreads();
  if (tid == 0 && s_max != 0u) {
    atomicMax(&ctrl[CTRL_DEPTH_MAX], s_max);
    atomicMax(&ctrl[CTRL_DEPTH_NEGMIN], s_negmin);
    // num_rendered = sum of tiles_touched (the reference reads it off the end of its prefix sum,
    // rasterizer_impl.cu:280-284); the host needs it before the instance list is laid out in depth order
    if (s_tiles) atomicAdd(&ctrl[CTRL_NUM_RENDERED], s_tiles);
  }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    uint8_t* __restrict__ present) {
  // checkFrustum, RAST/cuda_rasterizer/rasterizer_impl.cu:54-66: present = p_view.z > 0.2
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float x = means3D[3 * (size_t)idx], y = means3D[3 * (size_t)idx + 1], z = means3D[3 * (size_t)idx + 2];
  present[idx] = tp_row(viewmatrix, 2, x, y, z) > 0.2f ? 1 : 0;
}

}  // namespace

void launch_preprocess_fwd(const lsr_frame& f, const lsr_inputs& in, float focal_x, float focal_y, int32_t* radii,
                           GeomView g, uint32_t* ctrl, cudaStream_t s) {
  const int gx = (f.width + TILE_X - 1) / TILE_X, gy = (f.height + TILE_Y - 1) / TILE_Y;
  const int blocks = (f.P + PRE_THREADS - 1) / PRE_THREADS;
  if (blocks == 0) return;
  // fast path: the usual [P,16,3] SH tensor, 16-byte aligned -> rows staged through shared memory
  const bool sh16 = in.shs != nullptr && f.M == 16 && (((uintptr_t)in.shs) & 15) == 0;
  static DeviceOnce attr;  // per device: a process may drive several GPUs
  attr.run([] {
    cudaFuncSetAttribute(preprocess_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PRE_SH_SMEM);
    cudaFuncSetAttribute(preprocess_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PRE_SH_SMEM);
  });
#define LSR_PRE_ARGS                                                                                                  \
  f.P, f.D, f.M, in.means3D, in.scales, f.scale_modifier, in.rotations, in.opacities, in.shs, in.cov3D_precomp,       \
      in.colors_precomp, f.include_feature ? in.language_feature : nullptr, f.viewmatrix, f.projmatrix, f.campos,     \
      f.width, f.height, f.tanfovx, f.tanfovy, focal_x, focal_y, gx, gy, f.prefiltered, radii, g.rec, g.depth, g.tiles, \
      g.cov3D, g.clamped, ctrl
  if (sh16) preprocess_fwd_kernel<true><<<blocks, PRE_THREADS, PRE_SH_SMEM, s>>>(LSR_PRE_ARGS);
  else preprocess_fwd_kernel<false><<<blocks, PRE_THREADS, PRE_SH_SMEM, s>>>(LSR_PRE_ARGS);
#undef LSR_PRE_ARGS
  count_launch();
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s) {
  if (P <= 0) return;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
  count_launch();
}

}  // namespace lsr
