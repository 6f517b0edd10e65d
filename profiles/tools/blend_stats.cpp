// blend_stats.cpp — analysis helper (not product, not oracle): replays the blend's accept decisions to count how many
// (footprint, instance) pairs a kernel with a given warp footprint examines / has work for.  Drives design choices for
// render_fwd / render_bwd.  Build: g++ -O3 -fopenmp -shared -fPIC -o libblendstats.so blend_stats.cpp
#include <cmath>
#include <cstdint>
#include <algorithm>
extern "C" void blend_stats(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* means2D,
                            const float* conic_opacity, const uint32_t* n_contrib, int fw, int fh, int64_t* out) {
  const int gx = (W + 15) / 16, gy = (H + 15) / 16;
  int64_t fetched = 0, exam_tile = 0, exam_fp = 0, work_fp = 0, lanes = 0, bbox_fp = 0;
  int64_t hist[5] = {0, 0, 0, 0, 0};  // contributing pixels per working pair: 1-2, 3-4, 5-8, 9-16, 17+
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : fetched, exam_tile, exam_fp, work_fp, lanes, bbox_fp, hist[:5])
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t rb = ranges[2 * tile], re = ranges[2 * tile + 1];
    uint32_t tmax = 0;
    for (int ly = 0; ly < 16; ly++)
      for (int lx = 0; lx < 16; lx++) {
        const int px = tx * 16 + lx, py = ty * 16 + ly;
        if (px < W && py < H) tmax = std::max(tmax, n_contrib[(size_t)py * W + px]);
      }
    fetched += tmax;
    const int nfx = 16 / fw, nfy = 16 / fh;
    exam_tile += (int64_t)tmax * nfx * nfy;
    for (int fy = 0; fy < nfy; fy++)
      for (int fx = 0; fx < nfx; fx++) {
        uint32_t fmax = 0;
        for (int ly = 0; ly < fh; ly++)
          for (int lx = 0; lx < fw; lx++) {
            const int px = tx * 16 + fx * fw + lx, py = ty * 16 + fy * fh + ly;
            if (px < W && py < H) fmax = std::max(fmax, n_contrib[(size_t)py * W + px]);
          }
        exam_fp += fmax;
        for (uint32_t k = 0; k < fmax && rb + k < re; k++) {
          const uint32_t id = point_list[rb + k];
          const float* co = conic_opacity + 4 * (size_t)id;
          const float mx = means2D[2 * (size_t)id], my = means2D[2 * (size_t)id + 1];
          int cnt = 0;
          for (int ly = 0; ly < fh; ly++)
            for (int lx = 0; lx < fw; lx++) {
              const int px = tx * 16 + fx * fw + lx, py = ty * 16 + fy * fh + ly;
              if (px >= W || py >= H) continue;
              if (k >= n_contrib[(size_t)py * W + px]) continue;
              const float dx = mx - px, dy = my - py;
              const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
              if (power > 0) continue;
              const float alpha = std::min(0.99f, co[3] * std::exp(power));
              if (alpha < 1.0f / 255.0f) continue;
              cnt++;
            }
          if (cnt) {
            work_fp++;
            lanes += cnt;
            hist[cnt <= 2 ? 0 : cnt <= 4 ? 1 : cnt <= 8 ? 2 : cnt <= 16 ? 3 : 4]++;
          }
        }
      }
  }
  out[0] = fetched; out[1] = exam_tile; out[2] = exam_fp; out[3] = work_fp; out[4] = lanes;
  for (int i = 0; i < 5; i++) out[5 + i] = hist[i];
}
