// rows.cuh -- row source shared by the MLP kernels: row g = ray*S + s -> sample point + view direction.
#pragma once
#include "engine.h"

namespace nerf {

// ---- positional encoding of one MLP row --------------------------------------------------------
// Row g = ray*S + s.  Points are either explicit or o + d*t (volume_renderer.py:63,267).
__device__ __forceinline__ void row_point(const RowSrc& src, int64_t g, float p[3], float vd[3]) {
  const int64_t ray = g / src.S;
  if (src.pts) {
    p[0] = __ldg(src.pts + 3 * g); p[1] = __ldg(src.pts + 3 * g + 1); p[2] = __ldg(src.pts + 3 * g + 2);
  } else {
    const float t = __ldg(src.t + g);
#pragma unroll
    for (int c = 0; c < 3; ++c) p[c] = ray_point(__ldg(src.rays_o + 3 * ray + c), __ldg(src.rays_d + 3 * ray + c), t);
  }
  if (src.viewdirs) {
#pragma unroll
    for (int c = 0; c < 3; ++c) vd[c] = __ldg(src.viewdirs + 3 * ray + c);
  } else {  // viewdirs = rays_d / ||rays_d||  (volume_renderer.py:314)
    const float x = __ldg(src.rays_d + 3 * ray), y = __ldg(src.rays_d + 3 * ray + 1), z = __ldg(src.rays_d + 3 * ray + 2);
    const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)));
    vd[0] = __fdiv_rn(x, nrm); vd[1] = __fdiv_rn(y, nrm); vd[2] = __fdiv_rn(z, nrm);
  }
}

}  // namespace nerf
