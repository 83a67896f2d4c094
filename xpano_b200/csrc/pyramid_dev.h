// Device-side helpers shared by pyramid.cu and pyramid_fused.cu.
#pragma once

#include "pyramid_math.h"

namespace xp {

__device__ __forceinline__ void expand_row4(const float* __restrict__ r, int i,
                                            int xs, float e[4]) {
  const float2 p = __ldg(reinterpret_cast<const float2*>(r + 2 * i));
  const float2 q = __ldg(reinterpret_cast<const float2*>(r + 2 * i + 2));
  const float a[5] = {p.x, p.y, q.x, q.y, __ldg(r + 2 * i + 4)};
  pm::expand_h4(a, xs, e);
}

// expand(lo) at hi samples (4i .. 4i+3, y)
__device__ __forceinline__ void expand4(const float* __restrict__ lo,
                                        int lo_pitch, int xs, int ys, int i,
                                        int y, float e[4]) {
  if (2 * i + 4 < lo_pitch) {
    const int t = y + ys;
    const float* r1 = lo + (size_t)((t >> 1) + 1) * lo_pitch;
    float e0[4] = {0.f, 0.f, 0.f, 0.f}, e1[4], e2[4];
    expand_row4(r1, i, xs, e1);
    expand_row4(r1 + lo_pitch, i, xs, e2);
    const int even = (t & 1) == 0;
    if (even) expand_row4(r1 - lo_pitch, i, xs, e0);
#pragma unroll
    for (int k = 0; k < 4; ++k) e[k] = pm::expand_v(even, e0[k], e1[k], e2[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      e[k] = pm::expand_sample(lo, lo_pitch, xs, ys, 4 * i + k, y);
  }
}

}  // namespace xp
